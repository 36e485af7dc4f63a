"""Variable-update strategies: what happens to the gradients of a step.

The reference has eight ``VariableMgr`` classes (T/variable_mgr.py:28-831) plus the
Horovod and KungFu hooks; all of them are data parallelism. On one
NVSwitch box they collapse onto five behaviours, each a single fused kernel per
gradient bucket (``csrc/optim_comm.cu``):

==========================  =========================================================
variable_update             behaviour
==========================  =========================================================
independent                 local fused optimizer, no exchange
replicated, horovod,        all-reduce **sum** fused with unscale + update
 *_all_reduce                (ResNet divides its LR by num_gpus to compensate,
                             T/models/resnet_model_legacy.py:311-328)
parameter_server,           all-reduce **mean** fused with the update (reduce-to-owner +
 distributed_replicated      broadcast collapse to an all-reduce on NVSwitch)
kungfu sync_sgd             all-reduce mean (KungFu SynchronousSGDOptimizer)
kungfu sma                  all-reduce-average of the *weights*, blend, local step
kungfu async_sgd            pull one random peer's model, blend 1/2, local step (AD-PSGD)
kungfu ada_sgd              sma until --ada_sgd_switch_step, then sync_sgd
==========================  =========================================================

Buckets are offset ranges of the flat arenas (``nn/params.py``); the CUDA executor reports
ranges whose gradients are final during backward and every full bucket is launched on a
high-priority stream at once, overlapping the exchange with the rest of backward.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Tuple

import numpy as np
import torch

from ..nn.params import ParameterStore
from ..ops import native
from ..ops.native import f32, i32, i64, ptr
from ..ops import optim_ref
from .comm import Comm, P2PComm, TorchDistComm

OPT_IDS = {"sgd": 0, "momentum": 1, "rmsprop": 2, "adam": 3}


class PushTarget(ctypes.Structure):
    """Mirror of ``kfb_push_target`` (include/kfb200.h)."""
    _fields_ = [("nranks", ctypes.c_int), ("per", ctypes.c_longlong), ("rel", ctypes.c_longlong),
                ("dst", ctypes.c_void_p * 8)]


class Hyper(ctypes.Structure):
    """Mirror of ``kfb_hyper`` (include/kfb200.h)."""
    _fields_ = [("opt", ctypes.c_int), ("lr", ctypes.c_float), ("momentum", ctypes.c_float),
                ("decay", ctypes.c_float), ("epsilon", ctypes.c_float),
                ("weight_decay", ctypes.c_float), ("grad_scale", ctypes.c_float),
                ("clip", ctypes.c_float), ("adam_t", ctypes.c_float),
                ("decay_mask", ctypes.c_void_p)]


def reduction_semantics(params) -> Tuple[str, bool]:
    """(algorithm, average) for a parameter set: algorithm in
    {none, allreduce, sma, pair}."""
    vu = params.variable_update
    if vu == "independent":
        return "none", False
    if vu in ("replicated", "horovod", "distributed_all_reduce", "collective_all_reduce"):
        return "allreduce", False
    if vu in ("parameter_server", "distributed_replicated"):
        return "allreduce", True
    if vu == "kungfu":
        opt = params.kungfu_option
        if opt == "sync_sgd":
            return "allreduce", True
        if opt == "sma":
            return "sma", True
        if opt == "async_sgd":
            return "pair", False
        if opt == "ada_sgd":
            return "ada", True
    raise ValueError("unsupported variable_update=%s" % vu)


def plan_buckets(total: int, bucket_elems: int, num_buckets: int = 0) -> List[Tuple[int, int]]:
    """Split [0, total) into buckets: `num_buckets` equal parts (gradient_repacking) or
    size-based. Boundaries are multiples of 128 elements (arena slot alignment)."""
    if total <= 0:
        return []
    if num_buckets and num_buckets > 0:
        per = -(-total // num_buckets)
    else:
        per = max(bucket_elems, 128)
    per = (per + 127) // 128 * 128
    out = []
    lo = 0
    while lo < total:
        hi = min(lo + per, total)
        out.append((lo, hi))
        lo = hi
    return out


class UpdateEngine:
    """Applies one step's gradients to the model according to the strategy."""

    def __init__(self, params, store: ParameterStore, comm: Comm, model, seed: int = 1234):
        self.params = params
        self.store = store
        self.comm = comm
        self.model = model
        self.algo, self.average = reduction_semantics(params)
        if comm.size == 1 and self.algo != "none":
            self.algo = "none"  # a single replica: nothing to exchange
        self.device = store.device
        self.cuda = self.device.type == "cuda"
        self.opt = OPT_IDS[params.optimizer]
        self.step_count = 0
        n = store.num_trainable_padded
        self.total = n
        self.m = store.opt_slot("m") if self.opt != 0 else None
        self.v = store.opt_slot("v") if self.opt in (2, 3) else None
        if self.opt == 2 and float(self.v.abs().sum()) == 0.0:
            # tf.train.RMSPropOptimizer creates its "rms" slot filled with ones
            # (T/benchmark_cnn.py:1179-1185 uses it with momentum / decay / epsilon flags)
            self.v.fill_(1.0)
        self.auto_loss_scale = bool(params.use_fp16 and params.fp16_enable_auto_loss_scale)
        # Automatic loss scaling keeps its state machine on the device (T/variable_mgr_util.py:
        # 91-139 takes the skip / apply decision inside the graph with tf.cond): ls_state =
        # {scale, clean steps, skipped steps, -}; the loss kernels read the scale from it, the
        # update kernels read grad_scale from dyn[1] and the skip decision from skip_flag, and
        # kfb_loss_scale_update moves the state - no device->host read in the step. The host
        # path (CPU replicas, torch.distributed, SMA / PairAveraging) keeps the Python mirror.
        self.device_ls = bool(self.auto_loss_scale and self.device.type == "cuda" and
                              self.algo in ("none", "allreduce") and
                              not isinstance(comm, TorchDistComm) and
                              params.variable_consistency != "relaxed" and
                              not params.staged_vars)
        self.ls_state = None
        self._loss_scale = self._initial_loss_scale()
        self._ls_normal_steps = 0
        self._skipped_steps = 0
        if self.device_ls:
            self.ls_state = torch.tensor([self._loss_scale, 0.0, 0.0, 0.0], dtype=torch.float32,
                                         device=self.device)
        # bucket size: --bucket_bytes, but never more than ~12 buckets for very large models (each
        # bucket costs two flag rounds across the ranks)
        bucket_elems = max(int(params.bucket_bytes) // 4, n // 12, 128)
        self.buckets = plan_buckets(n, bucket_elems, params.gradient_repacking)
        self._ready_hi = 0
        self._launched = 0
        self._rng = np.random.RandomState(seed + 7919 * comm.rank)
        # variables a model excludes from the L2 term (SSD: batch norm): one byte per 128-element
        # arena block, read by the update kernels; None when everything is decayed
        self.decay_mask = None
        flt = getattr(model, "l2_decay_filter", None)
        if flt is not None and hasattr(store, "order") and (params.weight_decay or 0.0) > 0:
            keep = torch.ones((n + 127) // 128, dtype=torch.uint8)
            any_off = False
            for slot in store.order:
                if not flt(slot.spec.name):
                    keep[slot.offset // 128:(slot.offset + slot.numel + 127) // 128] = 0
                    any_off = True
            if any_off:
                self.decay_mask = keep.to(self.device)
        # --staged_vars (T/variable_mgr_util.py:236-393) software-pipelines the parameter-server
        # traffic by one step through StagingAreas: the step computes on variables that miss
        # the newest update. The deferred-exchange path below has the same one-step staleness
        # and the same overlap (exchange of step k under the forward pass of step k+1).
        self.relaxed = params.variable_consistency == "relaxed" or bool(params.staged_vars)
        if self.cuda:
            self.skip_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        # device-resident {lr, grad_scale, adam_t, -}: lets a captured CUDA graph see per-step
        # values (the kernels read it instead of the by-value fields when given)
        self.dyn = None
        self._dyn_host = None
        # p2p_fused: weight-gradient kernels push their tiles to the owning rank (fused GEMM +
        # reduce-scatter); only on the overlapped synchronous all-reduce path
        self.fused_push = bool(
            getattr(params, "comm_impl", "p2p") == "p2p_fused" and self.cuda and
            isinstance(comm, P2PComm) and self.algo == "allreduce" and
            not self.auto_loss_scale and not self.relaxed)
        self._push_targets = {}
        if self.fused_push:
            self.buckets = self._plan_buckets_on_slots(bucket_elems)
        self._setup_comm_buffers()

    # ------------------------------------------------------------------ loss-scale state
    def _ls_get(self, idx: int, host_value):
        if self.ls_state is None:
            return host_value
        return float(self.ls_state[idx].item())      # synchronises; not used inside the step

    def _ls_set(self, idx: int, value: float) -> None:
        if self.ls_state is not None:
            self.ls_state[idx] = float(value)
            if idx == 0 and self.dyn is not None:
                self.dyn[1] = self._base_grad_scale() / float(value)

    @property
    def loss_scale(self) -> float:
        return self._ls_get(0, self._loss_scale)

    @loss_scale.setter
    def loss_scale(self, value: float) -> None:
        self._loss_scale = float(value)
        self._ls_set(0, value)

    @property
    def loss_scale_normal_steps(self) -> int:
        return int(self._ls_get(1, self._ls_normal_steps))

    @loss_scale_normal_steps.setter
    def loss_scale_normal_steps(self, value: int) -> None:
        self._ls_normal_steps = int(value)
        self._ls_set(1, value)

    @property
    def skipped_steps(self) -> int:
        return int(self._ls_get(2, self._skipped_steps))

    @skipped_steps.setter
    def skipped_steps(self, value: int) -> None:
        self._skipped_steps = int(value)
        self._ls_set(2, value)

    def loss_scale_kwargs(self) -> dict:
        """loss_scale / dyn_scale arguments of ``executor.loss_backward``: with the
        device-resident state the loss kernels multiply by the scale they read from
        ``ls_state``."""
        if self.ls_state is not None:
            return dict(loss_scale=1.0, dyn_scale=self.ls_state[0:1])
        return dict(loss_scale=self._loss_scale, dyn_scale=None)

    def _base_grad_scale(self) -> float:
        return (1.0 / self.comm.size) if (self.algo == "allreduce" and self.average) else 1.0

    # ------------------------------------------------------------------ setup
    def _initial_loss_scale(self) -> float:
        p = self.params
        if not p.use_fp16 and self.device.type != "cuda":
            return 1.0
        if p.fp16_loss_scale is not None:
            return float(p.fp16_loss_scale)
        return float(self.model.get_fp16_loss_scale()) if (p.use_fp16 or self.cuda) else 1.0

    def _plan_buckets_on_slots(self, bucket_elems: int) -> List[Tuple[int, int]]:
        """Buckets that only cut between variables, so every weight gradient lies in exactly
        one bucket and its owner map is a single (lo, per) pair."""
        out, lo, hi = [], 0, 0
        for slot in self.store.order:
            hi = slot.offset + (slot.numel + 127) // 128 * 128
            if hi - lo >= bucket_elems:
                out.append((lo, hi))
                lo = hi
        if hi > lo:
            out.append((lo, hi))
        return out

    def register_rs_params(self, names) -> None:
        """Called by the executor with the variables whose gradient kernels can push (conv
        kernels, FC weights). Everything else in a bucket is pushed by ``kfb_push_ranges``."""
        if not self.fused_push:
            return
        c, st = self.comm, self.store
        n = c.size
        names = set(names)
        self.stage = c.symm_alloc(n * self.total, torch.float32)
        stage_peers = c.peers(self.stage)
        slot_bytes = self.total * 4
        self._slot_base = (ctypes.c_void_p * 8)()
        for o in range(n):
            self._slot_base[o] = stage_peers[o] + c.rank * slot_bytes
        self._bucket_push = []
        slots = list(st.order)
        for (lo, hi) in self.buckets:
            per = ((hi - lo + n - 1) // n + 3) & ~3
            ranges = []
            for s in slots:
                if not (lo <= s.offset < hi):
                    continue
                if s.spec.name in names:
                    t = PushTarget()
                    t.nranks, t.per, t.rel = n, per, s.offset - lo
                    for o in range(n):
                        t.dst[o] = stage_peers[o] + c.rank * slot_bytes + s.offset * 4
                    self._push_targets[s.spec.name] = t
                else:
                    ranges.append((s.offset, s.offset + (s.numel + 3) // 4 * 4))
            flat = [v for r in ranges for v in r]
            dev = (torch.tensor(flat, dtype=torch.int64, device=self.device) if flat else None)
            longest = max((b - a for a, b in ranges), default=0)
            self._bucket_push.append((per, dev, len(ranges), longest))

    def push_target(self, name):
        """ctypes descriptor for the fused reduce-scatter of variable ``name`` (or None)."""
        return self._push_targets.get(name)

    def _setup_comm_buffers(self) -> None:
        c = self.comm
        self.reduced = None
        self.nvls = False
        if isinstance(c, P2PComm) and self.algo != "none":
            st = self.store
            self.reduced = c.symm_alloc(self.total, torch.float32)
            self._grad_ptrs = c.ptr_array(st.grad)
            self._w_ptrs = c.ptr_array(st.w32)
            self._red_ptrs = c.ptr_array(self.reduced)
            self._flag_ptrs = c.ptr_array(c.flags)
            # --comm_impl=nvls: multicast aliases of the same arenas (in-switch reduce / bcast)
            self.nvls = bool(getattr(c, "nvls", False))
            if self.nvls:
                self._grad_mc = ctypes.c_void_p(c.mc_ptr(st.grad))
                self._w_mc = ctypes.c_void_p(c.mc_ptr(st.w32))
                self._red_mc = ctypes.c_void_p(c.mc_ptr(self.reduced))
            # second-generation kernel (kfb_allreduce_update2); KFB_AR_V2=0 keeps the first
            self.ar_v2 = os.environ.get("KFB_AR_V2", "1") != "0"
            # --compact_gradient_transfer (T/variable_mgr.py:311-313: whenever use_fp16): the
            # gradients cross NVLink as fp16 (summed in fp32); symmetric fp16 arenas for the
            # sources and the reduced shards
            p = self.params
            self.compact = bool(self.ar_v2 and p.use_fp16 and p.compact_gradient_transfer and
                                self.algo in ("allreduce", "ada") and not self.relaxed and
                                not self.fused_push)
            if self.compact:
                self.g16 = c.symm_alloc(self.total, torch.float16)
                self.r16 = c.symm_alloc(self.total, torch.float16)
                self._g16_ptrs = c.ptr_array(self.g16)
                self._r16_ptrs = c.ptr_array(self.r16)
                if self.nvls:
                    self._g16_mc = ctypes.c_void_p(c.mc_ptr(self.g16))
                    self._r16_mc = ctypes.c_void_p(c.mc_ptr(self.r16))
            self.ready_event = torch.cuda.Event()
            self.done_event = torch.cuda.Event()
            if self.algo in ("pair", "ada"):
                pass
            if self.algo == "pair":
                self.pub = c.symm_alloc(3 * self.total, torch.float32)
                self.ver = c.symm_alloc(64, torch.int32)
                self._pub_peers = c.peers(self.pub)
                self._ver_peers = c.peers(self.ver)
                self.torn = torch.zeros(1, dtype=torch.int32, device=self.device)
                self.pull_buf = torch.empty(self.total, dtype=torch.float32, device=self.device)
                self.pair_sync = torch.zeros(4, dtype=torch.int32, device=self.device)
                self.my_version = 0

    # ------------------------------------------------------------------ per-step
    def learning_rate(self, global_step: int, examples_per_step: int, num_examples: int) -> float:
        from .. import lr as lr_mod
        return lr_mod.get_learning_rate(self.params, global_step, num_examples, self.model,
                                        examples_per_step)

    def _hyper(self, lr: float) -> Hyper:
        p = self.params
        h = Hyper()
        h.opt = self.opt
        h.lr = lr
        if self.opt == 1:
            h.momentum = p.momentum
        elif self.opt == 2:
            h.momentum, h.decay, h.epsilon = p.rmsprop_momentum, p.rmsprop_decay, p.rmsprop_epsilon
        elif self.opt == 3:
            h.momentum, h.decay, h.epsilon = p.adam_beta1, p.adam_beta2, p.adam_epsilon
        # The reference adds num_towers * wd * l2_loss to the loss of ONE tower
        # (T/benchmark_cnn.py:3078-3099), so the L2 term follows the reduction: summed
        # gradients (replicated / horovod / *_all_reduce) carry N * wd * w, averaged ones
        # (parameter_server, KungFu S-SGD) and local steps carry wd * w.
        h.weight_decay = p.weight_decay or 0.0
        if self._effective_algo() == "allreduce" and not self.average:
            h.weight_decay *= self.comm.size
        scale = 1.0 / self._loss_scale      # device-resident scaling: dyn[1] overrides this
        if self._effective_algo() == "allreduce" and self.average:
            scale /= self.comm.size
        h.grad_scale = scale
        h.clip = p.gradient_clip if p.gradient_clip else 0.0
        h.adam_t = float(self.step_count + 1)
        h.decay_mask = self.decay_mask.data_ptr() if (self.decay_mask is not None and
                                                      self.cuda) else None
        return h

    def enable_dynamic_hyper(self) -> None:
        if self.dyn is None:
            self.dyn = torch.zeros(4, dtype=torch.float32, device=self.device)
            if self.ls_state is not None:
                self.dyn[1] = self._base_grad_scale() / self._loss_scale

    def push_dynamic_hyper(self, lr: float) -> None:
        """This step's learning rate / gradient scale / Adam step count into the device-resident
        hyper-parameters (graph replay, device-resident loss scaling). A one-thread kernel takes
        the values by argument, so the host may run any number of steps ahead of the device;
        under automatic loss scaling grad_scale belongs to kfb_loss_scale_update and is left
        alone."""
        h = self._hyper(lr)
        mask = 5 if self.ls_state is not None else 7
        native.call("set_hyper_dyn", ptr(self.dyn), f32(h.lr), f32(h.grad_scale), f32(h.adam_t),
                    i32(mask), self._s())

    def begin_step(self, lr: float, push_dyn: bool = True) -> None:
        """Called before backward: arms the overlap scheduler. ``push_dyn=False`` while a CUDA
        graph is being captured (the per-step values are pushed before every replay instead)."""
        self._h = self._hyper(lr)
        if self.relaxed and self.cuda and getattr(self, "_relaxed_pending", False) and \
                isinstance(self.comm, P2PComm):
            # the deferred all-reduce reads the gradient arena: it must be done before this
            # step's backward starts overwriting it (it had the whole forward pass to finish)
            torch.cuda.current_stream(self.device).wait_event(self.done_event)
        self._ready_hi = 0
        self._launched = 0
        self._sma_prefetched = False
        if self._effective_algo() == "sma" and self._sma_can_prefetch():
            # SynchronousAveraging exchanges the *weights*, which are final now: their average
            # is taken on the exchange stream while forward and backward run
            c = self.comm
            self.ready_event.record(torch.cuda.current_stream(self.device))
            c.comm_stream.wait_event(self.ready_event)
            self._allreduce_call(3, 0, self.total, self._h, None, c.comm_stream)
            self.done_event.record(c.comm_stream)
            self._sma_prefetched = True
        if self.device_ls:
            # the per-step values the update kernels read from device memory (the skip flag was
            # cleared by the previous kfb_loss_scale_update)
            self.enable_dynamic_hyper()
            if push_dyn:
                self.push_dynamic_hyper(lr)
        elif self.cuda and self.auto_loss_scale:
            self.skip_flag.zero_()

    def _effective_algo(self) -> str:
        if self.algo != "ada":
            return self.algo
        return "sma" if self.step_count < self.params.ada_sgd_switch_step else "allreduce"

    def on_grads_ready(self, lo: int, hi: int) -> None:
        """Backward-completion callback: gradients of arena range [lo, hi) are final. Ranges
        arrive in increasing offset order (the arena is laid out that way)."""
        self._ready_hi = max(self._ready_hi, hi)
        if not self._can_overlap():
            return
        while self._launched < len(self.buckets) and \
                self.buckets[self._launched][1] <= self._ready_hi:
            self._launch_bucket(self.buckets[self._launched], overlapped=True)
            self._launched += 1

    def _sma_can_prefetch(self) -> bool:
        return (self.cuda and isinstance(self.comm, P2PComm) and self.ar_v2 and
                not self.auto_loss_scale and not self.relaxed and self.total % 8 == 0 and
                os.environ.get("KFB_SMA_OVERLAP", "1") != "0")

    def _can_overlap(self) -> bool:
        return (self.cuda and isinstance(self.comm, P2PComm) and
                self._effective_algo() == "allreduce" and
                (not self.auto_loss_scale or self.device_ls) and not self.relaxed)

    def finish_step(self) -> None:
        """Called after backward: runs whatever has not been overlapped and makes the compute
        stream wait for the update."""
        algo = self._effective_algo()
        if self.device_ls:
            self._finish_device_loss_scale(algo)
            self.step_count += 1
            return
        if self.auto_loss_scale:
            self._check_finite()
        if self.relaxed and algo == "allreduce":
            self._finish_relaxed()
            self.step_count += 1
            return
        if not self.cuda or isinstance(self.comm, TorchDistComm):
            self._finish_torch(algo)
        elif algo == "none":
            self._local_update(0, self.total)
        elif algo == "pair":
            self._pair_update()
        elif algo == "sma" and getattr(self, "_sma_prefetched", False):
            st = self.store
            torch.cuda.current_stream(self.device).wait_event(self.done_event)
            native.call("optimizer_step_blend", ptr(st.w32), ptr(self.reduced), ptr(st.grad),
                        ptr(self.m), ptr(self.v), ptr(st.w16), i64(self.total),
                        ctypes.byref(self._h), ptr(self.dyn), f32(self.params.sma_alpha),
                        self._s())
        else:
            if self._launched == 0 and not self.fused_push:
                # nothing was overlapped (SMA, auto loss scale, the all-reduce micro-benchmark,
                # gradients that all became ready at once): one launch over the whole arena -
                # every extra bucket costs ~55 us of flag rounds and ramp (2 B200s, 102 MB:
                # 4 buckets 520 us, 1 launch 260 us)
                self._launch_bucket((0, self.total), overlapped=self._can_overlap())
                self._launched = len(self.buckets)
            while self._launched < len(self.buckets):
                self._launch_bucket(self.buckets[self._launched], overlapped=self._can_overlap())
                self._launched += 1
            if self._can_overlap():
                self.done_event.record(self.comm.comm_stream)
                torch.cuda.current_stream(self.device).wait_event(self.done_event)
        if self.auto_loss_scale:
            self._update_loss_scale()
        self.step_count += 1

    def _finish_device_loss_scale(self, algo: str) -> None:
        """Automatic loss scaling without a host round trip. The buckets were summed (reduce
        only) and checked for inf / nan on the exchange stream while backward ran - a non-finite
        gradient on any rank makes the sum non-finite on every rank, so all ranks take the same
        decision without exchanging it; the whole-arena update is predicated on the flag and
        the state machine moves on the device."""
        st, c = self.store, self.comm
        flag = ptr(self.skip_flag)
        if algo == "none" or not isinstance(c, P2PComm):
            native.call("check_finite", ptr(st.grad), i64(self.total), flag, self._s())
            src = st.grad
        else:
            if self._launched == 0:
                self._launch_bucket((0, self.total), overlapped=True)
                self._launched = len(self.buckets)
            while self._launched < len(self.buckets):
                self._launch_bucket(self.buckets[self._launched], overlapped=True)
                self._launched += 1
            self.done_event.record(c.comm_stream)
            torch.cuda.current_stream(self.device).wait_event(self.done_event)
            src = self.reduced
        native.call("optimizer_step", ptr(st.w32), ptr(src), ptr(self.m), ptr(self.v),
                    ptr(st.w16), i64(self.total), ctypes.byref(self._h), ptr(self.dyn), flag,
                    self._s())
        native.call("loss_scale_update", ptr(self.ls_state), flag,
                    i32(int(self.params.fp16_inc_loss_scale_every_n)), ptr(self.dyn),
                    f32(self._base_grad_scale()), self._s())

    def _finish_relaxed(self) -> None:
        """variable_consistency=relaxed (T/batch_allreduce.py:353-388): apply the all-reduced
        gradients of the PREVIOUS step now, then start reducing this step's gradients so the
        exchange overlaps the next forward pass."""
        st, c = self.store, self.comm
        if isinstance(c, P2PComm):
            if getattr(self, "_relaxed_pending", False):
                native.call("optimizer_step", ptr(st.w32), ptr(self.reduced), ptr(self.m),
                            ptr(self.v), ptr(st.w16), i64(self.total), ctypes.byref(self._h), None,
                            None, self._s())
            self.ready_event.record(torch.cuda.current_stream(self.device))
            c.comm_stream.wait_event(self.ready_event)
            self._allreduce_call(2, 0, self.total, self._h, None, c.comm_stream)
            self.done_event.record(c.comm_stream)
            self._relaxed_pending = True
        else:
            prev = getattr(self, "_relaxed_prev", None)
            if prev is not None:
                if self.cuda:
                    native.call("optimizer_step", ptr(st.w32), ptr(prev), ptr(self.m),
                                ptr(self.v), ptr(st.w16), i64(self.total), ctypes.byref(self._h),
                                None, None, self._s())
                else:
                    optim_ref.apply_update(self._h, st.w32, prev, self.m, self.v, st.w16,
                                           decay_mask=self.decay_mask)
            cur = st.grad.clone()
            c.all_reduce_(cur)
            self._relaxed_prev = cur

    # ------------------------------------------------------------------ CUDA paths
    def _s(self, stream=None):
        return ctypes.c_void_p((stream or torch.cuda.current_stream(self.device)).cuda_stream)

    def _ptr_at(self, t: Optional[torch.Tensor], off: int):
        if t is None:
            return None
        return ctypes.c_void_p(t.data_ptr() + off * t.element_size())

    def _local_update(self, lo: int, hi: int) -> None:
        st = self.store
        h = self._h
        if lo and h.decay_mask:
            # the kernel indexes the mask like its (offset) arrays: advance it with them
            h = Hyper.from_buffer_copy(h)
            h.decay_mask = self._h.decay_mask + lo // 128
        native.call("optimizer_step", self._ptr_at(st.w32, lo), self._ptr_at(st.grad, lo),
                    self._ptr_at(self.m, lo), self._ptr_at(self.v, lo), self._ptr_at(st.w16, lo),
                    i64(hi - lo), ctypes.byref(h), ptr(self.dyn),
                    ptr(self.skip_flag) if self.auto_loss_scale else None, self._s())

    def _launch_bucket(self, bucket, overlapped: bool) -> None:
        lo, hi = bucket
        st, c = self.store, self.comm
        algo = self._effective_algo()
        stream = None
        if overlapped:
            self.ready_event.record(torch.cuda.current_stream(self.device))
            c.comm_stream.wait_event(self.ready_event)
            stream = c.comm_stream
        mode = 1 if algo == "sma" else 0
        h = self._h
        if self.device_ls:
            # reduce only + finite check of the sum; the update follows once every bucket is in
            self._allreduce_call(2, lo, hi - lo, h, None, stream)
            native.call("check_finite", self._ptr_at(self.reduced, lo), i64(hi - lo),
                        ptr(self.skip_flag), self._s(stream))
            return
        if self.fused_push and self._push_targets:
            per, ranges_dev, nranges, longest = self._bucket_push[self.buckets.index(bucket)]
            if nranges:
                native.call("push_ranges", ptr(st.grad), ptr(ranges_dev), i32(nranges),
                            i64(longest), i64(lo), i64(per), i32(c.size), self._slot_base,
                            self._s(stream))
            native.call("allreduce_update_staged", i32(c.rank), i32(c.size), self._red_ptrs,
                        self._flag_ptrs, i64(lo), i64(hi - lo), ctypes.c_uint(c.next_seq()),
                        ptr(st.w32), ptr(self.m), ptr(self.v), ptr(st.w16), ctypes.byref(h),
                        None, None, ptr(self.stage), i64(self.total), i32(0), self._s(stream))
            return
        self._allreduce_call(mode, lo, hi - lo, h,
                             ptr(self.skip_flag) if self.auto_loss_scale else None, stream)

    def _allreduce_call(self, mode: int, lo: int, count: int, h, skip_flag, stream) -> None:
        """One fused all-reduce(+update) over [lo, lo+count): mode 0 gradients + optimizer,
        1 SMA (source = weights), 2 reduce only. Pull kernel, or its NVLS form. ``self.dyn``
        (null unless a CUDA graph is in use) carries the per-step learning rate / scale."""
        st, c = self.store, self.comm
        tail = (i64(lo), i64(count), ctypes.c_uint(c.next_seq()), ptr(st.w32), ptr(st.grad),
                ptr(self.m), ptr(self.v), ptr(st.w16), ctypes.byref(h), ptr(self.dyn), skip_flag,
                f32(self.params.sma_alpha if mode in (1, 3) else 0.0), i32(0), self._s(stream))
        if self.ar_v2 and lo % 8 == 0 and count % 8 == 0:
            half = self.compact and mode == 0
            if half:
                src, red, red_local = self._g16_ptrs, self._r16_ptrs, ptr(self.r16)
                src_mc = self._g16_mc if self.nvls else None
                red_mc = self._r16_mc if self.nvls else None
            else:
                src = self._w_ptrs if mode in (1, 3) else self._grad_ptrs
                red, red_local = self._red_ptrs, ptr(self.reduced)
                src_mc = (self._w_mc if mode in (1, 3) else self._grad_mc) if self.nvls else None
                red_mc = self._red_mc if self.nvls else None
            native.call("allreduce_update2", i32(mode), i32(c.rank), i32(c.size),
                        None if self.nvls else src, None if self.nvls else red, src_mc, red_mc,
                        red_local, self._flag_ptrs, i32(int(half)),
                        ptr(st.grad) if half else None, ptr(self.g16) if half else None, *tail)
        elif self.nvls:
            native.call("allreduce_update_nvls", i32(mode), i32(c.rank), i32(c.size),
                        self._w_mc if mode == 1 else self._grad_mc, self._red_mc,
                        ptr(self.reduced), self._flag_ptrs, *tail)
        else:
            native.call("allreduce_update", i32(mode), i32(c.rank), i32(c.size),
                        self._w_ptrs if mode == 1 else self._grad_ptrs, self._red_ptrs,
                        self._flag_ptrs, *tail)

    def _pair_update(self) -> None:
        c, st = self.comm, self.store
        peer = int(self._rng.randint(0, c.size - 1))
        if peer >= c.rank:
            peer += 1
        self.my_version += 1
        native.call("pair_average_update", ctypes.c_void_p(self._pub_peers[peer]),
                    ctypes.c_void_p(self._ver_peers[peer]), ptr(self.pub), ptr(self.ver),
                    i64(self.total), i64(self.total), ptr(st.w32), ptr(st.grad), ptr(self.m),
                    ptr(self.v), ptr(st.w16), ctypes.byref(self._h), None,
                    ctypes.c_uint(self.my_version), i32(1), ptr(self.torn), ptr(self.pull_buf),
                    ptr(self.pair_sync), i32(0), self._s())

    def _check_finite(self) -> None:
        st = self.store
        if self.cuda:
            native.call("check_finite", ptr(st.grad), i64(self.total), ptr(self.skip_flag),
                        self._s())
            if isinstance(self.comm, P2PComm):
                self.comm.flag_or(self.skip_flag)
            elif isinstance(self.comm, TorchDistComm):
                self.comm.all_reduce_(self.skip_flag)
        else:
            bad = 0 if bool(torch.isfinite(st.grad).all()) else 1
            if isinstance(self.comm, TorchDistComm):
                t = torch.tensor([bad], dtype=torch.int32)
                self.comm.all_reduce_(t)
                bad = int(t.item())
            self._cpu_skip = bad != 0

    def _update_loss_scale(self) -> None:
        """Host mirror of the state machine (T/variable_mgr_util.py:51-88) for the paths without
        the device-resident form (CPU replicas, torch.distributed, SMA / PairAveraging): the
        decision needs the flag on the host, one small D2H per step."""
        skipped = bool(self.skip_flag.item()) if self.cuda else self._cpu_skip
        self._loss_scale, self._ls_normal_steps = next_loss_scale(
            self._loss_scale, self._ls_normal_steps, skipped,
            self.params.fp16_inc_loss_scale_every_n)
        if skipped:
            self._skipped_steps += 1

    # ------------------------------------------------------------------ torch paths
    def _finish_torch(self, algo: str) -> None:
        """CPU replicas (gloo) and the NCCL comparison baseline: collective, then update."""
        st, c = self.store, self.comm
        h = self._h
        skip = False
        if self.auto_loss_scale:
            skip = bool(self.skip_flag.item()) if self.cuda else self._cpu_skip
        if algo == "allreduce":
            c.all_reduce_(st.grad)
        blend_src = None
        if algo == "sma":
            avg = st.w32.clone()
            c.all_reduce_(avg)
            avg /= c.size
            blend_src = (avg, self.params.sma_alpha)
        elif algo == "pair":
            blend_src = (self._pair_pull_torch(), 0.5)
        if skip:
            return
        if blend_src is not None:
            other, a = blend_src
            st.w32.mul_(1.0 - a).add_(other, alpha=a)
        if self.cuda:
            self._local_update(0, self.total)
        else:
            optim_ref.apply_update(h, st.w32, st.grad, self.m, self.v, st.w16,
                                   decay_mask=self.decay_mask)

    def _pair_pull_torch(self) -> torch.Tensor:
        """PairAveraging over torch.distributed: every rank draws a peer from a shared
        permutation-free schedule; implemented as an all-gather of the models (baseline
        quality - the product path reads the peer's HBM directly)."""
        c, st = self.comm, self.store
        peer = int(self._rng.randint(0, c.size - 1))
        if peer >= c.rank:
            peer += 1
        models = [torch.empty_like(st.w32) for _ in range(c.size)]
        c.dist.all_gather(models, st.w32)
        return models[peer]

    # ------------------------------------------------------------------ init-time collectives
    def broadcast_initial_variables(self) -> None:
        """Rank 0's variables (and BN statistics) become everyone's (T/variable_mgr.py:342-356,
        hvd.broadcast_global_variables / kungfu broadcast, T/benchmark_cnn.py:2094-2100)."""
        c, st = self.comm, self.store
        if c.size == 1:
            return
        if isinstance(c, P2PComm):
            w_ptrs = c.ptr_array(st.w32)
            native.call("broadcast_pull", i32(c.rank), i32(c.size), w_ptrs, c.ptr_array(c.flags),
                        i32(0), i64(0), i64(self.total), ctypes.c_uint(c.next_seq()),
                        ptr(st.w32), ptr(st.w16), i32(0), self._s())
            torch.cuda.synchronize(self.device)
            # the (small) non-trainable state goes through the bootstrap group
            state = st.state.cpu()
            c.dist.broadcast(state, src=0)
            st.state.copy_(state)
        else:
            c.broadcast_(st.w32, 0)
            c.broadcast_(st.state, 0)
            st.sync_shadow()
        c.barrier()


def next_loss_scale(scale: float, normal_steps: int, skipped: bool, inc_every_n: int):
    """One transition of the automatic loss-scale state machine."""
    if skipped:
        return scale / 2.0, 0
    if normal_steps < inc_every_n:
        return scale, normal_steps + 1
    return scale * 2.0, 0
