"""CUDA executor: runs a ``Graph`` forward and backward with the hand-written sm_100a
kernels of ``csrc/`` (no autograd, no cuDNN / cuBLAS, no Triton).

Design (B200-first, not a translation of the TF graph executor the reference leans on):

* every activation, gradient and scratch buffer is allocated once, up front - the step is a
  fixed sequence of kernel launches on static addresses, so it can be captured in a CUDA
  graph and nothing touches the allocator inside the timed loop (180 GB of HBM3e makes
  keeping all activations resident the cheap option);
* activations are NHWC fp16 ``[rows][C]``; flat tensors are rows padded to 8 columns so
  every row is a 16-byte multiple (TMA requirement); pad columns stay zero;
* convolutions / FC layers go through the tcgen05 implicit-GEMM kernel (fprop, dgrad,
  wgrad); weight gradients land directly in the fp32 gradient arena of the
  ``ParameterStore`` at the variable's offset, so the gradient exchange needs no packing;
* backward visits nodes in reverse order; after each node the arena range whose gradients
  just became final is reported through ``on_grads_ready(lo, hi)`` - the hook the
  bucket scheduler of ``parallel/`` uses to overlap the fused all-reduce+SGD kernel with
  the rest of the backward pass (reference: TF scheduling NcclAllReduce ops as their
  inputs become ready).
"""
from __future__ import annotations

import ctypes
import os
from typing import Callable, Dict, List, Optional

import torch

from ..ops import native
from ..ops.native import ConvDesc, f32, i32, i64, ptr
from .graph import Graph, Node, TensorRef
from .params import ParameterStore

u64 = ctypes.c_ulonglong


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


_TORCH_DTYPE = {"f16": torch.float16, "i32": torch.int32, "f32": torch.float32}


class _Buf:
    """A 2-D fp16 view [rows][cols] with row stride ld (elements)."""
    __slots__ = ("t", "rows", "cols", "ld")

    def __init__(self, t: torch.Tensor, rows: int, cols: int, ld: int):
        self.t, self.rows, self.cols, self.ld = t, rows, cols, ld

    @property
    def p(self):
        return ctypes.c_void_p(self.t.data_ptr())


class CudaExecutor:
    def __init__(self, graph: Graph, store: ParameterStore, nclass: int, aux_weight: float = 0.4,
                 dropout_seed: int = 4321, need_input_grad: bool = False):
        if not torch.cuda.is_available():
            raise RuntimeError("CudaExecutor needs a CUDA device")
        native.load()  # fail loudly if the extension is missing
        self.graph = graph
        self.store = store
        self.device = store.device
        self.nclass = nclass
        self.aux_weight = aux_weight
        self.dropout_seed = int(dropout_seed)
        self.step_index = 0
        self._stream = None
        self._nvtx = os.environ.get("KFB_NVTX") == "1"
        # tools/node_timing.py: when a list, every node appends (tag, node, start, end) events
        self.node_events: Optional[list] = None
        # set by the benchmark driver when --comm_impl=p2p_fused: name -> kfb_push_target
        self.push_lookup: Optional[Callable[[str], object]] = None
        dev = self.device
        n = graph.batch_size

        # ---- buffers ------------------------------------------------------------
        self.buf: Dict[int, _Buf] = {}
        self.gbuf: Dict[int, _Buf] = {}
        self._tmp: Dict[tuple, _Buf] = {}
        self._alias: Dict[int, int] = {}
        # per-step inputs: primary input, extra inputs, labels (in the input source's order);
        # host batches are prefetched into a second set on a copy stream (double buffering)
        self._in_refs = [graph.input] + list(graph.extra_inputs)
        specs = [(t.shape, t.dtype) for t in self._in_refs] + list(graph.label_specs)
        first = [torch.zeros(tuple(shape), dtype=_TORCH_DTYPE[dt], device=dev)
                 for shape, dt in specs]
        self._in_sets = [first, None]
        self._cur = 0
        self._bind_inputs(first)
        self._staged = False
        self.static_inputs = False       # set once a CUDA graph has captured the input addresses
        self._copy_stream = None
        self._ready = [torch.cuda.Event(), torch.cuda.Event()]
        self._free = [torch.cuda.Event(), torch.cuda.Event()]
        self.needs_grad = self._mark_needs_grad(need_input_grad)
        self._slice_of: Dict[int, tuple] = {}
        self._plan_concat_slices()
        for node in graph.nodes:
            self._alloc_output(node)
        # reductions / split-K scratch (sized below), loss outputs
        self._scratch_need = 0
        self._ws_need = 0
        # single-launch batch norm (csrc/bn_fused.cu); KFB_BN_FUSED=0 selects the three-kernel
        # forms of nn_kernels.cu (then the conv epilogue does not produce the statistics either)
        self.bn_fused = os.environ.get("KFB_BN_FUSED", "1") != "0"
        self._bn_max_grid = int(native.load().kfb_bn_fused_max_grid())
        # backward of layers smaller than this (bytes of one activation) and without a packed
        # mask use the three-kernel form: below ~60 MB the single launch loses to it (its two
        # CTAs per SM ramp up more slowly than the 4-8 of the separate kernels; measured on
        # the headline step: 0 -> 11 051, 30 MB -> 11 163, 60 MB -> 11 197, 120 MB -> 11 194 img/s)
        self.bn_fused_bwd_min_bytes = int(os.environ.get("KFB_BN_FUSED_BWD_MIN_BYTES",
                                                         "60000000"))
        # forward with conv-epilogue statistics: layers smaller than this run finalize + apply as
        # two launches instead of the single launch. Measured on the headline step (CUDA-graph
        # replay): threshold 0 / 30 MB / 60 MB / 120 MB / 250 MB / all layers -> 11 773 / 11 945 /
        # 12 034 / 12 144 / 12 180 / 12 237 img/s, so the two launches are the default for every
        # size; the single launch remains for BN layers that reduce their own statistics
        self.bn_fused_fwd_min_bytes = int(os.environ.get("KFB_BN_FUSED_FWD_MIN_BYTES",
                                                         str(1 << 62)))
        self._stats_need = 0
        self.fuse_stats_min_k = int(os.environ.get("KFB_FUSE_STATS_MIN_K", "0"))
        self._stats_for = None
        # fused residual gradient (kfb_conv_dgrad_residual): needs the packed masks of the
        # single-launch BN and the TMA-store epilogue
        self.fuse_res_grad = (self.bn_fused and os.environ.get("KFB_TMA_STORE", "1") != "0" and
                              os.environ.get("KFB_FUSE_RESIDUAL_GRAD", "1") != "0")
        # device-resident training-step counter (mirrors step_index): dropout / drop-path take the
        # step part of their Philox offset from it, so that a captured CUDA graph draws fresh
        # masks in every replay; drop-path keep probabilities (a schedule over the step) live
        # in a device array refreshed before each step
        self.step_dev = torch.zeros((1,), dtype=torch.int64, device=dev)
        self._dp_nodes = [nd for nd in graph.nodes if nd.op == "drop_path"]
        self.dp_keep = torch.ones((max(len(self._dp_nodes), 1),), dtype=torch.float32, device=dev)
        self._dp_slot = {id(nd): i for i, nd in enumerate(self._dp_nodes)}
        self.loss_rows = torch.zeros((2, n), dtype=torch.float32, device=dev)
        self.top = torch.zeros((4, n), dtype=torch.int32, device=dev)
        self.loss_out = torch.zeros((8,), dtype=torch.float32, device=dev)
        self.loss_scale_dev = torch.ones((1,), dtype=torch.float32, device=dev)
        self._per_node: Dict[int, dict] = {}
        for node in graph.nodes:
            self._prepare(node)
        self.scratch = torch.zeros((max(self._scratch_need, 1024),), dtype=torch.float32,
                                   device=dev)
        self.stats_buf = torch.zeros((max(self._stats_need, 1024),), dtype=torch.float32,
                                     device=dev)
        self.bn_sync = torch.zeros((2,), dtype=torch.int32, device=dev)
        self.workspace = torch.zeros((max(self._ws_need // 4, 1024),), dtype=torch.float32,
                                     device=dev)
        self._plan_grad_writes()
        self._plan_residual_grad_fusion()
        if graph.loss["kind"] not in ("softmax_xent", "ncf"):
            from . import losses_cuda
            losses_cuda.prepare(self)

    # ------------------------------------------------------------------ helpers
    def _s(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _mark_needs_grad(self, need_input_grad: bool) -> Dict[int, bool]:
        need = {self.graph.input.id: need_input_grad}
        for t in self.graph.extra_inputs:
            need[t.id] = False
        for node in self.graph.nodes:
            has_param = (any(p.trainable for p in node.params.values()) or
                         node.op == "embedding")  # shared tables: the owner node holds the spec
            need[node.output.id] = has_param or any(need[t.id] for t in node.inputs)
        return need

    def _plan_concat_slices(self) -> None:
        """Concat inputs that can be produced in place: a 4-D tensor written by a batch norm or a
        pooling layer (kernels that take the row stride of their output and of the incoming
        gradient) whose only consumer is the concat becomes a column slice of the concat
        buffer, forward and backward - no copy2d in either direction (the reference's concat is
        a real copy too, T/convnet_builder_legacy.py inception modules -> tf.concat; here the
        branch simply stores into its channel slice). KFB_CONCAT_INPLACE=0 keeps the copies."""
        if os.environ.get("KFB_CONCAT_INPLACE", "1") == "0":
            return
        producer = {node.output.id: node for node in self.graph.nodes}
        for node in self.graph.nodes:
            if node.op != "concat" or len(node.output.shape) != 4 or \
                    node.output.id in self._slice_of:
                continue
            off = 0
            for t in node.inputs:
                src = producer.get(t.id)
                if (src is not None and src.op in ("bn", "pool") and len(t.consumers) == 1 and
                        len(t.shape) == 4 and t.shape[-1] % 8 == 0 and off % 8 == 0 and
                        t is not self.graph.logits and t is not self.graph.aux_logits and
                        node.inputs.count(t) == 1):
                    self._slice_of[t.id] = (node.output.id, off)
                off += t.shape[-1]

    def _concat_buffer(self, tid: int) -> "_Buf":
        if tid not in self.buf:
            t = self.graph.tensors[tid]
            rows, cols = t.rows, t.shape[-1]
            self.buf[tid] = _Buf(torch.zeros((rows, cols), dtype=torch.float16,
                                             device=self.device), rows, cols, cols)
        return self.buf[tid]

    def _alloc_output(self, node: Node) -> None:
        t = node.output
        dev = self.device
        if t.id in self._slice_of:
            parent, off = self._slice_of[t.id]
            y = self._concat_buffer(parent)
            cols = t.shape[-1]
            self.buf[t.id] = _Buf(y.t[:, off:off + cols], y.rows, cols, y.ld)
            return
        if node.op == "concat" and t.id in self.buf:
            return                       # allocated early for its in-place inputs
        if len(t.shape) == 4:
            rows, cols = t.rows, t.shape[-1]
            if cols % 8 != 0:
                raise NotImplementedError(
                    f"{node.name}: {cols} channels - the sm_100a kernels need channel counts "
                    "that are multiples of 8 (use --device=cpu for this model)")
            ld = cols
        else:
            rows, cols = t.rows, t.shape[-1]
            ld = _pad8(cols)
        if node.op == "reshape":
            src = self.buf[node.inputs[0].id]
            if src.t.numel() == rows * ld and cols == ld:
                # pure view of the same storage
                self.buf[t.id] = _Buf(src.t.view(rows, ld), rows, cols, ld)
                self._alias[t.id] = node.inputs[0].id
                return
        is_logits = t is self.graph.logits or t is self.graph.aux_logits
        dt = torch.float32 if is_logits else torch.float16
        if is_logits and node.op != "affine":
            raise NotImplementedError("logits must come from an affine layer")
        self.buf[t.id] = _Buf(torch.zeros((rows, ld), dtype=dt, device=dev), rows, cols, ld)

    def _grad(self, t: TensorRef) -> _Buf:
        """Gradient buffer of tensor t (allocated lazily, aliases follow reshape views)."""
        if t.id in self._alias:
            base = self._grad(self.graph.tensors[self._alias[t.id]])
            b = self.buf[t.id]
            g = _Buf(base.t.view(b.rows, b.ld), b.rows, b.cols, b.ld)
            self.gbuf[t.id] = g
            return g
        if t.id in self._slice_of:
            if t.id not in self.gbuf:
                parent, off = self._slice_of[t.id]
                base = self._grad(self.graph.tensors[parent])
                b = self.buf[t.id]
                self.gbuf[t.id] = _Buf(base.t[:, off:off + b.cols], b.rows, b.cols, base.ld)
            return self.gbuf[t.id]
        if t.id not in self.gbuf:
            b = self.buf[t.id]
            self.gbuf[t.id] = _Buf(torch.zeros((b.rows, b.ld), dtype=torch.float16,
                                               device=self.device), b.rows, b.cols, b.ld)
        return self.gbuf[t.id]

    def _root(self, tid: int) -> int:
        while tid in self._alias:
            tid = self._alias[tid]
        return tid

    def _plan_grad_writes(self) -> None:
        """Static schedule of which backward contribution writes and which accumulates."""
        written = set()
        self._first: Dict[tuple, bool] = {}
        for node in reversed(self.graph.nodes):
            for idx, t in enumerate(node.inputs):
                if not self.needs_grad.get(t.id, False):
                    continue
                root = self._root(t.id)
                self._first[(id(node), idx)] = root not in written
                written.add(root)
        for t in self.graph.tensors:
            if self.needs_grad.get(t.id, False) and (t.consumers or t is self.graph.logits or
                                                     t is self.graph.aux_logits):
                self._grad(t)

    def _plan_residual_grad_fusion(self) -> None:
        """Identity residual blocks: y = relu(bn(conv_k(...)) + r) with r also feeding the block's
        first convolution. In backward the shortcut gradient dy * (y > 0) would be written by the
        BN backward and then read-modify-written by that convolution's data gradient; instead
        the convolution's dgrad epilogue adds dy * mask itself (kfb_conv_dgrad_residual) and the
        tensor is written once. Conditions: r has exactly these two consumers, the convolution
        is unit-stride with >= 64 input channels, the BN wrote a packed mask."""
        if not self.fuse_res_grad:
            return
        order = {id(n): i for i, n in enumerate(self.graph.nodes)}
        for node in self.graph.nodes:
            info = self._per_node[id(node)]
            if node.op != "bn" or "bitmask" not in info or len(node.inputs) < 2:
                continue
            r = node.inputs[1]
            if not self.needs_grad.get(r.id, False) or r.id in self._alias or \
                    len(r.consumers) != 2:
                continue
            others = [c for c in r.consumers if c is not node]
            if len(others) != 1 or others[0].op != "conv" or others[0].inputs[0] is not r:
                continue
            conv = others[0]
            cinfo = self._per_node[id(conv)]
            if cinfo.get("flat") or order[id(conv)] > order[id(node)] or not bool(
                    native.load().kfb_conv_dgrad_residual_ok(ctypes.byref(cinfo["desc"]))):
                continue
            info["res_grad_fused"] = True
            cinfo["res_from"] = node

    def _acc(self, node: Node, idx: int) -> int:
        return 0 if self._first[(id(node), idx)] else 1

    def _need_scratch(self, rows: int, c: int) -> None:
        nb = native.load().kfb_reduce_blocks(ctypes.c_longlong(rows), ctypes.c_int(c))
        self._scratch_need = max(self._scratch_need, nb * 2 * c)

    # ------------------------------------------------------------------ per-node setup
    def _prepare(self, node: Node) -> None:
        info: dict = {}
        a = node.attrs
        st = self.store
        dev = self.device
        if node.op == "conv":
            x = node.inputs[0]
            n, h, w, c = x.shape
            k = node.output.shape[-1]
            spec = node.params["kernel"]
            d = ConvDesc()
            d.n, d.h, d.w, d.c = n, h, w, c
            d.k, d.r, d.s = k, a["kh"], a["kw"]
            d.pad_h_lo, d.pad_h_hi = a["pad_h"]
            d.pad_w_lo, d.pad_w_hi = a["pad_w"]
            d.stride_h, d.stride_w = a["sh"], a["sw"]
            d.p, d.q = node.output.shape[1], node.output.shape[2]
            info["desc"] = d
            info["dgrad_zero_fill"] = bool(
                native.load().kfb_conv_dgrad_needs_zero_fill(ctypes.byref(d)))
            info["flat"] = spec.kind == "conv_kernel_flat"
            info["s2d"] = None
            if info["flat"]:
                from ..ops import stem
                kpad = spec.shape[1]
                info["kpad"] = kpad
                if stem.eligible(d) and not self.needs_grad.get(x.id, False):
                    # 7x7/2 stems: space-to-depth + overlapping-pixel TMA view (ops/stem.py)
                    # instead of the explicit column buffer (976 MB at batch 256)
                    pl = stem.plan(d, kpad)
                    info["s2d"] = pl
                    info["s2d_x"] = torch.zeros(pl.s2d_shape(), dtype=torch.float16, device=dev)
                    info["s2d_w"] = torch.zeros((k, pl.ta, 64), dtype=torch.float16, device=dev)
                    info["s2d_dw"] = torch.zeros((k, pl.ta, 64), dtype=torch.float32, device=dev)
                    fn = native.load().kfb_conv_wgrad_workspace_bytes
                    fn.restype = ctypes.c_longlong
                    self._ws_need = max(self._ws_need, int(fn(ctypes.byref(pl.desc))))
                else:
                    info["col"] = torch.zeros((n * d.p * d.q, kpad), dtype=torch.float16,
                                              device=dev)
                    self._ws_need = max(self._ws_need, self._nt_ws(k, kpad))
            else:
                fn = native.load().kfb_conv_wgrad_workspace_bytes
                fn.restype = ctypes.c_longlong
                self._ws_need = max(self._ws_need, int(fn(ctypes.byref(d))))
            if a["activation"] == "tanh":
                raise NotImplementedError("tanh epilogue is not built for the CUDA executor")
            self._need_scratch(node.output.rows, k)
            # conv -> bn adjacency: the conv epilogue can produce the BN statistics
            idx = self.graph.nodes.index(node)
            nxt = self.graph.nodes[idx + 1] if idx + 1 < len(self.graph.nodes) else None
            # Measured on B200 (profiles/README.md): with the per-lane store epilogue the fused
            # statistics cost 0.4-0.7 % end to end (the epilogue warps were the bottleneck of the
            # store-bound layers); with the TMA-store epilogue they gain 1.4 % on top of its
            # 0.9 %, so they are on for every conv -> bn pair (KFB_FUSE_STATS_MIN_K=<K> restricts
            # them to reductions >= K).
            k_reduction = a["kh"] * a["kw"] * c
            info["fuse_stats"] = bool(
                self.bn_fused and
                nxt is not None and nxt.op == "bn" and nxt.inputs[0] is node.output and
                len(node.output.consumers) == 1 and not a["has_bias"] and
                a["activation"] == "linear" and k_reduction >= self.fuse_stats_min_k)
            if info["fuse_stats"]:
                from ..ops import gemm_conv
                num_kb = ((info["kpad"] + 63) // 64) if info["flat"] else \
                    a["kh"] * a["kw"] * ((c + 63) // 64)
                if info["s2d"] is not None:
                    rows_, bn_ = gemm_conv.conv_stats_layout(info["s2d"].desc)
                elif info["flat"]:
                    rows_, bn_ = gemm_conv.stats_layout(node.output.rows, k, num_kb)
                else:
                    rows_, bn_ = gemm_conv.conv_stats_layout(d)
                info["stats_layout"] = (rows_, bn_)
                self._stats_need = max(self._stats_need, ((k + bn_ - 1) // bn_) * rows_ * 2 * bn_)
        elif node.op == "dwconv":
            x = node.inputs[0]
            n, h, w, c = x.shape
            d = ConvDesc()
            d.n, d.h, d.w, d.c = n, h, w, c
            d.k, d.r, d.s = c, a["kh"], a["kw"]
            d.pad_h_lo, d.pad_h_hi = a["pad_h"]
            d.pad_w_lo, d.pad_w_hi = a["pad_w"]
            d.stride_h, d.stride_w = a["sh"], a["sw"]
            d.p, d.q = node.output.shape[1], node.output.shape[2]
            info["desc"] = d
            fn = native.load().kfb_dwconv_wgrad_workspace_bytes
            fn.restype = ctypes.c_longlong
            self._ws_need = max(self._ws_need, int(fn(ctypes.byref(d))))
        elif node.op == "bn":
            c = node.output.shape[-1]
            info["mean_rstd"] = torch.zeros((2 * c,), dtype=torch.float32, device=dev)
            info["scale_shift"] = torch.zeros((2 * c,), dtype=torch.float32, device=dev)
            info["coef"] = torch.zeros((3 * c,), dtype=torch.float32, device=dev)
            self._need_scratch(node.output.rows, c)
            self._scratch_need = max(self._scratch_need, self._bn_max_grid * 2 * c)
            if self.bn_fused and a["relu"] and len(node.inputs) > 1:
                # packed ReLU mask of the block output: the backward reads one bit per element
                # instead of the fp16 output (twice)
                info["bitmask"] = torch.zeros((node.output.rows * (c // 8),), dtype=torch.uint8,
                                              device=dev)
        elif node.op == "affine":
            spec = node.params["weights"]
            out, in_pad = spec.shape
            self._ws_need = max(self._ws_need, self._nt_ws(out, in_pad))
            self._need_scratch(node.output.shape[0], _pad8(out))
        elif node.op == "pool":
            if a["kind"] == "max":
                o = node.output
                info["argmax"] = torch.zeros((o.rows, o.shape[-1]), dtype=torch.uint8, device=dev)
        elif node.op == "pad_channels":
            if a["before"] % 8 or a["after"] % 8:
                raise NotImplementedError("pad_channels: pads must be multiples of 8")
        elif node.op == "lstm":
            n, t_steps, _ = node.inputs[0].shape
            hid = a["hidden"]
            for d in a["dirs"]:
                info[d] = dict(
                    gates=torch.zeros((n * t_steps, 4 * hid), dtype=torch.float16, device=dev),
                    c=torch.zeros((n * t_steps, hid), dtype=torch.float32, device=dev),
                    hprev=torch.zeros((n * t_steps, hid), dtype=torch.float16, device=dev))
            info["gh"] = torch.zeros((n, 4 * hid), dtype=torch.float16, device=dev)
            info["dh"] = torch.zeros((n, hid), dtype=torch.float16, device=dev)
            info["dc"] = torch.zeros((n, hid), dtype=torch.float32, device=dev)
            info["h0"] = torch.zeros((n, hid), dtype=torch.float16, device=dev)
            for key, spec in node.params.items():
                if key.endswith("_b"):
                    continue
                self._ws_need = max(self._ws_need, self._nt_ws(spec.shape[0], spec.shape[1]))
            self._need_scratch(n * t_steps, 4 * hid)
        elif node.op == "embedding":
            if node.params:
                # dense gradient of the table is scatter-accumulated: zero it every backward
                self._scatter_tables = getattr(self, "_scatter_tables", []) + \
                    [node.params["table"].name]
        elif node.op == "relayout":
            info["index"] = torch.tensor(a["index"], dtype=torch.int32, device=dev)
            info["inverse"] = torch.tensor(a["inverse"], dtype=torch.int32, device=dev)
        elif node.op == "lrn" or node.op == "dropout":
            pass
        self._per_node[id(node)] = info

    @staticmethod
    def _nt_ws(m: int, n: int) -> int:
        bn = 64 if n <= 64 else (128 if n <= 128 else 256)
        tiles = ((m + 127) // 128) * ((n + bn - 1) // bn)
        splits = max(1, (2 * 148) // tiles)
        return m * n * 4 * splits

    # ------------------------------------------------------------------ inputs
    def _bind_inputs(self, tensors) -> None:
        ni = len(self._in_refs)
        for ref, t in zip(self._in_refs, tensors[:ni]):
            if len(ref.shape) >= 2:
                cols = ref.shape[-1]
                rows = t.numel() // cols
            else:
                rows, cols = ref.shape[0], 1
            self.buf[ref.id] = _Buf(t, rows, cols, cols)
        self.images = tensors[0]
        self.label_tensors = list(tensors[ni:])
        self.labels = self.label_tensors[0]

    def set_inputs(self, *tensors: torch.Tensor) -> None:
        """Makes a batch (primary input, *extra inputs, *labels) the input of the next step.
        Device tensors of the right type are used in place (the synthetic device source hands
        over the same cached buffers every step); host tensors are copied on the compute stream
        (see stage_inputs for the overlapped path)."""
        mine = self._in_sets[self._cur]
        if len(tensors) != len(mine):
            raise ValueError("expected %d input tensors, got %d" % (len(mine), len(tensors)))
        if all(t.is_cuda and t.dtype == m.dtype and t.is_contiguous() and t.shape == m.shape
               for t, m in zip(tensors, mine)):
            if self.static_inputs:
                # a captured CUDA graph reads the buffers bound at capture time: device batches
                # that live elsewhere (a ring of device-side preprocessing outputs) are copied in
                bound = [self.buf[r.id].t for r in self._in_refs] + self.label_tensors
                for t, b in zip(tensors, bound):
                    if t.data_ptr() != b.data_ptr():
                        b.copy_(t, non_blocking=True)
                return
            self._bind_inputs(list(tensors))
            return
        for t, m in zip(tensors, mine):
            m.copy_(t, non_blocking=True)
        self._bind_inputs(mine)

    def set_primary_input(self, images: torch.Tensor) -> None:
        """Only the primary input changes (evaluation of models whose labels are not fed)."""
        mine = self._in_sets[self._cur]
        mine[0].copy_(images, non_blocking=True)
        self._bind_inputs(mine)

    def stage_inputs(self, *tensors: torch.Tensor) -> None:
        """Starts the host->device copy of the NEXT step's batch on a copy stream into the back
        buffer set; it overlaps the kernels of the step that is currently enqueued."""
        b = 1 - self._cur
        if self._in_sets[b] is None:
            self._in_sets[b] = [torch.zeros_like(t) for t in self._in_sets[self._cur]]
            self._copy_stream = torch.cuda.Stream(device=self.device)
        cs = self._copy_stream
        cs.wait_event(self._free[b])  # the forward pass that last read this buffer is done
        with torch.cuda.stream(cs):
            for t, m in zip(tensors, self._in_sets[b]):
                m.copy_(t, non_blocking=True)
            self._ready[b].record(cs)
        self._staged = True

    def has_staged_inputs(self) -> bool:
        return self._staged

    def adopt_staged_inputs(self) -> None:
        """CUDA-graph mode: the captured kernels read fixed input buffers, so the staged batch
        is copied device-to-device into the current set instead of swapping the sets."""
        b = 1 - self._cur
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._ready[b])
        for m, t in zip(self._in_sets[self._cur], self._in_sets[b]):
            m.copy_(t, non_blocking=True)
        self._free[b].record(cur)       # the back set may be overwritten once this copy ran
        self._staged = False
        self._bind_inputs(self._in_sets[self._cur])

    def use_staged_inputs(self) -> None:
        b = 1 - self._cur
        torch.cuda.current_stream(self.device).wait_event(self._ready[b])
        self._cur = b
        self._staged = False
        self._bind_inputs(self._in_sets[b])

    # ------------------------------------------------------------------ forward
    def forward(self, train: bool = True) -> None:
        s = self._s()
        if train and self._dp_nodes and not torch.cuda.is_current_stream_capturing():
            self.push_step_values()
        if self.node_events is not None:
            for node in self.graph.nodes:
                self._timed_node("fwd", node, getattr(self, "_fwd_" + node.op), node, train, s)
        elif self._nvtx:
            # KFB_NVTX=1: one NVTX range per node so that profilers attribute kernels to layers
            # (ncu --nvtx --nvtx-include "fwd/<node name>")
            for node in self.graph.nodes:
                torch.cuda.nvtx.range_push("fwd/" + node.name)
                getattr(self, "_fwd_" + node.op)(node, train, s)
                torch.cuda.nvtx.range_pop()
        else:
            for node in self.graph.nodes:
                getattr(self, "_fwd_" + node.op)(node, train, s)
        # the image buffer is only read by the forward pass (backward uses saved activations)
        self._free[self._cur].record(torch.cuda.current_stream(self.device))

    @staticmethod
    def _zero(t: torch.Tensor, s) -> None:
        native.call("fill_zero", ptr(t), i64(t.numel() * t.element_size()), s)

    def _timed_node(self, tag, node, fn, *args):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        native.current_tag = tag + "/" + node.name
        e0.record()
        fn(*args)
        e1.record()
        self.node_events.append((tag, node, e0, e1))

    def _w16(self, name):
        return ctypes.c_void_p(self.store.view(name, "w16").data_ptr())

    def _w32(self, name):
        return ctypes.c_void_p(self.store.view(name, "w32").data_ptr())

    def _g32(self, name):
        return ctypes.c_void_p(self.store.view(name, "grad").data_ptr())

    def _fwd_conv(self, node, train, s):
        info = self._per_node[id(node)]
        a = node.attrs
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        bias = self._w32(node.params["biases"].name) if a["has_bias"] else None
        relu = i32(1 if a["activation"] == "relu" else 0)
        d = info["desc"]
        kname = node.params["kernel"].name
        # the batch-norm that immediately follows gets its statistics from the conv epilogue
        stats = ptr(self.stats_buf) if (train and info.get("fuse_stats")) else None
        if info["s2d"] is not None:
            from ..ops import stem
            pl = info["s2d"]
            stem.regroup(pl, x.t, info["s2d_x"], s)
            stem.pack_weights(pl, self.store.view(kname, "w16"), info["s2d_w"], s)
            native.call("conv_fprop", ctypes.byref(pl.desc), ptr(info["s2d_x"]),
                        ptr(info["s2d_w"]), y.p, i64(y.ld), bias, relu, stats, s)
        elif info["flat"]:
            col = info["col"]
            native.call("im2col_small_c", x.p, ptr(col), i32(d.n), i32(d.h), i32(d.w), i32(d.c),
                        i32(d.p), i32(d.q), i32(d.r), i32(d.s), i32(d.stride_h), i32(d.stride_w),
                        i32(d.pad_h_lo), i32(d.pad_w_lo), i32(info["kpad"]), s)
            native.call("gemm_tn", ptr(col), self._w16(kname), y.p, i32(y.rows), i32(d.k),
                        i32(info["kpad"]), i64(info["kpad"]), i64(info["kpad"]), i64(y.ld), bias,
                        relu, i32(0), stats, s)
        else:
            native.call("conv_fprop", ctypes.byref(d), x.p, self._w16(kname), y.p, i64(y.ld),
                        bias, relu, stats, s)
        self._stats_for = (node.output.id,) + info["stats_layout"] if stats is not None else None

    def _fwd_bn(self, node, train, s):
        info = self._per_node[id(node)]
        a = node.attrs
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        res = self.buf[node.inputs[1].id] if len(node.inputs) > 1 else None
        gamma = self._w32(node.params["gamma"].name) if a["scale"] else None
        beta = self._w32(node.params["beta"].name)
        mm = ptr(self.store.view(node.params["moving_mean"].name))
        mv = ptr(self.store.view(node.params["moving_variance"].name))
        c = y.cols
        if train and self.bn_fused:
            sf, self._stats_for = self._stats_for, None
            fused = sf is not None and sf[0] == node.inputs[0].id
            if fused and x.rows * c * 2 < self.bn_fused_fwd_min_bytes:
                native.call("bn_forward_train_stats", x.p, i64(x.ld), i64(x.rows), i32(c), gamma,
                            beta, mm, mv, f32(a["epsilon"]), f32(a["decay"]), i32(1),
                            ptr(info["mean_rstd"]), ptr(info["scale_shift"]),
                            ptr(self.stats_buf), i32(sf[1]), i32(sf[2]), res.p if res else None,
                            i64(res.ld if res else 0), i32(int(a["relu"])), y.p, i64(y.ld),
                            ptr(info.get("bitmask")), s)
                return
            native.call("bn_forward_train_fused", x.p, i64(x.ld), i64(x.rows), i32(c), gamma, beta,
                        mm, mv, f32(a["epsilon"]), f32(a["decay"]), i32(1),
                        ptr(info["mean_rstd"]), ptr(info["scale_shift"]),
                        ptr(self.stats_buf) if fused else None, i32(sf[1] if fused else 0),
                        i32(sf[2] if fused else 0), ptr(self.scratch), res.p if res else None,
                        i64(res.ld if res else 0), i32(int(a["relu"])), y.p, i64(y.ld),
                        ptr(info.get("bitmask")), ptr(self.bn_sync), s)
        elif train:
            native.call("bn_forward_train", x.p, i64(x.ld), i64(x.rows), i32(c), gamma, beta, mm,
                        mv, f32(a["epsilon"]), f32(a["decay"]), i32(1), ptr(info["mean_rstd"]),
                        ptr(info["scale_shift"]), ptr(self.scratch), i32(0),
                        res.p if res else None, i64(res.ld if res else 0), i32(int(a["relu"])),
                        y.p, i64(y.ld), s)
        else:
            native.call("bn_forward_inference", x.p, i64(x.ld), i64(x.rows), i32(c), gamma, beta,
                        mm, mv, f32(a["epsilon"]), ptr(info["scale_shift"]),
                        res.p if res else None, i64(res.ld if res else 0), i32(int(a["relu"])),
                        y.p, i64(y.ld), s)

    _ACT_KIND = {"relu": 1, "relu6": 2, "tanh": 3, "sigmoid": 4}

    def _fwd_act(self, node, train, s):
        kind = self._ACT_KIND[node.attrs["kind"]]
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        if kind == 1:
            native.call("add_act", x.p, i64(x.ld), None, i64(0), y.p, i64(y.ld), i64(x.rows),
                        i32(_pad8(x.cols)), i32(1), s)
        else:
            native.call("act_forward", x.p, i64(x.ld), y.p, i64(y.ld), i64(x.rows),
                        i32(_pad8(x.cols)), i32(kind), s)

    def _fwd_dwconv(self, node, train, s):
        d = self._per_node[id(node)]["desc"]
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        native.call("dwconv_fprop", ctypes.byref(d), x.p, i64(x.ld),
                    self._w16(node.params["kernel"].name), y.p, i64(y.ld), s)

    def push_step_values(self) -> None:
        """Per-step host scalars into device memory (before the step's first launch, outside a
        captured graph): the drop-path keep probabilities of this step."""
        if self._dp_nodes:
            vals = (ctypes.c_float * len(self._dp_nodes))(
                *[float(nd.attrs["keep_prob_fn"](self.step_index)) for nd in self._dp_nodes])
            native.call("set_floats", ptr(self.dp_keep), vals, i32(len(self._dp_nodes)),
                        self._s())

    def _drop_path_args(self, node):
        kp = float(node.attrs["keep_prob_fn"](self.step_index))
        idx = self.graph.nodes.index(node)
        b = self.buf[node.output.id]
        keep = ctypes.c_void_p(self.dp_keep.data_ptr() + 4 * self._dp_slot[id(node)])
        return (i64(b.rows), i64(b.rows // self.graph.batch_size), i32(_pad8(b.cols)), f32(kp),
                u64(self.dropout_seed), u64(idx * 65536), ptr(self.step_dev), i32(28), keep)

    def _fwd_drop_path(self, node, train, s):
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        if not train:
            native.call("copy2d", x.p, i64(x.ld), y.p, i64(y.ld), i64(x.rows), i32(x.cols),
                        i32(0), s)
            return
        native.call("drop_path_dev", x.p, i64(x.ld), y.p, i64(y.ld), i32(0),
                    *self._drop_path_args(node), s)

    def _fwd_add(self, node, train, s):
        a, b = self.buf[node.inputs[0].id], self.buf[node.inputs[1].id]
        y = self.buf[node.output.id]
        native.call("add_act", a.p, i64(a.ld), b.p, i64(b.ld), y.p, i64(y.ld), i64(a.rows),
                    i32(_pad8(a.cols)), i32(int(node.attrs["relu"])), s)

    def _pool_args(self, node):
        a = node.attrs
        n, h, w, c = node.inputs[0].shape
        p, q = node.output.shape[1], node.output.shape[2]
        return (i32(n), i32(h), i32(w), i32(c), i32(p), i32(q), i32(a["kh"]), i32(a["kw"]),
                i32(a["sh"]), i32(a["sw"]), i32(a["pad_h"][0]), i32(a["pad_w"][0]),
                i32(1 if a["kind"] == "max" else 0))

    def _fwd_pool(self, node, train, s):
        info = self._per_node[id(node)]
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        native.call("pool_forward", x.p, i64(x.ld), y.p, i64(y.ld), ptr(info.get("argmax")),
                    *self._pool_args(node), s)

    def _fwd_spatial_mean(self, node, train, s):
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        n, h, w, c = node.inputs[0].shape
        native.call("spatial_mean_forward", x.p, i64(x.ld), y.p, i64(y.ld), i32(n), i32(h * w),
                    i32(c), s)

    def _fwd_reshape(self, node, train, s):
        if node.output.id in self._alias:
            return
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        per_row = y.cols
        native.call("copy_rows", x.p, i64(per_row), y.p, i64(y.ld), i64(y.rows), i64(per_row),
                    i32(0), s)

    def _fwd_affine(self, node, train, s):
        a = node.attrs
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        spec = node.params["weights"]
        out, in_pad = spec.shape
        native.call("gemm_tn", x.p, self._w16(spec.name), y.p, i32(x.rows), i32(out),
                    i32(spec.logical_cols), i64(x.ld), i64(in_pad), i64(y.ld),
                    self._w32(node.params["biases"].name),
                    i32(1 if a["activation"] == "relu" else 0),
                    i32(int(y.t.dtype == torch.float32)), None, s)

    def _dropout_offset(self, node) -> int:
        idx = self.graph.nodes.index(node)
        return idx * (1 << 32)      # the step part (step << 44) comes from step_dev

    def _fwd_dropout(self, node, train, s):
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        kp = node.attrs["keep_prob"] if train else 1.0
        native.call("dropout_dev", x.p, y.p, i64(x.rows * x.ld), f32(kp), u64(self.dropout_seed),
                    u64(self._dropout_offset(node)), ptr(self.step_dev), i32(44), s)

    def _fwd_lrn(self, node, train, s):
        a = node.attrs
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        native.call("lrn_forward", x.p, y.p, i64(x.rows), i32(x.cols), i32(a["radius"]),
                    f32(a["bias"]), f32(a["alpha"]), f32(a["beta"]), s)

    def _fwd_concat(self, node, train, s):
        y = self.buf[node.output.id]
        off = 0
        for t in node.inputs:
            x = self.buf[t.id]
            if t.id not in self._slice_of:       # in-place inputs already sit in their slice
                native.call("copy2d", x.p, i64(x.ld),
                            ctypes.c_void_p(y.t.data_ptr() + off * 2), i64(y.ld), i64(x.rows),
                            i32(x.cols), i32(0), s)
            off += x.cols

    # LSTM layer = one big input-projection GEMM per direction, then per time step a small
    # recurrent GEMM (batch x 4H x H) and the pointwise cell kernel. Batch-major [N, T, *]
    # buffers are sliced per step through the row stride (ld = T * width), so the two
    # directions write straight into their halves of the concatenated output.
    def _fwd_lstm(self, node, train, s):
        info = self._per_node[id(node)]
        a = node.attrs
        hid = a["hidden"]
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        n, t_steps, _ = node.inputs[0].shape
        e2 = 2  # bytes per half
        for di, d in enumerate(a["dirs"]):
            st = info[d]
            wx, wh, b = (node.params[d + k] for k in ("_wx", "_wh", "_b"))
            gates, cbuf, hprev = st["gates"], st["c"], st["hprev"]
            native.call("gemm_tn", x.p, self._w16(wx.name), ptr(gates), i32(x.rows), i32(4 * hid),
                        i32(wx.logical_cols), i64(x.ld), i64(wx.shape[1]), i64(4 * hid), None,
                        i32(0), i32(0), None, s)
            order = range(t_steps) if d == "fw" else range(t_steps - 1, -1, -1)
            step = 1 if d == "fw" else -1
            ybase = y.t.data_ptr() + di * hid * e2
            for t in order:
                tp = t - step                       # previous step in processing order
                first = tp < 0 or tp >= t_steps
                if first:
                    hp_ptr, hp_ld = info["h0"].data_ptr(), hid
                else:
                    hp_ptr, hp_ld = ybase + tp * y.ld * e2, t_steps * y.ld
                native.call("gemm_tn", ctypes.c_void_p(hp_ptr), self._w16(wh.name),
                            ptr(info["gh"]), i32(n), i32(4 * hid), i32(hid), i64(hp_ld),
                            i64(wh.shape[1]), i64(4 * hid), None, i32(0), i32(0), None, s)
                tn = t + step
                has_next = 0 <= tn < t_steps
                native.call(
                    "lstm_cell_forward",
                    ctypes.c_void_p(gates.data_ptr() + t * 4 * hid * e2), i64(t_steps * 4 * hid),
                    ptr(info["gh"]), i64(4 * hid), self._w32(b.name),
                    None if first else ctypes.c_void_p(cbuf.data_ptr() + tp * hid * 4),
                    i64(t_steps * hid),
                    ctypes.c_void_p(cbuf.data_ptr() + t * hid * 4), i64(t_steps * hid),
                    ctypes.c_void_p(ybase + t * y.ld * e2), i64(t_steps * y.ld),
                    ctypes.c_void_p(hprev.data_ptr() + tn * hid * e2) if has_next else None,
                    i64(t_steps * hid), i32(n), i32(hid), f32(a["forget_bias"]), s)

    def _bwd_lstm(self, node, s):
        info = self._per_node[id(node)]
        a = node.attrs
        hid = a["hidden"]
        xin = node.inputs[0]
        x = self.buf[xin.id]
        gy = self._grad(node.output)
        n, t_steps, _ = xin.shape
        e2 = 2
        need_dx = self.needs_grad[xin.id]
        acc = self._acc(node, 0) if need_dx else 0
        for di, d in enumerate(a["dirs"]):
            st = info[d]
            wx, wh, b = (node.params[d + k] for k in ("_wx", "_wh", "_b"))
            gates, cbuf, hprev = st["gates"], st["c"], st["hprev"]
            self._zero(info["dc"], s)
            step = 1 if d == "fw" else -1
            order = range(t_steps - 1, -1, -1) if d == "fw" else range(t_steps)
            gybase = gy.t.data_ptr() + di * hid * e2
            for k, t in enumerate(order):
                tp = t - step
                first = tp < 0 or tp >= t_steps
                native.call(
                    "lstm_cell_backward",
                    ctypes.c_void_p(gates.data_ptr() + t * 4 * hid * e2), i64(t_steps * 4 * hid),
                    ctypes.c_void_p(gybase + t * gy.ld * e2), i64(t_steps * gy.ld),
                    ptr(info["dh"]) if k > 0 else None, i64(hid),
                    ctypes.c_void_p(cbuf.data_ptr() + t * hid * 4), i64(t_steps * hid),
                    None if first else ctypes.c_void_p(cbuf.data_ptr() + tp * hid * 4),
                    i64(t_steps * hid), ptr(info["dc"]), i32(n), i32(hid), s)
                if not first:
                    # dh_{t-1} = dgates_t . Wh
                    native.call("gemm_nn",
                                ctypes.c_void_p(gates.data_ptr() + t * 4 * hid * e2),
                                self._w16(wh.name), ptr(info["dh"]), i32(n), i32(wh.shape[1]),
                                i32(4 * hid), i64(t_steps * 4 * hid), i64(wh.shape[1]), i64(hid),
                                i32(0), i32(0), s)
            rows = n * t_steps
            # bias gradient = column sums of the gate gradients
            native.call("bias_act_backward", ptr(gates), i64(4 * hid), None, i64(0), i64(rows),
                        i32(4 * hid), None, i64(0), self._g32(b.name), i32(0), f32(1.0),
                        ptr(self.scratch), s)
            ws, wsb = ptr(self.workspace), i64(self.workspace.numel() * 4)
            native.call("gemm_nt", ptr(gates), ptr(hprev), self._g32(wh.name), i32(4 * hid),
                        i32(wh.shape[1]), i32(rows), i64(4 * hid), i64(hid), i64(wh.shape[1]),
                        ws, wsb, i32(0), f32(1.0), s)
            native.call("gemm_nt", ptr(gates), x.p, self._g32(wx.name), i32(4 * hid),
                        i32(wx.shape[1]), i32(rows), i64(4 * hid), i64(x.ld), i64(wx.shape[1]),
                        ws, wsb, i32(0), f32(1.0), s)
            if need_dx:
                gx = self._grad(xin)
                native.call("gemm_nn", ptr(gates), self._w16(wx.name), gx.p, i32(rows),
                            i32(wx.shape[1]), i32(4 * hid), i64(4 * hid), i64(wx.shape[1]),
                            i64(gx.ld), i32(0), i32(1 if (acc or di > 0) else 0), s)

    def _fwd_embedding(self, node, train, s):
        a = node.attrs
        ids, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        native.call("embedding_gather", self._w16(a["table"]), i64(a["dim"]), i32(a["col_lo"]),
                    ids.p, i64(a["num_rows"]), y.p, i64(y.ld), i64(y.rows), i32(y.cols), s)

    def _bwd_embedding(self, node, s):
        a = node.attrs
        ids, gy = self.buf[node.inputs[0].id], self._grad(node.output)
        native.call("embedding_scatter_add", gy.p, i64(gy.ld), ids.p, i64(a["num_rows"]),
                    self._g32(a["table"]), i64(a["dim"]), i32(a["col_lo"]), i64(gy.rows),
                    i32(gy.cols), s)

    def _fwd_mul(self, node, train, s):
        a, b = self.buf[node.inputs[0].id], self.buf[node.inputs[1].id]
        y = self.buf[node.output.id]
        native.call("mul_forward", a.p, i64(a.ld), b.p, i64(b.ld), y.p, i64(y.ld), i64(a.rows),
                    i32(_pad8(a.cols)), s)

    def _bwd_mul(self, node, s):
        gy = self._grad(node.output)
        for idx, t in enumerate(node.inputs):
            if not self.needs_grad[t.id]:
                continue
            other = self.buf[node.inputs[1 - idx].id]
            gx = self._grad(t)
            native.call("mul_backward", gy.p, i64(gy.ld), other.p, i64(other.ld), gx.p,
                        i64(gx.ld), i32(self._acc(node, idx)), i64(gy.rows),
                        i32(_pad8(gy.cols)), s)

    def _fwd_relayout(self, node, train, s):
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        native.call("channel_gather", x.p, i64(x.ld), y.p, i64(y.ld), i32(0), i64(x.rows),
                    i32(y.cols), ptr(self._per_node[id(node)]["index"]), s)

    def _bwd_relayout(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        native.call("channel_gather", gy.p, i64(gy.ld), gx.p, i64(gx.ld),
                    i32(self._acc(node, 0)), i64(gy.rows), i32(gx.cols),
                    ptr(self._per_node[id(node)]["inverse"]), s)

    def _fwd_pad_channels(self, node, train, s):
        x, y = self.buf[node.inputs[0].id], self.buf[node.output.id]
        off = node.attrs["before"]
        native.call("copy2d", x.p, i64(x.ld), ctypes.c_void_p(y.t.data_ptr() + off * 2), i64(y.ld),
                    i64(x.rows), i32(x.cols), i32(0), s)

    # ------------------------------------------------------------------ loss + backward
    def loss_backward(self, loss_scale: float = 1.0, backward: bool = True,
                      on_grads_ready: Optional[Callable[[int, int], None]] = None,
                      dyn_scale: Optional[torch.Tensor] = None) -> None:
        """Softmax-xent (+0.4 aux) and the whole backward pass. Gradients of
        ``loss * loss_scale`` land in ``store.grad``; ``dyn_scale`` (device scalar) multiplies
        loss_scale for device-resident automatic loss scaling."""
        s = self._s()
        g = self.graph
        n = g.batch_size
        kind = g.loss["kind"]
        heads = [(g.logits, 1.0, 0)] if kind == "softmax_xent" else []
        if kind == "ncf":
            lb = self.buf[g.logits.id]
            gb = self._grad(g.logits) if backward else None
            native.call("ncf_loss", lb.p, i64(lb.ld), ptr(self.labels), i32(n),
                        f32(loss_scale / n), ptr(dyn_scale), gb.p if gb else None,
                        i64(gb.ld if gb else 0), ptr(self.loss_out), s)
        elif kind != "softmax_xent":
            from . import losses_cuda
            losses_cuda.run(self, kind, loss_scale, dyn_scale, backward, s)
        if g.aux_logits is not None:
            heads.append((g.aux_logits, self.aux_weight, 1))
        for t, weight, slot in heads:
            lb = self.buf[t.id]
            gb = self._grad(t) if backward else None
            native.call("softmax_xent", lb.p, i64(lb.ld), ptr(self.labels), i32(n),
                        i32(self.nclass), f32(loss_scale * weight / n), ptr(dyn_scale),
                        gb.p if gb else None, i64(lb.ld), ptr(self.loss_rows[slot]),
                        ptr(self.top[2 * slot]), ptr(self.top[2 * slot + 1]),
                        ctypes.c_void_p(self.loss_out.data_ptr() + 16 * slot), s)
        if not backward:
            return
        for name in getattr(self, "_scatter_tables", []):
            self._zero(self.store.view(name, "grad"), s)
        for node in reversed(g.nodes):
            if not self.needs_grad[node.output.id]:
                continue
            if self.node_events is not None:
                self._timed_node("bwd", node, getattr(self, "_bwd_" + node.op), node, s)
            elif self._nvtx:
                torch.cuda.nvtx.range_push("bwd/" + node.name)
                getattr(self, "_bwd_" + node.op)(node, s)
                torch.cuda.nvtx.range_pop()
            else:
                getattr(self, "_bwd_" + node.op)(node, s)
            if on_grads_ready is not None and node.params:
                lo, hi = self.store.span([p.name for p in node.params.values() if p.trainable])
                if hi > lo:
                    on_grads_ready(lo, hi)
        native.call("step_advance", ptr(self.step_dev), s)
        self.step_index += 1

    def read_loss(self):
        """(loss, top1 hits, top5 hits) of the last step; synchronises the stream."""
        v = self.loss_out.tolist()
        loss = v[0] + (self.aux_weight * v[4] if self.graph.aux_logits is not None else 0.0)
        return loss, int(v[1]), int(v[2])

    def _bias_act_bwd(self, node, gy: _Buf, y: _Buf, relu: bool, bias_name: Optional[str], s,
                      cols: int) -> None:
        """In place: gy <- gy * (y > 0); dbias <- column sums."""
        if not relu and bias_name is None:
            return
        native.call("bias_act_backward", gy.p, i64(gy.ld), y.p if relu else None, i64(y.ld),
                    i64(gy.rows), i32(cols), gy.p if relu else None, i64(gy.ld),
                    self._g32(bias_name) if bias_name else None, i32(0), f32(1.0),
                    ptr(self.scratch), s)

    def _bwd_conv(self, node, s):
        info = self._per_node[id(node)]
        a = node.attrs
        xin = node.inputs[0]
        x, y = self.buf[xin.id], self.buf[node.output.id]
        gy = self._grad(node.output)
        d = info["desc"]
        kname = node.params["kernel"].name
        self._bias_act_bwd(node, gy, y, a["activation"] == "relu",
                           node.params["biases"].name if a["has_bias"] else None, s, y.cols)
        if info["flat"]:
            if self.needs_grad[xin.id]:
                raise NotImplementedError("input gradient of a tiny-C stem conv is not built")
            if info["s2d"] is not None:
                from ..ops import stem
                pl = info["s2d"]
                native.call("conv_wgrad", ctypes.byref(pl.desc), ptr(info["s2d_x"]), gy.p,
                            i64(gy.ld), ptr(info["s2d_dw"]), ptr(self.workspace),
                            i64(self.workspace.numel() * 4), i32(0), f32(1.0), s)
                stem.unpack_grad(pl, info["s2d_dw"], self.store.view(kname, "grad"), False, 1.0, s)
                return
            col = info["col"]
            push = self.push_lookup(kname) if self.push_lookup else None
            if push is not None:
                native.call("gemm_nt_rs", gy.p, ptr(col), i32(d.k), i32(info["kpad"]),
                            i32(gy.rows), i64(gy.ld), i64(info["kpad"]), ptr(self.workspace),
                            i64(self.workspace.numel() * 4), f32(1.0), ctypes.byref(push), s)
                return
            native.call("gemm_nt", gy.p, ptr(col), self._g32(kname), i32(d.k), i32(info["kpad"]),
                        i32(gy.rows), i64(gy.ld), i64(info["kpad"]), i64(info["kpad"]),
                        ptr(self.workspace), i64(self.workspace.numel() * 4), i32(0), f32(1.0), s)
            return
        if self.needs_grad[xin.id] and info.get("res_from") is not None:
            # dx = dgrad + (gradient of the block output) * (packed ReLU mask of the block output)
            bn = info["res_from"]
            gout = self._grad(bn.output)
            bits = self._per_node[id(bn)]["bitmask"]
            gx = self._grad(xin)
            native.call("conv_dgrad_residual", ctypes.byref(d), gy.p, self._w16(kname), gx.p,
                        gout.p, i64(gout.ld), ptr(bits), i64(gout.cols // 8), s)
        elif self.needs_grad[xin.id]:
            gx = self._grad(xin)
            acc = self._acc(node, 0)
            if info["dgrad_zero_fill"] and not acc:
                self._zero(gx.t, s)
            native.call("conv_dgrad", ctypes.byref(d), gy.p, self._w16(kname), gx.p, i32(acc), s)
        push = self.push_lookup(kname) if self.push_lookup else None
        if push is not None:
            # fused GEMM + reduce-scatter: tiles go straight to the owning rank over NVLink
            native.call("conv_wgrad_rs", ctypes.byref(d), x.p, gy.p, i64(gy.ld),
                        ptr(self.workspace), i64(self.workspace.numel() * 4), f32(1.0),
                        ctypes.byref(push), s)
            return
        native.call("conv_wgrad", ctypes.byref(d), x.p, gy.p, i64(gy.ld), self._g32(kname),
                    ptr(self.workspace), i64(self.workspace.numel() * 4), i32(0), f32(1.0), s)

    def _bwd_bn(self, node, s):
        info = self._per_node[id(node)]
        a = node.attrs
        zin = node.inputs[0]
        z, y = self.buf[zin.id], self.buf[node.output.id]
        gy = self._grad(node.output)
        gamma = self._w32(node.params["gamma"].name) if a["scale"] else None
        dgamma = self._g32(node.params["gamma"].name) if a["scale"] else None
        dbeta = self._g32(node.params["beta"].name)
        has_res = len(node.inputs) > 1
        gres = self._grad(node.inputs[1]) if has_res and self.needs_grad[node.inputs[1].id] else None
        if info.get("res_grad_fused"):
            gres = None        # the block's first convolution adds the shortcut gradient itself
        res_acc = self._acc(node, 1) if gres is not None else 0
        need_dz = self.needs_grad[zin.id]
        if need_dz and self._acc(node, 0):
            gz = self._tmp_like(z)
        else:
            gz = self._grad(zin) if need_dz else self._tmp_like(z)
        # plain BN+ReLU: the mask is recomputed from z (no need to re-read y); with a residual
        # the stored output is the only source of the mask
        mask = y.p if (a["relu"] and has_res) else None
        affine = ptr(info["scale_shift"]) if (a["relu"] and not has_res) else None
        bits = info.get("bitmask")
        if self.bn_fused and (bits is not None or
                              z.rows * z.cols * 2 >= self.bn_fused_bwd_min_bytes):
            native.call("bn_backward_fused", gy.p, i64(gy.ld), None if bits is not None else mask,
                        i64(y.ld), ptr(bits), affine, z.p, i64(z.ld), i64(z.rows), i32(z.cols),
                        gamma, ptr(info["mean_rstd"]), dgamma, dbeta, i32(0), f32(1.0),
                        ptr(info["coef"]), ptr(self.scratch), gz.p, i64(gz.ld),
                        gres.p if gres else None, i64(gres.ld if gres else 0), i32(res_acc),
                        ptr(self.bn_sync), s)
        else:
            native.call("bn_backward", gy.p, i64(gy.ld), mask, i64(y.ld), affine, z.p,
                        i64(z.ld), i64(z.rows), i32(z.cols), gamma, ptr(info["mean_rstd"]),
                        dgamma, dbeta, i32(0), f32(1.0), ptr(info["coef"]), ptr(self.scratch),
                        gz.p, i64(gz.ld), gres.p if gres else None,
                        i64(gres.ld if gres else 0), i32(res_acc), s)
        if need_dz and self._acc(node, 0):
            gdst = self._grad(zin)
            native.call("copy2d", gz.p, i64(gz.ld), gdst.p, i64(gdst.ld), i64(gz.rows),
                        i32(gz.cols), i32(1), s)

    def _tmp_like(self, b: _Buf) -> _Buf:
        """Scratch activation-gradient buffer (one per distinct shape, owned by this executor)."""
        key = (b.rows, b.ld)
        if key not in self._tmp:
            self._tmp[key] = _Buf(torch.zeros((b.rows, b.ld), dtype=torch.float16,
                                              device=self.device), b.rows, b.cols, b.ld)
        t = self._tmp[key]
        return _Buf(t.t, b.rows, b.cols, b.ld)

    def _bwd_act(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        y, gy, gx = self.buf[node.output.id], self._grad(node.output), self._grad(xin)
        kind = self._ACT_KIND[node.attrs["kind"]]
        if kind == 1:
            native.call("relu_backward", gy.p, i64(gy.ld), y.p, i64(y.ld), gx.p, i64(gx.ld),
                        i32(self._acc(node, 0)), i64(gy.rows), i32(_pad8(gy.cols)), s)
        else:
            native.call("act_backward", gy.p, i64(gy.ld), y.p, i64(y.ld), gx.p, i64(gx.ld),
                        i32(self._acc(node, 0)), i64(gy.rows), i32(_pad8(gy.cols)), i32(kind), s)

    def _bwd_dwconv(self, node, s):
        d = self._per_node[id(node)]["desc"]
        xin = node.inputs[0]
        x, gy = self.buf[xin.id], self._grad(node.output)
        kname = node.params["kernel"].name
        if self.needs_grad[xin.id]:
            gx = self._grad(xin)
            native.call("dwconv_dgrad", ctypes.byref(d), gy.p, i64(gy.ld), self._w16(kname), gx.p,
                        i64(gx.ld), i32(self._acc(node, 0)), s)
        native.call("dwconv_wgrad", ctypes.byref(d), x.p, i64(x.ld), gy.p, i64(gy.ld),
                    self._g32(kname), ptr(self.workspace), i64(self.workspace.numel() * 4),
                    i32(0), f32(1.0), s)

    def _bwd_drop_path(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        native.call("drop_path_dev", gy.p, i64(gy.ld), gx.p, i64(gx.ld), i32(self._acc(node, 0)),
                    *self._drop_path_args(node), s)

    def _bwd_add(self, node, s):
        y, gy = self.buf[node.output.id], self._grad(node.output)
        for idx, t in enumerate(node.inputs):
            if not self.needs_grad[t.id]:
                continue
            gx = self._grad(t)
            native.call("relu_backward", gy.p, i64(gy.ld), y.p if node.attrs["relu"] else None,
                        i64(y.ld), gx.p, i64(gx.ld), i32(self._acc(node, idx)), i64(gy.rows),
                        i32(_pad8(gy.cols)), s)

    def _bwd_pool(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        info = self._per_node[id(node)]
        gy, gx = self._grad(node.output), self._grad(xin)
        args = self._pool_args(node)
        native.call("pool_backward", gy.p, i64(gy.ld), ptr(info.get("argmax")), gx.p, i64(gx.ld),
                    i32(self._acc(node, 0)), *args, s)

    def _bwd_spatial_mean(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        n, h, w, c = xin.shape
        native.call("spatial_mean_backward", gy.p, i64(gy.ld), gx.p, i64(gx.ld),
                    i32(self._acc(node, 0)), i32(n), i32(h * w), i32(c), s)

    def _bwd_reshape(self, node, s):
        xin = node.inputs[0]
        if node.output.id in self._alias or not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        native.call("copy_rows", gy.p, i64(gy.ld), gx.p, i64(gy.cols), i64(gy.rows), i64(gy.cols),
                    i32(self._acc(node, 0)), s)

    def _bwd_affine(self, node, s):
        a = node.attrs
        xin = node.inputs[0]
        x, y = self.buf[xin.id], self.buf[node.output.id]
        gy = self._grad(node.output)
        spec = node.params["weights"]
        out, in_pad = spec.shape
        relu = a["activation"] == "relu"
        self._bias_act_bwd(node, gy, y, relu, node.params["biases"].name, s, gy.ld)
        if self.needs_grad[xin.id]:
            gx = self._grad(xin)
            native.call("gemm_nn", gy.p, self._w16(spec.name), gx.p, i32(gy.rows), i32(in_pad),
                        i32(out), i64(gy.ld), i64(in_pad), i64(gx.ld), i32(0),
                        i32(self._acc(node, 0)), s)
        push = self.push_lookup(spec.name) if self.push_lookup else None
        if push is not None:
            native.call("gemm_nt_rs", gy.p, x.p, i32(out), i32(in_pad), i32(gy.rows), i64(gy.ld),
                        i64(x.ld), ptr(self.workspace), i64(self.workspace.numel() * 4), f32(1.0),
                        ctypes.byref(push), s)
        else:
            native.call("gemm_nt", gy.p, x.p, self._g32(spec.name), i32(out), i32(in_pad),
                        i32(gy.rows), i64(gy.ld), i64(x.ld), i64(in_pad), ptr(self.workspace),
                        i64(self.workspace.numel() * 4), i32(0), f32(1.0), s)

    def _bwd_dropout(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        dst = gx if not self._acc(node, 0) else self._tmp_like(gy)
        native.call("dropout_dev", gy.p, dst.p, i64(gy.rows * gy.ld),
                    f32(node.attrs["keep_prob"]), u64(self.dropout_seed),
                    u64(self._dropout_offset(node)), ptr(self.step_dev), i32(44), s)
        if dst is not gx:
            native.call("copy2d", dst.p, i64(dst.ld), gx.p, i64(gx.ld), i64(gx.rows),
                        i32(_pad8(gx.cols)), i32(1), s)

    def _bwd_lrn(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        a = node.attrs
        x, gy, gx = self.buf[xin.id], self._grad(node.output), self._grad(xin)
        native.call("lrn_backward", x.p, gy.p, gx.p, i32(self._acc(node, 0)), i64(x.rows),
                    i32(x.cols), i32(a["radius"]), f32(a["bias"]), f32(a["alpha"]),
                    f32(a["beta"]), s)

    def _bwd_concat(self, node, s):
        gy = self._grad(node.output)
        off = 0
        for idx, t in enumerate(node.inputs):
            cols = self.buf[t.id].cols
            if self.needs_grad[t.id] and t.id not in self._slice_of:
                gx = self._grad(t)
                native.call("copy2d", ctypes.c_void_p(gy.t.data_ptr() + off * 2), i64(gy.ld),
                            gx.p, i64(gx.ld), i64(gx.rows), i32(cols),
                            i32(self._acc(node, idx)), s)
            off += cols

    def _bwd_pad_channels(self, node, s):
        xin = node.inputs[0]
        if not self.needs_grad[xin.id]:
            return
        gy, gx = self._grad(node.output), self._grad(xin)
        off = node.attrs["before"]
        native.call("copy2d", ctypes.c_void_p(gy.t.data_ptr() + off * 2), i64(gy.ld), gx.p,
                    i64(gx.ld), i64(gx.rows), i32(gx.cols), i32(self._acc(node, 0)), s)

    def rs_capable_params(self) -> List[str]:
        """Variables whose weight-gradient kernel can scatter its result to the owning ranks
        (kfb_conv_wgrad_rs / kfb_gemm_nt_rs): conv kernels and FC weights."""
        names = []
        for node in self.graph.nodes:
            if node.op == "conv":
                if self._per_node[id(node)].get("s2d") is None:   # the s2d stem unpacks locally
                    names.append(node.params["kernel"].name)
            elif node.op == "affine":
                names.append(node.params["weights"].name)
        return names

    # ------------------------------------------------------------------ inspection
    def logits(self) -> torch.Tensor:
        b = self.buf[self.graph.logits.id]
        return b.t[:, :b.cols].reshape(self.graph.logits.shape)

    def value(self, tensor_id: int) -> torch.Tensor:
        b = self.buf[tensor_id]
        t = self.graph.tensors[tensor_id]
        return b.t[:, :b.cols].reshape(t.shape)

    def grad_value(self, tensor_id: int) -> torch.Tensor:
        b = self.gbuf[tensor_id]
        t = self.graph.tensors[tensor_id]
        return b.t[:, :b.cols].reshape(t.shape)
