"""BenchmarkCNN: the benchmark driver (training loop, evaluation, logging, checkpoints).

Capabilities and text output follow the reference driver (T/benchmark_cnn.py:1233-2391):
the same flag-combination rules, header block, ``images/sec: m +/- u (jitter = j)`` step
lines, ``total images/sec`` footer and returned stats dict. What differs is the
machinery: one process per GPU, a static kernel schedule instead of ``sess.run``,
CUDA-event timing (max over ranks for the totals) instead of host wall-clock, and the
fused NVLink update kernels instead of TF's gradient-aggregation graph.
"""
from __future__ import annotations

import os
import time
from typing import Dict, List

import numpy as np
import torch

from . import ckpt as ckpt_lib
from . import datasets, lr as lr_mod
from .constants import BenchmarkMode
from .models import model_config
from .nn.params import ParameterStore
from .parallel import comm as comm_lib
from .parallel.strategy import UpdateEngine
from .params import IGNORED_ON_B200, UNSUPPORTED_WHEN_SET, make_params
from .utils import log_util
from .utils.log_util import log_fn  # noqa: F401  (re-exported, monkey-patchable via log_util)

_NUM_STEPS_TO_PROFILE = 10


def _log(msg: str) -> None:
    log_util.log_fn(msg)


def get_mode_from_params(params) -> str:
    if params.forward_only and params.eval:
        raise ValueError("Only one of forward_only and eval parameters is true")
    if params.eval:
        return BenchmarkMode.EVAL
    if params.forward_only:
        return BenchmarkMode.FORWARD_ONLY
    if (params.eval_during_training_every_n_steps or params.eval_during_training_every_n_epochs or
            params.eval_during_training_at_specified_steps or
            params.eval_during_training_at_specified_epochs):
        return BenchmarkMode.TRAIN_AND_EVAL
    return BenchmarkMode.TRAIN


def setup(params):
    """Process-level setup (T/benchmark_cnn.py:3356-3395): thread counts, unsupported-flag
    checks. Returns the (possibly adjusted) params."""
    for name, why in UNSUPPORTED_WHEN_SET.items():
        if getattr(params, name):
            raise ValueError("--%s: %s" % (name, why))
    if params.num_intra_threads:
        torch.set_num_threads(int(params.num_intra_threads))
    _apply_cluster_role(params)
    return params


def _apply_cluster_role(params) -> None:
    """TF cluster roles (T/benchmark_cnn.py:1354-1406, T/cnn_util.py:201-251) on a runtime that
    has one process per GPU and no parameter-server processes:

    * ``--job_name=worker --task_index=i --worker_hosts=h0:p0,h1:p1,...`` - this process is rank
      ``i`` of ``len(worker_hosts)``; the first worker's host:port is the rendezvous address.
      Nothing is changed when a launcher (torchrun / parallel.launcher) already set RANK.
    * ``--job_name=ps`` / ``controller`` - there is nothing for these roles to do (variables are
      replicated and updated by the fused all-reduce kernels; there is no separate graph-owning
      controller): the process logs that and exits 0, so reference launch scripts keep working.
    """
    role = params.job_name
    if not role:
        return
    if role in ("ps", "controller"):
        _log("--job_name=%s: this runtime has no %s processes (variables live with the workers); "
             "nothing to do, exiting." % (role, "parameter-server" if role == "ps"
                                          else "controller"))
        raise SystemExit(0)
    hosts = [h for h in (params.worker_hosts or "").split(",") if h]
    if not hosts:
        raise ValueError("--job_name=worker needs --worker_hosts")
    if not 0 <= params.task_index < len(hosts):
        raise ValueError("--task_index=%d out of range for %d worker_hosts" %
                         (params.task_index, len(hosts)))
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ:
        return
    host, _, port = hosts[0].partition(":")
    os.environ["RANK"] = str(params.task_index)
    os.environ["LOCAL_RANK"] = str(params.task_index)
    os.environ["WORLD_SIZE"] = str(len(hosts))
    os.environ.setdefault("MASTER_ADDR", host or "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", port or "29500")


class BenchmarkCNN:
    def __init__(self, params, dataset=None, model=None):
        self.params = params
        p = params
        self.dataset = dataset or datasets.create_dataset(p.data_dir, p.data_name)
        self.model = model or model_config.get_model_config(p.model, self.dataset, p)
        a = p.autotune_threshold if p.autotune_threshold else 1
        self.num_warmup_batches = (p.num_warmup_batches if p.num_warmup_batches is not None
                                   else max(10, 5 * a * a))
        self.num_gpus = p.num_gpus
        self.device_kind = p.device.lower()
        self._validate()
        if p.batch_size > 0:
            self.model.set_batch_size(p.batch_size)
        self.mode = get_mode_from_params(p)
        # ---- process model: one rank per device --------------------------------
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        if p.variable_update in ("kungfu", "horovod") or p.job_name == "worker":
            # one GPU per worker process (KungFu / Horovod / TF-cluster style launches)
            if p.job_name == "worker" and p.num_gpus != 1:
                raise ValueError("--job_name=worker: start one worker process per GPU "
                                 "(--num_gpus=1) on this runtime")
            self.num_workers = self.world      # num_gpus == 1 per worker
            self.replicas_per_worker = 1
        else:
            self.num_workers = 1
            self.replicas_per_worker = max(self.world, 1)
            if self.world > 1 and self.world != p.num_gpus:
                raise ValueError("launched %d ranks but --num_gpus=%d" % (self.world, p.num_gpus))
        self.batch_size_per_device = self.model.get_batch_size()
        self.batch_size = self.batch_size_per_device * self.num_gpus  # per worker, as reference
        if self.device_kind == "gpu":
            if not p.use_fp16:
                # The reference's default precision (T/benchmark_cnn.py:464-466) has no device
                # path here: every sm_100a kernel stores activations as fp16 and feeds the tensor
                # cores fp16 operands (kind::f16). Running that under an fp32 label would make
                # the numbers look like something they are not, so the combination is refused.
                raise ValueError(
                    "--use_fp16=false is not available with --device=gpu: the B200 kernels keep "
                    "activations and tensor-core operands in fp16 (fp32 accumulation, fp32 master "
                    "weights, optimizer state and batch-norm statistics). Pass --use_fp16, or "
                    "--device=cpu --data_format=NHWC for full fp32 arithmetic.")
            if not torch.cuda.is_available():
                raise RuntimeError("--device=gpu but no CUDA device is visible "
                                   "(use --device=cpu --data_format=NHWC for the CPU path)")
            idx = self.local_rank
            if p.gpu_indices:
                idx = [int(x) for x in p.gpu_indices.split(",")][self.local_rank]
            self.device = torch.device("cuda", idx)
            torch.cuda.set_device(self.device)
        else:
            self.device = torch.device("cpu")
        self.enable_auto_loss_scale = p.use_fp16 and p.fp16_enable_auto_loss_scale
        self.num_batches, self.num_epochs = lr_mod.get_num_batches_and_epochs(
            p, self.batch_size * self.num_workers,
            self.dataset.num_examples_per_epoch("train"))
        self.eval_steps = self._eval_schedule()
        self.comm = None
        self.store = None
        self.executor = None
        self.engine = None
        self.global_step = 0
        self.seed = p.tf_random_seed
        if p.variable_update in ("kungfu", "horovod"):
            self.seed += self.rank  # T/benchmark_cnn.py:2625-2636

    # ------------------------------------------------------------------ validation
    def _validate(self) -> None:
        """Flag-combination rules of T/benchmark_cnn.py:1268-1352."""
        p = self.params
        if p.device.lower() == "cpu" and p.data_format == "NCHW" and not p.mkl:
            raise ValueError("device=cpu requires that data_format=NHWC")
        if ((p.num_epochs_per_decay or p.learning_rate_decay_factor) and
                not (p.init_learning_rate is not None and p.num_epochs_per_decay and
                     p.learning_rate_decay_factor)):
            raise ValueError("If one of num_epochs_per_decay or learning_rate_decay_factor is "
                             "set, both must be set and learning_rate must be set")
        if (p.minimum_learning_rate and
                not (p.init_learning_rate is not None and p.num_epochs_per_decay and
                     p.learning_rate_decay_factor)):
            raise ValueError("minimum_learning_rate requires learning_rate, "
                             "num_epochs_per_decay, and learning_rate_decay_factor to be set")
        if (p.use_fp16 and p.fp16_vars and "replicated" in p.variable_update and
                p.all_reduce_spec and "nccl" in p.all_reduce_spec):
            raise ValueError("fp16 variables are not supported with NCCL")
        if p.use_fp16 and p.fp16_vars and p.gradient_repacking:
            raise ValueError("--fp16_vars cannot be used with --gradient_repacking")
        if p.variable_update == "horovod" and p.num_gpus > 1:
            raise ValueError("Horovod benchmarks require num_gpus=1 on each worker")
        if p.variable_update == "horovod" and p.job_name:
            raise ValueError("job_name should not be specified for Horovod.")
        if p.variable_update == "kungfu" and p.num_gpus > 1:
            raise ValueError("KungFu benchmarks require num_gpus=1 on each worker")
        if p.variable_update == "kungfu" and p.job_name:
            raise ValueError("job_name should not be specified for KungFu.")
        if p.use_fp16 and p.fp16_enable_auto_loss_scale:
            if p.all_reduce_spec and "nccl" in p.all_reduce_spec:
                raise ValueError("Automatic loss scaling is not supported with NCCL.")
            if p.variable_update not in ("parameter_server", "replicated", "independent"):
                raise ValueError("Automatic loss scaling is not supported with "
                                 "variable_update=%s." % p.variable_update)
            if p.staged_vars:
                raise ValueError("Automatic loss scaling is not supported with staged_vars.")
            if p.variable_consistency == "relaxed":
                # the deferred update would apply a sum whose finite check and loss scale
                # belong to another step
                raise ValueError("Automatic loss scaling is not supported with "
                                 "variable_consistency=relaxed.")
        if p.debugger is not None and p.debugger != "cli" and ":" not in p.debugger:
            raise ValueError('--debugger must be "cli" or in the form host:port')
        if p.hierarchical_copy and p.num_gpus <= 1:
            raise ValueError("--hierarchical_copy requires --num_gpus to be greater than 1")
        if p.save_model_secs and p.save_model_steps:
            raise ValueError("At most one of --save_model_secs and --save_model_steps can be "
                             "specified")
        ev = [bool(p.eval_during_training_every_n_steps),
              bool(p.eval_during_training_every_n_epochs),
              bool(p.eval_during_training_at_specified_steps),
              bool(p.eval_during_training_at_specified_epochs)]
        if ev.count(True) > 1:
            raise ValueError("At most one flag with --eval_during_training_* prefix must be "
                             "specified.")
        if any(ev):
            if p.eval:
                raise ValueError("At most one of --eval and --eval_during_training_* must be "
                                 "specified")
            if p.forward_only:
                raise ValueError("At most one of --forward_only and --eval_during_training_* "
                                 "must be specified")
            if p.job_name:
                raise ValueError("--eval_during_training_* is not yet supported in distributed "
                                 "mode.")
            if p.staged_vars:
                raise ValueError("--eval_during_training_* is not currently compatible with "
                                 "staged_vars")
        if p.stop_at_top_1_accuracy and not any(ev):
            raise ValueError("--stop_at_top_1_accuracy is only supported with "
                             "--eval_during_training_*")
        if p.forward_only and p.freeze_when_forward_only:
            if p.train_dir is not None:
                raise ValueError("In forward_only mode, when --freeze_when_forward_only is True, "
                                 "--train_dir should not be specified")
        elif p.trt_mode:
            raise ValueError("--trt_mode should not be specified if one of --forward_only and "
                             "--freeze_when_forward_only is set to False")
        if p.staged_vars and p.variable_update != "parameter_server":
            raise ValueError("staged_vars for now is only supported with "
                             "variable_update=parameter_server")
        if p.all_reduce_spec:
            from .parallel import allreduce_spec
            allreduce_spec.parse_all_reduce_spec(p.all_reduce_spec)  # syntax check

    def _eval_schedule(self):
        p = self.params
        per_epoch = self.dataset.num_examples_per_epoch("train") / float(
            self.batch_size * max(self.num_workers, 1))
        if p.eval_during_training_every_n_steps:
            n = p.eval_during_training_every_n_steps
            return set(range(n, self.num_batches + 1, n))
        if p.eval_during_training_every_n_epochs:
            n = max(int(p.eval_during_training_every_n_epochs * per_epoch), 1)
            return set(range(n, self.num_batches + 1, n))
        if p.eval_during_training_at_specified_steps:
            return set(int(x) for x in p.eval_during_training_at_specified_steps)
        if p.eval_during_training_at_specified_epochs:
            return set(max(int(float(x) * per_epoch), 1)
                       for x in p.eval_during_training_at_specified_epochs)
        return set()

    # ------------------------------------------------------------------ info
    def print_info(self) -> None:
        """T/benchmark_cnn.py:1633-1666."""
        p = self.params
        if self.rank != 0:
            return
        _log("Model:       %s" % self.model.get_model_name())
        subset = "validation" if p.eval else "train"
        ds = self.dataset.name + (" (synthetic)" if self.dataset.use_synthetic_gpu_inputs() else "")
        _log("Dataset:     %s" % ds)
        _log("Mode:        %s" % self.mode)
        _log("SingleSess:  %s" % False)
        _log("Batch size:  %s global" % (self.batch_size * self.num_workers))
        _log("             %s per device" % self.batch_size_per_device)
        if p.batch_group_size > 1:
            _log("             %d batches per prepocessing group" % p.batch_group_size)
        _log("Num batches: %d" % self.num_batches)
        _log("Num epochs:  %.2f" % self.num_epochs)
        kind = "gpu" if self.device_kind == "gpu" else "cpu"
        _log("Devices:     %s" % ["/%s:%d" % (kind, i) for i in range(self.num_gpus)])
        _log("Data format: %s" % p.data_format)
        _log("Optimizer:   %s" % p.optimizer)
        _log("Variables:   %s" % p.variable_update)
        if p.variable_update in ("replicated", "distributed_all_reduce", "collective_all_reduce"):
            _log("AllReduce:   %s" % p.all_reduce_spec)
        if p.variable_update == "kungfu":
            _log("KungFu option: %s" % p.kungfu_option)
        _log("==========")
        for name in IGNORED_ON_B200:
            from . import flags
            if getattr(p, name) != flags.param_specs[name].default_value:
                _log("note: --%s has no effect on the B200-native runtime" % name)

    # ------------------------------------------------------------------ build
    def _build(self, phase_train: bool) -> None:
        p = self.params
        _log_rank0 = _log if self.rank == 0 else (lambda *_: None)
        _log_rank0("Generating training model" if phase_train else "Generating model")
        nclass = self.dataset.num_classes
        self.nclass = nclass
        self.graph = self.model.build_graph(nclass, phase_train=phase_train,
                                            batch_size=self.batch_size_per_device)
        if self.device.type == "cuda":
            # odd channel counts (DenseNet k=12, NASNet, SSD heads) get zero phantom channels so
            # every tensor meets the 16-byte granularity of the TMA / vectorised kernels
            from .nn.legalize import legalize
            self.graph = legalize(self.graph)
        self.comm = comm_lib.create_comm(p, self.device)
        self.store = ParameterStore(self.graph, self.device,
                                    fp16_shadow=self.device.type == "cuda",
                                    allocator=self.comm.allocator())
        _log_rank0("Initializing graph")
        self.store.initialize(self.seed)
        if self.device.type == "cuda":
            from .nn.executor_cuda import CudaExecutor
            self.executor = CudaExecutor(self.graph, self.store, nclass,
                                         aux_weight=self.model.aux_loss_weight,
                                         dropout_seed=4321 + self.rank)
        else:
            from .nn.executor_ref import RefExecutor
            self.executor = RefExecutor(self.graph, self.store, nclass,
                                        aux_weight=self.model.aux_loss_weight,
                                        dropout_seed=4321 + self.rank)
        self.engine = UpdateEngine(p, self.store, self.comm, self.model, seed=self.seed)
        if getattr(self.engine, "fused_push", False) and hasattr(self.executor,
                                                                 "rs_capable_params"):
            self.engine.register_rs_params(self.executor.rs_capable_params())
            self.executor.push_lookup = self.engine.push_target
        self.input_source = self._make_input_source(phase_train)
        self._dump_graph_files()

    def _make_input_source(self, train: bool):
        from .data import inputs
        return inputs.create_input_source(self.params, self.dataset, self.model,
                                          self.batch_size_per_device, self.device, self.nclass,
                                          seed=self.seed, rank=self.rank,
                                          world=max(self.world, 1), train=train)

    # ------------------------------------------------------------------ run
    def run(self) -> Dict:
        """Runs the benchmark and returns the stats dict (T/benchmark_cnn.py:1726-1755)."""
        try:
            if self.params.eval:
                return self._run_eval()
            return self._benchmark_train()
        finally:
            self._close_source(getattr(self, "input_source", None))
            if self.comm is not None:
                self.comm.close()

    @staticmethod
    def _close_source(src) -> None:
        """Real-data sources own a producer thread and a decode pool: stop them."""
        close = getattr(src, "close", None)
        if close is not None:
            try:
                close()
            except Exception:  # pragma: no cover - shutting down must not mask the result
                pass

    def _examples_per_step(self) -> int:
        """Batch size the LR schedule sees (T/benchmark_cnn.py:2757-2765)."""
        if self.params.variable_update in ("kungfu", "horovod", "collective_all_reduce"):
            return self.batch_size * self.num_workers
        return self.batch_size

    def _maybe_restore(self) -> None:
        p = self.params
        if not p.train_dir:
            return
        try:
            path, step = ckpt_lib.latest_checkpoint(p.train_dir)
        except ckpt_lib.CheckpointNotFoundException:
            return
        arrays = ckpt_lib.load_arrays(path)
        ckpt_lib.restore_into(self.store, arrays, ckpt_lib.variable_prefix(p.variable_update))
        self.global_step = int(arrays.get("global_step", step))
        self._restore_optimizer_state(arrays)
        if self.rank == 0:
            _log("Restored checkpoint %s (global_step %d)" % (path, self.global_step))

    def _restore_optimizer_state(self, arrays) -> None:
        """Optimizer slots (``<var>/Momentum`` ...) and the loss-scale state are saved
        variables in the reference too, so a resumed run continues with them."""
        from .nn.params import _tf_to_native
        eng, st = self.engine, self.store
        suffixes = {"momentum": {"m": "Momentum"}, "rmsprop": {"v": "RMSProp", "m": "RMSProp_1"},
                    "adam": {"m": "Adam", "v": "Adam_1"}, "sgd": {}}[self.params.optimizer]
        for key, suffix in suffixes.items():
            arena = st.opt_slots.get(key)
            if arena is None:
                continue
            for spec in st.graph.params.values():
                if not spec.trainable:
                    continue
                for prefix in ("v0/", "v/"):
                    name = "%s%s/%s" % (prefix, spec.name, suffix)
                    if name in arrays:
                        slot = st.slots[spec.name]
                        val = _tf_to_native(spec, torch.from_numpy(
                            np.asarray(arrays[name], dtype=np.float32)))
                        arena[slot.offset:slot.offset + slot.numel].copy_(val.reshape(-1))
                        break
        if eng is not None:
            # Adam's bias correction (TF saves beta1_power / beta2_power) and the ada_sgd switch
            # count optimizer steps from the start of training, not of this process
            eng.step_count = int(self.global_step)
        if "loss_scale" in arrays and eng is not None:
            eng.loss_scale = float(arrays["loss_scale"])
            eng.loss_scale_normal_steps = int(arrays.get("loss_scale_normal_steps", 0))

    def _maybe_load_backbone(self) -> None:
        """--backbone_model_path (T/benchmark_cnn.py:2204-2205, T/models/ssd_model.py:256-297):
        initialise the backbone variables from a checkpoint before training starts."""
        p = self.params
        if not p.backbone_model_path or self.global_step > 0:
            return
        loader = getattr(self.model, "load_backbone_model", None)
        if loader is None:
            raise ValueError("--backbone_model_path: model %s has no backbone to load" % p.model)
        n = loader(self.store, p.backbone_model_path)
        if self.rank == 0:
            _log("Loaded %d backbone variables from %s" % (n, p.backbone_model_path))

    def _dump_graph_files(self) -> None:
        """--graph_file / --partitioned_graph_file_prefix (T/benchmark_cnn.py:2142-2148,
        869-883): the layer graph as text (or JSON for a .json name), and per rank the graph
        the executor actually runs (after channel legalisation) with the bucket plan."""
        p = self.params
        import json as _json

        def dump(graph, path, extra=None):
            if path.endswith(".json"):
                doc = {"nodes": [{"op": n.op, "name": n.name, "inputs": [t.id for t in n.inputs],
                                  "output": n.output.id, "shape": list(n.output.shape),
                                  "params": {k: v.name for k, v in n.params.items()}}
                                 for n in graph.nodes],
                       "variables": {v.name: list(v.tf_shape) for v in graph.params.values()}}
                doc.update(extra or {})
                with open(path, "w") as f:
                    _json.dump(doc, f, indent=1)
            else:
                with open(path, "w") as f:
                    f.write(graph.describe() + "\n")
                    for k, v in (extra or {}).items():
                        f.write("%s: %s\n" % (k, v))
        if p.graph_file and self.rank == 0:
            dump(getattr(self.graph, "logical", self.graph), p.graph_file)
        if p.partitioned_graph_file_prefix:
            ext = ".json" if (p.graph_file or "").endswith(".json") else ".txt"
            dump(self.graph, "%s%d%s" % (p.partitioned_graph_file_prefix, self.rank, ext),
                 {"rank": self.rank, "buckets": [list(b) for b in self.engine.buckets]})

    def _save(self) -> None:
        p = self.params
        if not p.train_dir or self.rank != 0:
            return
        extra = {"global_step": np.asarray(self.global_step, dtype=np.int64)}
        if self.engine is not None and (self.engine.auto_loss_scale or self.engine.loss_scale != 1):
            extra["loss_scale"] = np.asarray(self.engine.loss_scale, dtype=np.float32)
            extra["loss_scale_normal_steps"] = np.asarray(self.engine.loss_scale_normal_steps,
                                                          dtype=np.int64)
        path = ckpt_lib.save_checkpoint(p.train_dir, self.global_step, self.store, extra,
                                        max_to_keep=p.max_ckpts_to_keep,
                                        opt_slots=self._opt_slot_arrays(),
                                        fmt=p.checkpoint_format,
                                        prefix=ckpt_lib.variable_prefix(p.variable_update))
        _log("Saved checkpoint %s" % path)

    def _opt_slot_arrays(self):
        names = {"momentum": {"m": "Momentum"}, "rmsprop": {"v": "RMSProp", "m": "RMSProp_1"},
                 "adam": {"m": "Adam", "v": "Adam_1"}, "sgd": {}}[self.params.optimizer]
        out = {}
        for key, suffix in names.items():
            if key in self.store.opt_slots:
                out[suffix] = self.store.opt_slots[key]
        return out

    def _maybe_inject_fault(self) -> None:
        """Test hook: KFB_FAULT_INJECT="<rank>:<global_step>:<exit|hang>" makes
        one rank die or stall at a step so the launcher's kill-the-group path can be exercised."""
        spec = os.environ.get("KFB_FAULT_INJECT")
        if not spec:
            return
        rank, step, kind = spec.split(":")
        if int(rank) == self.rank and int(step) == self.global_step:
            if kind == "exit":
                os._exit(17)
            time.sleep(10 ** 6)

    def _graph_eligible(self) -> bool:
        """--cuda_graph: the whole step (forward, loss, backward, fused update) is captured once
        and replayed, with device-resident inputs. Several ranks can do so when the gradient
        exchange is the overlapped P2P all-reduce: the kernels then take part of their per-call
        sequence number from a device counter that the captured step advances
        (P2PComm.enable_device_seq; checked on hardware against the eager step,
        tests/test_gpu_multi.py). KFB_MULTI_RANK_GRAPH=0 switches the multi-rank form off."""
        want = self.params.cuda_graph
        if want is False or self.device.type != "cuda" or \
                os.environ.get("KFB_CUDA_GRAPH", "1") == "0":
            return False
        eng = self.engine
        if eng.auto_loss_scale and not eng.device_ls:
            return False                        # the host mirror reads the skip flag every step
        if want is None and not self._graph_safe_to_freeze():
            return False
        if self.world <= 1:
            return True
        from .parallel.comm import P2PComm
        return (os.environ.get("KFB_MULTI_RANK_GRAPH", "1") != "0" and isinstance(self.comm, P2PComm)
                and not eng.fused_push and
                (eng._can_overlap() or
                 (eng._effective_algo() == "sma" and eng._sma_can_prefetch())))

    def _graph_safe_to_freeze(self) -> bool:
        """Automatic mode only captures steps whose replay is exactly the eager step. Dropout /
        drop-path qualify: the step part of their mask offsets and the drop-path keep
        probabilities are read from device memory (executor.step_dev, dp_keep). Only the
        softmax cross-entropy head is known to be free of per-step host work."""
        cached = getattr(self, "_graph_safe", None)
        if cached is None:
            g = getattr(self.executor, "graph", None)
            cached = bool(g is not None and g.loss["kind"] == "softmax_xent" and
                          not any(n.op in ("lstm", "embedding") for n in g.nodes))
            self._graph_safe = cached
        return cached

    def _step_graphed(self) -> None:
        ex, eng = self.executor, self.engine
        lr = eng.learning_rate(self.global_step, self._examples_per_step(),
                               self.dataset.num_examples_per_epoch("train"))
        src = self.input_source
        host = getattr(src, "host_resident", False) and hasattr(ex, "stage_inputs")
        if host and ex.has_staged_inputs():
            # the batch staged on the copy stream during the previous step moves into the input
            # buffers the captured kernels read (device-to-device, ~0.1 % of the step)
            ex.adopt_staged_inputs()
        else:
            ex.set_inputs(*src.next_batch())
        if getattr(self, "_graph", None) is None:
            eng.enable_dynamic_hyper()
            self._graph_warm = getattr(self, "_graph_warm", 0) + 1
            eng.push_dynamic_hyper(lr)
            if self._graph_warm <= 2:            # eager steps: lazy inits happen outside capture
                self._last_lr = lr
                eng.begin_step(lr)
                ex.forward(True)
                ex.loss_backward(on_grads_ready=eng.on_grads_ready, **eng.loss_scale_kwargs())
                eng.finish_step()
                self.global_step += 1
                if host:
                    ex.stage_inputs(*src.next_batch())
                return
            multi = self.world > 1
            counter = self.comm.enable_device_seq() if multi else None
            torch.cuda.synchronize(self.device)
            if multi:
                self.comm.barrier()               # every rank starts capturing at the same step
            seq0 = self.comm.seq if multi else 0
            g = torch.cuda.CUDAGraph()
            from .ops import native as _native
            launches0 = _native.launch_count
            # A device-side input pipeline (--gpu_preprocess) has a producer thread that
            # allocates, copies and launches (nvJPEG, the transform kernel) on its own stream
            # while this thread captures: in the default global mode its calls would fail and
            # invalidate the capture, so the capture is thread-local then. Every other source
            # keeps the global mode (the one the multi-rank captures were validated with).
            busy_producer = getattr(self.input_source, "_gpu", None) is not None
            mode = {"capture_error_mode": "thread_local"} if busy_producer else {}
            with torch.cuda.graph(g, **mode):
                eng.begin_step(lr, push_dyn=False)
                ex.forward(True)
                ex.loss_backward(on_grads_ready=eng.on_grads_ready, **eng.loss_scale_kwargs())
                eng.finish_step()
                if multi:
                    # last node: move the device part of the sequence numbers past this step
                    self._graph_collectives = self.comm.seq - seq0
                    from .ops import native
                    import ctypes
                    native.call("seq_advance", native.ptr(counter),
                                ctypes.c_uint(self._graph_collectives), native.stream_ptr())
            if multi:
                self.comm.seq = seq0              # nothing ran during capture
            eng.step_count -= 1                   # finish_step counted the captured (not run) step
            ex.static_inputs = True               # later device batches are copied, not rebound
            self._graph_launches = _native.launch_count - launches0   # kernels per replay
            _native.launch_count = launches0
            self._graph = g
            _log("Captured the training step in a CUDA graph")
        eng.push_dynamic_hyper(lr)
        ex.push_step_values()                     # drop-path keep probabilities of this step
        self._last_lr = lr
        self._graph.replay()
        ex.step_index += 1                        # host mirror of the device step counter
        from .ops import native as _native
        _native.launch_count += getattr(self, "_graph_launches", 0)
        eng.step_count += 1                       # Adam's bias correction reads it (dyn[2])
        if self.world > 1:
            self.comm.graph_replayed(self._graph_collectives)
        self.global_step += 1
        if host:
            ex.stage_inputs(*src.next_batch())

    def _step(self, train: bool = True, backward: bool = True) -> None:
        """One step on this rank: inputs -> forward -> loss/backward -> exchange + update."""
        self._maybe_inject_fault()
        if train and backward and self._graph_eligible():
            return self._step_graphed()
        ex, eng = self.executor, self.engine
        src = self.input_source
        prefetch = getattr(src, "host_resident", False) and hasattr(ex, "stage_inputs")
        if prefetch and ex.has_staged_inputs():
            ex.use_staged_inputs()
        else:
            ex.set_inputs(*src.next_batch())
        ex.forward(train)
        if not backward:
            ex.loss_backward(backward=False)
            if prefetch:
                ex.stage_inputs(*src.next_batch())
            return
        lr = eng.learning_rate(self.global_step, self._examples_per_step(),
                               self.dataset.num_examples_per_epoch("train"))
        self._last_lr = lr
        eng.begin_step(lr)
        ex.loss_backward(on_grads_ready=eng.on_grads_ready, **eng.loss_scale_kwargs())
        eng.finish_step()
        self.global_step += 1
        if prefetch:
            # the next batch's host->device copy overlaps the kernels enqueued above
            ex.stage_inputs(*src.next_batch())

    def _benchmark_train(self) -> Dict:
        """Timed training loop; a device-side failure (a trapped flag / mbarrier wait - the
        kernels bound every wait - or any other CUDA error) is reported with the rank, the step
        and the state of the gradient exchange before it propagates, so a stuck peer is
        identifiable from the logs of the surviving ranks."""
        try:
            return self._benchmark_train_impl()
        except RuntimeError as e:
            text = str(e)
            if "CUDA" in text or "device-side" in text or "launch failure" in text or \
                    "kfb_" in text:
                eng = self.engine
                _log("FATAL rank %d/%d: device error at global step %d (%s). Gradient exchange: "
                     "algo=%s, buckets=%d, launched this step=%s, flag sequence=%s. A trap in a "
                     "peer wait means another rank stopped signalling: check its log." %
                     (self.rank, max(self.world, 1), self.global_step, text.splitlines()[0][:200],
                      getattr(eng, "algo", "?"), len(getattr(eng, "buckets", []) or []),
                      getattr(eng, "_launched", "?"), getattr(self.comm, "seq", "?")))
            raise

    def _benchmark_train_impl(self) -> Dict:
        p = self.params
        forward_only = bool(p.forward_only)
        self._build(phase_train=not forward_only)
        self._maybe_restore()
        self._maybe_load_backbone()
        self.engine.broadcast_initial_variables()
        rank0 = self.rank == 0
        cuda = self.device.type == "cuda"
        profiler = None
        if p.tfprof_file or p.trace_file:
            from .utils import profile as profile_lib
            profiler = profile_lib.StepProfiler(p, self.device,
                                                graph=getattr(self.executor, "graph", None))
        from .utils.summaries import SummaryWriter
        summaries = SummaryWriter(p, p.train_dir) if rank0 else None
        if rank0:
            _log("Running warm up")
        step_train_times: List[float] = []
        loop_start = None
        display = max(int(p.display_every), 1)
        accuracy_cols = bool(p.print_training_accuracy or forward_only)
        last_save_time = time.time()
        ev_start = ev_end = None
        pending = []  # (event_start, event_end) of steps whose time has not been read yet
        n_done = 0
        stop_early = False
        # --num_batches counts from global step 0: a run resumed from a checkpoint only does the
        # remaining steps (T/benchmark_cnn.py:2242-2250); warm-up steps run (and train) anyway
        end_local_step = max(self.num_batches - self.global_step, 0)
        for local_step in range(-self.num_warmup_batches, end_local_step):
            if local_step == 0:
                self._sync()
                self.comm.barrier()
                if rank0:
                    _log("Done warm up")
                    _log(log_util.header_line(accuracy_cols, p.loss_type_to_report))
                loop_start = time.time()
            if profiler is not None:
                profiler.before_step(local_step)
            if cuda:
                ev_start, ev_end = torch.cuda.Event(enable_timing=True), \
                    torch.cuda.Event(enable_timing=True)
                ev_start.record()
            else:
                t0 = time.time()
            self._step(train=not forward_only, backward=not forward_only)
            if cuda:
                ev_end.record()
                if p.sync_on_finish:
                    ev_end.synchronize()
            else:
                step_time = time.time() - t0
            if profiler is not None:
                profiler.after_step(local_step)
            if local_step < 0:
                continue
            n_done += 1
            if cuda:
                pending.append((ev_start, ev_end))
            else:
                step_train_times.append(step_time)
            step1 = local_step + 1
            if step1 == 1 or step1 % display == 0 or step1 == self.num_batches:
                if cuda:
                    ev_end.synchronize()
                    step_train_times.extend(s.elapsed_time(e) / 1e3 for s, e in pending)
                    pending = []
                if (step1 == 1 or step1 % display == 0) and rank0:
                    loss, t1, t5 = self.executor.read_loss()
                    loss = self._reported_loss(loss)
                    n = self.batch_size_per_device
                    _log(log_util.step_line(
                        step1, self.batch_size, step_train_times, loss,
                        t1 / n if accuracy_cols else None, t5 / n if accuracy_cols else None))
                    if p.benchmark_log_dir:
                        from .utils import benchmark_logger
                        speed = log_util.get_perf_timing(self.batch_size, step_train_times)[0]
                        benchmark_logger.log_metric(p, "current_examples_per_sec", speed, step1,
                                                    "examples/sec")
                        if accuracy_cols:
                            benchmark_logger.log_metric(p, "top_1_accuracy", t1 / n, step1)
                            benchmark_logger.log_metric(p, "top_5_accuracy", t5 / n, step1)
            if summaries is not None and summaries.path is not None and \
                    self.global_step % max(summaries.every, 1) == 0:
                loss_s = self._reported_loss(self.executor.read_loss()[0])
                summaries.maybe_write(self.global_step, getattr(self, "_last_lr", 0.0), loss_s,
                                      self.engine, self.store)
            if p.save_model_steps and step1 % p.save_model_steps == 0:
                self._save()
            elif p.save_model_secs and time.time() - last_save_time >= p.save_model_secs:
                self._save()
                last_save_time = time.time()
            if step1 in self.eval_steps:
                acc = self._eval_during_training()
                if p.stop_at_top_1_accuracy and acc["top_1_accuracy"] >= p.stop_at_top_1_accuracy:
                    if rank0:
                        _log("Stopping, as eval accuracy at least %s was reached" %
                             p.stop_at_top_1_accuracy)
                    stop_early = True
            if stop_early:
                break
        self._sync()
        if cuda and pending:
            step_train_times.extend(s.elapsed_time(e) / 1e3 for s, e in pending)
        # device-timed total: sum of this rank's step times, max over ranks
        local_elapsed = float(np.sum(step_train_times)) if step_train_times else 0.0
        elapsed = self.comm.max_over_ranks(local_elapsed)
        num_steps = n_done
        total_replicas = self.num_workers * self.replicas_per_worker \
            if p.variable_update not in ("kungfu", "horovod") else self.num_workers
        images_per_sec = (total_replicas * num_steps * self.batch_size_per_device /
                          elapsed) if elapsed > 0 else 0.0
        loss, t1, t5 = self.executor.read_loss()
        loss = self._reported_loss(loss)
        if rank0:
            _log("-" * 64)
            _log("total images/sec: %.2f" % images_per_sec)
            _log("-" * 64)
        if profiler is not None:
            profiler.finish()
        if p.train_dir and not forward_only:
            self._save()
        if self.engine.algo == "pair" and hasattr(self.engine, "torn") and rank0:
            torn = int(self.engine.torn.item())
            if torn:
                _log("PairAveraging: %d model pulls raced with a publish" % torn)
        stats = {
            "num_workers": self.num_workers, "num_steps": num_steps,
            "average_wall_time": elapsed / max(num_steps, 1),
            "images_per_sec": images_per_sec, "last_average_loss": loss,
            "loss_scale": self.engine.loss_scale, "skipped_steps": self.engine.skipped_steps,
        }
        if p.benchmark_log_dir and rank0:
            from .utils import benchmark_logger
            benchmark_logger.write(p, stats, self)
        return stats

    def _reported_loss(self, base_loss: float) -> float:
        """total_loss adds the L2 term (T/benchmark_cnn.py:3068-3099); it is only evaluated
        when a loss is reported, never inside the timed kernels."""
        p = self.params
        if p.loss_type_to_report == "base_loss" or not p.weight_decay or p.forward_only:
            return base_loss
        w = self.store.w32.double() ** 2
        mask = getattr(self.engine, "decay_mask", None)
        if mask is not None:        # variables the model keeps out of the L2 term (SSD: BN)
            w = w * mask.to(w.dtype).repeat_interleave(128)[:w.numel()]
        l2 = 0.5 * float(w.sum().item())
        return base_loss + p.weight_decay * l2

    def _sync(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ eval
    def _eval_batches(self) -> int:
        p = self.params
        if p.num_eval_batches and p.num_eval_epochs:
            raise ValueError("At most one of --num_eval_batches and --num_eval_epochs may be "
                             "specified.")
        if p.num_eval_epochs:
            return max(int(p.num_eval_epochs * self.dataset.num_examples_per_epoch("validation") /
                           (self.batch_size * self.num_workers)), 1)
        return p.num_eval_batches or p.num_batches or 100

    def _eval_once(self, num_batches: int, source=None) -> Dict:
        """T/benchmark_cnn.py:1864-1923."""
        src = source or self.input_source
        top1 = top5 = 0
        total = 0
        loss_sum = 0.0
        self._sync()
        start = time.time()
        custom = getattr(self.model, "eval_step", None)
        for step in range(num_batches):
            try:
                batch = src.next_batch()
            except StopIteration:
                break
            if custom is not None:
                # models whose metric is not top-k accuracy (SSD: mAP, DeepSpeech2: CER / WER)
                # run their own forward + bookkeeping and report at the end (eval_finish)
                loss, t1, t5 = custom(self.executor, batch, src), 0, 0
            else:
                self.executor.set_inputs(*batch)
                self.executor.forward(False)
                self.executor.loss_backward(backward=False)
                loss, t1, t5 = self.executor.read_loss()
            top1 += t1
            top5 += t5
            loss_sum += loss
            total += self.batch_size_per_device
            if (step + 1) % max(self.params.display_every, 1) == 0 and self.rank == 0:
                dt = time.time() - start
                _log("%i\t%.1f examples/sec" % (step + 1, total / max(dt, 1e-9)))
        dt = time.time() - start
        acc1, acc5 = top1 / max(total, 1), top5 / max(total, 1)
        if self.rank == 0:
            _log("Accuracy @ 1 = %.4f Accuracy @ 5 = %.4f [%d examples]" % (acc1, acc5, total))
        result = {"top_1_accuracy": acc1, "top_5_accuracy": acc5, "num_examples": total,
                  "eval_average_loss": loss_sum / max(num_batches, 1),
                  "images_per_sec": total / max(dt, 1e-9)}
        if self.rank == 0 and self.params.benchmark_log_dir:
            from .utils import benchmark_logger
            benchmark_logger.log_evaluation_result(self.params, result, self.global_step)
        finish = getattr(self.model, "eval_finish", None)
        if finish is not None:
            extra = finish(self.params) or {}
            for k, v in extra.items():
                if self.rank == 0:
                    _log("%s = %s" % (k, v))
            result.update(extra)
        return result

    def _eval_during_training(self) -> Dict:
        from .data import inputs
        src = inputs.create_input_source(self.params, self.dataset, self.model,
                                         self.batch_size_per_device, self.device, self.nclass,
                                         seed=self.seed + 1, rank=self.rank,
                                         world=max(self.world, 1), train=False)
        try:
            return self._eval_once(self._eval_batches(), src)
        finally:
            self._close_source(src)

    def _write_eval_summary(self, step: int, result: Dict) -> None:
        """--eval_dir: one JSON line per evaluated checkpoint in eval_dir/summaries.jsonl (the
        reference writes the same scalars as a TensorBoard event, T/benchmark_cnn.py:1771,
        1895-1904)."""
        d = self.params.eval_dir
        if not d:
            return
        import json
        try:
            os.makedirs(d, exist_ok=True)
            rec = _eval_summary_record(step, result)
            with open(os.path.join(d, "summaries.jsonl"), "a") as f:
                f.write(json.dumps(rec) + "\n")
            from .utils import tfevents
            if getattr(self, "_eval_events", None) is None:
                self._eval_events = tfevents.EventFileWriter(d)
            self._eval_events.add(step, {k: v for k, v in rec.items() if k != "step"})
        except OSError as e:
            _log("note: could not write the eval summary to %s: %s" % (d, e))

    def _run_eval(self) -> Dict:
        """T/benchmark_cnn.py:1757-1794."""
        p = self.params
        self._build(phase_train=False)
        result = {}
        while True:
            path, step = ckpt_lib.latest_checkpoint(p.train_dir)
            ckpt_lib.restore_into(self.store, ckpt_lib.load_arrays(path),
                                  ckpt_lib.variable_prefix(p.variable_update))
            self.global_step = step
            result = self._eval_once(self._eval_batches())
            if self.rank == 0:
                _log("-" * 64)
                _log("total images/sec: %.2f" % result["images_per_sec"])
                _log("-" * 64)
                self._write_eval_summary(step, result)
            if p.eval_interval_secs <= 0:
                break
            time.sleep(p.eval_interval_secs)
        return result


def _eval_summary_record(step: int, result: Dict) -> Dict:
    """The scalars the reference adds to its eval summary (T/benchmark_cnn.py:1895-1904):
    eval/Accuracy@1, eval/Accuracy@5 and every model result marked as a simple value."""
    from . import constants
    rec = {"step": int(step), "eval/Accuracy@1": result["top_1_accuracy"],
           "eval/Accuracy@5": result["top_5_accuracy"]}
    prefix = constants.SIMPLE_VALUE_RESULT_PREFIX
    for k, v in result.items():
        if isinstance(k, str) and k.startswith(prefix):
            rec["eval/" + k[len(prefix):]] = float(v)
    return rec


def run_from_kwargs(**kwargs) -> Dict:
    """Library entry: ``run_from_kwargs(model='resnet50', batch_size=64, ...)``."""
    params = setup(make_params(**kwargs))
    bench = BenchmarkCNN(params)
    bench.print_info()
    return bench.run()
