"""The command-line / library parameter surface.

All 129 flags of the reference (T/benchmark_cnn.py:114-636) are accepted under the same
names, types, defaults, bounds and enum values, so existing launch scripts keep working.
Flags that configure TensorFlow / cuDNN / XLA / TensorRT / MKL internals have no meaning
for the B200-native runtime: they are accepted and listed in ``IGNORED_ON_B200`` (a note
is logged when they are set to a non-default value). B200-specific additions are at the
end of the file (``--kungfu_strategy`` alias, ``--comm_impl`` ...).

``make_params(**kwargs)`` / ``make_params_from_flags(argv)`` / ``validate_params`` mirror
T/benchmark_cnn.py:959-1034.
"""
from __future__ import annotations

import sys
from collections import namedtuple
from typing import Optional, Sequence

from . import flags
from .constants import NetworkTopology

_D = flags

# ---- model / mode ------------------------------------------------------------------
_D.DEFINE_string("model", "trivial", "Name of the model to run (see models/model_config.py).")
_D.DEFINE_boolean("eval", False, "Evaluate checkpoints from --train_dir instead of training.")
_D.DEFINE_integer("eval_interval_secs", 0, "With --eval: seconds between evaluations of the "
                  "latest checkpoint; 0 evaluates once.")
_D.DEFINE_integer("eval_during_training_every_n_steps", None,
                  "Run an evaluation every N training steps.", lower_bound=1)
_D.DEFINE_float("eval_during_training_every_n_epochs", None,
                "Run an evaluation every N training epochs.")
_D.DEFINE_list("eval_during_training_at_specified_steps", [],
               "Comma-separated training steps after which to evaluate.")
_D.DEFINE_list("eval_during_training_at_specified_epochs", [],
               "Comma-separated training epochs after which to evaluate.")
_D.DEFINE_boolean("forward_only", False, "Only run the forward pass (no loss gradient, no update).")
_D.DEFINE_boolean("freeze_when_forward_only", False,
                  "Freeze variables into constants in forward-only mode (inference graph).")
_D.DEFINE_boolean("print_training_accuracy", False, "Also log top-1/top-5 during training.")
# ---- sizes / steps ------------------------------------------------------------------
_D.DEFINE_integer("batch_size", 0, "Per-device batch size; 0 uses the model default.")
_D.DEFINE_integer("batch_group_size", 1, "Batches produced per input-pipeline wake-up.")
_D.DEFINE_integer("num_batches", None, "Timed steps to run (default 100). Exclusive with "
                  "--num_epochs.")
_D.DEFINE_integer("num_eval_batches", None, "Evaluation batches. Exclusive with --num_eval_epochs.")
_D.DEFINE_float("num_epochs", None, "Training epochs to run. Exclusive with --num_batches.")
_D.DEFINE_float("num_eval_epochs", None, "Evaluation epochs. Exclusive with --num_eval_batches.")
_D.DEFINE_float("stop_at_top_1_accuracy", None,
                "Stop training once an evaluation reaches this top-1 accuracy.")
_D.DEFINE_integer("num_warmup_batches", None, "Untimed warm-up steps (default max(10, 5*a^2)).")
_D.DEFINE_integer("autotune_threshold", None, "Accepted for compatibility (cuDNN autotune).")
_D.DEFINE_integer("num_gpus", 1, "GPUs (= ranks) on this machine.")
_D.DEFINE_string("gpu_indices", "", "Comma-separated device ordinals to use.")
_D.DEFINE_integer("display_every", 10, "Log throughput every N steps.")
# ---- data ---------------------------------------------------------------------------
_D.DEFINE_string("data_dir", None, "Dataset directory; unset means synthetic data.")
_D.DEFINE_string("data_name", None, "imagenet | cifar10 (inferred from --data_dir if unset).")
_D.DEFINE_string("resize_method", "bilinear",
                 "round_robin | nearest | bilinear | bicubic | area.")
_D.DEFINE_boolean("distortions", True, "Random crops / flips / colour distortions in training.")
_D.DEFINE_boolean("use_datasets", True, "Use the batched dataset pipeline for real data.")
_D.DEFINE_string("input_preprocessor", "default", "Named input preprocessor of the dataset.")
_D.DEFINE_string("gpu_thread_mode", "gpu_private", "Accepted for compatibility (TF GPU threads).")
_D.DEFINE_integer("per_gpu_thread_count", 0, "Accepted for compatibility (TF GPU threads).")
_D.DEFINE_boolean("hierarchical_copy", False, "Reference: two-level copy reduce on DGX-1. Here: "
                  "selects the NVSwitch P2P all-reduce like every other algorithm.")
_D.DEFINE_enum("network_topology", NetworkTopology.DGX1,
               (NetworkTopology.DGX1, NetworkTopology.GCP_V100),
               "Accepted for compatibility; NVSwitch gives every GPU pair full bandwidth.")
_D.DEFINE_integer("gradient_repacking", 0, "Number of equal gradient buckets (0: size-based "
                  "buckets).", lower_bound=0)
_D.DEFINE_boolean("compact_gradient_transfer", True,
                  "Move gradients between GPUs in fp16 when computing in fp16.")
_D.DEFINE_enum("variable_consistency", "strong", ("strong", "relaxed"),
               "relaxed: apply the all-reduced gradients of the previous step (one step stale).")
_D.DEFINE_boolean("datasets_repeat_cached_sample", False,
                  "Repeat one cached sample forever (removes input IO).")
_D.DEFINE_enum("local_parameter_device", "gpu", ("cpu", "gpu", "CPU", "GPU"),
               "Where parameter_server mode keeps variables.")
_D.DEFINE_enum("device", "gpu", ("cpu", "gpu", "CPU", "GPU"), "Compute device.")
_D.DEFINE_enum("data_format", "NCHW", ("NHWC", "NCHW"),
               "Accepted for compatibility: activations are always NHWC on B200.")
_D.DEFINE_integer("num_intra_threads", None, "CPU threads for intra-op parallelism.")
_D.DEFINE_integer("num_inter_threads", 0, "Accepted for compatibility.")
_D.DEFINE_string("trace_file", "", "Write a Chrome trace of one warm-up step to this file.")
_D.DEFINE_boolean("use_chrome_trace_format", True, "Chrome trace JSON (else a plain-text table).")
_D.DEFINE_string("tfprof_file", None, "Write a per-kernel profile of the first 10 timed steps.")
_D.DEFINE_string("graph_file", None, "Write the layer graph to this file (.txt / .json).")
_D.DEFINE_string("partitioned_graph_file_prefix", None,
                 "Write per-device layer graphs with this prefix.")
# ---- optimizer / LR -----------------------------------------------------------------
_D.DEFINE_enum("optimizer", "sgd", ("momentum", "sgd", "rmsprop", "adam"), "Optimizer.")
_D.DEFINE_float("init_learning_rate", None, "Initial learning rate (else the model's schedule).")
_D.DEFINE_string("piecewise_learning_rate_schedule", None,
                 "'LR0;E1;LR1;...;En;LRn': LRi applies from epoch Ei.")
_D.DEFINE_float("num_epochs_per_decay", 0, "Epochs between staircase decays.")
_D.DEFINE_float("learning_rate_decay_factor", 0, "Staircase decay factor.")
_D.DEFINE_float("num_learning_rate_warmup_epochs", 0, "Linear warm-up epochs.")
_D.DEFINE_float("minimum_learning_rate", 0, "Lower bound of the decayed learning rate.")
_D.DEFINE_float("momentum", 0.9, "Momentum (Nesterov).")
_D.DEFINE_float("rmsprop_decay", 0.9, "RMSProp decay.")
_D.DEFINE_float("rmsprop_momentum", 0.9, "RMSProp momentum.")
_D.DEFINE_float("rmsprop_epsilon", 1.0, "RMSProp epsilon.")
_D.DEFINE_float("adam_beta1", 0.9, "Adam beta1.")
_D.DEFINE_float("adam_beta2", 0.999, "Adam beta2.")
_D.DEFINE_float("adam_epsilon", 1e-8, "Adam epsilon.")
_D.DEFINE_float("gradient_clip", None, "Clip reduced gradients elementwise to [-c, c].")
_D.DEFINE_float("weight_decay", 0.00004, "L2 regularisation coefficient.")
_D.DEFINE_float("gpu_memory_frac_for_testing", 0, "Accepted for compatibility.",
                lower_bound=0.0, upper_bound=1.0)
_D.DEFINE_boolean("use_unified_memory", False, "Accepted for compatibility.")
_D.DEFINE_boolean("use_tf_layers", True, "Accepted for compatibility.")
_D.DEFINE_integer("tf_random_seed", 1234, "Seed for weight init and synthetic data.")
_D.DEFINE_string("debugger", None, "Accepted for compatibility (tfdbg).")
_D.DEFINE_boolean("use_python32_barrier", False, "Accepted for compatibility.")
_D.DEFINE_boolean("datasets_use_prefetch", True, "Prefetch real-data batches to the device.")
_D.DEFINE_integer("datasets_prefetch_buffer_size", 1, "Prefetch depth.")
_D.DEFINE_integer("datasets_num_private_threads", None, "Input pipeline worker threads.")
_D.DEFINE_boolean("datasets_use_caching", False, "Cache decoded records in host memory.")
_D.DEFINE_integer("datasets_parallel_interleave_cycle_length", None, "Files read in parallel.")
_D.DEFINE_boolean("datasets_sloppy_parallel_interleave", False, "Allow out-of-order records.")
_D.DEFINE_integer("datasets_parallel_interleave_prefetch", None, "Per-file prefetch.")
_D.DEFINE_boolean("use_multi_device_iterator", True, "Accepted for compatibility.")
_D.DEFINE_integer("multi_device_iterator_max_buffer_size", 1, "Accepted for compatibility.")
_D.DEFINE_boolean("winograd_nonfused", True, "Accepted for compatibility (cuDNN).")
_D.DEFINE_boolean("batchnorm_persistent", True, "Accepted for compatibility (cuDNN).")
_D.DEFINE_boolean("sync_on_finish", False, "Synchronise the device at the end of every step.")
_D.DEFINE_boolean("staged_vars", False, "parameter_server only: stage variable reads (one step "
                  "stale).")
_D.DEFINE_boolean("force_gpu_compatible", False, "Accepted for compatibility.")
_D.DEFINE_boolean("allow_growth", None, "Accepted for compatibility.")
_D.DEFINE_boolean("xla", False, "Not supported on this build (no tracing compiler).")
_D.DEFINE_boolean("xla_compile", False, "Not supported on this build (no tracing compiler).")
_D.DEFINE_boolean("fuse_decode_and_crop", True, "Decode only the cropped JPEG region.")
_D.DEFINE_boolean("distort_color_in_yiq", True, "Colour distortion in YIQ space.")
_D.DEFINE_boolean("enable_optimizations", True, "Accepted for compatibility (Grappler).")
_D.DEFINE_string("rewriter_config", None, "Accepted for compatibility (Grappler).")
_D.DEFINE_enum("loss_type_to_report", "total_loss", ("base_loss", "total_loss"),
               "Report the loss with or without the L2 term.")
_D.DEFINE_boolean("single_l2_loss_op", False, "Accepted: the L2 term is always one fused pass.")
_D.DEFINE_boolean("use_resource_vars", False, "Accepted for compatibility.")
_D.DEFINE_boolean("compute_lr_on_cpu", False, "Accepted: the schedule is always host-side.")
_D.DEFINE_boolean("sparse_to_dense_grads", False, "Accepted: all gradients are dense.")
_D.DEFINE_boolean("mkl", False, "Accepted for compatibility.")
_D.DEFINE_integer("kmp_blocktime", 0, "Accepted for compatibility.")
_D.DEFINE_string("kmp_affinity", "granularity=fine,verbose,compact,1,0",
                 "Accepted for compatibility.")
_D.DEFINE_integer("kmp_settings", 1, "Accepted for compatibility.")
# ---- fp16 -------------------------------------------------------------------------
_D.DEFINE_boolean("use_fp16", False, "fp16 compute with fp32 master weights.")
_D.DEFINE_float("fp16_loss_scale", None, "Static loss scale (default: the model's, 128).")
_D.DEFINE_boolean("fp16_vars", False, "Keep variables in fp16 too.")
_D.DEFINE_boolean("fp16_enable_auto_loss_scale", False, "Automatic loss scaling.")
_D.DEFINE_integer("fp16_inc_loss_scale_every_n", 1000,
                  "Clean steps before the automatic loss scale doubles.")
# ---- gradient path ----------------------------------------------------------------
_D.DEFINE_enum("variable_update", "parameter_server",
               ("parameter_server", "replicated", "distributed_replicated", "independent",
                "distributed_all_reduce", "collective_all_reduce", "horovod", "kungfu"),
               "How replicas exchange gradients / variables.")
_D.DEFINE_enum("kungfu_option", "sync_sgd", ("async_sgd", "sync_sgd", "ada_sgd", "sma"),
               "KungFu optimizer: sync_sgd (S-SGD), async_sgd (PairAveraging / AD-PSGD), "
               "sma (SynchronousAveraging), ada_sgd (SMA then S-SGD).")
_D.DEFINE_string("all_reduce_spec", None, "Reference all-reduce grammar (alg[#shards][:limit:...]);"
                 " every algorithm maps onto the NVSwitch P2P kernel family.")
_D.DEFINE_integer("agg_small_grads_max_bytes", 0, "Fold tensors up to this size into one bucket.")
_D.DEFINE_integer("agg_small_grads_max_group", 10, "Most small tensors per folded bucket.")
_D.DEFINE_integer("allreduce_merge_scope", 1, "Accepted for compatibility.")
_D.DEFINE_enum("job_name", "", ("ps", "worker", "controller", ""),
               "TF-cluster role; not supported on this build (single NVSwitch box).")
_D.DEFINE_string("ps_hosts", "", "Comma-separated list of parameter-server hosts (accepted; "
                 "the ps role has no work on this runtime).")
_D.DEFINE_string("worker_hosts", "", "Comma-separated host:port list of the workers; with "
                 "--job_name=worker it defines world size and the rendezvous address.")
_D.DEFINE_string("controller_host", None, "Accepted for compatibility.")
_D.DEFINE_integer("task_index", 0, "Index of this task within --job_name (the worker's rank).")
_D.DEFINE_string("server_protocol", "grpc", "Accepted for compatibility.")
_D.DEFINE_boolean("cross_replica_sync", True, "Accepted for compatibility.")
_D.DEFINE_string("horovod_device", "", "Accepted for compatibility.")
# ---- observability / checkpoints ------------------------------------------------------
_D.DEFINE_integer("summary_verbosity", 0, "0 none .. 3 histograms (JSONL summaries).")
_D.DEFINE_integer("save_summaries_steps", 0, "Steps between summaries.")
_D.DEFINE_integer("save_model_secs", 0, "Seconds between checkpoints.")
_D.DEFINE_integer("save_model_steps", None, "Steps between checkpoints.")
_D.DEFINE_integer("max_ckpts_to_keep", 5, "Checkpoints kept in --train_dir.")
_D.DEFINE_string("train_dir", None, "Directory for checkpoints and summaries.")
_D.DEFINE_string("eval_dir", "/tmp/tf_cnn_benchmarks/eval", "Directory for evaluation summaries.")
_D.DEFINE_string("backbone_model_path", None, "Checkpoint of a backbone to load before training.")
_D.DEFINE_enum("trt_mode", "", ["", "FP32", "FP16", "INT8"], "Not supported on this build.")
_D.DEFINE_integer("trt_max_workspace_size_bytes", 4 << 30, "Accepted for compatibility.")
_D.DEFINE_string("benchmark_log_dir", None, "Write benchmark_run.log / metric.log JSON here.")
_D.DEFINE_string("benchmark_test_id", None, "Identifier stored in the benchmark log.")

NUM_REFERENCE_FLAGS = len(flags.param_specs)

# ---- B200-native additions ------------------------------------------------------------
_D.DEFINE_alias("kungfu_strategy", "kungfu_option")
_D.DEFINE_enum("checkpoint_format", "tf", ("tf", "npz"),
               "tf (default): a TensorFlow V2 checkpoint bundle (model.ckpt-N.index + .data-*), "
               "what the reference's Saver writes and can restore. npz: one .npz per checkpoint "
               "keyed by the same TF variable names. Both are read back transparently.")
_D.DEFINE_enum("comm_impl", "auto", ("auto", "p2p", "p2p_fused", "nvls", "nccl", "gloo"),
               "auto (default): nvls from 4 ranks up where the NVSwitch multicast bind succeeds, "
               "p2p otherwise (measured: 8 ranks 304 us vs 347 us per ResNet-50 all-reduce + "
               "update, 2 ranks 305 us vs 248 us). "
               "p2p: fused NVLink all-reduce+update kernels (the product). p2p_fused: in "
               "addition the weight-gradient GEMMs scatter their tiles to the owning ranks from "
               "their own epilogue (GEMM + reduce-scatter in one kernel). nvls: the p2p kernels with "
               "the reduction and the broadcast done inside the NVSwitch (multimem.ld_reduce / "
               "multimem.st on multicast-bound arenas). nccl: "
               "torch.distributed NCCL + separate update (comparison baseline). gloo: CPU ranks.")
_D.DEFINE_float("sma_alpha", 0.1, "Blend factor of SynchronousAveraging.")
_D.DEFINE_integer("bucket_bytes", 8 << 20, "Target gradient bucket size for overlap.")
_D.DEFINE_boolean("cuda_graph", None,
                  "Capture the training step in a CUDA graph and replay it. Unset = automatic: "
                  "on whenever the step is a fixed kernel sequence (GPU, no dropout / drop-path "
                  "whose masks must change per step); true / false force it.")
_D.DEFINE_boolean("gpu_preprocess", False,
                  "Real-data input: the host threads only decode; crop / resize / mirror / colour "
                  "jitter / normalise / fp16 run as one CUDA kernel per batch on the raw uploaded "
                  "images (bilinear, nearest and crop_and_resize; other resize methods stay on "
                  "the host).")
_D.DEFINE_integer("ada_sgd_switch_step", 0, "ada_sgd: step at which SMA hands over to S-SGD.")

# Reference flags that configure TF/cuDNN/XLA/TRT/MKL internals: accepted, no effect here.
IGNORED_ON_B200 = (
    "autotune_threshold", "gpu_thread_mode", "per_gpu_thread_count", "network_topology",
    "num_inter_threads", "gpu_memory_frac_for_testing", "use_unified_memory", "use_tf_layers",
    "debugger", "use_python32_barrier", "use_multi_device_iterator",
    "multi_device_iterator_max_buffer_size", "winograd_nonfused", "batchnorm_persistent",
    "force_gpu_compatible", "allow_growth", "enable_optimizations", "rewriter_config",
    "single_l2_loss_op", "use_resource_vars", "compute_lr_on_cpu", "sparse_to_dense_grads", "mkl",
    "kmp_blocktime", "kmp_affinity", "kmp_settings", "allreduce_merge_scope", "server_protocol",
    "cross_replica_sync", "horovod_device", "trt_max_workspace_size_bytes",
    # one input pipeline serves both settings of --use_datasets; prefetching is the pinned ring
    # plus the copy stream; Pillow decodes whole images; saturation and hue commute in HSV, so
    # the two --distort_color_in_yiq orders coincide
    "use_datasets", "datasets_use_prefetch", "datasets_prefetch_buffer_size",
    "datasets_parallel_interleave_prefetch", "fuse_decode_and_crop", "distort_color_in_yiq",
    # gradients lie back to back in one fp32 arena: nothing to aggregate or compact, no
    # parameter device to choose; there is no controller process
    "compact_gradient_transfer", "agg_small_grads_max_bytes", "agg_small_grads_max_group",
    "local_parameter_device", "controller_host",
)
# Reference features this build rejects with an explicit error when requested.
UNSUPPORTED_WHEN_SET = {
    "xla": "XLA is not available (no tracing compiler on the hot path)",
    "xla_compile": "XLA is not available (no tracing compiler on the hot path)",
    "trt_mode": "TensorRT conversion is not available",
}

Params = namedtuple("Params", list(flags.param_specs.keys()))  # type: ignore


def validate_params(params) -> None:
    for name, value in params._asdict().items():
        flags.validate_value(name, value)


def make_params(**kwargs):
    """Params from keyword overrides (library entry, T/benchmark_cnn.py:993-1010)."""
    values = {name: spec.default_value for name, spec in flags.param_specs.items()}
    for k, v in kwargs.items():
        k = flags.aliases.get(k, k)
        if k not in values:
            raise ValueError("Unknown parameter: %s" % k)
        values[k] = v
    params = Params(**values)
    validate_params(params)
    return params


def make_params_from_flags(argv: Optional[Sequence[str]] = None):
    given = flags.parse_argv(sys.argv[1:] if argv is None else argv)
    return make_params(**given)


def remove_param_fields(params, fields_to_remove):
    """T/benchmark_cnn.py:1026-1034."""
    d = params._asdict()
    for f in fields_to_remove:
        assert f in d, "Invalid Params field: " + f
    kept = {k: v for k, v in d.items() if k not in fields_to_remove}
    return namedtuple("Params", kept.keys())(**kept)
