"""All-reduce micro-benchmark with the gradient shapes of a model.

Counterpart of the reference's ``all_reduce_benchmark.py`` (T/all_reduce_benchmark.py:61-265):
it takes the variable shapes of ``--model``, fills per-rank gradient arenas with random values
and times ``--iters_per_step`` chained all-reduces per step for ``--num_batches`` steps after
``--num_warmup_batches`` warm-up steps, then prints ``Average time per step: <seconds>`` (the
line the reference's test greps for). Here the all-reduce is the fused P2P all-reduce+update
kernel (``--comm_impl=p2p``) or NCCL through torch.distributed (``--comm_impl=nccl``), timed on
the device (CUDA events, max over ranks), and the achieved fraction of the NVLink roofline is
reported next to it. Run one rank per GPU:

    python -m kf_benchmarks_b200.parallel.launcher -np 8 -- -m \\
        kf_benchmarks_b200.all_reduce_benchmark --model=resnet50 --variable_update=replicated
"""
from __future__ import annotations

import os
import sys

import torch

from . import datasets, flags, params as params_lib
from .models import model_config
from .nn.params import ParameterStore
from .parallel import comm as comm_lib
from .parallel.strategy import UpdateEngine
from .utils import log_util

DEFAULT_ITERS_PER_STEP = 5  # --iters_per_step: all-reduce iterations chained in one timed step

NVLINK_GBS_PER_DIRECTION = 770.0  # measured peer-copy bandwidth (B200_PROFILING.md)


def get_var_shapes(model, nclass):
    """Shapes of the trainable variables of the model (T/all_reduce_benchmark.py:61-67)."""
    g = model.build_graph(nclass, batch_size=2)
    return g, [p.tf_shape for p in g.trainable_params()]


def run_benchmark(params) -> float:
    if params.variable_update not in ("replicated", "kungfu", "horovod", "parameter_server"):
        raise ValueError("all_reduce_benchmark needs an all-reduce based --variable_update")
    if params.variable_consistency == "relaxed":
        raise ValueError("variable_consistency=relaxed is not supported by this benchmark")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    dev = torch.device("cuda", local) if params.device.lower() == "gpu" else torch.device("cpu")
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    ds = datasets.create_dataset(params.data_dir, params.data_name)
    model = model_config.get_model_config(params.model, ds, params)
    graph, shapes = get_var_shapes(model, ds.num_classes)
    comm = comm_lib.create_comm(params, dev)
    store = ParameterStore(graph, dev, fp16_shadow=dev.type == "cuda", allocator=comm.allocator())
    store.initialize(params.tf_random_seed)
    engine = UpdateEngine(params, store, comm, model)
    gen = torch.Generator(device=dev).manual_seed(1 + rank)
    store.grad.copy_(torch.randn(store.grad.numel(), device=dev, generator=gen))
    iters = params.iters_per_step
    warm = params.num_warmup_batches if params.num_warmup_batches is not None else 5
    steps = params.num_batches or 20
    nbytes = store.num_trainable_padded * 4

    def one_step():
        for _ in range(iters):
            engine.begin_step(1e-6)
            engine.finish_step()

    for _ in range(warm):
        one_step()
    comm.barrier()
    if dev.type == "cuda":
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            one_step()
        e1.record()
        e1.synchronize()
        secs = e0.elapsed_time(e1) / 1e3 / steps
    else:
        import time
        t0 = time.time()
        for _ in range(steps):
            one_step()
        secs = (time.time() - t0) / steps
    secs = comm.max_over_ranks(secs)
    if rank == 0:
        log_util.log_fn("Model: %s  variables: %d tensors, %.1f MiB fp32  ranks: %d  impl: %s" %
                        (params.model, len(shapes), nbytes / 2 ** 20, world, comm.kind))
        log_util.log_fn("Average time per step: %s" % secs)
        if world > 1:
            per_iter = secs / iters
            link = 2.0 * (world - 1) / world * nbytes / NVLINK_GBS_PER_DIRECTION / 1e9
            hbm = store.num_trainable_padded * 20 / 6545.6e9
            roof = max(link, hbm)
            log_util.log_fn("per all-reduce+update: %.1f us; roofline max(link %.1f us, HBM %.1f "
                            "us) -> %.0f%% of roofline" %
                            (per_iter * 1e6, link * 1e6, hbm * 1e6, 100 * roof / per_iter))
    comm.close()
    return secs


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    iters = DEFAULT_ITERS_PER_STEP
    rest = []
    i = 0
    while i < len(argv):
        a = argv[i]
        if a.startswith("--iters_per_step"):
            if "=" in a:
                iters = int(a.split("=", 1)[1])
            else:
                i += 1
                iters = int(argv[i])
        else:
            rest.append(a)
        i += 1
    try:
        given = flags.parse_argv(rest)
    except flags.FlagError as e:
        print("FATAL Flags parsing error: %s" % e, file=sys.stderr)
        return 2
    given.setdefault("variable_update", "replicated")
    base = params_lib.make_params(**given)
    P2 = type("P", (), dict(base._asdict(), iters_per_step=iters))
    run_benchmark(P2)
    return 0


if __name__ == "__main__":
    sys.exit(main())
