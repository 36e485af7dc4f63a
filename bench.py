#!/usr/bin/env python
"""Headline benchmark: ResNet-50 fp16 synthetic-ImageNet training images/sec on N B200s.

Contract (driver): ``python bench.py --gpus N --steps K --warmup W`` (under torchrun for
N > 1) prints ONE JSON line from rank 0. The metric/config are BASELINE.json's: ResNet-50
v1, fp16 compute / fp32 master weights, batch 256 per GPU, synthetic ImageNet,
``variable_update=replicated`` (fused P2P all-reduce + Nesterov-momentum SGD), device-timed,
max over ranks. ``--impl reference`` reports that the TensorFlow/KungFu reference cannot be
installed on this image.

Defaults: 200 timed steps after 10 warm-up steps (~4 s per timed region, so the number is not a
burst at boost clocks; profiles/sustained_1gpu.jsonl has a 600-step run).

Timed region = exactly K full training steps (forward, loss, backward, gradient exchange,
optimizer update, BN statistics) bracketed by barrier + synchronize on both sides and CUDA
events on the launching stream. Every step touches > 5 GB of activations, far more than the
126 MB L2, so no explicit L2 flush is needed between iterations.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# reference numbers that exist: RTX 3090, fp32, parameter_server
_PUBLISHED = {1: 416.43, 2: 713.23}


def _reference_arm(args) -> int:
    """The reference is a TF 2.5 (compat.v1) + KungFu/Horovod program; neither TensorFlow nor
    KungFu nor Horovod wheels exist in /opt/wheelhouse and the tree has no setup.py /
    pyproject.toml, so `pip install /root/reference` cannot succeed."""
    why = ("reference (tf_cnn_benchmarks fork) needs TensorFlow 2.5 + KungFu/Horovod; none are "
           "installable offline on this image and the tree has no setup.py/pyproject.toml")
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    try:
        sys.path.insert(0, ref_dir)
        import tensorflow  # noqa: F401
        why = "tensorflow imported but the reference run path is not wired on this image"
    except Exception:
        pass
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def _timed_steps(bench, steps, torch, e2e: bool):
    """K steps between barrier+sync on both sides; returns device ms (this rank)."""
    dev = bench.device
    torch.cuda.synchronize(dev)
    bench.comm.barrier()
    torch.cuda.synchronize(dev)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    start.record()
    last = None
    for _ in range(steps):
        bench._step(train=True, backward=True)
        if e2e:
            last = bench.executor.read_loss()  # device->host read of the step's result
    end.record()
    end.synchronize()
    torch.cuda.synchronize(dev)
    bench.comm.barrier()
    wall = time.time() - t0
    return start.elapsed_time(end), wall, last


def _run_standin(args, torch, dev, rank, world, group=None):
    """Times the PyTorch stand-in (kf_benchmarks_b200/baseline/torch_standin.py: cuDNN / cuBLAS
    fp16 model + fp32 master, NCCL all-reduce of a flat fp16 gradient buffer, CUDA-graphed
    forward + backward) on the same config with the same harness; returns a dict for rank 0.
    It is neither the reference nor the product."""
    import torch.distributed as dist
    from kf_benchmarks_b200 import datasets, lr as lr_mod, params as params_lib
    from kf_benchmarks_b200.baseline.torch_standin import StandinTrainer
    from kf_benchmarks_b200.models import model_config
    kungfu = args.variable_update in ("kungfu", "horovod")
    p = params_lib.make_params(model=args.model, batch_size=args.batch_size, use_fp16=True,
                               variable_update=args.variable_update,
                               kungfu_option=args.kungfu_option,
                               num_gpus=1 if kungfu else args.gpus, optimizer=args.optimizer,
                               data_format="NHWC")
    ds = datasets.create_dataset(None, None)
    model = model_config.get_model_config(args.model, ds, p)
    model.set_batch_size(args.batch_size)
    graph = model.build_graph(ds.num_classes, batch_size=args.batch_size)
    tr = StandinTrainer(p, model, graph, ds.num_classes, dev, world, rank, group=group)
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    h = model.image_size
    images = (torch.randn(args.batch_size, h, h, 3, device=dev, generator=gen) * 60 + 127).half()
    labels = torch.randint(0, ds.num_classes - 1, (args.batch_size,), device=dev, generator=gen)
    warmup = max(args.warmup, 3)
    lr = lr_mod.get_learning_rate(p, 1000, ds.num_examples_per_epoch("train"), model,
                                  args.batch_size * (world if kungfu else args.gpus))
    for _ in range(warmup + 6):  # cudnn.benchmark, then graph capture, then replays
        tr.step(images, labels, lr)

    def sync():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier(group=group)
        torch.cuda.synchronize(dev)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        tr.step(images, labels, lr)
    e1.record()
    e1.synchronize()
    sync()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        ms = float(t.item())
    gb = args.batch_size * world
    return {"impl": "eager-graphed torch cuDNN/cuBLAS + NCCL stand-in - NOT the reference",
            "value": gb * args.steps / (ms / 1e3), "unit": "images/sec",
            "ms_per_step": ms / args.steps, "steps": args.steps, "n_gpus": world,
            "final_loss": float(tr.last_loss)}


def _standin_arm(args, torch) -> int:
    """--impl standin: the stand-in alone, as its own JSON line."""
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    out = _run_standin(args, torch, dev, rank, world)
    if rank == 0:
        out.update({"metric": "ResNet-50 fp16 synthetic-ImageNet training images/sec "
                              "(device-timed, max over ranks)", "higher_is_better": True,
                    "scaling": "weak", "dtype": "fp16", "data": "synthetic",
                    "warmup": max(args.warmup, 3),
                    "config": {"model": args.model, "global_batch": args.batch_size * world,
                               "variable_update": args.variable_update,
                               "optimizer": args.optimizer}})
        out["impl"] = "standin"
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


def _replica_checksums_equal(bench, torch):
    """64-bit checksum of every rank's fp32 master weights; True when all ranks hold the same
    bits (what the synchronous modes guarantee), None for the asynchronous ones."""
    eng = bench.engine
    if getattr(eng, "algo", "none") in ("pair", "sma", "ada", "none") and bench.world > 1:
        return None
    w = bench.store.w32
    chk = int(w.view(torch.int32).to(torch.int64).sum().item())
    sums = bench.comm.allgather_object(chk) if bench.world > 1 else [chk]
    return all(s == sums[0] for s in sums)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "standin"])
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch_size", type=int, default=256)
    ap.add_argument("--variable_update", default="replicated")
    ap.add_argument("--kungfu_option", default="sync_sgd")
    ap.add_argument("--optimizer", default="momentum")
    ap.add_argument("--comm_impl", default="auto",
                    help="auto (NVLS from 4 ranks up where the multicast bind succeeds, else the "
                         "pull kernel), p2p, nvls, p2p_fused, nccl")
    ap.add_argument("--bucket_bytes", type=int, default=None,
                    help="gradient bucket size of the overlapped exchange (default: the flag's)")
    ap.add_argument("--no_e2e", action="store_true")
    ap.add_argument("--no_standin", action="store_true",
                    help="skip the same-box PyTorch stand-in that is timed after the product")
    ap.add_argument("--cuda_graph", type=int, default=-1,
                    help="1 / 0 force the CUDA-graph replay of the step on / off; -1 (default) is "
                         "the framework's automatic choice (on for ResNet-50)")
    args = ap.parse_args()
    if args.impl == "reference":
        return _reference_arm(args)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        from kf_benchmarks_b200.parallel import launcher
        return launcher.launch_local([os.path.abspath(__file__)] + sys.argv[1:], args.gpus)

    import torch
    from kf_benchmarks_b200 import benchmark, params as params_lib
    from kf_benchmarks_b200.data import inputs
    from kf_benchmarks_b200.ops import native

    if args.impl == "standin":
        return _standin_arm(args, torch)
    warmup = max(args.warmup, 3)
    kungfu = args.variable_update in ("kungfu", "horovod")
    p = params_lib.make_params(
        model=args.model, batch_size=args.batch_size, use_fp16=True,
        variable_update=args.variable_update, kungfu_option=args.kungfu_option,
        num_gpus=1 if kungfu else args.gpus, optimizer=args.optimizer,
        comm_impl=args.comm_impl, num_batches=args.steps, num_warmup_batches=warmup,
        cuda_graph=None if args.cuda_graph < 0 else bool(args.cuda_graph),
        data_format="NHWC", **({"bucket_bytes": args.bucket_bytes} if args.bucket_bytes else {}))
    p = benchmark.setup(p)
    bench = benchmark.BenchmarkCNN(p)
    rank = bench.rank
    bench._build(phase_train=True)
    bench.engine.broadcast_initial_variables()
    for _ in range(warmup):
        bench._step()
    sampler = ClockSampler(bench.device.index or 0)
    if rank == 0:
        sampler.start()
    launches0 = native.launch_count
    ms, wall, _ = _timed_steps(bench, args.steps, torch, e2e=False)
    launches = native.launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_max = bench.comm.max_over_ranks(ms)
    n_ranks = max(bench.world, 1)
    global_batch = args.batch_size * n_ranks
    value = global_batch * args.steps / (ms_max / 1e3)
    loss = bench.executor.read_loss()[0]

    e2e = None
    if not args.no_e2e:
        # same metric through the public step API with per-step H2D of the inputs from pinned
        # host memory and a D2H read of the loss
        src = inputs.SyntheticHostSource(args.batch_size, bench.model.image_size, 3, bench.nclass,
                                         bench.device, seed=77 + rank, ring=3)
        bench.input_source = src
        for _ in range(2):
            bench._step()
        ms_e, _, last = _timed_steps(bench, args.steps, torch, e2e=True)
        ms_e = bench.comm.max_over_ranks(ms_e)
        e2e = {"value": global_batch * args.steps / (ms_e / 1e3), "unit": "images/sec",
               "h2d_bytes_per_step": src.bytes_per_step * n_ranks,
               "d2h_bytes_per_step": 32 * n_ranks, "ms_per_step": ms_e / args.steps}
    replicas_equal = _replica_checksums_equal(bench, torch)
    standin = None
    if not args.no_standin:
        # the same-box comparison point, timed after the product's own regions with the same
        # harness (the reference itself cannot be installed here, see --impl reference)
        try:
            group = None
            if n_ranks > 1:
                import torch.distributed as dist
                group = dist.new_group(backend="nccl")
            standin = _run_standin(args, torch, bench.device, rank, n_ranks, group=group)
        except Exception as e:  # the product's numbers stand on their own
            standin = {"error": repr(e)[:300]}
    if rank == 0:
        out = {
            "metric": "ResNet-50 fp16 synthetic-ImageNet training images/sec "
                      "(device-timed, max over ranks)",
            "value": value, "unit": "images/sec", "n_gpus": n_ranks, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak",
            # BASELINE.json publishes no number for this config (the reference's only figures are
            # RTX 3090, fp32, parameter_server with a nan loss: 416 / 713 img/s on 1 / 2 GPUs)
            "vs_baseline": None,
            "dtype": "fp16", "data": "synthetic",
            "config": {"model": args.model, "global_batch": global_batch, "seq_len": None,
                       "image_size": bench.model.image_size,
                       "parallelism": "dp%d" % n_ranks,
                       "variable_update": args.variable_update, "optimizer": args.optimizer,
                       "kungfu_option": args.kungfu_option if args.variable_update == "kungfu"
                       else "",
                       "comm_impl": args.comm_impl if n_ranks > 1 else "none",
                       "l2_hygiene": "per-step working set (>5 GB activations) exceeds the "
                                     "126 MB L2; no explicit flush",
                       "cuda_graph": bool(getattr(bench, "_graph", None) is not None),
                       "published_context": _PUBLISHED},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "replica_checksum_equal": replicas_equal, "standin": standin,
            "final_loss": loss, "wall_s_timed": wall,
        }
    bench.comm.close()
    if n_ranks > 1:
        # tear the process groups down before the result line goes out, so that nothing (NCCL
        # shutdown warnings of other ranks, for instance) follows it on the job's output
        try:
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
        except Exception:
            pass
    if rank == 0:
        sys.stdout.flush()
        time.sleep(0.5 if n_ranks > 1 else 0.0)
        print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
