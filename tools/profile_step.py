#!/usr/bin/env python
"""Run a few training steps of one model (for ncu launch lists / kernel captures).

    python tools/profile_step.py --model resnet50 --batch_size 256 --steps 3
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch_size", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--optimizer", default="momentum")
    a = ap.parse_args()
    import torch
    from kf_benchmarks_b200 import benchmark, params as params_lib
    from kf_benchmarks_b200.ops import native
    p = params_lib.make_params(model=a.model, batch_size=a.batch_size, use_fp16=True,
                               variable_update="replicated", optimizer=a.optimizer,
                               data_format="NHWC")
    bench = benchmark.BenchmarkCNN(benchmark.setup(p))
    bench._build(phase_train=True)
    torch.cuda.synchronize()
    n0 = native.launch_count
    for i in range(a.steps):
        bench._step()
        torch.cuda.synchronize()
        if i == 0:
            print("launches per step (approx): %d" % (native.launch_count - n0))
    print("loss", bench.executor.read_loss()[0])


if __name__ == "__main__":
    main()
