#!/usr/bin/env python
"""Device time of the batch-norm kernels in isolation: three-kernel form (nn_kernels.cu) against
the single-launch form (bn_fused.cu), forward and backward, for the layer shapes of ResNet-50 at
batch 256. Back-to-back launches on one stream, CUDA events around 30 iterations, inputs rotated
over enough buffers to exceed the L2 for the large shapes. Prints microseconds per call and the
bytes-over-copy-bandwidth roof of each form.

    python tools/bn_microbench.py [--iters 30]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--phases", action="store_true",
                    help="also print the phase times of the single-launch backward (CTA 0)")
    ap.add_argument("--shape", default=None, help="only this shape, e.g. 50176x256")
    a = ap.parse_args()
    import torch
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.ops.native import f32, i32, i64, ptr
    from kf_benchmarks_b200.utils import roofline
    bw, _ = roofline.peaks()
    lib = native.load()
    s = native.stream_ptr()
    shapes = [(1024, 64), (256 * 49, 512), (256 * 196, 256), (256 * 49, 2048), (256 * 784, 128),
              (256 * 196, 1024), (256 * 3136, 64), (256 * 784, 512), (256 * 3136, 256)]
    if a.shape:
        shapes = [tuple(int(v) for v in a.shape.split("x"))]
    print("%-16s %-10s %10s %10s %10s" % ("rows x C", "pass", "3-kernel", "1-launch", "roof us"))
    for rows, c in shapes:
        nbuf = max(1, min(6, int(400e6 // (rows * c * 2))))
        xs = [torch.randn(rows, c, device="cuda").half() for _ in range(nbuf)]
        dys = [torch.randn(rows, c, device="cuda").half() for _ in range(nbuf)]
        y = torch.empty_like(xs[0])
        dz = torch.empty_like(xs[0])
        gamma, beta = torch.ones(c, device="cuda"), torch.zeros(c, device="cuda")
        mm, mv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
        mr, ss, coef = (torch.zeros(2 * c, device="cuda"), torch.zeros(2 * c, device="cuda"),
                        torch.zeros(3 * c, device="cuda"))
        dg, db = torch.zeros(c, device="cuda"), torch.zeros(c, device="cuda")
        nb = lib.kfb_reduce_blocks(i64(rows), i32(c))
        scratch = torch.zeros(max(nb, int(lib.kfb_bn_fused_max_grid())) * 2 * c, device="cuda")
        sync = torch.zeros(2, dtype=torch.int32, device="cuda")

        def fwd3(i):
            native.call("bn_forward_train", ptr(xs[i % nbuf]), i64(c), i64(rows), i32(c),
                        ptr(gamma), ptr(beta), ptr(mm), ptr(mv), f32(1e-5), f32(0.9), i32(1),
                        ptr(mr), ptr(ss), ptr(scratch), i32(0), None, i64(0), i32(1), ptr(y),
                        i64(c), s)

        def fwd1(i):
            native.call("bn_forward_train_fused", ptr(xs[i % nbuf]), i64(c), i64(rows), i32(c),
                        ptr(gamma), ptr(beta), ptr(mm), ptr(mv), f32(1e-5), f32(0.9), i32(1),
                        ptr(mr), ptr(ss), None, i32(0), i32(0), ptr(scratch), None, i64(0),
                        i32(1), ptr(y), i64(c), None, ptr(sync), s)

        def bwd3(i):
            native.call("bn_backward", ptr(dys[i % nbuf]), i64(c), None, i64(c), ptr(ss),
                        ptr(xs[i % nbuf]), i64(c), i64(rows), i32(c), ptr(gamma), ptr(mr),
                        ptr(dg), ptr(db), i32(0), f32(1.0), ptr(coef), ptr(scratch), ptr(dz),
                        i64(c), None, i64(0), i32(0), s)

        def bwd1(i):
            native.call("bn_backward_fused", ptr(dys[i % nbuf]), i64(c), None, i64(c), None,
                        ptr(ss), ptr(xs[i % nbuf]), i64(c), i64(rows), i32(c), ptr(gamma),
                        ptr(mr), ptr(dg), ptr(db), i32(0), f32(1.0), ptr(coef), ptr(scratch),
                        ptr(dz), i64(c), None, i64(0), i32(0), ptr(sync), s)

        def timed(fn):
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(a.iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1e3 / a.iters

        t = rows * c * 2
        for name, f3, f1, nbytes in (("fwd", fwd3, fwd1, 3 * t), ("bwd", bwd3, bwd1, 5 * t)):
            print("%-16s %-10s %10.1f %10.1f %10.1f" % ("%dx%d" % (rows, c), name, timed(f3),
                                                        timed(f1), nbytes / bw * 1e6))
        if a.phases:
            dbg = torch.zeros(8, dtype=torch.int64, device="cuda")
            lib.kfb_bn_fused_debug_stamps(ptr(dbg))
            for i in range(4):
                bwd1(i)
            torch.cuda.synchronize()
            lib.kfb_bn_fused_debug_stamps(None)
            t = dbg.tolist()
            print("   bwd phases of CTA 0 (us): reduce %.1f | barrier %.1f | coefficients %.1f | "
                  "barrier %.1f | apply %.1f" % tuple((t[i + 1] - t[i]) / 1e3 for i in range(5)))
        del xs, dys


if __name__ == "__main__":
    main()
