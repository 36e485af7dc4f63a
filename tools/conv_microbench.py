#!/usr/bin/env python
"""Device time of the convolution kernels on ResNet-50's 3x3 layers (batch 256), one line per
(layer, pass). Inputs are larger than L2 for stages 1-2 and rotated over four buffers otherwise.

    KFB_SV3=0 python tools/conv_microbench.py      # nine im2col boxes per 3x3 tile
    KFB_SV3=1 python tools/conv_microbench.py      # three (shifted-view rows)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    from kf_benchmarks_b200.ops import gemm_conv as gc
    n = int(os.environ.get("BATCH", "256"))
    shapes = [(56, 64), (28, 128), (14, 256), (7, 512)]
    only = os.environ.get("LAYERS")            # e.g. LAYERS=56 PASSES=fprop for a profiler run
    if only:
        shapes = [sh for sh in shapes if str(sh[0]) in only.split(",")]
    only_pass = os.environ.get("PASSES")
    iters = int(os.environ.get("ITERS", "20"))
    for hw, c in shapes:
        xs = [torch.randn(n, hw, hw, c, device="cuda").half() for _ in range(4)]
        dys = [torch.randn(n, hw, hw, c, device="cuda").half() for _ in range(4)]
        w = (torch.randn(c, 3, 3, c, device="cuda") * (9 * c) ** -0.5).half()
        d = gc.make_desc(xs[0].shape, w.shape, (1, 1), (1, 1), (1, 1))
        out = torch.empty(n, hw, hw, c, device="cuda", dtype=torch.float16)
        srows, sbn = gc.conv_stats_layout(d)
        stats = torch.empty(((c + sbn - 1) // sbn, srows, 2, sbn), device="cuda")
        passes = {
            "fprop": lambda i: gc.conv_fprop(xs[i % 4], w, d, out=out),
            "fprop+stats": lambda i: gc.conv_fprop(xs[i % 4], w, d, out=out, stats=stats),
            "dgrad": lambda i: gc.conv_dgrad(dys[i % 4], w, d, out=out),
            "wgrad": lambda i: gc.conv_wgrad(xs[i % 4], dys[i % 4], d),
        }
        flops = 2.0 * n * hw * hw * c * c * 9
        for name, fn in passes.items():
            if only_pass and name not in only_pass.split(","):
                continue
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            print(json.dumps({"layer": f"3x3_{hw}_c{c}", "pass": name, "ms": round(ms, 4),
                              "tflops": round(flops / ms / 1e9, 1),
                              "sv3": os.environ.get("KFB_SV3", "1")}), flush=True)


def residual_dgrads():
    """The fused dgrad + shortcut gradient of the first 1x1 convolution of ResNet-50's blocks."""
    import ctypes
    from kf_benchmarks_b200.ops import gemm_conv as gc
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.ops.native import i64, ptr
    n = int(os.environ.get("BATCH", "256"))
    for hw, c, k in [(56, 256, 64), (28, 512, 128), (14, 1024, 256), (7, 2048, 512)]:
        dys = [torch.randn(n, hw, hw, k, device="cuda").half() for _ in range(2)]
        srcs = [torch.randn(n, hw, hw, c, device="cuda").half() for _ in range(2)]
        bits = torch.randint(0, 256, (n * hw * hw, c // 8), device="cuda", dtype=torch.uint8)
        w = (torch.randn(k, 1, 1, c, device="cuda") * c ** -0.5).half()
        d = gc.make_desc((n, hw, hw, c), w.shape, (1, 1), (0, 0), (0, 0))
        out = torch.empty(n, hw, hw, c, device="cuda", dtype=torch.float16)
        passes = {
            "dgrad": lambda i: gc.conv_dgrad(dys[i % 2], w, d, out=out),
            "dgrad+acc": lambda i: gc.conv_dgrad(dys[i % 2], w, d, out=out, accumulate=True),
            "dgrad+residual": lambda i: native.call(
                "conv_dgrad_residual", ctypes.byref(d), ptr(dys[i % 2]), ptr(w), ptr(out),
                ptr(srcs[i % 2]), i64(c), ptr(bits), i64(c // 8), native.stream_ptr()),
        }
        nbytes = {"dgrad": n * hw * hw * (k + c) * 2, "dgrad+acc": n * hw * hw * (k + 2 * c) * 2,
                  "dgrad+residual": n * hw * hw * (k + 2 * c) * 2 + bits.numel()}
        for name, fn in passes.items():
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(20):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            print(json.dumps({"layer": f"1x1_{hw}_c{c}_k{k}", "pass": name, "ms": round(ms, 4),
                              "GBps": round(nbytes[name] / ms / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "residual":
        residual_dgrads()
    else:
        main()
