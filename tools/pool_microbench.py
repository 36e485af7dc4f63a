#!/usr/bin/env python
"""Device time of the max-pool kernels on the ResNet stem shape (256 x 112 x 112 x 64, 3x3/2 SAME)
for the CTA cap given by KFB_POOL_CTAS_PER_SM (read once per process)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from kf_benchmarks_b200.ops import native  # noqa: E402
from kf_benchmarks_b200.ops.native import i32, i64, ptr  # noqa: E402

n, h, c, p = 256, 112, 64, 56
x = torch.randn(n, h, h, c, device="cuda").half()
y = torch.empty(n, p, p, c, device="cuda", dtype=torch.float16)
arg = torch.zeros(n, p, p, c, device="cuda", dtype=torch.uint8)
dy = torch.randn(n, p, p, c, device="cuda").half()
dx = torch.empty_like(x)
geom = (i32(n), i32(h), i32(h), i32(c), i32(p), i32(p), i32(3), i32(3), i32(2), i32(2), i32(0),
        i32(0), i32(1))
s = native.stream_ptr()


def fwd():
    native.call("pool_forward", ptr(x), i64(c), ptr(y), i64(c), ptr(arg), *geom, s)


def bwd():
    native.call("pool_backward", ptr(dy), i64(c), ptr(arg), ptr(dx), i64(c), i32(0), *geom, s)


for name, fn in (("fwd", fwd), ("bwd", bwd)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print("KFB_POOL_CTAS_PER_SM=%s max pool %s: %.1f us" %
          (os.environ.get("KFB_POOL_CTAS_PER_SM", "0"), name, e0.elapsed_time(e1) * 50))
