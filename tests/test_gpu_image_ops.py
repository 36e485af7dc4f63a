"""Device-side batch image transform (csrc/image_ops.cu) against the host library
(csrc/host/image_ops.cpp): same descriptors, outputs equal to one fp16 ulp."""
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ulp_diff(a: torch.Tensor, b: torch.Tensor) -> int:
    """largest distance in fp16 ulps (monotone integer mapping of the bit patterns)"""
    def key(t):
        i = t.view(torch.int16).to(torch.int32)
        return torch.where(i < 0, -(i & 0x7FFF), i)
    return int((key(a) - key(b)).abs().max())


def test_batch_transform_matches_the_host_library():
    from kf_benchmarks_b200.data import gpu_image, native_image as ni
    if not ni.available():
        pytest.skip("host library not built")
    rng = random.Random(5)
    nprng = np.random.RandomState(5)
    images, ops = [], []
    for k in range(24):
        ih, iw = rng.randint(60, 400), rng.randint(60, 500)
        img = nprng.randint(0, 256, (ih, iw, 3)).astype(np.uint8)
        distort = None
        if k % 3:
            distort = (1 + k % 2, rng.uniform(-32, 32), rng.uniform(0.5, 1.5),
                       rng.uniform(-0.2, 0.2), rng.uniform(0.5, 1.5))
        if k % 4 == 3:      # SSD style: crop_and_resize of a window that may leave the image
            box = (rng.uniform(-0.1, 0.3), rng.uniform(-0.1, 0.3), rng.uniform(0.6, 1.1),
                   rng.uniform(0.6, 1.1))
            op = ni.make_crop_and_resize_op(img, box, (64, 64), flip=bool(k % 2), distort=distort,
                                            scale3=(0.017, 0.018, 0.019), bias3=(-2.1, -2.0, -1.8),
                                            clamp=bool(k % 8 == 3), out_f16=True)
        elif k % 4 == 2:    # evaluation geometry: resize, then a central window
            op = ni.make_op(img, None, (73, 80), (4, 8, 64, 64), "bilinear", False, None,
                            out_f16=True)
        else:               # training: random crop, mirror, jitter
            ch, cw = rng.randint(20, ih), rng.randint(20, iw)
            crop = (rng.randint(0, ih - ch), rng.randint(0, iw - cw), ch, cw)
            op = ni.make_op(img, crop, (64, 64), None, "nearest" if k % 4 == 1 else "bilinear",
                            bool(k % 2), distort, out_f16=True)
        images.append(img)
        ops.append(op)
    batcher = gpu_image.GpuImageBatcher("cuda", len(images), 64, 64)
    for _ in range(2):      # second call reuses the other slot's buffers
        got = batcher.run(images, ops)
    torch.cuda.synchronize()
    assert got.shape == (len(images), 64, 64, 3) and got.dtype == torch.float16
    worst = 0
    for k, (img, op) in enumerate(zip(images, ops)):
        want = torch.from_numpy(ni.transform(img, op))
        d = _ulp_diff(got[k].cpu(), want)
        worst = max(worst, d)
        assert d <= 1, (k, int(op.method), int(op.distort_order), d)
    assert batcher.h2d_bytes_last < sum(i.size for i in images) + 64 * 1024
