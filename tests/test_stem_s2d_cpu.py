"""Geometry of the space-to-depth stem (ops/stem.py) on the CPU: regroup + weight pack + a
ceil(R/2) x 4 unit-stride convolution must equal the original stride-2 convolution for every
padding parity, odd sizes and window shapes the plan accepts (T/models/resnet_model.py:283-286 is
the layer this serves; the CUDA kernels are compared with the same restatements on the GPU)."""
import pytest
import torch
import torch.nn.functional as F

from kf_benchmarks_b200.ops import stem
from kf_benchmarks_b200.ops.native import ConvDesc


def _desc(n, h, w, c, k, r, s, pad_h, pad_w):
    d = ConvDesc()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s = n, h, w, c, k, r, s
    d.pad_h_lo, d.pad_h_hi = pad_h
    d.pad_w_lo, d.pad_w_hi = pad_w
    d.stride_h = d.stride_w = 2
    d.p = (h + pad_h[0] + pad_h[1] - r) // 2 + 1
    d.q = (w + pad_w[0] + pad_w[1] - s) // 2 + 1
    return d


CASES = [
    (2, 32, 32, 3, 8, 7, 7, (3, 3), (3, 3)),      # SAME_RESNET
    (1, 32, 32, 3, 8, 7, 7, (2, 3), (2, 3)),      # SAME at an even size: even low padding
    (2, 23, 27, 3, 8, 7, 7, (3, 3), (3, 3)),      # odd sizes
    (1, 20, 20, 4, 8, 5, 7, (0, 0), (0, 0)),      # VALID, 3 vertical taps, 4 channels
    (1, 18, 22, 1, 8, 8, 8, (3, 4), (3, 4)),      # even window
    (1, 17, 19, 2, 8, 3, 7, (1, 1), (3, 3)),      # 2 vertical taps
]


@pytest.mark.parametrize("n,h,w,c,k,r,s,pad_h,pad_w", CASES)
def test_regrouped_convolution_equals_the_original(n, h, w, c, k, r, s, pad_h, pad_w):
    d = _desc(n, h, w, c, k, r, s, pad_h, pad_w)
    assert stem.eligible(d)
    kpad = (r * s * c + 7) // 8 * 8
    pl = stem.plan(d, kpad)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(n, h, w, c, generator=g)
    wf = torch.zeros(k, kpad)
    wf[:, :r * s * c] = torch.randn(k, r * s * c, generator=g)
    w_oihw = wf[:, :r * s * c].reshape(k, r, s, c).permute(0, 3, 1, 2)
    xp = F.pad(x.permute(0, 3, 1, 2), (pad_w[0], pad_w[1], pad_h[0], pad_h[1]))
    want = F.conv2d(xp, w_oihw, stride=2).permute(0, 2, 3, 1)
    got = stem.forward_ref(pl, x, wf)
    assert got.shape == want.shape
    assert torch.allclose(got, want, atol=1e-4, rtol=1e-4)
    # the virtual descriptor is a consistent convolution (what check_conv verifies in C)
    v = pl.desc
    assert (v.h + v.pad_h_lo + v.pad_h_hi - v.r) + 1 == v.p and v.w == v.q and v.c == 64
    assert v.x_pix_stride == 16 and v.x_row_stride == pl.w2p * 16
    # every regrouped pixel a virtual pixel reads lies inside its row
    assert pl.w2p == d.q + 3


def test_weight_pack_and_gradient_unpack_are_inverse_gathers():
    d = _desc(1, 32, 32, 3, 16, 7, 7, (3, 3), (3, 3))
    pl = stem.plan(d, 152)
    wf = torch.zeros(16, 152)
    wf[:, :147] = torch.arange(16 * 147, dtype=torch.float32).reshape(16, 147)
    w4 = stem.pack_weights_ref(pl, wf)
    assert w4.shape == (16, 4, 4, 16)
    assert int((w4 != 0).sum()) == int((wf != 0).sum())          # nothing lost, nothing doubled
    assert torch.equal(stem.unpack_grad_ref(pl, w4), wf)
    # taps beyond the window and the 4 pad channels stay zero
    assert float(w4[:, 3, :, 6:12].abs().sum()) == 0.0            # r = 7 does not exist
    assert float(w4[..., 12:].abs().sum()) == 0.0


def test_eligibility():
    assert not stem.eligible(_desc(1, 32, 32, 3, 8, 3, 3, (1, 1), (1, 1)))   # 2 horizontal taps
    assert not stem.eligible(_desc(1, 32, 32, 8, 8, 7, 7, (3, 3), (3, 3)))   # 8 channels
    d = _desc(1, 32, 32, 3, 8, 7, 7, (3, 3), (3, 3))
    d.stride_h = d.stride_w = 1
    assert not stem.eligible(d)
