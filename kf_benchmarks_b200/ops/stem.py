"""Space-to-depth execution plan of a tiny-C stride-2 stem convolution (csrc/stem_s2d.cu).

ResNet's first layer (7x7 / 2 on 3 channels, T/models/resnet_model.py:283-286) fills 3 of the 64
channels of an implicit-GEMM K block. Regrouped into 2x2 pixel blocks it is a 4x4 unit-stride
convolution on 16 channels, and its four horizontal taps are 64 contiguous halves of the
regrouped image - one dense K block for a TMA im2col map whose pixels overlap (pixel stride 16,
pixel size 64). The convolution then runs as a ``ceil(R/2) x 1`` convolution on the ordinary
tcgen05 FPROP / WGRAD paths; this module holds the geometry and the three small kernels around
it (regroup, weight pack, gradient unpack).
"""
import ctypes
import os
from dataclasses import dataclass

import torch

from . import native
from .native import ConvDesc, f32, i32, ptr, stream_ptr


@dataclass
class StemPlan:
    n: int
    h: int
    w: int
    c: int
    k: int
    r: int
    s: int
    kpad: int
    p: int
    q: int
    eh: int          # parity of the low padding (rows / columns): input i = 2A + b - e
    ew: int
    ta: int          # vertical taps of the regrouped convolution, ceil(R / 2)
    h2: int          # rows of the regrouped image
    w2p: int         # columns of the regrouped image including the materialised zero padding
    lw: int          # zero columns on the left
    expand: bool     # True: the 64-channel pixels are materialised (no overlapping strides)
    desc: ConvDesc   # the ceil(R/2) x 1 convolution over 64-channel pixels

    def s2d_shape(self):
        if self.expand:
            return (self.n, self.h2, self.q, 64)
        return (self.n, self.h2, self.w2p, 16)


def eligible(d: ConvDesc) -> bool:
    """Stride 2x2, at most 4 input channels and exactly four horizontal block taps."""
    return (d.stride_h == 2 and d.stride_w == 2 and 1 <= d.c <= 4 and (d.s + 1) // 2 == 4 and
            d.k % 8 == 0 and os.environ.get("KFB_STEM_S2D", "1") != "0")


def plan(d: ConvDesc, kpad: int) -> StemPlan:
    """``d`` describes the original convolution (its ``c`` may be < 8: never passed to C)."""
    assert eligible(d)
    eh, ew = d.pad_h_lo & 1, d.pad_w_lo & 1
    lh, lw = (d.pad_h_lo - eh) // 2, (d.pad_w_lo - ew) // 2
    ta = (d.r + 1) // 2
    h2 = (d.h + eh + 1) // 2
    w2p = d.q + 3
    expand = os.environ.get("KFB_STEM_EXPAND", "0") == "1"
    v = ConvDesc()
    v.n, v.h, v.w, v.c = d.n, h2, d.q, 64
    v.k, v.r, v.s = d.k, ta, 1
    v.pad_h_lo, v.pad_h_hi = lh, d.p - 1 + ta - lh - h2
    v.pad_w_lo, v.pad_w_hi = 0, 0
    v.stride_h, v.stride_w = 1, 1
    v.p, v.q = d.p, d.q
    if not expand:
        v.x_pix_stride, v.x_row_stride = 16, w2p * 16
    return StemPlan(d.n, d.h, d.w, d.c, d.k, d.r, d.s, kpad, d.p, d.q, eh, ew, ta, h2, w2p, lw,
                    expand, v)


def regroup(pl: StemPlan, x, out, stream=None):
    """x[n, h, w, c] fp16 -> out (``pl.s2d_shape()``)."""
    native.call("stem_s2d", ptr(x), ptr(out), i32(pl.n), i32(pl.h), i32(pl.w), i32(pl.c),
                i32(pl.h2), i32(pl.w2p), i32(pl.eh), i32(pl.ew), i32(pl.lw),
                i32(int(pl.expand)), i32(pl.q), stream or stream_ptr())
    return out


def pack_weights(pl: StemPlan, w_flat, w4, stream=None):
    """w_flat[k, kpad] fp16 -> w4[k, ta, 64] fp16."""
    native.call("stem_pack_w", ptr(w_flat), ptr(w4), i32(pl.k), i32(pl.kpad), i32(pl.r),
                i32(pl.s), i32(pl.c), i32(pl.ta), stream or stream_ptr())
    return w4


def unpack_grad(pl: StemPlan, dw4, dw_flat, accumulate=False, scale=1.0, stream=None):
    """dw4[k, ta, 64] fp32 -> dw_flat[k, kpad] fp32 (+)= scale * ..."""
    native.call("stem_unpack_dw", ptr(dw4), ptr(dw_flat), i32(pl.k), i32(pl.kpad), i32(pl.r),
                i32(pl.s), i32(pl.c), i32(pl.ta), i32(int(accumulate)), f32(scale),
                stream or stream_ptr())
    return dw_flat


def forward(pl: StemPlan, x, w_flat, bias=None, relu=False):
    """Convenience wrapper for tests / tools: returns (y, s2d)."""
    from . import gemm_conv
    s2d = torch.zeros(pl.s2d_shape(), device=x.device, dtype=torch.float16)
    w4 = torch.empty((pl.k, pl.ta, 1, 64), device=x.device, dtype=torch.float16)
    regroup(pl, x, s2d)
    pack_weights(pl, w_flat, w4)
    y = gemm_conv.conv_fprop(s2d, w4, pl.desc, bias=bias, relu=relu)
    return y, s2d


def weight_grad(pl: StemPlan, s2d, dy):
    """dW[k, kpad] fp32 of the original convolution."""
    from . import gemm_conv
    dw4 = gemm_conv.conv_wgrad(s2d, dy, pl.desc)
    out = torch.empty((pl.k, pl.kpad), device=dy.device, dtype=torch.float32)
    return unpack_grad(pl, dw4, out)


# ---------------------------------------------------------------- PyTorch restatements
# The index arithmetic of csrc/stem_s2d.cu in plain tensor ops: the CPU tests check the plan's
# geometry with them (regroup + pack + a ceil(R/2) x 4 convolution == the original convolution)
# and the GPU tests compare the kernels against them.
def regroup_ref(pl: StemPlan, x: torch.Tensor) -> torch.Tensor:
    """x[n, h, w, c] -> the padded regrouped image [n, h2, w2p, 16] (same dtype)."""
    out = torch.zeros((pl.n, pl.h2, pl.w2p, 16), dtype=x.dtype, device=x.device)
    for a in range(pl.h2):
        for bh in range(2):
            i = 2 * a - pl.eh + bh
            if not 0 <= i < pl.h:
                continue
            for jp in range(pl.w2p):
                for bw in range(2):
                    j = 2 * (jp - pl.lw) - pl.ew + bw
                    if 0 <= j < pl.w:
                        c0 = (bh * 2 + bw) * pl.c
                        out[:, a, jp, c0:c0 + pl.c] = x[:, i, j, :]
    return out


def pack_weights_ref(pl: StemPlan, w_flat: torch.Tensor) -> torch.Tensor:
    """w_flat[k, kpad] (columns (r*S + s)*C + c) -> [k, ta, 4, 16] (vertical tap, horizontal
    tap, regrouped channel)."""
    w4 = torch.zeros((pl.k, pl.ta, 4, 16), dtype=w_flat.dtype, device=w_flat.device)
    for r in range(pl.r):
        for s in range(pl.s):
            ch = ((r & 1) * 2 + (s & 1)) * pl.c
            col = (r * pl.s + s) * pl.c
            w4[:, r >> 1, s >> 1, ch:ch + pl.c] = w_flat[:, col:col + pl.c]
    return w4


def unpack_grad_ref(pl: StemPlan, dw4: torch.Tensor) -> torch.Tensor:
    """Inverse gather of pack_weights_ref: [k, ta, 4, 16] -> [k, kpad] (pad columns zero)."""
    dw4 = dw4.reshape(pl.k, pl.ta, 4, 16)
    out = torch.zeros((pl.k, pl.kpad), dtype=dw4.dtype, device=dw4.device)
    for r in range(pl.r):
        for s in range(pl.s):
            ch = ((r & 1) * 2 + (s & 1)) * pl.c
            col = (r * pl.s + s) * pl.c
            out[:, col:col + pl.c] = dw4[:, r >> 1, s >> 1, ch:ch + pl.c]
    return out


def forward_ref(pl: StemPlan, x: torch.Tensor, w_flat: torch.Tensor) -> torch.Tensor:
    """The convolution as the kernels run it, in fp32 tensor ops: returns y[n, p, q, k]."""
    import torch.nn.functional as F
    s2d = regroup_ref(pl, x.float())                       # [n, h2, w2p, 16]
    w4 = pack_weights_ref(pl, w_flat.float())              # [k, ta, 4, 16]
    d = pl.desc
    xin = F.pad(s2d.permute(0, 3, 1, 2), (0, 0, d.pad_h_lo, max(d.pad_h_hi, 0)))
    if d.pad_h_hi < 0:
        xin = xin[:, :, :xin.shape[2] + d.pad_h_hi, :]
    y = F.conv2d(xin, w4.permute(0, 3, 1, 2))              # 4 horizontal taps = 4 pixels
    return y.permute(0, 2, 3, 1)[:, :pl.p, :pl.q, :]


__all__ = ["StemPlan", "eligible", "plan", "regroup", "pack_weights", "unpack_grad", "forward",
           "weight_grad", "regroup_ref", "pack_weights_ref", "unpack_grad_ref", "forward_ref"]
_ = ctypes
