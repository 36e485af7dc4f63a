"""Raw wrappers over the tcgen05 GEMM / implicit-GEMM convolution entry points.

These are the thin, allocation-explicit layer: NHWC fp16 activations, OHWI
fp16 filters, fp32 weight gradients. The autograd-facing ops live in
``ops/functional.py``. Reference behaviour: ``ConvNetBuilder.conv`` /
``affine`` (T/convnet_builder_legacy.py:143-226, 288-325).
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import native
from .native import ConvDesc, i32, i64, f32, ptr, stream_ptr


def conv_out_size(size: int, k: int, stride: int, lo: int, hi: int) -> int:
    return (size + lo + hi - k) // stride + 1


def resolve_padding(mode: str, in_size: int, k: int, stride: int) -> Tuple[int, int]:
    """(lo, hi) padding for TF-style padding modes.

    SAME: TF semantics (extra pixel goes to the end). VALID: none.
    SAME_RESNET: explicit pad (k-1)//2 at the start, rest at the end, then VALID
    (T/convnet_builder_legacy.py:182-199).
    """
    mode = mode.upper()
    if mode == "VALID":
        return 0, 0
    if mode == "SAME":
        out = -(-in_size // stride)
        total = max((out - 1) * stride + k - in_size, 0)
        return total // 2, total - total // 2
    if mode == "SAME_RESNET":
        if stride == 1:
            total = k - 1
            return total // 2, total - total // 2
        total = k - 1
        lo = total // 2
        return lo, total - lo
    raise ValueError(f"unknown padding mode {mode}")


def make_desc(x_shape, w_shape, stride, pad_h, pad_w) -> ConvDesc:
    n, h, w, c = x_shape
    k, r, s, c2 = w_shape
    assert c == c2, (x_shape, w_shape)
    sh, sw = stride
    d = ConvDesc()
    d.n, d.h, d.w, d.c = n, h, w, c
    d.k, d.r, d.s = k, r, s
    d.pad_h_lo, d.pad_h_hi = pad_h
    d.pad_w_lo, d.pad_w_hi = pad_w
    d.stride_h, d.stride_w = sh, sw
    d.p = conv_out_size(h, r, sh, *pad_h)
    d.q = conv_out_size(w, s, sw, *pad_w)
    return d


def _chk(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda or t.dtype != dtype or not t.is_contiguous():
        raise ValueError(f"{name}: expected contiguous CUDA {dtype}, got {t.dtype} "
                         f"cuda={t.is_cuda} contiguous={t.is_contiguous()}")


def gemm_tn(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None,
            relu: bool = False, out_dtype=torch.float16,
            out: Optional[torch.Tensor] = None,
            stats: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a[M,K] @ b[N,K].T`` (+bias) (ReLU)."""
    _chk(a, torch.float16, "a")
    _chk(b, torch.float16, "b")
    m, k = a.shape
    n, k2 = b.shape
    assert k == k2
    if out is None:
        out = torch.empty((m, n), device=a.device, dtype=out_dtype)
    native.call("gemm_tn", ptr(a), ptr(b), ptr(out), i32(m), i32(n), i32(k), i64(a.stride(0)),
                i64(b.stride(0)), i64(out.stride(0)), ptr(bias), i32(int(relu)),
                i32(int(out.dtype == torch.float32)), ptr(stats), stream_ptr())
    return out


def gemm_nt(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
            accumulate: bool = False, scale: float = 1.0,
            workspace: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a[K,M].T @ b[K,N]`` -> fp32 [M,N] (split-K through a workspace)."""
    _chk(a, torch.float16, "a")
    _chk(b, torch.float16, "b")
    k, m = a.shape
    k2, n = b.shape
    assert k == k2
    if out is None:
        out = torch.empty((m, n), device=a.device, dtype=torch.float32)
        accumulate = False
    if workspace is None:
        workspace = get_workspace(a.device, gemm_nt_workspace_bytes(m, n))
    native.call("gemm_nt", ptr(a), ptr(b), ptr(out), i32(m), i32(n), i32(k), i64(a.stride(0)),
                i64(b.stride(0)), i64(out.stride(0)), ptr(workspace),
                i64(workspace.numel() * workspace.element_size()), i32(int(accumulate)),
                f32(scale), stream_ptr())
    return out


def gemm_nt_workspace_bytes(m: int, n: int) -> int:
    bn = 64 if n <= 64 else (128 if n <= 128 else 256)
    tiles = ((m + 127) // 128) * ((n + bn - 1) // bn)
    splits = max(1, (2 * 148) // tiles)
    return m * n * 4 * splits


_workspaces = {}


def get_workspace(device, nbytes: int) -> torch.Tensor:
    """Grow-only fp32 scratch buffer per device (split-K partial sums)."""
    key = (device.type, device.index)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() * 4 < nbytes:
        ws = torch.empty(((nbytes + 3) // 4,), device=device, dtype=torch.float32)
        _workspaces[key] = ws
    return ws


def stats_layout(m: int, n: int, num_kb: int):
    """(rows, bn) of the fused BN statistics a GEMM of that shape writes: for every bn-wide
    column block, `rows` partial rows of [2][bn] floats (sum, sum of squares)."""
    rows, bn = ctypes.c_int(0), ctypes.c_int(0)
    native.check(native.load().kfb_gemm_stats_layout(i32(m), i32(n), i32(num_kb),
                                                     ctypes.byref(rows), ctypes.byref(bn)),
                 "gemm_stats_layout")
    return rows.value, bn.value


def conv_stats_layout(d: ConvDesc, ldy=None):
    """(rows, bn) of the statistics conv_fprop writes for `d` into rows of `ldy` elements."""
    rows, bn = ctypes.c_int(0), ctypes.c_int(0)
    native.check(native.load().kfb_conv_stats_layout(ctypes.byref(d), i64(ldy or d.k),
                                                     ctypes.byref(rows), ctypes.byref(bn)),
                 "conv_stats_layout")
    return rows.value, bn.value


def stats_buffer(m: int, n: int, num_kb: int, device) -> torch.Tensor:
    """Buffer for the fused statistics, shaped [column blocks][rows][2][bn]."""
    rows, bn = stats_layout(m, n, num_kb)
    return torch.empty(((n + bn - 1) // bn, rows, 2, bn), device=device, dtype=torch.float32)


def fold_stats(stats: torch.Tensor, n: int) -> torch.Tensor:
    """[2][n] totals of a statistics buffer (what the fused BN forward kernel folds)."""
    t = stats.sum(1)                                   # [blocks][2][bn]
    return t.permute(1, 0, 2).reshape(2, -1)[:, :n]


def conv_fprop(x, w, d: ConvDesc, bias=None, relu=False, out=None, ldy=None, stats=None):
    _chk(x, torch.float16, "x")
    _chk(w, torch.float16, "w")
    if out is None:
        out = torch.empty((d.n, d.p, d.q, d.k), device=x.device, dtype=torch.float16)
        ldy = d.k
    elif ldy is None:
        ldy = out.stride(-2)
    native.call("conv_fprop", ctypes.byref(d), ptr(x), ptr(w), ptr(out), i64(ldy), ptr(bias),
                i32(int(relu)), ptr(stats), stream_ptr())
    return out


def conv_dgrad(dy, w, d: ConvDesc, out=None, accumulate=False):
    _chk(dy, torch.float16, "dy")
    _chk(w, torch.float16, "w")
    zero_fill = bool(native.load().kfb_conv_dgrad_needs_zero_fill(ctypes.byref(d)))
    if out is None:
        alloc = torch.zeros if zero_fill else torch.empty
        out = alloc((d.n, d.h, d.w, d.c), device=dy.device, dtype=torch.float16)
        accumulate = False
    elif zero_fill and not accumulate:
        out.zero_()
    native.call("conv_dgrad", ctypes.byref(d), ptr(dy), ptr(w), ptr(out), i32(int(accumulate)),
                stream_ptr())
    return out


def gemm_nn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None,
            out_dtype=torch.float16, accumulate: bool = False, k: Optional[int] = None,
            n: Optional[int] = None) -> torch.Tensor:
    """``a[M,K] @ b[K,N]`` with b read in place (MN-major operand)."""
    _chk2d(a, "a")
    _chk2d(b, "b")
    m = a.shape[0]
    k = a.shape[1] if k is None else k
    n = b.shape[1] if n is None else n
    if out is None:
        out = torch.empty((m, n), device=a.device, dtype=out_dtype)
        accumulate = False
    native.call("gemm_nn", ptr(a), ptr(b), ptr(out), i32(m), i32(n), i32(k), i64(a.stride(0)),
                i64(b.stride(0)), i64(out.stride(0)), i32(int(out.dtype == torch.float32)),
                i32(int(accumulate)), stream_ptr())
    return out


def _chk2d(t: torch.Tensor, name: str):
    if not t.is_cuda or t.dtype != torch.float16 or t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name}: expected 2-D CUDA fp16 with unit inner stride")


def conv_wgrad(x, dy, d: ConvDesc, out=None, accumulate=False, scale=1.0, workspace=None):
    _chk(x, torch.float16, "x")
    _chk(dy, torch.float16, "dy")
    if out is None:
        out = torch.empty((d.k, d.r, d.s, d.c), device=x.device, dtype=torch.float32)
        accumulate = False
    if workspace is None:
        nbytes = native.load().kfb_conv_wgrad_workspace_bytes
        nbytes.restype = ctypes.c_longlong
        workspace = get_workspace(x.device, int(nbytes(ctypes.byref(d))))
    native.call("conv_wgrad", ctypes.byref(d), ptr(x), ptr(dy), i64(d.k), ptr(out),
                ptr(workspace), i64(workspace.numel() * 4), i32(int(accumulate)), f32(scale),
                stream_ptr())
    return out
