"""Kernels of the wider zoo (csrc/extra_ops.cu) vs plain PyTorch fp32 restatements."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from kf_benchmarks_b200.ops import native  # noqa: E402
from kf_benchmarks_b200.ops.native import ConvDesc, f32, i32, i64, ptr  # noqa: E402

u64 = ctypes.c_ulonglong


def _s():
    return native.stream_ptr()


@pytest.mark.parametrize("kind,name", [(2, "relu6"), (3, "tanh"), (4, "sigmoid")])
def test_activation_forward_backward(kind, name):
    g = torch.Generator(device="cuda").manual_seed(kind)
    rows, c = 333, 40
    x = (torch.randn(rows, c, device="cuda", generator=g) * 4).half()
    dy = torch.randn(rows, c, device="cuda", generator=g).half()
    y, dx = torch.empty_like(x), torch.empty_like(x)
    native.call("act_forward", ptr(x), i64(c), ptr(y), i64(c), i64(rows), i32(c), i32(kind), _s())
    native.call("act_backward", ptr(dy), i64(c), ptr(y), i64(c), ptr(dx), i64(c), i32(0),
                i64(rows), i32(c), i32(kind), _s())
    xf = x.float().requires_grad_(True)
    yr = {"relu6": F.relu6, "tanh": torch.tanh, "sigmoid": torch.sigmoid}[name](xf)
    yr.backward(dy.float())
    assert torch.allclose(y.float(), yr, atol=2e-3, rtol=2e-3)
    assert torch.allclose(dx.float(), xf.grad, atol=1e-2, rtol=1e-2)
    # accumulate
    native.call("act_backward", ptr(dy), i64(c), ptr(y), i64(c), ptr(dx), i64(c), i32(1),
                i64(rows), i32(c), i32(kind), _s())
    assert torch.allclose(dx.float(), 2 * xf.grad, atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("n,h,w,c,r,stride,pad", [
    (2, 14, 14, 32, 3, 1, (1, 1)), (3, 15, 15, 96, 3, 2, (1, 1)), (2, 12, 12, 48, 5, 2, (1, 2)),
    (1, 9, 9, 264, 7, 1, (3, 3)), (2, 8, 8, 8, 3, 2, (0, 1))])
def test_depthwise_conv(n, h, w, c, r, stride, pad):
    g = torch.Generator(device="cuda").manual_seed(c)
    x = torch.randn(n, h, w, c, device="cuda", generator=g).half()
    wt = (torch.randn(r, r, c, device="cuda", generator=g) * 0.3).half()
    p = (h + pad[0] + pad[1] - r) // stride + 1
    d = ConvDesc()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s = n, h, w, c, c, r, r
    d.pad_h_lo, d.pad_h_hi, d.pad_w_lo, d.pad_w_hi = pad[0], pad[1], pad[0], pad[1]
    d.stride_h = d.stride_w = stride
    d.p = d.q = p
    y = torch.empty(n, p, p, c, device="cuda", dtype=torch.float16)
    native.call("dwconv_fprop", ctypes.byref(d), ptr(x), i64(c), ptr(wt), ptr(y), i64(c), _s())
    xf = x.float().requires_grad_(True)
    wf = wt.float().requires_grad_(True)
    xp = F.pad(xf.permute(0, 3, 1, 2), (pad[0], pad[1], pad[0], pad[1]))
    yr = F.conv2d(xp, wf.permute(2, 0, 1).unsqueeze(1), stride=stride, groups=c)
    yr = yr.permute(0, 2, 3, 1)
    assert yr.shape == y.shape
    assert torch.allclose(y.float(), yr, atol=2e-2, rtol=2e-2)
    dy = torch.randn(y.shape, device="cuda", generator=g).half()
    yr.backward(dy.float())
    dx = torch.empty_like(x)
    native.call("dwconv_dgrad", ctypes.byref(d), ptr(dy), i64(c), ptr(wt), ptr(dx), i64(c), i32(0),
                _s())
    assert torch.allclose(dx.float(), xf.grad, atol=2e-2, rtol=2e-2)
    fn = native.load().kfb_dwconv_wgrad_workspace_bytes
    fn.restype = ctypes.c_longlong
    ws = torch.zeros(int(fn(ctypes.byref(d))) // 4 + 16, device="cuda")
    dw = torch.zeros(r, r, c, device="cuda")
    native.call("dwconv_wgrad", ctypes.byref(d), ptr(x), i64(c), ptr(dy), i64(c), ptr(dw), ptr(ws),
                i64(ws.numel() * 4), i32(0), f32(1.0), _s())
    ref = wf.grad
    assert (dw - ref).abs().max() <= 2e-2 * ref.abs().max() + 1e-2


def test_drop_path_is_per_sample_and_replayable():
    n, hw, c = 64, 9, 16
    x = torch.ones(n * hw, c, device="cuda", dtype=torch.float16)
    y, y2 = torch.empty_like(x), torch.empty_like(x)
    args = (i64(n * hw), i64(hw), i32(c), f32(0.75), u64(99), u64(12345))
    native.call("drop_path", ptr(x), i64(c), ptr(y), i64(c), i32(0), *args, _s())
    native.call("drop_path", ptr(x), i64(c), ptr(y2), i64(c), i32(0), *args, _s())
    assert torch.equal(y, y2)
    per = y.float().view(n, hw * c)
    assert bool(((per == 0).all(1) | torch.isclose(per, torch.tensor(1 / 0.75, device="cuda"),
                                                    atol=1e-3).all(1)).all())
    kept = int((per[:, 0] > 0).sum())
    assert 32 <= kept <= 60


def test_channel_gather_and_inverse():
    rows, cin, cout = 77, 32, 24
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(rows, cin, device="cuda", generator=g).half()
    index = [-1] * cout
    for dst, src in zip((0, 1, 2, 5, 9, 10, 11, 17, 23), (3, 2, 1, 30, 16, 17, 18, 0, 31)):
        index[dst] = src
    idx = torch.tensor(index, dtype=torch.int32, device="cuda")
    y = torch.empty(rows, cout, device="cuda", dtype=torch.float16)
    native.call("channel_gather", ptr(x), i64(cin), ptr(y), i64(cout), i32(0), i64(rows),
                i32(cout), ptr(idx), _s())
    li = idx.long()
    ref = x[:, li.clamp(min=0)] * (li >= 0)
    assert torch.equal(y, ref.half())
    native.call("channel_gather", ptr(x), i64(cin), ptr(y), i64(cout), i32(1), i64(rows),
                i32(cout), ptr(idx), _s())
    assert torch.allclose(y.float(), 2 * ref.float(), atol=1e-2)


def test_embedding_gather_scatter_and_mul():
    g = torch.Generator(device="cuda").manual_seed(11)
    rows_t, dim, n, lo, cols = 500, 48, 1000, 16, 32
    table = torch.randn(rows_t, dim, device="cuda", generator=g).half()
    ids = torch.randint(0, rows_t, (n,), device="cuda", generator=g, dtype=torch.int32)
    y = torch.empty(n, cols, device="cuda", dtype=torch.float16)
    native.call("embedding_gather", ptr(table), i64(dim), i32(lo), ptr(ids), i64(rows_t), ptr(y),
                i64(cols), i64(n), i32(cols), _s())
    assert torch.equal(y, table[ids.long(), lo:lo + cols])
    dy = torch.randn(n, cols, device="cuda", generator=g).half()
    dt = torch.zeros(rows_t, dim, device="cuda")
    native.call("embedding_scatter_add", ptr(dy), i64(cols), ptr(ids), i64(rows_t), ptr(dt),
                i64(dim), i32(lo), i64(n), i32(cols), _s())
    ref = torch.zeros(rows_t, dim, device="cuda")
    ref[:, lo:lo + cols].index_add_(0, ids.long(), dy.float())
    assert torch.allclose(dt, ref, atol=1e-3, rtol=1e-3)
    a = torch.randn(n, cols, device="cuda", generator=g).half()
    b = torch.randn(n, cols, device="cuda", generator=g).half()
    out, da = torch.empty_like(a), torch.empty_like(a)
    native.call("mul_forward", ptr(a), i64(cols), ptr(b), i64(cols), ptr(out), i64(cols), i64(n),
                i32(cols), _s())
    assert torch.allclose(out.float(), a.float() * b.float(), atol=1e-2, rtol=1e-2)
    native.call("mul_backward", ptr(dy), i64(cols), ptr(b), i64(cols), ptr(da), i64(cols), i32(0),
                i64(n), i32(cols), _s())
    assert torch.allclose(da.float(), dy.float() * b.float(), atol=1e-2, rtol=1e-2)


def test_ncf_loss_matches_two_class_softmax():
    g = torch.Generator(device="cuda").manual_seed(3)
    n = 3000
    z = torch.randn(n, 8, device="cuda", generator=g) * 3
    labels = torch.randint(0, 2, (n,), device="cuda", generator=g, dtype=torch.int32)
    dz = torch.zeros(n, 8, device="cuda", dtype=torch.float16)
    out = torch.zeros(8, device="cuda")
    native.call("ncf_loss", ptr(z), i64(8), ptr(labels), i32(n), f32(64.0 / n), None, ptr(dz),
                i64(8), ptr(out), _s())
    zf = z[:, 0].clone().requires_grad_(True)
    two = torch.stack([torch.ones_like(zf), zf], dim=1)
    loss = F.cross_entropy(two, labels.long())
    (loss * 64.0).backward()
    assert abs(float(out[0]) - float(loss)) < 1e-4
    assert torch.allclose(dz[:, 0].float(), zf.grad, atol=2e-4, rtol=1e-2)
    assert int(out[1]) == int(((z[:, 0] > 1.0) == (labels != 0)).sum())


def test_ssd_multibox_loss_and_gradients():
    """csrc/loss_kernels.cu vs nn/losses_ref.ssd_loss on small heads with real positives."""
    from kf_benchmarks_b200.nn import losses_ref
    g = torch.Generator(device="cuda").manual_seed(21)
    n, L = 3, 11
    sizes, ndef = (6, 4, 3, 2, 1, 1), (4, 6, 6, 6, 4, 4)
    locs = [torch.randn(n, s, s, d * 4, device="cuda", generator=g).half()
            for s, d in zip(sizes, ndef)]
    lds = [((d * L + 7) // 8) * 8 for d in ndef]                      # padded rows (phantom lanes)
    confs = [(torch.randn(n, s, s, ld, device="cuda", generator=g) * 2).half()
             for s, ld in zip(sizes, lds)]
    A = sum(s * s * d for s, d in zip(sizes, ndef))
    gt_loc = torch.randn(n, A, 4, device="cuda", generator=g) * 1.5
    gt_label = (torch.rand(n, A, device="cuda", generator=g) < 0.1).float() * \
        torch.randint(1, L, (n, A), device="cuda", generator=g).float()
    nm = torch.tensor([3.0, 1.7, 9.2], device="cuda")
    dlocs = [torch.full_like(t, 7.0) for t in locs]
    dconfs = [torch.zeros_like(t) for t in confs]
    fn = native.load().kfb_ssd_loss_workspace_floats
    fn.restype = ctypes.c_longlong
    ws = torch.zeros(int(fn(i32(n), i32(A))), device="cuda")
    out = torch.zeros(8, device="cuda")

    def parr(ts):
        return (ctypes.c_void_p * 6)(*[t.data_ptr() for t in ts])
    scale = 128.0
    native.call("ssd_loss", i32(n), i32(L), (ctypes.c_int * 6)(*sizes), (ctypes.c_int * 6)(*ndef),
                parr(locs), (ctypes.c_longlong * 6)(*[d * 4 for d in ndef]), parr(confs),
                (ctypes.c_longlong * 6)(*lds), parr(dlocs), parr(dconfs), ptr(gt_loc),
                ptr(gt_label), ptr(nm), i32(3), f32(scale / n), None, ptr(ws), i64(ws.numel()),
                ptr(out), _s())
    # oracle
    lf = [t.float().requires_grad_(True) for t in locs]
    cf = [t.float().requires_grad_(True) for t in confs]
    pl = torch.cat([t.reshape(n, s, s, d, 4).permute(0, 3, 1, 2, 4).reshape(n, -1, 4)
                    for t, s, d in zip(lf, sizes, ndef)], 1)
    pc = torch.cat([t[..., :d * L].reshape(n, s, s, d, L).permute(0, 3, 1, 2, 4).reshape(n, -1, L)
                    for t, s, d in zip(cf, sizes, ndef)], 1)
    total, cls, box = losses_ref.ssd_loss(pl, pc, gt_loc, gt_label, nm, 3)
    (total * scale).backward()
    assert abs(float(out[0]) - float(total)) < 2e-3 * float(total)
    assert abs(float(out[4]) - float(cls)) < 2e-3 * float(cls)
    assert abs(float(out[5]) - float(box)) < 2e-3 * float(box) + 1e-4
    for got, ref in zip(dlocs + dconfs, lf + cf):
        assert torch.allclose(got.float(), ref.grad, atol=2e-2, rtol=2e-2)


def test_ctc_loss_matches_reference_formulation():
    from kf_benchmarks_b200.nn import ctc_ref
    g = torch.Generator(device="cuda").manual_seed(8)
    n, T, C, lmax = 5, 40, 29, 12
    z = torch.randn(n * T, 32, device="cuda", generator=g) * 2
    labels = torch.randint(0, C - 1, (n, lmax), device="cuda", generator=g, dtype=torch.int32)
    labels[1, 3] = labels[1, 2]                      # a repeated symbol needs a blank in between
    in_len = torch.tensor([40, 33, 40, 9, 25], dtype=torch.int32, device="cuda")
    lab_len = torch.tensor([12, 7, 1, 12, 12], dtype=torch.int32, device="cuda")   # sample 3 infeasible
    fn = native.load().kfb_ctc_loss_workspace_floats
    fn.restype = ctypes.c_longlong
    ws = torch.zeros(int(fn(i32(n), i32(T), i32(lmax))), device="cuda")
    dz = torch.full((n * T, 32), 5.0, device="cuda", dtype=torch.float16)
    out = torch.zeros(8, device="cuda")
    scale = 64.0
    native.call("ctc_loss", ptr(z), i64(32), i32(n), i32(T), i32(C), ptr(labels), i32(lmax),
                ptr(in_len), ptr(lab_len), f32(1e-7), f32(scale / n), None, ptr(dz), i64(32),
                ptr(ws), i64(ws.numel()), ptr(out), _s())
    zf = z[:, :C].reshape(n, T, C).clone().requires_grad_(True)
    loss = ctc_ref.ctc_loss(zf, labels, in_len, lab_len)
    (loss * scale).backward()
    assert abs(float(out[0]) - float(loss)) < 1e-3 * float(loss)
    got = dz[:, :C].float().reshape(n, T, C)
    assert torch.allclose(got, zf.grad, atol=3e-3, rtol=2e-2)
    assert float(got[3].abs().max()) == 0.0           # ignored sample
