"""Memory-bound layer kernels vs plain PyTorch fp32 restatements of the same op."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from kf_benchmarks_b200.ops import native  # noqa: E402
from kf_benchmarks_b200.ops.native import f32, i32, i64, ptr  # noqa: E402

u64 = ctypes.c_ulonglong


def _s():
    return native.stream_ptr()


def _scratch(rows, c):
    nb = native.load().kfb_reduce_blocks(ctypes.c_longlong(rows), ctypes.c_int(c))
    return torch.zeros(nb * 2 * c, device="cuda")


@pytest.mark.parametrize("form", ["three_kernels", "single_launch", "single_launch_bitmask"])
@pytest.mark.parametrize("rows,c,relu,res,scale", [(4 * 7 * 7, 64, True, False, True),
                                                   (3 * 14 * 14, 256, True, True, True),
                                                   (1000, 2048, False, False, False),
                                                   (777, 24, True, True, True),
                                                   (64 * 28 * 28, 128, True, True, True),
                                                   (37, 4096, True, False, True)])
def test_bn_forward_backward(rows, c, relu, res, scale, form):
    """Both forms of the training batch norm (nn_kernels.cu: reduce / finalize / apply kernels;
    bn_fused.cu: one persistent kernel per direction with grid barriers, optionally with the
    packed ReLU mask) against the fp32 autograd restatement."""
    fused = form != "three_kernels"
    use_bits = form == "single_launch_bitmask" and relu
    if form == "single_launch_bitmask" and not relu:
        pytest.skip("no ReLU, no mask")
    g = torch.Generator(device="cuda").manual_seed(0)
    x = (torch.randn(rows, c, device="cuda", generator=g) * 2 + 0.5).half()
    r = torch.randn(rows, c, device="cuda", generator=g).half() if res else None
    gamma = (torch.rand(c, device="cuda", generator=g) + 0.5) if scale else None
    beta = torch.randn(c, device="cuda", generator=g)
    mm, mv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    mean_rstd, ss, coef = (torch.zeros(2 * c, device="cuda"), torch.zeros(2 * c, device="cuda"),
                           torch.zeros(3 * c, device="cuda"))
    y = torch.empty_like(x)
    eps, decay = 1e-5, 0.9
    scratch = _scratch(rows, c)
    sync = torch.zeros(2, dtype=torch.int32, device="cuda")
    bits = torch.zeros(rows * (c // 8), dtype=torch.uint8, device="cuda") if use_bits else None
    if fused:
        scratch = torch.zeros(int(native.load().kfb_bn_fused_max_grid()) * 2 * c, device="cuda")
        native.call("bn_forward_train_fused", ptr(x), i64(c), i64(rows), i32(c), ptr(gamma),
                    ptr(beta), ptr(mm), ptr(mv), f32(eps), f32(decay), i32(1), ptr(mean_rstd),
                    ptr(ss), None, i32(0), i32(0), ptr(scratch), ptr(r), i64(c), i32(int(relu)),
                    ptr(y), i64(c), ptr(bits), ptr(sync), _s())
    else:
        native.call("bn_forward_train", ptr(x), i64(c), i64(rows), i32(c), ptr(gamma), ptr(beta),
                    ptr(mm), ptr(mv), f32(eps), f32(decay), i32(1), ptr(mean_rstd), ptr(ss),
                    ptr(scratch), i32(0), ptr(r), i64(c), i32(int(relu)), ptr(y), i64(c), _s())
    # oracle
    xf = x.float().requires_grad_(True)
    gf = gamma.clone().requires_grad_(True) if scale else None
    bf = beta.clone().requires_grad_(True)
    rf = r.float().requires_grad_(True) if res else None
    mean, var = xf.mean(0), xf.var(0, unbiased=False)
    yr = (xf - mean) * torch.rsqrt(var + eps) * (gf if scale else 1.0) + bf
    if res:
        yr = yr + rf
    if relu:
        yr = F.relu(yr)
    assert torch.allclose(y.float(), yr, atol=2e-2, rtol=2e-2)
    assert torch.allclose(mm, mean.detach() * (1 - decay), atol=1e-3, rtol=1e-3)
    unb = var.detach() * rows / (rows - 1)
    assert torch.allclose(mv, decay + unb * (1 - decay), atol=1e-3, rtol=1e-3)
    # backward
    dy = torch.randn(rows, c, device="cuda", generator=g).half()
    yr.backward(dy.float())
    dz = torch.empty_like(x)
    dres = torch.empty_like(x) if res else None
    dgamma = torch.zeros(c, device="cuda") if scale else None
    dbeta = torch.zeros(c, device="cuda")
    use_affine = relu and not res       # mask recomputed from x through the forward affine
    if fused:
        mask = None if use_bits else (ptr(y) if (relu and res) else None)
        native.call("bn_backward_fused", ptr(dy), i64(c), mask, i64(c), ptr(bits),
                    ptr(ss) if (use_affine and not use_bits) else None, ptr(x), i64(c), i64(rows),
                    i32(c), ptr(gamma), ptr(mean_rstd), ptr(dgamma), ptr(dbeta), i32(0), f32(1.0),
                    ptr(coef), ptr(scratch), ptr(dz), i64(c), ptr(dres), i64(c), i32(0),
                    ptr(sync), _s())
    else:
        native.call("bn_backward", ptr(dy), i64(c), ptr(y) if (relu and res) else None, i64(c),
                    ptr(ss) if use_affine else None, ptr(x), i64(c),
                    i64(rows), i32(c), ptr(gamma), ptr(mean_rstd), ptr(dgamma), ptr(dbeta), i32(0),
                    f32(1.0), ptr(coef), ptr(scratch), ptr(dz), i64(c), ptr(dres), i64(c), i32(0),
                    _s())
    torch.cuda.synchronize()
    assert int(sync[0]) == 0      # the grid barrier leaves its arrival count at zero
    # the ReLU mask comes from the fp16 output; elements within rounding of zero may differ
    tol = dict(atol=3e-2, rtol=3e-2)
    assert torch.allclose(dbeta, bf.grad, atol=0.15, rtol=2e-2)
    if scale:
        assert torch.allclose(dgamma, gf.grad, atol=0.3, rtol=3e-2)
    frac_bad = ((dz.float() - xf.grad).abs() > 3e-2 + 3e-2 * xf.grad.abs()).float().mean()
    assert frac_bad < 2e-3
    if res:
        frac_bad = ((dres.float() - rf.grad).abs() > 3e-2).float().mean()
        assert frac_bad < 2e-3


@pytest.mark.parametrize("kind,k,s,pad", [("max", 3, 2, "SAME"), ("max", 2, 2, "VALID"),
                                          ("avg", 3, 1, "SAME"), ("avg", 2, 2, "VALID"),
                                          ("max", 3, 2, "VALID"), ("avg", 1, 2, "VALID"),
                                          ("max", 3, 1, "SAME")])
@pytest.mark.parametrize("hw", [(15, 15), (16, 22)])
def test_pool(kind, k, s, pad, hw):
    from kf_benchmarks_b200.nn.builder import out_size, resolve_padding
    n, (h, w), c = 2, hw, 24
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(n, h, w, c, device="cuda", generator=g).half()
    ph, pw = resolve_padding(pad, h, k, s), resolve_padding(pad, w, k, s)
    p, q = out_size(h, k, s, ph), out_size(w, k, s, pw)
    y = torch.empty(n, p, q, c, device="cuda", dtype=torch.float16)
    arg = torch.zeros(n, p, q, c, device="cuda", dtype=torch.uint8)
    is_max = int(kind == "max")
    geom = (i32(n), i32(h), i32(w), i32(c), i32(p), i32(q), i32(k), i32(k), i32(s), i32(s),
            i32(ph[0]), i32(pw[0]), i32(is_max))
    native.call("pool_forward", ptr(x), i64(c), ptr(y), i64(c), ptr(arg), *geom, _s())
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    if kind == "max":
        yr = F.max_pool2d(F.pad(xf, (pw[0], pw[1], ph[0], ph[1]), value=float("-inf")), k, s)
    else:
        num = F.avg_pool2d(F.pad(xf, (pw[0], pw[1], ph[0], ph[1])), k, s, divisor_override=1)
        cnt = F.avg_pool2d(F.pad(torch.ones_like(xf[:1, :1]), (pw[0], pw[1], ph[0], ph[1])), k, s,
                           divisor_override=1)
        yr = num / cnt
    assert torch.allclose(y.float(), yr.permute(0, 2, 3, 1), atol=2e-3, rtol=2e-3)
    dy = torch.randn(n, p, q, c, device="cuda", generator=g).half()
    yr.backward(dy.float().permute(0, 3, 1, 2))
    dx = torch.empty_like(x)
    native.call("pool_backward", ptr(dy), i64(c), ptr(arg), ptr(dx), i64(c), i32(0), *geom, _s())
    assert torch.allclose(dx.float(), xf.grad.permute(0, 2, 3, 1), atol=5e-3, rtol=5e-3)
    # accumulate
    native.call("pool_backward", ptr(dy), i64(c), ptr(arg), ptr(dx), i64(c), i32(1), *geom, _s())
    assert torch.allclose(dx.float(), 2 * xf.grad.permute(0, 2, 3, 1), atol=1e-2, rtol=1e-2)


def test_spatial_mean_and_copy_and_add():
    n, hw, c = 3, 49, 64
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn(n * hw, c, device="cuda", generator=g).half()
    y = torch.empty(n, c, device="cuda", dtype=torch.float16)
    native.call("spatial_mean_forward", ptr(x), i64(c), ptr(y), i64(c), i32(n), i32(hw), i32(c),
                _s())
    assert torch.allclose(y.float(), x.float().view(n, hw, c).mean(1), atol=2e-3)
    dy = torch.randn(n, c, device="cuda", generator=g).half()
    dx = torch.empty_like(x)
    native.call("spatial_mean_backward", ptr(dy), i64(c), ptr(dx), i64(c), i32(0), i32(n), i32(hw),
                i32(c), _s())
    assert torch.allclose(dx.float().view(n, hw, c), (dy.float() / hw)[:, None, :].expand(n, hw, c),
                          atol=1e-3)
    # strided copy into a wider buffer + accumulate
    wide = torch.zeros(n * hw, 2 * c, device="cuda", dtype=torch.float16)
    native.call("copy2d", ptr(x), i64(c), ctypes.c_void_p(wide.data_ptr() + c * 2), i64(2 * c),
                i64(n * hw), i32(c), i32(0), _s())
    assert torch.equal(wide[:, c:], x) and float(wide[:, :c].abs().sum()) == 0
    native.call("copy2d", ptr(x), i64(c), ctypes.c_void_p(wide.data_ptr() + c * 2), i64(2 * c),
                i64(n * hw), i32(c), i32(1), _s())
    assert torch.allclose(wide[:, c:].float(), 2 * x.float(), atol=1e-2)
    out = torch.empty_like(x)
    native.call("add_act", ptr(x), i64(c), ptr(x), i64(c), ptr(out), i64(c), i64(n * hw), i32(c),
                i32(1), _s())
    assert torch.allclose(out.float(), F.relu(2 * x.float()), atol=1e-2)


@pytest.mark.parametrize("rows,classes", [(8, 1001), (256, 1001), (33, 11)])
def test_softmax_xent(rows, classes):
    g = torch.Generator(device="cuda").manual_seed(3)
    ld = (classes + 7) // 8 * 8
    logits = torch.zeros(rows, ld, device="cuda")
    logits[:, :classes] = torch.randn(rows, classes, device="cuda", generator=g) * 3
    labels = torch.randint(0, classes, (rows,), device="cuda", generator=g, dtype=torch.int32)
    dl = torch.empty(rows, ld, device="cuda", dtype=torch.float16)
    loss_rows = torch.empty(rows, device="cuda")
    t1 = torch.empty(rows, device="cuda", dtype=torch.int32)
    t5 = torch.empty(rows, device="cuda", dtype=torch.int32)
    out3 = torch.zeros(4, device="cuda")
    scale = 128.0
    native.call("softmax_xent", ptr(logits), i64(ld), ptr(labels), i32(rows), i32(classes),
                f32(scale / rows), None, ptr(dl), i64(ld), ptr(loss_rows), ptr(t1), ptr(t5),
                ptr(out3), _s())
    lf = logits[:, :classes].clone().requires_grad_(True)
    loss = F.cross_entropy(lf, labels.long())
    (loss * scale).backward()
    assert abs(float(out3[0]) - float(loss)) < 1e-4
    assert torch.allclose(dl[:, :classes].float(), lf.grad, atol=2e-3, rtol=2e-2)
    assert float(dl[:, classes:].abs().sum()) == 0
    top5 = lf.detach().topk(5, dim=1).indices
    assert int(out3[1]) == int((top5[:, 0] == labels).sum())
    assert int(out3[2]) == int((top5 == labels[:, None].long()).any(1).sum())


def test_bias_act_backward_and_im2col():
    rows, c = 1000, 72
    g = torch.Generator(device="cuda").manual_seed(4)
    dy = torch.randn(rows, c, device="cuda", generator=g).half()
    y = torch.randn(rows, c, device="cuda", generator=g).half()
    dz = torch.empty_like(dy)
    db = torch.zeros(c, device="cuda")
    native.call("bias_act_backward", ptr(dy), i64(c), ptr(y), i64(c), i64(rows), i32(c), ptr(dz),
                i64(c), ptr(db), i32(0), f32(1.0), ptr(_scratch(rows, c)), _s())
    ref = dy.float() * (y.float() > 0)
    assert torch.allclose(dz.float(), ref, atol=1e-3)
    assert torch.allclose(db, ref.sum(0), atol=5e-2, rtol=1e-3)
    # explicit im2col of a 3-channel image
    n, h, w, cc, k, s = 2, 12, 12, 3, 5, 2
    x = torch.randn(n, h, w, cc, device="cuda", generator=g).half()
    pad = 2
    p = (h + 2 * pad - k) // s + 1
    kpad = (k * k * cc + 7) // 8 * 8
    col = torch.zeros(n * p * p, kpad, device="cuda", dtype=torch.float16)
    native.call("im2col_small_c", ptr(x), ptr(col), i32(n), i32(h), i32(w), i32(cc), i32(p), i32(p),
                i32(k), i32(k), i32(s), i32(s), i32(pad), i32(pad), i32(kpad), _s())
    unf = F.unfold(x.float().permute(0, 3, 1, 2), k, padding=pad, stride=s)  # [n, c*k*k, L]
    unf = unf.view(n, cc, k * k, -1).permute(0, 3, 2, 1).reshape(n * p * p, k * k * cc)
    assert torch.allclose(col[:, :k * k * cc].float(), unf, atol=1e-3)


def test_dropout_and_lrn_and_synth():
    n = 1 << 16
    x = torch.ones(n, device="cuda", dtype=torch.float16)
    y = torch.empty_like(x)
    native.call("dropout", ptr(x), ptr(y), i64(n), f32(0.5), u64(7), u64(123), _s())
    kept = float((y > 0).float().mean())
    assert abs(kept - 0.5) < 0.02 and float(y.max()) == 2.0
    y2 = torch.empty_like(x)
    native.call("dropout", ptr(x), ptr(y2), i64(n), f32(0.5), u64(7), u64(123), _s())
    assert torch.equal(y, y2)  # same (seed, offset) => same mask (backward relies on it)
    rows, c = 50, 64
    g = torch.Generator(device="cuda").manual_seed(5)
    a = torch.randn(rows, c, device="cuda", generator=g).half()
    out = torch.empty_like(a)
    native.call("lrn_forward", ptr(a), ptr(out), i64(rows), i32(c), i32(4), f32(1.0),
                f32(0.001 / 9), f32(0.75), _s())
    af = a.float().requires_grad_(True)
    sq = F.pad((af * af).unsqueeze(1), (4, 4)).unfold(-1, 9, 1).sum(-1).squeeze(1)
    ref = af / (1.0 + 0.001 / 9 * sq) ** 0.75
    assert torch.allclose(out.float(), ref, atol=2e-3, rtol=2e-3)
    dy = torch.randn(rows, c, device="cuda", generator=g).half()
    ref.backward(dy.float())
    dx = torch.empty_like(a)
    native.call("lrn_backward", ptr(a), ptr(dy), ptr(dx), i32(0), i64(rows), i32(c), i32(4),
                f32(1.0), f32(0.001 / 9), f32(0.75), _s())
    assert torch.allclose(dx.float(), af.grad, atol=3e-3, rtol=3e-3)
    img = torch.empty(1 << 18, device="cuda", dtype=torch.float16)
    native.call("synth_images", ptr(img), i64(img.numel()), f32(127.0), f32(60.0), u64(99), _s())
    f = img.float()
    assert abs(float(f.mean()) - 127) < 1.0 and 50 < float(f.std()) < 56  # truncated at 2 sigma
    assert float(f.min()) >= 127 - 120.5 and float(f.max()) <= 127 + 120.5
    lab = torch.empty(4096, device="cuda", dtype=torch.int32)
    native.call("synth_labels", ptr(lab), i32(4096), i32(0), i32(1000), u64(5), _s())
    assert int(lab.min()) >= 0 and int(lab.max()) < 1000 and lab.float().std() > 200
