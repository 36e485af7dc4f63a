"""tcgen05 GEMM / implicit-GEMM convolution kernels vs fp32 PyTorch restatements."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

import kernel_check  # noqa: E402


@pytest.mark.parametrize("name", sorted(kernel_check.CASES))
def test_case(name):
    r = kernel_check.run_case(name)
    assert r["finite"] and r["ok"], r


def test_accumulate_and_gemm_nn():
    import torch
    from kf_benchmarks_b200.ops import gemm_conv as gc
    g = torch.Generator(device="cuda").manual_seed(0)
    a = torch.randn(300, 1001, device="cuda", generator=g).half()
    apad = torch.zeros(300, 1008, device="cuda", dtype=torch.float16)
    apad[:, :1001] = a
    b = (torch.randn(1001, 2048, device="cuda", generator=g) * 0.03).half()
    out = gc.gemm_nn(apad, b, k=1001)
    ref = a.float() @ b.float()
    assert torch.allclose(out.float(), ref, atol=2e-2, rtol=2e-2)
    out2 = out.clone()
    gc.gemm_nn(apad, b, out=out2, accumulate=True, k=1001)
    assert torch.allclose(out2.float(), 2 * ref, atol=4e-2, rtol=2e-2)
    # dgrad accumulation into an existing gradient
    x = torch.randn(2, 14, 14, 64, device="cuda", generator=g).half()
    w = (torch.randn(128, 3, 3, 64, device="cuda", generator=g) * 0.05).half()
    d = gc.make_desc(x.shape, w.shape, (1, 1), (1, 1), (1, 1))
    dy = torch.randn(2, 14, 14, 128, device="cuda", generator=g).half()
    dx = gc.conv_dgrad(dy, w, d)
    dx2 = dx.clone()
    gc.conv_dgrad(dy, w, d, out=dx2, accumulate=True)
    assert torch.allclose(dx2.float(), 2 * dx.float(), atol=2e-2, rtol=1e-2)
    # weight gradients are deterministic run to run (fixed-order split-K reduction)
    g1 = gc.conv_wgrad(x, dy, d)
    g2 = gc.conv_wgrad(x, dy, d)
    assert torch.equal(g1, g2)


def test_fused_bn_statistics_from_conv_epilogue():
    import torch
    from kf_benchmarks_b200.ops import gemm_conv as gc
    g = torch.Generator(device="cuda").manual_seed(3)
    for (n, h, c, k, r) in [(4, 28, 64, 256, 1), (3, 14, 128, 64, 3), (2, 9, 64, 320, 3), (64, 28, 64, 128, 1),
                            (40, 56, 64, 64, 1)]:
        x = torch.randn(n, h, h, c, device="cuda", generator=g).half()
        w = (torch.randn(k, r, r, c, device="cuda", generator=g) * (r * r * c) ** -0.5).half()
        pad = (r // 2, r // 2)
        d = gc.make_desc(x.shape, w.shape, (1, 1), pad, pad)
        stats = gc.stats_buffer(d.n * d.p * d.q, k, r * r * ((c + 63) // 64), "cuda")
        stats.fill_(7.0)   # every row the fold reads is written by the kernel, none accumulated
        y = gc.conv_fprop(x, w, d, stats=stats)
        torch.cuda.synchronize()
        yf = y.float().reshape(-1, k)
        tot = gc.fold_stats(stats, k)
        assert torch.allclose(tot[0], yf.sum(0), atol=0.05, rtol=1e-3)
        assert torch.allclose(tot[1], (yf * yf).sum(0), atol=0.05, rtol=1e-3)
        # the single-launch BN forward consumes the buffer in place of its own statistics pass
        from kf_benchmarks_b200.ops import native
        from kf_benchmarks_b200.ops.native import f32, i32, i64, ptr
        rows = yf.shape[0]
        srows, sbn = gc.conv_stats_layout(d)
        gamma = torch.rand(k, device="cuda", generator=g) + 0.5
        beta = torch.randn(k, device="cuda", generator=g)
        mm, mv = torch.zeros(k, device="cuda"), torch.ones(k, device="cuda")
        mean_rstd, ss = torch.zeros(2 * k, device="cuda"), torch.zeros(2 * k, device="cuda")
        sync = torch.zeros(2, dtype=torch.int32, device="cuda")
        z = torch.empty_like(y)
        native.call("bn_forward_train_fused", ptr(y), i64(k), i64(rows), i32(k), ptr(gamma),
                    ptr(beta), ptr(mm), ptr(mv), f32(1e-5), f32(0.9), i32(1), ptr(mean_rstd),
                    ptr(ss), ptr(stats), i32(srows), i32(sbn), None, None, i64(0), i32(1), ptr(z),
                    i64(k), None, ptr(sync), native.stream_ptr())
        torch.cuda.synchronize()
        ref = torch.relu((yf - yf.mean(0)) * torch.rsqrt(yf.var(0, unbiased=False) + 1e-5) * gamma
                         + beta)
        assert torch.allclose(z.float().reshape(-1, k), ref, atol=2e-2, rtol=2e-2)
        # ... and so does the two-launch form (finalize + apply), the default in training; with a
        # residual and the packed ReLU mask here
        res = torch.randn_like(y)
        z2 = torch.empty_like(y)
        bits = torch.zeros(rows * (k // 8), dtype=torch.uint8, device="cuda")
        mm2, mv2 = torch.zeros(k, device="cuda"), torch.ones(k, device="cuda")
        native.call("bn_forward_train_stats", ptr(y), i64(k), i64(rows), i32(k), ptr(gamma),
                    ptr(beta), ptr(mm2), ptr(mv2), f32(1e-5), f32(0.9), i32(1), ptr(mean_rstd),
                    ptr(ss), ptr(stats), i32(srows), i32(sbn), ptr(res), i64(k), i32(1), ptr(z2),
                    i64(k), ptr(bits), native.stream_ptr())
        torch.cuda.synchronize()
        pre = (yf - yf.mean(0)) * torch.rsqrt(yf.var(0, unbiased=False) + 1e-5) * gamma + beta \
            + res.float().reshape(-1, k)
        assert torch.allclose(z2.float().reshape(-1, k), torch.relu(pre), atol=2e-2, rtol=2e-2)
        assert torch.allclose(mm2, mm) and torch.allclose(mv2, mv)
        shifts = torch.arange(8, device="cuda", dtype=torch.uint8)
        got_mask = ((bits.reshape(rows, k // 8, 1) >> shifts) & 1).reshape(rows, k).bool()
        assert (got_mask != (z2.reshape(-1, k) > 0)).float().mean() < 1e-4


def test_per_lane_store_epilogue_matches_the_references():
    """The TMA-store epilogue is the default (fp16 outputs of GEMM_TN / FPROP / unit-stride DGRAD
    leave through cp.async.bulk.tensor stores) and is what the other tests of this file run;
    here every GEMM / convolution case runs again with KFB_TMA_STORE=0, the per-lane 16-byte
    store epilogue that remains for accumulating / scattering / fp32 outputs."""
    import subprocess
    env = dict(os.environ, KFB_TMA_STORE="0", PYTHONPATH=ROOT)
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "kernel_check.py")],
                         cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)
    tail = res.stdout.strip().splitlines()[-1] if res.stdout.strip() else ""
    assert tail.startswith("SUMMARY"), res.stdout[-2000:] + res.stderr[-2000:]
    n_ok, n_all = tail.split()[1].split("/")
    assert n_ok == n_all, res.stdout[-4000:]
