#!/usr/bin/env python
"""On-GPU numerical sweep of the tcgen05 GEMM / conv kernels against fp32 oracles.

Each case runs in its own subprocess (a device fault in one case must not take
the others down) under a timeout. Results go to stdout and to
``gpurun_out/kernel_check.json``. Usage:

    python tools/kernel_check.py            # run every case
    python tools/kernel_check.py --case fprop_3x3_c64   # run one case in-process
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# name -> (kind, kwargs)
CASES = {}


def case(name, kind, **kw):
    CASES[name] = (kind, kw)


case("gemm_tn_128x64x64", "gemm_tn", m=128, n=64, k=64)
case("gemm_tn_256x128x128", "gemm_tn", m=256, n=128, k=128)
case("gemm_tn_512x256x512", "gemm_tn", m=512, n=256, k=512)
case("gemm_tn_tail_1000x1001x520_bias_relu_f32", "gemm_tn", m=1000, n=1001, k=520, bias=True,
     relu=True, f32=True)
case("gemm_tn_4096x512x2048", "gemm_tn", m=4096, n=512, k=2048)
case("gemm_nt_128x64x64", "gemm_nt", m=128, n=64, k=64)
case("gemm_nt_256x256x4096", "gemm_nt", m=256, n=256, k=4096)
case("gemm_nt_tail_1001x2048x250", "gemm_nt", m=1001 // 8 * 8, n=2048, k=250)
for mode in ("fprop", "dgrad", "wgrad"):
    case(f"{mode}_1x1_c64_k256", mode, n=4, h=14, w=14, c=64, k=256, r=1, s=1, stride=1, pad="SAME")
    case(f"{mode}_3x3_c64_k64", mode, n=2, h=56, w=56, c=64, k=64, r=3, s=3, stride=1, pad="SAME")
    case(f"{mode}_3x3_c128_k128_tail", mode, n=3, h=13, w=13, c=128, k=128, r=3, s=3, stride=1,
         pad="SAME")
    case(f"{mode}_1x1_s2_c256_k512", mode, n=4, h=28, w=28, c=256, k=512, r=1, s=1, stride=2,
         pad="VALID")
    case(f"{mode}_3x3_valid_c64_k192", mode, n=2, h=17, w=17, c=64, k=192, r=3, s=3, stride=1,
         pad="VALID")
    case(f"{mode}_1x7_c128_k128", mode, n=2, h=17, w=17, c=128, k=128, r=1, s=7, stride=1,
         pad="SAME")
    case(f"{mode}_3x3_c32_k32", mode, n=2, h=16, w=16, c=32, k=32, r=3, s=3, stride=1, pad="SAME")
case("fprop_3x3_s2_c64_k64", "fprop", n=2, h=56, w=56, c=64, k=64, r=3, s=3, stride=2,
     pad="SAME_RESNET")
case("wgrad_3x3_s2_c64_k64", "wgrad", n=2, h=56, w=56, c=64, k=64, r=3, s=3, stride=2,
     pad="SAME_RESNET")
case("dgrad_3x3_s2_c64_k64", "dgrad", n=2, h=56, w=56, c=64, k=64, r=3, s=3, stride=2,
     pad="SAME_RESNET")
case("dgrad_3x3_s2_valid_c64_k96_odd", "dgrad", n=2, h=35, w=35, c=64, k=96, r=3, s=3, stride=2,
     pad="VALID")
case("dgrad_3x3_s2_same_c32_k32", "dgrad", n=3, h=16, w=16, c=32, k=32, r=3, s=3, stride=2,
     pad="SAME")
case("dgrad_5x5_s2_same_c64_k64", "dgrad", n=2, h=15, w=15, c=64, k=64, r=5, s=5, stride=2,
     pad="SAME")
case("fprop_7x7_s2_c8_k64", "fprop", n=2, h=64, w=64, c=8, k=64, r=7, s=7, stride=2,
     pad="SAME_RESNET")
case("fprop_big_3x3_c256", "fprop", n=32, h=14, w=14, c=256, k=256, r=3, s=3, stride=1, pad="SAME")
case("wgrad_big_3x3_c256", "wgrad", n=32, h=14, w=14, c=256, k=256, r=3, s=3, stride=1, pad="SAME")
case("dgrad_big_3x3_c256", "dgrad", n=32, h=14, w=14, c=256, k=256, r=3, s=3, stride=1, pad="SAME")
# dgrad accumulating into an existing gradient: TMA-prefetched old tiles (unit stride) and the
# per-lane read-modify-write together with the row scatter (strided)
case("dgrad_acc_1x1_c256_k64", "dgrad", n=4, h=14, w=14, c=256, k=64, r=1, s=1, stride=1,
     pad="SAME", acc=True)
case("dgrad_acc_1x1_c512_k128_tail", "dgrad", n=3, h=13, w=13, c=512, k=128, r=1, s=1, stride=1,
     pad="SAME", acc=True)
case("dgrad_acc_3x3_c128_k128", "dgrad", n=2, h=28, w=28, c=128, k=128, r=3, s=3, stride=1,
     pad="SAME", acc=True)
case("dgrad_acc_1x1_s2_c256_k512_scatter", "dgrad", n=4, h=28, w=28, c=256, k=512, r=1, s=1,
     stride=2, pad="VALID", acc=True)
case("dgrad_acc_3x3_s2_c64_k64_scatter", "dgrad", n=2, h=56, w=56, c=64, k=64, r=3, s=3,
     stride=2, pad="SAME_RESNET", acc=True)
# dgrad + shortcut gradient * packed ReLU mask in one kernel (kfb_conv_dgrad_residual)
case("dgrad_res_1x1_c256_k64", "dgrad", n=4, h=14, w=14, c=256, k=64, r=1, s=1, stride=1,
     pad="SAME", res=True)
case("dgrad_res_1x1_c512_k128_tail", "dgrad", n=3, h=13, w=13, c=512, k=128, r=1, s=1, stride=1,
     pad="SAME", res=True)
case("dgrad_res_3x3_c64_k64", "dgrad", n=2, h=28, w=28, c=64, k=64, r=3, s=3, stride=1,
     pad="SAME", res=True)
# space-to-depth stem (ops/stem.py): forward and weight gradient against conv2d
case("stem_s2d_fprop_7x7_s2_c3_k64", "stem_s2d", n=2, h=64, w=64, c=3, k=64, r=7, s=7, stride=2,
     pad="SAME_RESNET")
case("stem_s2d_wgrad_7x7_s2_c3_k64", "stem_s2d", n=2, h=64, w=64, c=3, k=64, r=7, s=7, stride=2,
     pad="SAME_RESNET", wgrad=True)
case("stem_s2d_fprop_7x7_s2_c3_k32_odd_same", "stem_s2d", n=3, h=45, w=51, c=3, k=32, r=7, s=7,
     stride=2, pad="SAME")
case("stem_s2d_wgrad_7x7_s2_c3_k32_odd_same", "stem_s2d", n=3, h=45, w=51, c=3, k=32, r=7, s=7,
     stride=2, pad="SAME", wgrad=True)
case("stem_s2d_fprop_7x7_s2_c3_k64_same_even_pad", "stem_s2d", n=2, h=32, w=32, c=3, k=64, r=7,
     s=7, stride=2, pad="SAME")
case("stem_s2d_wgrad_7x7_s2_c3_k64_same_even_pad", "stem_s2d", n=2, h=32, w=32, c=3, k=64, r=7,
     s=7, stride=2, pad="SAME", wgrad=True)
case("stem_s2d_wgrad_5x7_s2_c4_k64_valid", "stem_s2d", n=2, h=40, w=40, c=4, k=64, r=5, s=7,
     stride=2, pad="VALID", wgrad=True)
# headline shapes (ResNet-50, batch 256; `--headline` runs them, they need ~6 GB)
HEADLINE = []


def hcase(name, kind, **kw):
    CASES[name] = (kind, kw)
    HEADLINE.append(name)


for mode in ("fprop", "dgrad", "wgrad"):
    hcase(f"{mode}_head_1x1_56_c64_k256", mode, n=256, h=56, w=56, c=64, k=256, r=1, s=1,
          stride=1, pad="SAME")
    hcase(f"{mode}_head_3x3_56_c64_k64", mode, n=256, h=56, w=56, c=64, k=64, r=3, s=3, stride=1,
          pad="SAME")
    hcase(f"{mode}_head_1x1_s2_56_c256_k512", mode, n=256, h=56, w=56, c=256, k=512, r=1, s=1,
          stride=2, pad="VALID")
hcase("dgrad_head_acc_1x1_56_c256_k64", "dgrad", n=256, h=56, w=56, c=256, k=64, r=1, s=1,
      stride=1, pad="SAME", acc=True)
hcase("dgrad_head_acc_1x1_s2_56_c256_k512_scatter", "dgrad", n=256, h=56, w=56, c=256, k=512,
      r=1, s=1, stride=2, pad="VALID", acc=True)
hcase("dgrad_head_res_1x1_56_c256_k64", "dgrad", n=256, h=56, w=56, c=256, k=64, r=1, s=1,
      stride=1, pad="SAME", res=True)
hcase("stem_head_7x7_s2_c3_k64", "stem", n=256, h=224, w=224, c=3, k=64, r=7, s=7, stride=2,
      pad="SAME_RESNET")
hcase("stem_s2d_head_fprop", "stem_s2d", n=256, h=224, w=224, c=3, k=64, r=7, s=7, stride=2,
      pad="SAME_RESNET")
hcase("stem_s2d_head_wgrad", "stem_s2d", n=256, h=224, w=224, c=3, k=64, r=7, s=7, stride=2,
      pad="SAME_RESNET", wgrad=True)


def run_case(name: str) -> dict:
    import torch
    import torch.nn.functional as F
    from kf_benchmarks_b200.ops import gemm_conv as gc

    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    kind, kw = CASES[name]
    g = torch.Generator(device="cuda").manual_seed(1234)
    dev = "cuda"

    def rnd(*shape, scale=1.0):
        return (torch.randn(*shape, device=dev, generator=g) * scale).half()

    if kind == "gemm_tn":
        a, b = rnd(kw["m"], kw["k"]), rnd(kw["n"], kw["k"])
        bias = torch.randn(kw["n"], device=dev, generator=g) if kw.get("bias") else None
        out = gc.gemm_tn(a, b, bias=bias, relu=kw.get("relu", False),
                         out_dtype=torch.float32 if kw.get("f32") else torch.float16)
        ref = a.float() @ b.float().t()
        if bias is not None:
            ref = ref + bias
        if kw.get("relu"):
            ref = ref.relu()
    elif kind == "gemm_nt":
        a, b = rnd(kw["k"], kw["m"]), rnd(kw["k"], kw["n"])
        out = gc.gemm_nt(a, b)
        ref = a.float().t() @ b.float()
    elif kind == "stem_s2d":
        from kf_benchmarks_b200.ops import stem
        n, h, w, c, k, r, s = (kw[x] for x in ("n", "h", "w", "c", "k", "r", "s"))
        st = kw["stride"]
        ph = gc.resolve_padding(kw["pad"], h, r, st)
        pw = gc.resolve_padding(kw["pad"], w, s, st)
        x = rnd(n, h, w, c)
        kpad = (r * s * c + 7) // 8 * 8
        wt = rnd(k, kpad, scale=(r * s * c) ** -0.5)
        wt[:, r * s * c:] = 0
        d = gc.ConvDesc()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s = n, h, w, c, k, r, s
        d.pad_h_lo, d.pad_h_hi, d.pad_w_lo, d.pad_w_hi = ph[0], ph[1], pw[0], pw[1]
        d.stride_h = d.stride_w = st
        d.p, d.q = (h + ph[0] + ph[1] - r) // st + 1, (w + pw[0] + pw[1] - s) // st + 1
        pl = stem.plan(d, kpad)
        y, s2d = stem.forward(pl, x, wt)
        w_oihw = wt[:, :r * s * c].float().reshape(k, r, s, c).permute(0, 3, 1, 2).contiguous()
        w_oihw.requires_grad_(True)
        xp = F.pad(x.float().permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1]))
        y_ref = F.conv2d(xp, w_oihw, stride=st)
        if kw.get("wgrad"):
            dy = rnd(n, d.p, d.q, k)
            y_ref.backward(dy.float().permute(0, 3, 1, 2))
            out = stem.weight_grad(pl, s2d, dy)[:, :r * s * c]
            ref = w_oihw.grad.permute(0, 2, 3, 1).reshape(k, r * s * c)
        else:
            out = y
            ref = y_ref.detach().permute(0, 2, 3, 1)
    elif kind == "stem":
        # the tiny-C stem: explicit im2col (kfb_im2col_small_c) + GEMM, as the executor runs it
        from kf_benchmarks_b200.ops import native
        from kf_benchmarks_b200.ops.native import i32, ptr
        n, h, w, c, k, r, s = (kw[x] for x in ("n", "h", "w", "c", "k", "r", "s"))
        st = kw["stride"]
        ph = gc.resolve_padding(kw["pad"], h, r, st)
        pw = gc.resolve_padding(kw["pad"], w, s, st)
        p_, q_ = (h + ph[0] + ph[1] - r) // st + 1, (w + pw[0] + pw[1] - s) // st + 1
        x = rnd(n, h, w, c)
        kpad = (r * s * c + 7) // 8 * 8
        wt = rnd(k, kpad, scale=(r * s * c) ** -0.5)
        wt[:, r * s * c:] = 0
        col = torch.zeros(n * p_ * q_, kpad, device=dev, dtype=torch.float16)
        native.call("im2col_small_c", ptr(x), ptr(col), i32(n), i32(h), i32(w), i32(c), i32(p_),
                    i32(q_), i32(r), i32(s), i32(st), i32(st), i32(ph[0]), i32(pw[0]), i32(kpad),
                    native.stream_ptr())
        out = gc.gemm_tn(col, wt).reshape(n, p_, q_, k)
        w_oihw = wt[:, :r * s * c].float().reshape(k, r, s, c).permute(0, 3, 1, 2)
        xp = F.pad(x.float().permute(0, 3, 1, 2), (pw[0], pw[1], ph[0], ph[1]))
        ref = F.conv2d(xp, w_oihw, stride=st).permute(0, 2, 3, 1)
    else:
        n, h, w, c, k, r, s = (kw[x] for x in ("n", "h", "w", "c", "k", "r", "s"))
        st = kw["stride"]
        ph = gc.resolve_padding(kw["pad"], h, r, st)
        pw = gc.resolve_padding(kw["pad"], w, s, st)
        x = rnd(n, h, w, c)
        wt = rnd(k, r, s, c, scale=(r * s * c) ** -0.5)
        d = gc.make_desc(x.shape, wt.shape, (st, st), ph, pw)
        x_nchw = x.float().permute(0, 3, 1, 2)
        w_oihw = wt.float().permute(0, 3, 1, 2)
        xp = F.pad(x_nchw, (pw[0], pw[1], ph[0], ph[1])).requires_grad_(kind != "fprop")
        wv = w_oihw.clone().requires_grad_(kind != "fprop")
        y_ref = F.conv2d(xp, wv, stride=st)
        if kind == "fprop":
            out = gc.conv_fprop(x, wt, d)
            ref = y_ref.permute(0, 2, 3, 1)
        else:
            dy = rnd(n, d.p, d.q, k)
            y_ref.backward(dy.float().permute(0, 3, 1, 2))
            if kind == "dgrad":
                gx = xp.grad[:, :, ph[0]:ph[0] + h, pw[0]:pw[0] + w]
                ref = gx.permute(0, 2, 3, 1)
                if kw.get("res"):
                    import ctypes
                    from kf_benchmarks_b200.ops import native
                    from kf_benchmarks_b200.ops.native import i64, ptr
                    src = rnd(n, h, w, c)
                    bits = torch.randint(0, 256, (n * h * w, c // 8), device=dev,
                                         dtype=torch.uint8, generator=g)
                    out = torch.full((n, h, w, c), 7.0, device=dev, dtype=torch.float16)
                    native.call("conv_dgrad_residual", ctypes.byref(d), ptr(dy), ptr(wt),
                                ptr(out), ptr(src), i64(c), ptr(bits), i64(c // 8),
                                native.stream_ptr())
                    shifts = torch.arange(8, device=dev, dtype=torch.uint8)
                    mask = ((bits[:, :, None] >> shifts) & 1).reshape(n, h, w, c).float()
                    ref = ref + src.float() * mask
                elif kw.get("acc"):
                    prev = rnd(n, h, w, c)
                    out = prev.clone()
                    gc.conv_dgrad(dy, wt, d, out=out, accumulate=True)
                    ref = ref + prev.float()
                else:
                    out = gc.conv_dgrad(dy, wt, d)
            else:
                out = gc.conv_wgrad(x, dy, d)
                ref = wv.grad.permute(0, 2, 3, 1)
    torch.cuda.synchronize()
    out = out.float()
    ref = ref.contiguous().float()
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    bad = int(((out - ref).abs() > 2e-2 * scale).sum().item())
    return {"case": name, "max_abs_err": err, "ref_max": scale, "rel": err / scale,
            "mismatches": bad, "numel": out.numel(), "ok": bool(err / scale < 2e-2),
            "finite": bool(torch.isfinite(out).all().item())}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", help="comma-separated cases to run in this process")
    ap.add_argument("--filter", default="")
    ap.add_argument("--headline", action="store_true",
                    help="run the ResNet-50 batch-256 shapes (skipped by default: ~6 GB, slower)")
    ap.add_argument("--timeout", type=int, default=180)
    args = ap.parse_args()
    if args.case:
        for name in args.case.split(","):
            try:
                r = run_case(name)
            except Exception as e:  # a sticky CUDA error ends this batch
                r = {"case": name, "ok": False, "error": repr(e)[-1500:]}
                print("RESULT " + json.dumps(r), flush=True)
                sys.exit(3)
            print("RESULT " + json.dumps(r), flush=True)
        return
    todo = [n for n in CASES if (not args.filter or args.filter in n) and
            ((n in HEADLINE) == bool(args.headline))]
    if os.environ.get("KFB_TMA_STORE") == "0":
        # the fused residual gradient exists only in the TMA-store epilogue
        todo = [n for n in todo if not CASES[n][1].get("res")]
    results = []
    while todo:
        t0 = time.time()
        try:
            res = subprocess.run([sys.executable, __file__, "--case", ",".join(todo)],
                                 capture_output=True, text=True, timeout=args.timeout)
            out, err = res.stdout, res.stderr
        except subprocess.TimeoutExpired as e:
            out = (e.stdout or b"").decode() if isinstance(e.stdout, bytes) else (e.stdout or "")
            err = "timeout"
        got = [json.loads(l[7:]) for l in out.splitlines() if l.startswith("RESULT ")]
        for r in got:
            results.append(r)
            print(json.dumps(r), flush=True)
        done = {r["case"] for r in got}
        rest = [n for n in todo if n not in done]
        if rest and (not got or got[-1].get("ok") or "error" not in got[-1]):
            # the process died without reporting the case it was on
            r = {"case": rest[0], "ok": False, "error": "process died: " + err[-1200:]}
            results.append(r)
            print(json.dumps(r), flush=True)
            rest = rest[1:]
        todo = rest
        print(f"# batch took {time.time() - t0:.1f}s, {len(todo)} cases left", flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "kernel_check.json"), "w") as f:
        json.dump(results, f, indent=1)
    n_ok = sum(1 for r in results if r.get("ok"))
    print(f"SUMMARY {n_ok}/{len(results)} ok")


if __name__ == "__main__":
    main()
