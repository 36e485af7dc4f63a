"""GPU tests of the second half of round 2: the space-to-depth stem kernels against their PyTorch
restatements (ops/stem.py), in-place concat inputs against the copying form, and the shifted-view
3x3 convolutions against the nine-box form (same kernels, KFB_SV3 switch)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _stem_plan(n=2, h=40, w=44, c=3, k=16, r=7, s=7, pad=(3, 3)):
    from kf_benchmarks_b200.ops import stem
    from kf_benchmarks_b200.ops.native import ConvDesc
    d = ConvDesc()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s = n, h, w, c, k, r, s
    d.pad_h_lo, d.pad_h_hi = pad
    d.pad_w_lo, d.pad_w_hi = pad
    d.stride_h = d.stride_w = 2
    d.p = (h + pad[0] + pad[1] - r) // 2 + 1
    d.q = (w + pad[0] + pad[1] - s) // 2 + 1
    kpad = (r * s * c + 7) // 8 * 8
    return stem.plan(d, kpad), kpad


@pytest.mark.parametrize("pad", [(3, 3), (2, 3), (0, 0)])
def test_stem_kernels_equal_their_restatements(pad):
    from kf_benchmarks_b200.ops import stem
    pl, kpad = _stem_plan(pad=pad)
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(pl.n, pl.h, pl.w, pl.c, device="cuda", generator=g).half()
    out = torch.full(pl.s2d_shape(), 7.0, device="cuda", dtype=torch.float16)
    stem.regroup(pl, x, out)
    assert torch.equal(out, stem.regroup_ref(pl, x))
    wf = torch.zeros(pl.k, kpad, device="cuda", dtype=torch.float16)
    wf[:, :pl.r * pl.s * pl.c] = torch.randn(pl.k, pl.r * pl.s * pl.c, device="cuda",
                                             generator=g).half()
    w4 = torch.full((pl.k, pl.ta, 64), 7.0, device="cuda", dtype=torch.float16)
    stem.pack_weights(pl, wf, w4)
    assert torch.equal(w4.reshape(pl.k, pl.ta, 4, 16), stem.pack_weights_ref(pl, wf))
    dw4 = torch.randn(pl.k, pl.ta, 64, device="cuda", generator=g)
    dw = torch.full((pl.k, kpad), 3.0, device="cuda")
    stem.unpack_grad(pl, dw4, dw, accumulate=False, scale=0.5)
    assert torch.allclose(dw, 0.5 * stem.unpack_grad_ref(pl, dw4))
    stem.unpack_grad(pl, dw4, dw, accumulate=True, scale=1.0)
    assert torch.allclose(dw, 1.5 * stem.unpack_grad_ref(pl, dw4), atol=1e-6)


def test_stem_forward_and_weight_gradient_match_conv2d_and_the_expanded_form():
    import torch.nn.functional as F
    from kf_benchmarks_b200.ops import stem
    pl, kpad = _stem_plan(n=3, h=64, w=64, k=64)
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(pl.n, pl.h, pl.w, pl.c, device="cuda", generator=g).half()
    wf = torch.zeros(pl.k, kpad, device="cuda", dtype=torch.float16)
    wf[:, :147] = (torch.randn(pl.k, 147, device="cuda", generator=g) * 147 ** -0.5).half()
    y, s2d = stem.forward(pl, x, wf)
    want = stem.forward_ref(pl, x, wf)
    assert float((y.float() - want).abs().max()) < 2e-2 * float(want.abs().max())
    dy = torch.randn(pl.n, pl.p, pl.q, pl.k, device="cuda", generator=g).half()
    dw = stem.weight_grad(pl, s2d, dy)[:, :147]
    w_oihw = wf[:, :147].float().reshape(pl.k, 7, 7, 3).permute(0, 3, 1, 2).requires_grad_(True)
    xp = F.pad(x.float().permute(0, 3, 1, 2), (3, 3, 3, 3))
    F.conv2d(xp, w_oihw, stride=2).backward(dy.float().permute(0, 3, 1, 2))
    ref = w_oihw.grad.permute(0, 2, 3, 1).reshape(pl.k, 147)
    assert float((dw - ref).abs().max()) < 1e-2 * float(ref.abs().max())


def _run(code, **env):
    e = dict(os.environ)
    e.update({k: str(v) for k, v in env.items()})
    res = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=e, capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    return res.stdout


_STEP = """
import sys, torch
sys.path.insert(0, 'tools')
from kf_benchmarks_b200 import datasets
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.nn.executor_cuda import CudaExecutor
from kf_benchmarks_b200.nn.legalize import legalize
from kf_benchmarks_b200.nn.params import ParameterStore
ds = datasets.create_dataset(None, 'imagenet')
model = model_config.get_model_config('%s', ds, None)
g = legalize(model.build_graph(ds.num_classes, batch_size=%d))
for node in g.nodes:
    if node.op == 'dropout':
        node.attrs['keep_prob'] = 1.0
st = ParameterStore(g, torch.device('cuda'))
st.initialize(1234)
ex = CudaExecutor(g, st, ds.num_classes)
gen = torch.Generator().manual_seed(0)
h = model.image_size
ex.set_inputs((torch.randn(%d, h, h, 3, generator=gen) * 0.5).half().cuda(),
              torch.randint(0, 1000, (%d,), generator=gen).int().cuda())
ex.forward(True)
ex.loss_backward(loss_scale=1.0)
torch.cuda.synchronize()
print('SLICES', len(ex._slice_of))
print('LOGITS', float(ex.logits().double().abs().sum()))
print('GRADS', float(st.grad.double().abs().sum()), float(st.grad.double().sum()))
"""


def test_in_place_concat_inputs_change_nothing_but_the_copies():
    code = _STEP % ("inception3", 4, 4, 4)
    a = _run(code, KFB_CONCAT_INPLACE=0, KFB_CUDA_GRAPH=0)
    b = _run(code, KFB_CONCAT_INPLACE=1, KFB_CUDA_GRAPH=0)
    la = dict(line.split(" ", 1) for line in a.strip().splitlines() if " " in line)
    lb = dict(line.split(" ", 1) for line in b.strip().splitlines() if " " in line)
    assert int(la["SLICES"]) == 0 and int(lb["SLICES"]) > 20
    # the same kernels compute the same values; only where they are stored differs
    assert la["LOGITS"] == lb["LOGITS"] and la["GRADS"] == lb["GRADS"]


def test_shifted_view_convolutions_match_the_nine_box_form():
    code = _STEP % ("resnet50", 4, 4, 4)
    a = _run(code, KFB_SV3=0, KFB_CUDA_GRAPH=0)
    b = _run(code, KFB_SV3=1, KFB_CUDA_GRAPH=0)
    la = dict(line.split(" ", 1) for line in a.strip().splitlines() if " " in line)
    lb = dict(line.split(" ", 1) for line in b.strip().splitlines() if " " in line)
    # fp32 accumulation order inside a K block differs (taps are issued per filter row), so the
    # sums agree to rounding, not bit for bit
    for key in ("LOGITS",):
        x, y = float(la[key]), float(lb[key])
        assert abs(x - y) <= 5e-2 * abs(x), (key, x, y)
