#!/usr/bin/env python
"""Compare the CUDA executor against the PyTorch fp32 reference executor on one model.

    python tools/executor_check.py --model resnet50 --batch 4
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from kf_benchmarks_b200 import datasets  # noqa: E402
from kf_benchmarks_b200.models import model_config  # noqa: E402
from kf_benchmarks_b200.nn.executor_cuda import CudaExecutor  # noqa: E402
from kf_benchmarks_b200.nn.executor_ref import RefExecutor  # noqa: E402
from kf_benchmarks_b200.nn.params import ParameterStore  # noqa: E402


LOSS_SCALE = 1024.0


def compare(model_name, data_name, batch, image_size=None, seed=0, verbose=False, tweak=None):
    ds = datasets.create_dataset(None, data_name)
    model = model_config.get_model_config(model_name, ds, None)
    if image_size:
        model.image_size = image_size
    if tweak is not None:
        tweak(model)
    nclass = ds.num_classes
    from kf_benchmarks_b200.nn.legalize import legalize
    g = legalize(model.build_graph(nclass, batch_size=batch))
    for node in g.nodes:  # the two executors draw different random streams: compare without them
        if node.op == "dropout":
            node.attrs["keep_prob"] = 1.0
        elif node.op == "drop_path":
            node.attrs["keep_prob_fn"] = lambda step: 1.0
    dev = torch.device("cuda")
    st_gpu = ParameterStore(g, dev)
    st_gpu.initialize(1234)
    st_cpu = ParameterStore(g, "cuda", fp16_shadow=False)  # reference also on GPU (fp32 torch)
    st_cpu.w32.copy_(st_gpu.w16.float())  # identical (fp16-rounded) weights
    st_cpu.state.copy_(st_gpu.state)
    st_gpu.w32.copy_(st_cpu.w32)
    gen = torch.Generator().manual_seed(seed)
    if hasattr(model, "make_input_source"):
        model.set_batch_size(batch)
        inputs = list(model.make_input_source(batch, "cpu", seed).next_batch())
    else:
        h = model.image_size
        images = (torch.randn(batch, h, h, 3, generator=gen) * 0.5).half()
        labels = torch.randint(0, nclass - 1, (batch,), generator=gen)
        inputs = [images, labels.int()]
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ex = CudaExecutor(g, st_gpu, nclass)
    ex.set_inputs(*[t.cuda() for t in inputs])
    ex.forward(True)
    ex.loss_backward(loss_scale=LOSS_SCALE)
    torch.cuda.synchronize()
    ref = RefExecutor(g, st_cpu, nclass, emulate_fp16=True)
    ref.set_inputs(*[t.float() if t.dtype == torch.float16 else t for t in inputs])
    # dropout: both must draw the same mask -> disable by comparing only deterministic models
    ref.forward(True)
    if verbose:
        for v in ref._vals.values():
            if v.requires_grad:
                v.retain_grad()
    ref.loss_backward(LOSS_SCALE)
    if verbose:
        cs = torch.nn.functional.cosine_similarity
        for node in reversed(g.nodes):
            t = node.output
            rv = ref._vals[t.id]
            fa = ex.value(t.id).float()
            line = f"{node.op:12s} {node.name[-40:]:40s} act_rel {float((fa - rv).abs().max() / (rv.abs().max() + 1e-6)):.4f}"
            if t.id in ex.gbuf and rv.grad is not None:
                ga, gr_ = ex.grad_value(t.id).float().flatten(), rv.grad.flatten()
                line += f"  grad_cos {float(cs(ga, gr_, dim=0)):.4f} gmax {float(gr_.abs().max()):.3e}"
            print(line)
    torch.cuda.synchronize()
    out = {"model": model_name, "batch": batch}
    if g.logits is not None:
        lg, lr = ex.logits().float(), ref.logits().float()
        out["logits_rel"] = float((lg - lr).abs().max() / (lr.abs().max() + 1e-6))
    else:
        out["logits_rel"] = None
    out["loss_gpu"], out["loss_ref"] = ex.read_loss()[0], ref.read_loss()[0]
    gg, gr = st_gpu.grad, st_cpu.grad
    worst = (None, 0.0)
    for name, slot in st_gpu.slots.items():
        if slot.arena != "w":
            continue
        a = st_gpu.view(name, "grad").float()
        b = st_cpu.view(name, "grad").float()
        rel = float((a - b).abs().max() / (b.abs().max() + 1e-6))
        if verbose:
            print(f"{name:60s} rel {rel:.4f} refmax {float(b.abs().max()):.3e}")
        if rel > worst[1]:
            worst = (name, rel)
    out["grad_rel_global"] = float((gg - gr).abs().max() / (gr.abs().max() + 1e-6))
    cos = torch.nn.functional.cosine_similarity(gg.flatten(), gr.flatten(), dim=0)
    out["grad_cosine"] = float(cos)
    out["worst_param"], out["worst_rel"] = worst
    sm, sr = st_gpu.state, st_cpu.state
    out["moving_stats_rel"] = float((sm - sr).abs().max() / (sr.abs().max() + 1e-6))
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--data_name", default="imagenet")
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--image_size", type=int, default=None)
    ap.add_argument("-v", action="store_true")
    a = ap.parse_args()
    print("RESULT " + json.dumps(compare(a.model, a.data_name, a.batch, a.image_size, verbose=a.v)))
