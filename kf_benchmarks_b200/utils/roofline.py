"""Analytic roofline of one training step from the model graph (no GPU needed).

For every node it counts the FLOPs of the forward / dgrad / wgrad GEMMs and the HBM bytes each
kernel class has to move with the pass structure this framework uses (fp16 NHWC activations,
fp32 weight gradients, batch-norm statistics from the conv epilogue, one apply pass forward, a
reduce and an apply pass backward - the single-launch backward kernel is booked under "apply") and turns both into a time at
the measured peaks of MEASURED_PEAKS.json: roof = max(FLOPs / tensor peak, bytes / copy
bandwidth), summed kernel by kernel. ``CLASS_OF`` maps kernel names to those classes. Used by
``tools/roofline_model.py`` (against ncu launch lists) and by ``--tfprof_file`` (against the
profiler's per-kernel times). SURVEY.md 8 asks for the achieved fraction of the roof per row.
"""
from __future__ import annotations

import json
import os
import re
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

_FALLBACK_PEAKS = (6550.7e9, 1399.7e12)   # B200 measured here: copy bytes/s, 16-bit dense FLOP/s


def peaks():
    """(HBM copy bytes/s, sustained dense 16-bit FLOP/s) from MEASURED_PEAKS.json."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]) * 1e9, float(d["bf16_tflops_sustained"]) * 1e12
    except (OSError, KeyError, ValueError):
        return _FALLBACK_PEAKS


def numel(shape):
    n = 1
    for s in shape:
        n *= s
    return n


class _Costs(defaultdict):
    """class -> [flops, bytes, roof seconds]; the roof is summed kernel by kernel (a class that
    mixes tensor-bound and HBM-bound layers is bounded by the sum of the per-layer maxima)."""

    def __init__(self, bw, fl):
        super().__init__(lambda: [0.0, 0.0, 0.0])
        self.bw, self.fl = bw, fl
        self.layers = defaultdict(lambda: [0.0, 0.0, 0.0])   # "fwd/<node>" | "bwd/<node>"

    def add(self, cls, flops, nbytes, layer=None):
        roof = max(flops / self.fl, nbytes / self.bw)
        for e in ([self[cls]] + ([self.layers[layer]] if layer else [])):
            e[0] += flops
            e[1] += nbytes
            e[2] += roof


def model_costs(graph, bw, fl):
    """Costs of one training step (forward + backward) per kernel class."""
    c = _Costs(bw, fl)
    first = graph.nodes[0]
    for n in graph.nodes:
        x = numel(n.inputs[0].shape)
        y = numel(n.output.shape)
        if n.op == "conv":
            k = n.attrs["kh"] * n.attrs["kw"] * n.inputs[0].shape[-1]
            w = k * n.output.shape[-1]
            macs = (y // n.output.shape[-1]) * n.output.shape[-1] * k
            c.add("conv fprop (W1)", 2.0 * macs, 2 * x + 2 * w + 2 * y, "fwd/" + n.name)
            if n is not first:                       # no input gradient for the image
                c.add("conv dgrad (W2)", 2.0 * macs, 2 * y + 2 * w + 2 * x, "bwd/" + n.name)
            c.add("conv wgrad (W3)", 2.0 * macs, 2 * x + 2 * y + 4 * w, "bwd/" + n.name)
        elif n.op == "bn":
            res = len(n.inputs) > 1
            # forward: the statistics come from the conv epilogue (no pass over x); the apply
            # phase reads x (+ residual) and writes y (+ 1 bit per element of packed ReLU mask)
            c.add("BN fwd apply+ReLU+add (W5)", 0, 2 * x * (3 if res else 2) + (x / 8 if res else 0),
                  "fwd/" + n.name)
            # backward: the reduce phase reads dy, x (+ mask bits); the apply phase reads them
            # again and writes dx; the shortcut gradient of an identity block is added by the
            # block's first convolution instead of being written here
            c.add("BN bwd reduce (W5)", 0, 2 * x * 2 + (x / 8 if res else 0), "bwd/" + n.name)
            c.add("BN bwd apply (W5)", 0, 2 * x * 3 + (x / 8 if res else 0), "bwd/" + n.name)
        elif n.op == "pool":
            c.add("pool fwd+bwd (W6)", 0, 2 * (x + y), "fwd/" + n.name)
            c.add("pool fwd+bwd (W6)", 0, 2 * (y + x), "bwd/" + n.name)
        elif n.op == "spatial_mean":
            c.add("pool fwd+bwd (W6)", 0, 2 * (x + y), "fwd/" + n.name)
            c.add("pool fwd+bwd (W6)", 0, 2 * (x + y), "bwd/" + n.name)
        elif n.op == "affine":
            k, m = n.inputs[0].shape[-1], n.output.shape[-1]
            rows = x // k
            for cls, ph in (("conv fprop (W1)", "fwd/"), ("conv dgrad (W2)", "bwd/"),
                            ("conv wgrad (W3)", "bwd/")):
                c.add(cls, 2.0 * rows * k * m, 2 * x + 2 * k * m + 4 * y, ph + n.name)
    ntrain = sum(p.numel for p in graph.params.values() if p.trainable)
    # fused optimizer: read g, w, m (fp32), write w, m (fp32) and the fp16 shadow
    c.add("optimizer / all-reduce epilogue (W10/W11)", 0, ntrain * (4 * 5 + 2))
    return c


CLASS_OF = [
    (r"gemm_tcgen05_kernel<\d+, 0,", "conv fprop (W1)"),
    (r"gemm_tcgen05_kernel<\d+, 2,", "conv fprop (W1)"),
    (r"im2col_small_c", "conv fprop (W1)"),
    (r"gemm_tcgen05_kernel<\d+, 3,", "conv dgrad (W2)"),
    (r"gemm_tcgen05_kernel<\d+, 5,", "conv dgrad (W2)"),
    (r"gemm_tcgen05_kernel<\d+, [14],", "conv wgrad (W3)"),
    (r"splitk_reduce", "conv wgrad (W3)"),
    (r"channel_reduce_kernel<0>|bn_finalize", "BN fwd apply+ReLU+add (W5)"),
    (r"bn_apply_kernel|bn_fwd_fused", "BN fwd apply+ReLU+add (W5)"),
    (r"channel_reduce_kernel<1>|bn_bwd_finalize", "BN bwd reduce (W5)"),
    (r"bn_bwd_apply|bn_bwd_fused", "BN bwd apply (W5)"),
    (r"pool_|spatial_mean", "pool fwd+bwd (W6)"),
    (r"optimizer_step|allreduce_update", "optimizer / all-reduce epilogue (W10/W11)"),
]


def class_of(kernel_name: str):
    """Roofline class of a kernel name (None for kernels outside the table)."""
    for pat, cls in CLASS_OF:
        if re.search(pat, kernel_name):
            return cls
    return None
