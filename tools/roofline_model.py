#!/usr/bin/env python
"""Analytic roofline of one training step, per kernel class, against the measured launch list.

For every node of the model graph (built on the CPU, no GPU needed) it counts the FLOPs of the
forward / dgrad / wgrad GEMMs and the HBM bytes each kernel class *has* to move with the pass
structure this framework uses (fp16 NHWC activations, fp32 weight gradients, two-pass batch norm
forward and backward), turns both into a time at the measured peaks of MEASURED_PEAKS.json
(roofline = max(FLOPs / tensor peak, bytes / copy bandwidth)) and sets that against the device
time ncu measured for the same class (profiles/resnet50_bs256_launches_*.txt, written by
tools/summarize_launches.py). SURVEY.md 8 asks for the achieved fraction of the roof per row.

usage: python tools/roofline_model.py [--model resnet50] [--batch_size 256]
                                      [--launches profiles/resnet50_bs256_launches_r1e.txt]
"""
import argparse
import os
import re
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


from kf_benchmarks_b200.utils.roofline import CLASS_OF, model_costs, peaks  # noqa: E402


def measured(path):
    out = defaultdict(float)
    other = 0.0
    with open(path) as f:
        for ln in f:
            m = re.match(r"\s+([\d.]+)\s+[\d.]+%\s+\d+\s+[\d.]+\s+[\d.]+\s+[\d.]+%\s+(.*)$", ln)
            if not m:
                continue
            ms, name = float(m.group(1)), m.group(2)
            for pat, cls in CLASS_OF:
                if re.search(pat, name):
                    out[cls] += ms
                    break
            else:
                other += ms
    return out, other


def by_layer_report(costs, path, top=25):
    """Per-layer table from a launch list whose kernels were renamed after their NVTX range
    (tools/capture_profiles.sh step 2b): measured time of "fwd/<node>" / "bwd/<node>" against
    the roof of the same node and phase, largest absolute gap first."""
    meas = {}
    with open(path) as f:
        for ln in f:
            m = re.match(r"\s+([\d.]+)\s+[\d.]+%\s+\d+\s+[\d.]+\s+[\d.]+\s+[\d.]+%\s+((?:fwd|bwd)/\S+)",
                         ln)
            if m:
                meas[m.group(2).rstrip("/")] = meas.get(m.group(2).rstrip("/"), 0.0) + \
                    float(m.group(1))
    rows = []
    for layer, (flops, nbytes, roof_s) in costs.layers.items():
        t = meas.get(layer.rstrip("/"))
        if t is not None:
            rows.append((t - roof_s * 1e3, layer, flops, nbytes, roof_s * 1e3, t))
    rows.sort(reverse=True)
    print()
    print("## Layers farthest from their roof (%d of %d matched layers)" % (min(top, len(rows)),
                                                                            len(rows)))
    print()
    print("| layer / phase | GFLOP | GB | roof ms | measured ms | achieved | gap ms |")
    print("|---|---|---|---|---|---|---|")
    for gap, layer, flops, nbytes, roof, t in rows[:top]:
        print("| %s | %.1f | %.3f | %.3f | %.3f | %.0f %% | %.3f |" %
              (layer, flops / 1e9, nbytes / 1e9, roof, t, 100 * roof / t if t else 0.0, gap))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch_size", type=int, default=256)
    ap.add_argument("--launches",
                    default=os.path.join(ROOT, "profiles", "resnet50_bs256_launches_r1e.txt"))
    ap.add_argument("--by_layer", default=None,
                    help="launch list with NVTX-renamed kernels (per-layer table)")
    ap.add_argument("--list_layers", type=int, default=0,
                    help="also list the N layers / phases with the largest roof (analytic only)")
    ap.add_argument("--step_ms", type=float, default=None,
                    help="free-running step time to set the total against")
    args = ap.parse_args()
    from kf_benchmarks_b200 import datasets, params as params_lib
    from kf_benchmarks_b200.models import model_config
    p = params_lib.make_params(model=args.model, batch_size=args.batch_size, use_fp16=True,
                               data_format="NHWC")
    ds = datasets.create_dataset(None, None)
    model = model_config.get_model_config(args.model, ds, p)
    model.set_batch_size(args.batch_size)
    graph = model.build_graph(ds.num_classes, batch_size=args.batch_size)
    bw, fl = peaks()
    costs = model_costs(graph, bw, fl)
    meas, other = measured(args.launches) if os.path.exists(args.launches) else ({}, 0.0)
    print("# Roofline of one %s training step, batch %d (tools/roofline_model.py)" %
          (args.model, args.batch_size))
    print()
    print("Peaks: HBM copy %.1f GB/s, dense 16-bit GEMM %.1f TFLOP/s sustained "
          "(MEASURED_PEAKS.json). Roof = sum over kernels of max(FLOPs / peak, bytes / bandwidth) "
          "with the bytes this pass structure must move; measured = ncu device time of the class (%s)." %
          (bw / 1e9, fl / 1e12, os.path.relpath(args.launches, ROOT)))
    print()
    print("| class | GFLOP | GB accessed (HBM traffic if nothing stays in L2) | roof ms | bound | measured ms | achieved |")
    print("|---|---|---|---|---|---|---|")
    tot_roof = tot_meas = 0.0
    for cls, (flops, nbytes, roof_s) in costs.items():
        t_f, t_b = flops / fl * 1e3, nbytes / bw * 1e3
        roof = roof_s * 1e3
        m = meas.get(cls)
        tot_roof += roof
        tot_meas += m or 0.0
        print("| %s | %.0f | %.2f | %.3f | %s | %s | %s |" %
              (cls, flops / 1e9, nbytes / 1e9, roof, "tensor" if t_f > t_b else ("HBM" if flops == 0 or roof <= t_b * 1.001
                                                           else "mixed"),
               "%.3f" % m if m else "-", "%.0f %%" % (100 * roof / m) if m else "-"))
    if tot_meas + other > 0:
        print("| **sum** | | | **%.2f** | | **%.2f** (+ %.2f other) | **%.0f %%** |" %
              (tot_roof, tot_meas, other, 100 * tot_roof / (tot_meas + other)))
    else:
        print("| **sum** | | | **%.2f** | | - | - |" % tot_roof)
        print()
        print("No launch list given: at the roof this step would take %.2f ms = %.0f examples/sec." %
              (tot_roof, args.batch_size / (tot_roof * 1e-3)))
    if args.by_layer and os.path.exists(args.by_layer):
        by_layer_report(costs, args.by_layer)
    if args.list_layers > 0:
        print()
        print("## Largest layers by roof (analytic)")
        print()
        print("| layer / phase | GFLOP | GB | roof us | bound |")
        print("|---|---|---|---|---|")
        rows = sorted(costs.layers.items(), key=lambda kv: -kv[1][2])[:args.list_layers]
        for layer, (flops, nbytes, roof_s) in rows:
            print("| %s | %.1f | %.3f | %.1f | %s |" %
                  (layer, flops / 1e9, nbytes / 1e9, roof_s * 1e6,
                   "tensor" if flops / fl > nbytes / bw else "HBM"))
    if args.step_ms:
        print()
        print("Free-running step: %.2f ms -> %.0f %% of the summed roof." %
              (args.step_ms, 100 * tot_roof / args.step_ms))


if __name__ == "__main__":
    main()
