#!/usr/bin/env python
"""Per-node device time of one training step against the node's analytic roof.

    python tools/node_timing.py --model resnet50 --batch_size 256 [--top 40] [--out file.md]

Every node's forward and backward is bracketed by CUDA events on the launching stream (no
synchronisation inside the step, so the kernels run back to back as in training); the table
lists, per node and pass, the measured time (median over --steps steps), the roof of
``utils/roofline.py`` for that node (max of FLOPs / tensor peak and bytes / copy bandwidth at
the measured peaks) and the gap in microseconds, largest gap first. This is the table the
optimisation work is planned from (reference counterpart: --tfprof_file per-op times,
T/benchmark_cnn.py:1216-1260).
"""
import argparse
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50")
    ap.add_argument("--batch_size", type=int, default=256)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    from kf_benchmarks_b200 import benchmark, params as params_lib
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.utils import roofline
    # node by node: a graph replay would hide the calls the events bracket
    p = params_lib.make_params(model=a.model, batch_size=a.batch_size, use_fp16=True,
                               variable_update="replicated", optimizer="momentum",
                               data_format="NHWC", cuda_graph=False)
    bench = benchmark.BenchmarkCNN(benchmark.setup(p))
    bench._build(phase_train=True)
    ex = bench.executor
    for _ in range(3):
        bench._step()
    torch.cuda.synchronize()
    per = defaultdict(list)
    order = []
    step_ms = []
    for _ in range(a.steps):
        ex.node_events = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        bench._step()
        e1.record()
        torch.cuda.synchronize()
        step_ms.append(e0.elapsed_time(e1))
        for tag, node, s0, s1 in ex.node_events:
            key = tag + "/" + node.name
            if key not in per:
                order.append((key, node))
            per[key].append(s0.elapsed_time(s1) * 1e3)
        ex.node_events = None
    # second view: device time per C-ABI entry point inside each node (one more step)
    ex.node_events = []
    native.call_events = []
    bench._step()
    torch.cuda.synchronize()
    calls = list(native.call_events)
    native.call_events = None
    ex.node_events = None
    call_ms = [(n, e0.elapsed_time(e1) * 1e3) for n, _, e0, e1 in calls]
    per_call_total = defaultdict(float)
    for n, us in call_ms:
        per_call_total[n] += us
    detail = defaultdict(list)
    for (n, tag, _, _), (_, us) in zip(calls, call_ms):
        detail[tag].append("%s %.0f" % (n, us))
    bw, fl = roofline.peaks()
    costs = roofline.model_costs(ex.graph, bw, fl)
    rows = []
    tot_meas = tot_roof = 0.0
    by_class = defaultdict(lambda: [0.0, 0.0])
    for key, node in order:
        v = sorted(per[key])
        us = v[len(v) // 2]
        layer = costs.layers.get(key)
        roof = layer[2] * 1e6 if layer else 0.0
        shape = "x".join(str(d) for d in node.output.shape)
        extra = ""
        if node.op == "conv":
            extra = "k%dx%d s%d cin%d" % (node.attrs["kh"], node.attrs["kw"], node.attrs["sh"],
                                           node.inputs[0].shape[-1])
        rows.append((us - roof, key, node.op, shape, extra, us, roof))
        tot_meas += us
        tot_roof += roof
        c = by_class[key.split("/")[0] + " " + node.op]
        c[0] += us
        c[1] += roof
    rows.sort(reverse=True)
    step_ms.sort()
    lines = ["# Per-node device time vs analytic roof: %s, batch %d" % (a.model, a.batch_size), "",
             "Step (events around the whole step, median of %d): %.2f ms; sum over nodes %.2f ms; "
             "sum of node roofs %.2f ms. Peaks: %.0f GB/s copy, %.0f TFLOP/s." %
             (a.steps, step_ms[len(step_ms) // 2], tot_meas / 1e3, tot_roof / 1e3, bw / 1e9,
              fl / 1e12), "",
             "| pass + op | measured ms | roof ms | achieved |", "|---|---|---|---|"]
    for k in sorted(by_class, key=lambda k: -by_class[k][0]):
        m, r = by_class[k]
        lines.append("| %s | %.3f | %.3f | %s |" % (k, m / 1e3, r / 1e3,
                                                   ("%.0f %%" % (100 * r / m)) if r else "-"))
    lines += ["", "| node | op | output | geometry | measured us | roof us | gap us | calls (us) |",
              "|---|---|---|---|---|---|---|---|"]
    for gap, key, op, shape, extra, us, roof in rows[:a.top]:
        lines.append("| %s | %s | %s | %s | %.1f | %.1f | %.1f | %s |" %
                     (key, op, shape, extra, us, roof, gap, " + ".join(detail.get(key, []))))
    lines += ["", "| C-ABI entry point | calls | device ms (events around each call) |",
              "|---|---|---|"]
    cnt = defaultdict(int)
    for n, _ in call_ms:
        cnt[n] += 1
    for n in sorted(per_call_total, key=lambda k: -per_call_total[k]):
        lines.append("| %s | %d | %.3f |" % (n, cnt[n], per_call_total[n] / 1e3))
    text = "\n".join(lines) + "\n"
    if a.out:
        with open(a.out, "w") as f:
            f.write(text)
    print(text)


if __name__ == "__main__":
    main()
