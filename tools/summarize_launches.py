#!/usr/bin/env python
"""Aggregate an ncu launch list (`--metrics gpu__time_duration.sum[,dram__bytes_read.sum,
dram__bytes_write.sum] --csv`) by kernel name: launches, device time, share, DRAM traffic and the
achieved DRAM bandwidth against the measured copy peak of MEASURED_PEAKS.json."""
import csv
import json
import os
import re
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        for k in ("hbm_copy_gbs", "copy_bandwidth_gbs", "hbm_gbs"):
            if k in d:
                return float(d[k])
        for v in d.values():
            if isinstance(v, dict):
                for k, x in v.items():
                    if "gb" in k.lower() and isinstance(x, (int, float)):
                        return float(x)
    except Exception:
        pass
    return 6545.6


def short(name):
    if name.startswith(("fwd/", "bwd/")):   # renamed after its NVTX range (KFB_NVTX=1): a layer
        return name
    name = re.sub(r"^void ", "", name)
    m = re.match(r"([\w:]+(?:<[^(]*>)?)", name)
    return (m.group(1) if m else name).replace("kfb::", "")


def main(path, top=45):
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    per_launch = defaultdict(dict)
    for r in csv.DictReader(lines):
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        metric = r["Metric Name"]
        if metric == "gpu__time_duration.sum":
            val *= {"ns": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}.get(unit, 1.0)
        elif unit in ("Kbyte", "KB"):
            val *= 1e3
        elif unit in ("Mbyte", "MB"):
            val *= 1e6
        elif unit in ("Gbyte", "GB"):
            val *= 1e9
        d = per_launch[r["ID"]]
        d["name"] = short(r["Kernel Name"])
        d["grid"] = r.get("Grid Size", "")
        d[metric] = val
    agg = defaultdict(lambda: [0, 0.0, 0.0])
    total = 0.0
    for d in per_launch.values():
        t = d.get("gpu__time_duration.sum", 0.0)
        b = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        a = agg[d["name"]]
        a[0] += 1
        a[1] += t
        a[2] += b
        total += t
    peak = peak_gbs()
    print("total %.3f ms over %d launches (serialised under ncu, cold caches: compare shares); "
          "DRAM peak used for %%: %.1f GB/s (measured copy bandwidth)" %
          (total / 1e6, len(per_launch), peak))
    print("%9s %6s %6s %10s %9s %6s  %s" % ("time ms", "share", "count", "DRAM MB", "GB/s",
                                             "%peak", "kernel"))
    for name, (n, t, b) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        gbs = b / t if t else 0.0          # bytes per ns == GB/s
        print("%9.3f %5.1f%% %6d %10.1f %9.1f %5.1f%%  %s" %
              (t / 1e6, 100 * t / total, n, b / 1e6, gbs, 100 * gbs / peak, name[:90]))
    print("--- slowest single launches")
    rows = sorted(per_launch.values(), key=lambda d: -d.get("gpu__time_duration.sum", 0.0))[:20]
    for d in rows:
        t = d.get("gpu__time_duration.sum", 0.0)
        b = d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        print("%8.1f us  %8.1f MB %7.1f GB/s  grid %-16s %s" %
              (t / 1e3, b / 1e6, b / t if t else 0.0, d["grid"], d["name"][:80]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 45)
