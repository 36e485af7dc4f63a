#!/usr/bin/env python
"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict


def main(path, top=40):
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rd = csv.DictReader(lines)
    agg = defaultdict(lambda: [0, 0.0])
    total = 0.0
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = r["Kernel Name"]
        name = re.sub(r"\(.*$", "", name)
        val = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        if unit in ("us", "usecond"):
            val *= 1e3
        elif unit in ("ms", "msecond"):
            val *= 1e6
        grid = r.get("Grid Size", "")
        agg[name][0] += 1
        agg[name][1] += val
        total += val
        rows.append((name, val, grid))
    print("total %.3f ms over %d launches" % (total / 1e6, len(rows)))
    for name, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%8.3f ms %5.1f%% %5d  %s" % (t / 1e6, 100 * t / total, n, name[:110]))
    print("--- slowest single launches")
    for name, val, grid in sorted(rows, key=lambda r: -r[1])[:25]:
        print("%8.1f us  grid %-18s %s" % (val / 1e3, grid, name[:100]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
