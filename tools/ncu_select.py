#!/usr/bin/env python
"""Selected raw metrics of an ncu report (the `--page raw --csv` export) as a small CSV.

    ncu -i report.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_select.py raw.csv profiles/ncu_<tag>_<name>.csv
"""
import csv
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__pcsamp_warps_issue_stalled_long_scoreboard",
        "smsp__pcsamp_sample_count")


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(src)))
    if len(rows) < 3:
        print("no rows in", src)
        return 1
    hdr = rows[0]
    cols = [i for i, h in enumerate(hdr) if h in ("ID", "Kernel Name") or h in KEEP]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        for r in rows:
            w.writerow([r[i][:120] for i in cols if i < len(r)])
    print("%s: %d launches, %d metrics" % (dst, len(rows) - 2, len(cols)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
