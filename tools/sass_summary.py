#!/usr/bin/env python
"""Regenerates profiles/sass_summary.md: per-kernel counts of the SASS mnemonics that prove
the Blackwell-native path (B200_PROFILING.md) from `cuobjdump -sass` of the in-tree library."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "kf_benchmarks_b200", "_lib", "libkfb200.so")
KEYS = ("UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR",
        "UTCATOMSWS", "SYNCS", "HMMA", "MEMBAR", "LDGMC", "LDG", "STG", "LDS", "STS", "ATOM", "RED",
        "LD.E", "ST.E")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    fn = None
    counts = collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            fn = m.group(1)
            counts[fn] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
        if m and fn:
            op = m.group(1)
            counts[fn]["_total"] += 1
            for k in KEYS:
                if op.startswith(k):
                    # keep the informative suffixes (IM2COL, SYS scope, vector width)
                    keep = op if k in ("UTMALDG", "MEMBAR", "LDTM", "LDGMC") else k
                    counts[fn][keep] += 1
                    break
    out = ["# SASS evidence for libkfb200.so (cuobjdump -sass, sm_100a)", "",
           "Mnemonics that prove the Blackwell-native path (B200_PROFILING.md): UTCHMMA = "
           "tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA load (incl. .IM2COL), UTCBAR = "
           "tcgen05.commit, SYNCS = mbarrier, UTMASTG = TMA tensor store (the opt-in KFB_TMA_STORE "
           "epilogue of the FAST GEMM_TN / FPROP / DGRAD variants). In the weight-gradient variants (MODE 1 = GEMM_NT, "
           "MODE 4 = WGRAD) `MEMBAR.SC.SYS` is the system-scope fence that follows the epilogue's "
           "peer (NVLink) stores of the fused GEMM + reduce-scatter (--comm_impl=p2p_fused); no "
           "HMMA (legacy mma.sync) appears anywhere.", ""]
    for fn, c in counts.items():
        if not any(k.startswith(("UTC", "UTMA", "LDTM")) for k in c):
            continue
        out.append("## " + fn)
        out.append("  total instructions: %d" % c["_total"])
        out.append("  " + ", ".join("%s x%d" % (k, v) for k, v in sorted(c.items())
                                    if k != "_total"))
        out.append("")
    others = [fn for fn, c in counts.items()
              if not any(k.startswith(("UTC", "UTMA", "LDTM")) for k in c)]
    out.append("## other kernels (CUDA-core, bandwidth-bound): %d" % len(others))
    out.append("  " + ", ".join(re.sub(r"^_ZN3kfb\d+", "", f)[:40] for f in others))
    out.append("")
    out.append("## NVLS all-reduce (--comm_impl=nvls): multimem.ld_reduce = LDGMC.*.ADD, "
               "multimem.st = STG.E.*.STRONG.SYS on the multicast alias")
    for fn, c in counts.items():
        mc = {k: v for k, v in c.items() if k.startswith("LDGMC")}
        if mc:
            out.append("  %s: %s" % (fn, ", ".join("%s x%d" % kv for kv in sorted(mc.items()))))
    hm = sum(c["HMMA"] for c in counts.values())
    out.append("")
    out.append("HMMA (legacy tensor path) instructions in the whole library: %d" % hm)
    path = os.path.join(ROOT, "profiles", "sass_summary.md")
    with open(path, "w") as f:
        f.write("\n".join(out) + "\n")
    print("wrote", path, "kernels:", len(counts))


if __name__ == "__main__":
    sys.exit(main())
