"""Compares --comm_impl=p2p and p2p_fused on N GPUs: trained variables per rank and per impl.
Usage: python tools/fused_rs_check.py <model> <data> <batch> [lr] [ranks]"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(impl, model, data, batch, lr, n, out):
    kw = dict(model=model, data_name=data, batch_size=batch, use_fp16=True, optimizer="momentum",
              init_learning_rate=lr, num_batches=5, num_warmup_batches=0,
              variable_update="replicated", num_gpus=n, comm_impl=impl)
    env = dict(os.environ, PYTHONPATH=ROOT)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", str(n), "--",
           os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw), out]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    if res.returncode != 0:
        print(res.stdout[-3000:], res.stderr[-3000:])
        raise SystemExit(1)
    return [np.load(os.path.join(out, "rank%d.npz" % r)) for r in range(n)]


def main():
    model, data, batch = sys.argv[1], sys.argv[2], int(sys.argv[3])
    lr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.01
    n = int(sys.argv[5]) if len(sys.argv) > 5 else 2
    res = {}
    for impl in ("p2p", "p2p_fused"):
        with tempfile.TemporaryDirectory() as d:
            rs = run(impl, model, data, batch, lr, n, d)
            res[impl] = [r["w"].copy() for r in rs]
            print(impl, "loss", float(rs[0]["loss"]), "img/s", float(rs[0]["ips"]), "nan",
                  int(np.isnan(res[impl][0]).sum()), "ranks identical",
                  all(np.array_equal(res[impl][0], w, equal_nan=True) for w in res[impl]))
    a, b = res["p2p"][0], res["p2p_fused"][0]
    same = np.array_equal(a, b, equal_nan=True)
    print("FUSED_RS", "IDENTICAL" if same else "DIFFERENT", "max abs diff",
          float(np.nanmax(np.abs(a - b))))
    if not same:
        idx = np.flatnonzero(~((a == b) | (np.isnan(a) & np.isnan(b))))
        print("first differing indices", idx[:10], "count", idx.size)


if __name__ == "__main__":
    main()
