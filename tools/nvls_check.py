"""--comm_impl=nvls against --comm_impl=p2p on N GPUs (default 2): trained variables per rank and
per impl (replicas must be bit-identical; the two impls differ only by summation order), then the
isolated all-reduce micro-benchmark for p2p / nvls / nccl.

Usage: python tools/nvls_check.py [ranks] [model]      (needs an NVSwitch box with NVLS)"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(n, args, timeout=900):
    env = dict(os.environ, PYTHONPATH=ROOT)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", str(n), "--"] + args
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)


def main() -> int:
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    model = sys.argv[2] if len(sys.argv) > 2 else "resnet50"
    res = {}
    for impl in ("p2p", "nvls"):
        with tempfile.TemporaryDirectory() as d:
            kw = dict(model="resnet20", data_name="cifar10", batch_size=32, use_fp16=True,
                      optimizer="momentum", init_learning_rate=0.05, num_batches=6,
                      num_warmup_batches=0, variable_update="replicated", num_gpus=n,
                      comm_impl=impl)
            r = _launch(n, [os.path.join(ROOT, "tests", "helpers", "rank_worker.py"),
                            json.dumps(kw), d])
            if r.returncode != 0:
                print(impl, "FAILED\n", r.stdout[-3000:], r.stderr[-3000:])
                return 1
            res[impl] = [np.load(os.path.join(d, "rank%d.npz" % k))["w"].copy() for k in range(n)]
            print(impl, "replicas identical:",
                  all(np.array_equal(res[impl][0], w) for w in res[impl]))
    a, b = res["p2p"][0], res["nvls"][0]
    print("NVLS_CHECK max abs diff vs p2p %.3e, cosine %.8f" %
          (float(np.abs(a - b).max()),
           float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b)))))
    env_nvls = dict(os.environ, KFB_P2P_NVLS="1")
    os.environ.update(env_nvls)       # _launch copies os.environ
    r = _launch(n, [os.path.join(ROOT, "tools", "p2p_check.py")])
    os.environ.pop("KFB_P2P_NVLS", None)
    print("p2p_check under NVLS: rc", r.returncode,
          "ALL PASS" if "P2P_CHECK ALL PASS" in r.stdout else "FAILED")
    if r.returncode != 0:
        print(r.stdout[-3000:], r.stderr[-2000:])
    for impl in ("p2p", "nvls", "nccl"):
        r = _launch(n, ["-m", "kf_benchmarks_b200.all_reduce_benchmark", "--model=" + model,
                        "--variable_update=replicated", "--num_gpus=%d" % n, "--use_fp16",
                        "--optimizer=momentum", "--comm_impl=" + impl, "--num_batches=20",
                        "--num_warmup_batches=5"])
        print("==", impl, "rc", r.returncode)
        print("\n".join(l for l in r.stdout.splitlines()
                        if "Average" in l or "per all-reduce" in l or "Model:" in l))
        if r.returncode != 0:
            print(r.stderr[-2000:])
    return 0


if __name__ == "__main__":
    sys.exit(main())
