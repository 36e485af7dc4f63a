#!/usr/bin/env python
"""Test runner (counterpart of T/run_tests.py:58-100).

    python run_tests.py                      # CPU suite (what the driver runs every round)
    python run_tests.py --gpu                # GPU suite (needs a B200; multi-GPU tests need >= 2)
    python run_tests.py --full_tests         # everything that can run on this machine

The reference splits "distributed" tests into a second process group because TF grabs all GPU
memory; here multi-rank tests spawn their own rank processes (gloo on CPU, one process per GPU
on a GPU box), so a single pytest invocation is enough.
"""
import argparse
import subprocess
import sys


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--full_tests", action="store_true")
    ap.add_argument("--run_distributed_tests", action="store_true",
                    help="accepted for compatibility: multi-rank tests are part of both suites")
    a, rest = ap.parse_known_args()
    cmd = [sys.executable, "-m", "pytest", "tests", "-q"]
    if a.full_tests:
        pass
    elif a.gpu:
        cmd += ["-m", "gpu"]
    else:
        cmd += ["-m", "not gpu"]
    return subprocess.call(cmd + rest)


if __name__ == "__main__":
    sys.exit(main())
