"""Local launcher: N ranks of one command on one box (the role of ``kungfu-run -np N`` /
``mpirun -np N`` in the reference's launch scripts, T/README.md:97-114, T/run_kf.sh-bak:18-21).

Each rank gets RANK / LOCAL_RANK / WORLD_SIZE / LOCAL_WORLD_SIZE / MASTER_ADDR / MASTER_PORT. Rank 0's stdout is
passed through; other ranks are prefixed. The first non-zero exit kills the whole group
(failure mode of the reference: one rank dies, the launcher reports it and reaps the rest,
SURVEY.md 5.3).

``--native`` hands the job to the C++ launcher ``_lib/kfb-run`` (csrc/host/launcher/kfb_run.cpp,
same environment, prefixes and failure handling, plus ``-logdir`` / ``-timeout``), which can
also be used directly for any command: ``kfb-run -np 8 -- python -m kf_benchmarks_b200.cli ...``.
"""
from __future__ import annotations

import os
import signal
import socket
import subprocess
import sys
import threading
import time
from typing import List


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch_local(py_args: List[str], nproc: int, env_extra=None, timeout=None) -> int:
    port = _free_port()
    procs = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc),
                    "LOCAL_WORLD_SIZE": str(nproc), "MASTER_ADDR": "127.0.0.1",
                    "MASTER_PORT": str(port)})
        env.update(env_extra or {})
        out = None if r == 0 else subprocess.PIPE
        p = subprocess.Popen([sys.executable] + py_args, env=env, stdout=out,
                             stderr=subprocess.STDOUT if r else None, start_new_session=True)
        procs.append(p)
        if r:
            threading.Thread(target=_pump, args=(p, r), daemon=True).start()
    start = time.time()
    rc = 0
    alive = set(range(nproc))
    while alive:
        for r in list(alive):
            code = procs[r].poll()
            if code is None:
                continue
            alive.discard(r)
            if code != 0 and rc == 0:
                rc = code
                sys.stderr.write("[launcher] rank %d exited with %d; stopping the others\n" %
                                 (r, code))
                _kill_all(procs)
        if timeout and time.time() - start > timeout:
            sys.stderr.write("[launcher] timeout after %.0fs; stopping all ranks\n" % timeout)
            _kill_all(procs)
            rc = rc or 124
            break
        time.sleep(0.05)
    return rc


def _pump(p, rank):
    for line in iter(p.stdout.readline, b""):
        sys.stdout.write("[rank %d] %s" % (rank, line.decode("utf-8", "replace")))
        sys.stdout.flush()


def _kill_all(procs):
    for q in procs:
        if q.poll() is None:
            try:
                os.killpg(q.pid, signal.SIGTERM)
            except (ProcessLookupError, PermissionError):
                pass


def main(argv=None) -> int:
    """``python -m kf_benchmarks_b200.parallel.launcher -np 4 -- -m kf_benchmarks_b200.cli ...``"""
    argv = sys.argv[1:] if argv is None else argv
    if len(argv) < 3 or argv[0] != "-np":
        print("usage: launcher -np N [--] <python args>", file=sys.stderr)
        return 2
    n = int(argv[1])
    rest = argv[2:]
    native = bool(rest) and rest[0] == "--native"
    if native:
        rest = rest[1:]
    if rest and rest[0] == "--":
        rest = rest[1:]
    if native:
        from .. import build as build_lib
        exe = build_lib.build_launcher()
        os.execv(exe, [exe, "-np", str(n), "--", sys.executable] + rest)
    return launch_local(rest, n)


if __name__ == "__main__":
    sys.exit(main())
