"""Build the in-tree CUDA library for sm_100a with plain nvcc.

``python -m kf_benchmarks_b200.build`` compiles every ``csrc/*.cu`` into
``kf_benchmarks_b200/_lib/libkfb200.so``. nvcc cross-compiles without a GPU, so
this is also the CPU-side "does it build" check used by ``__graft_entry__.build``.
Objects are cached by source mtime under ``build/``.
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "kf_benchmarks_b200", "_lib")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(OUT_DIR, "libkfb200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include"),
]
# No --use_fast_math: division, rsqrt, exp and log in BN finalize, the optimizers, the losses
# and the LSTM cell are IEEE-rounded (the hot kernels are memory- or tensor-bound and do not
# depend on it; KFB_FAST_MATH=1 at build time restores the flag for an A/B measurement).
if os.environ.get("KFB_FAST_MATH") == "1":
    NVCC_FLAGS.insert(NVCC_FLAGS.index("-I"), "--use_fast_math")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    return "nvcc"


def _deps_mtime() -> float:
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    return max([os.path.getmtime(h) for h in hdrs] + [0.0])


def _compile_one(src: str, verbose: bool) -> str:
    obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
    newest = max(os.path.getmtime(src), _deps_mtime())
    if os.path.exists(obj) and os.path.getmtime(obj) >= newest:
        return obj
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
    if verbose:
        sys.stderr.write(res.stderr)
    return obj


HOST_LIB = os.path.join(OUT_DIR, "libkfb200_host.so")


def build_host(force: bool = False) -> str:
    """Host-side input runtime (csrc/host/*.cpp, plain g++): TFRecord streaming threads,
    CRC32C, tf.Example scanner. No CUDA involved, so it also serves the --device=cpu path."""
    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "host", "*.cpp")))
    if not srcs:
        raise RuntimeError("no host sources found under csrc/host/")
    newest = max(os.path.getmtime(s) for s in srcs)
    if not force and os.path.exists(HOST_LIB) and os.path.getmtime(HOST_LIB) >= newest:
        return HOST_LIB
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-msse4.2", "-mf16c", "-o", HOST_LIB] \
        + srcs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"host library build failed:\n{res.stdout}\n{res.stderr}")
    return HOST_LIB


LAUNCHER = os.path.join(OUT_DIR, "kfb-run")


def build_launcher(force: bool = False) -> str:
    """Native rank launcher (csrc/host/launcher/kfb_run.cpp): the `kungfu-run -np N` /
    `mpirun -np N` of this runtime as one small executable."""
    os.makedirs(OUT_DIR, exist_ok=True)
    src = os.path.join(CSRC, "host", "launcher", "kfb_run.cpp")
    if not force and os.path.exists(LAUNCHER) and os.path.getmtime(LAUNCHER) >= os.path.getmtime(src):
        return LAUNCHER
    cxx = os.environ.get("CXX", "g++")
    res = subprocess.run([cxx, "-O2", "-std=c++17", "-o", LAUNCHER, src], capture_output=True,
                         text=True)
    if res.returncode != 0:
        raise RuntimeError(f"launcher build failed:\n{res.stdout}\n{res.stderr}")
    return LAUNCHER


def build(force: bool = False, verbose: bool = False) -> str:
    build_host(force)
    build_launcher(force)
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    if not srcs:
        raise RuntimeError("no CUDA sources found under csrc/")
    if force:
        for o in glob.glob(os.path.join(OBJ_DIR, "*.o")):
            os.remove(o)
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose), srcs))
    newest_obj = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < newest_obj:
        # the arch at link time too: without it nvcc adds an (empty) sm_52 device-link stub
        cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
