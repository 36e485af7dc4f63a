"""ctypes binding to the in-tree CUDA library (``_lib/libkfb200.so``).

The library is built by ``kf_benchmarks_b200.build`` (plain ``nvcc`` for
``sm_100a``) and exposes a C ABI declared in ``include/kfb200.h``. Tensors are
handed over as raw device pointers plus the current CUDA stream, so there is no
dependency on torch's C++ headers and the ``.so`` is visible in the process's
map when it is used.

On a machine with a GPU a missing library is a hard error: the ops in
``kf_benchmarks_b200.ops`` never fall back to PyTorch math on CUDA tensors.
"""
from __future__ import annotations

import ctypes
import os
import threading

# KFB_LIB_NAME selects another build of the library in _lib/ (A/B measurements of build flags)
_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib",
                         os.environ.get("KFB_LIB_NAME", "libkfb200.so"))
_lib = None
_lock = threading.Lock()

# number of kernel launches issued through this binding (bench.py reports it)
launch_count = 0


class KfbError(RuntimeError):
    pass


class ConvDesc(ctypes.Structure):
    """Mirror of ``kfb_conv_desc`` in include/kfb200.h."""
    _fields_ = [(name, ctypes.c_int) for name in (
        "n", "h", "w", "c", "k", "r", "s", "pad_h_lo", "pad_h_hi", "pad_w_lo", "pad_w_hi",
        "stride_h", "stride_w", "p", "q", "x_pix_stride", "x_row_stride")]


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def load():
    """Load the shared library once; raise KfbError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_LIB_PATH):
            raise KfbError(
                f"native library not built: {_LIB_PATH} (run `python -m kf_benchmarks_b200.build`)")
        lib = ctypes.CDLL(_LIB_PATH)
        lib.kfb_last_error.restype = ctypes.c_char_p
        _lib = lib
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().kfb_last_error().decode("utf-8", "replace")
        raise KfbError(f"{what} failed (rc={rc}): {msg}")


# kernels launched per entry point (used for the launch count bench.py reports; entry points
# not listed launch one kernel; conv_dgrad launches one per stride phase and is counted once)
_LAUNCHES = {"gemm_nt": 2, "conv_wgrad": 2, "bn_forward_train": 3, "bn_forward_train_stats": 2, "fill_zero": 0, "step_advance": 1, "set_floats": 1, "bn_forward_inference": 2,
             "bn_backward": 3, "bias_act_backward": 2, "softmax_xent": 2, "dwconv_wgrad": 2, "ssd_loss": 4, "ctc_loss": 3, "gemm_nt_rs": 2, "conv_wgrad_rs": 2}
# allreduce_update2 launches a second (cast) kernel with compact gradient transfer; counted once


# tools/node_timing.py: when a list, every call appends (entry point, start event, end event)
call_events = None
current_tag = ""


def call(name: str, *args) -> None:
    """Call ``kfb_<name>`` and raise on a non-zero return code."""
    global launch_count
    fn = getattr(load(), "kfb_" + name)
    if call_events is not None:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = fn(*args)
        e1.record()
        call_events.append((name, current_tag, e0, e1))
    else:
        rc = fn(*args)
    launch_count += _LAUNCHES.get(name, 1)
    check(rc, name)


def ptr(t) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def stream_ptr() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


i32 = ctypes.c_int
i64 = ctypes.c_longlong
f32 = ctypes.c_float
