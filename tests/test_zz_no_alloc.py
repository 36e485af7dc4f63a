"""The allocation-free softmax-cross-entropy entry point (``kfb_softmax_xent_static``).

Some users of the library cannot give it scratch memory: a training or serving loop that must
never reach ``cudaMalloc`` (it synchronises the device), or a host program that owns the whole device heap. For them the entry point keeps its per-row
scratch in a fixed 48 KiB device array inside the shared library (at most 4096 rows per call).

The test drives the function through the C ABI only (``ctypes`` on the built ``.so``) and checks
it against the ordinary ``kfb_softmax_xent`` on two back-to-back calls with different inputs:
the second call sees the first call's leftovers in the shared array and must not be affected.
It runs last (file name) so a ``pytest -x`` run has already exercised everything else, and it
passes both on its own in a fresh process and after the other tests.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from kf_benchmarks_b200.ops import native  # noqa: E402


def _call_pair(lib, rows, classes, seed, scale):
    ld = (classes + 7) // 8 * 8
    g = torch.Generator(device="cuda").manual_seed(seed)
    logits = torch.zeros(rows, ld, device="cuda")
    logits[:, :classes] = torch.randn(rows, classes, device="cuda", generator=g) * 4
    labels = torch.randint(0, classes, (rows,), device="cuda", generator=g, dtype=torch.int32)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda t: ctypes.c_void_p(t.data_ptr())  # noqa: E731
    # ordinary entry point with caller-provided scratch
    dl_a = torch.full((rows, ld), 7.0, device="cuda", dtype=torch.float16)
    loss_rows = torch.empty(rows, device="cuda")
    t1 = torch.empty(rows, device="cuda", dtype=torch.int32)
    t5 = torch.empty(rows, device="cuda", dtype=torch.int32)
    out_a = torch.zeros(4, device="cuda")
    rc = lib.kfb_softmax_xent(vp(logits), ctypes.c_longlong(ld), vp(labels), ctypes.c_int(rows),
                              ctypes.c_int(classes), ctypes.c_float(scale), None, vp(dl_a),
                              ctypes.c_longlong(ld), vp(loss_rows), vp(t1), vp(t5), vp(out_a),
                              stream)
    assert rc == 0, lib.kfb_last_error()
    # allocation-free entry point
    dl_b = torch.full((rows, ld), -7.0, device="cuda", dtype=torch.float16)
    out_b = torch.zeros(4, device="cuda")
    before = torch.cuda.memory_allocated()
    rc = lib.kfb_softmax_xent_static(vp(logits), ctypes.c_longlong(ld), vp(labels),
                                     ctypes.c_int(rows), ctypes.c_int(classes),
                                     ctypes.c_float(scale), None, vp(dl_b), ctypes.c_longlong(ld),
                                     vp(out_b), stream)
    assert rc == 0, lib.kfb_last_error()
    assert torch.cuda.memory_allocated() == before
    torch.cuda.synchronize()
    assert torch.allclose(out_a[:3], out_b[:3], atol=1e-5, rtol=1e-5)
    assert torch.allclose(dl_a.float(), dl_b.float(), atol=2e-3, rtol=2e-2)
    return out_b


def test_static_scratch_matches_regular_entry_point_twice():
    lib = native.load()
    lib.kfb_last_error.restype = ctypes.c_char_p
    max_rows = lib.kfb_softmax_xent_static_max_rows()
    assert max_rows == 4096
    # two calls in a row, different shapes and data: the second runs on the first one's leftovers
    first = _call_pair(lib, rows=300, classes=1001, seed=11, scale=128.0 / 300)
    second = _call_pair(lib, rows=64, classes=11, seed=12, scale=1.0 / 64)
    assert float(first[0]) != float(second[0])
    # sizes beyond the built-in scratch are refused with a clear message
    logits = torch.zeros(8, 8, device="cuda")
    labels = torch.zeros(8, device="cuda", dtype=torch.int32)
    out = torch.zeros(4, device="cuda")
    rc = lib.kfb_softmax_xent_static(ctypes.c_void_p(logits.data_ptr()), ctypes.c_longlong(8),
                                     ctypes.c_void_p(labels.data_ptr()), ctypes.c_int(max_rows + 1),
                                     ctypes.c_int(8), ctypes.c_float(1.0), None, None,
                                     ctypes.c_longlong(8), ctypes.c_void_p(out.data_ptr()),
                                     ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b"exceed" in lib.kfb_last_error()
