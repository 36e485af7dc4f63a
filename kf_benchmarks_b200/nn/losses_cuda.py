"""CUDA side of the non-classification loss heads: marshals the executor's buffers into the
C ABI of csrc/loss_kernels.cu."""
from __future__ import annotations

import ctypes

import torch

from ..ops import native
from ..ops.native import f32, i32, i64, ptr


def prepare(ex) -> None:
    """Allocates everything the loss head needs up front (no allocation inside a step)."""
    g = ex.graph
    kind = g.loss["kind"]
    if kind == "ssd":
        spec = g.loss
        n = g.batch_size
        sizes = [g.tensors[t].shape[1] for t in spec["loc"]]
        anchors = sum(s * s * d for s, d in zip(sizes, spec["num_dboxes"]))
        fn = native.load().kfb_ssd_loss_workspace_floats
        fn.restype = ctypes.c_longlong
        ex._loss_ws = torch.zeros(int(fn(i32(n), i32(anchors))), dtype=torch.float32,
                                  device=ex.device)
        for tid in list(spec["loc"]) + list(spec["conf"]):
            ex._grad(g.tensors[tid])
        ex._ssd_sizes = (ctypes.c_int * 6)(*sizes)
        ex._ssd_ndef = (ctypes.c_int * 6)(*spec["num_dboxes"])
    elif kind == "ctc":
        from . import ctc_cuda
        ctc_cuda.prepare(ex)


def _ptr_array(bufs):
    return (ctypes.c_void_p * len(bufs))(*[b.t.data_ptr() for b in bufs])


def run(ex, kind, loss_scale, dyn_scale, backward, s) -> None:
    g = ex.graph
    if kind == "ssd":
        spec = g.loss
        n = g.batch_size
        loc = [ex.buf[t] for t in spec["loc"]]
        conf = [ex.buf[t] for t in spec["conf"]]
        dloc = [ex.gbuf[t] for t in spec["loc"]]
        dconf = [ex.gbuf[t] for t in spec["conf"]]
        gt_loc, gt_label, num_matched = ex.label_tensors
        native.call("ssd_loss", i32(n), i32(spec["num_classes"]), ex._ssd_sizes, ex._ssd_ndef,
                    _ptr_array(loc), (ctypes.c_longlong * 6)(*[b.ld for b in loc]),
                    _ptr_array(conf), (ctypes.c_longlong * 6)(*[b.ld for b in conf]),
                    _ptr_array(dloc) if backward else None,
                    _ptr_array(dconf) if backward else None,
                    ptr(gt_loc), ptr(gt_label), ptr(num_matched),
                    i32(spec.get("negs_per_positive", 3)), f32(loss_scale / n), ptr(dyn_scale),
                    ptr(ex._loss_ws), i64(ex._loss_ws.numel()), ptr(ex.loss_out), s)
    elif kind == "ctc":
        from . import ctc_cuda
        ctc_cuda.run(ex, loss_scale, dyn_scale, backward, s)
    else:
        raise NotImplementedError("loss head %r" % kind)
