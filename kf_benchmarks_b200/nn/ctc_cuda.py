"""CUDA side of the CTC loss head (csrc/loss_kernels.cu)."""
from __future__ import annotations

import ctypes

import torch

from ..ops import native
from ..ops.native import f32, i32, i64, ptr
from .ctc_ref import ctc_input_lengths


def prepare(ex) -> None:
    g = ex.graph
    n, t, c = g.logits.shape
    lmax = g.label_specs[0][0][1]
    fn = native.load().kfb_ctc_loss_workspace_floats
    fn.restype = ctypes.c_longlong
    ex._loss_ws = torch.zeros(int(fn(i32(n), i32(t), i32(lmax))), dtype=torch.float32,
                              device=ex.device)
    ex._ctc_len = torch.zeros((n,), dtype=torch.int32, device=ex.device)
    ex._grad(g.logits)


def run(ex, loss_scale, dyn_scale, backward, s) -> None:
    g = ex.graph
    n, t, c = g.logits.shape
    labels, in_len, lab_len = ex.label_tensors
    # sequence length seen by the CTC = floor(actual * T_ctc / T_input), a handful of ints
    ex._ctc_len.copy_(ctc_input_lengths(in_len, t, g.loss["max_time_steps"]))
    lb = ex.buf[g.logits.id]
    gb = ex._grad(g.logits) if backward else None
    native.call("ctc_loss", lb.p, i64(lb.ld), i32(n), i32(t), i32(c), ptr(labels),
                i32(labels.shape[1]), ptr(ex._ctc_len), ptr(lab_len), f32(g.loss.get("eps", 1e-7)),
                f32(loss_scale / n), ptr(dyn_scale), gb.p if gb else None,
                i64(gb.ld if gb else 0), ptr(ex._loss_ws), i64(ex._loss_ws.numel()),
                ptr(ex.loss_out), s)
