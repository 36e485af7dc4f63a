"""CTC loss oracle (PyTorch): the reference's formulation,
T/models/experimental/deepspeech.py:359-396 - emissions (softmax(z)+eps)/(1+C*eps), blank is
the last class, infeasible samples are ignored, mean over the batch."""
from __future__ import annotations

import torch
import torch.nn.functional as F


def ctc_input_lengths(actual_time_steps, ctc_time_steps: int, max_time_steps: int):
    """floor(actual * ctc_steps / max_steps) as the reference computes it in float32."""
    a = actual_time_steps.reshape(-1).to(torch.float32)
    return torch.floor(a * float(ctc_time_steps) / float(max_time_steps)).to(torch.int32)


def ctc_loss(logits, labels, input_len, label_len, eps=1e-7):
    """logits [N, T, C] fp32; returns the batch-mean loss."""
    n, t, c = logits.shape
    q = F.softmax(logits.float(), dim=-1)
    logp = torch.log((q + eps) / (1.0 + c * eps))
    losses = F.ctc_loss(logp.permute(1, 0, 2), labels.long(), input_len.long().reshape(-1),
                        label_len.long().reshape(-1), blank=c - 1, reduction="none",
                        zero_infinity=True)
    return losses.mean()


def loss(graph, vals, labels):
    spec = graph.loss
    logits = vals[graph.logits.id].float()
    lab, in_len, lab_len = labels
    t = logits.shape[1]
    ilen = ctc_input_lengths(in_len, t, spec["max_time_steps"])
    return ctc_loss(logits, lab, ilen, lab_len, spec.get("eps", 1e-7)), 0, 0
