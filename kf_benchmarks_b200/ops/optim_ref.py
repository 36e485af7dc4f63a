"""PyTorch restatement of the fused update rule of ``csrc/optim_comm.cu`` (apply_update).

Used on the CPU path and as the oracle of the optimizer / all-reduce kernel tests. The
rules are TensorFlow's (T/benchmark_cnn.py:1172-1190): plain SGD, Nesterov momentum as
``ApplyMomentum(use_nesterov=True)``, RMSProp with momentum, Adam with bias correction.
``g = grad * grad_scale + weight_decay * w`` then an optional elementwise clip.
"""
from __future__ import annotations

import torch


def apply_update(h, w, grad, m=None, v=None, w16=None, blend=None, decay_mask=None):
    """In-place update of the flat arena ``w``; ``h`` has the fields of ``kfb_hyper``.
    ``decay_mask`` (optional, one uint8 per 128-element block, 0 = no L2 term) is the torch form
    of ``kfb_hyper.decay_mask``."""
    wd = h.weight_decay
    if decay_mask is not None:
        wd = wd * decay_mask.to(w.dtype).repeat_interleave(128)[:w.numel()]
    g = grad * h.grad_scale + wd * w
    if h.clip > 0:
        g = g.clamp(-h.clip, h.clip)
    if h.opt == 0:
        w.sub_(h.lr * g)
    elif h.opt == 1:
        m.mul_(h.momentum).add_(g)
        w.sub_(h.lr * (g + h.momentum * m))
    elif h.opt == 2:
        v.mul_(h.decay).add_((1.0 - h.decay) * g * g)
        m.mul_(h.momentum).add_(h.lr * g * torch.rsqrt(v + h.epsilon))
        w.sub_(m)
    else:
        b1, b2 = h.momentum, h.decay
        m.mul_(b1).add_((1.0 - b1) * g)
        v.mul_(b2).add_((1.0 - b2) * g * g)
        lr_t = h.lr * (1.0 - b2 ** h.adam_t) ** 0.5 / (1.0 - b1 ** h.adam_t)
        w.sub_(lr_t * m / (v.sqrt() + h.epsilon))
    if w16 is not None:
        w16.copy_(w)
    return w
