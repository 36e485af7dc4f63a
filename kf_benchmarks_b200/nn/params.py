"""ParameterStore: all variables of a model in a few flat arenas.

The reference keeps one TF variable per tensor and, for the gradient exchange,
flattens / concatenates / casts / splits them every step (``_TensorPacker``,
T/batch_allreduce.py:391-481; small-tensor packing T/allreduce.py:514-588). Here
that packing is a property of the layout: trainable variables, their gradients and
optimizer slots live at the same offsets of contiguous fp32 arenas (plus an fp16
shadow of the weights for the tensor-core kernels), ordered by *backward
completion* (reverse creation order) so a "bucket" of the gradient exchange is
just an offset range that becomes final early in the backward pass.

Non-trainable state (batch-norm moving statistics) lives in a separate arena.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np
import torch

from .graph import Graph, ParamSpec

ALIGN = 128  # elements; keeps every slot 512-byte (fp32) / 256-byte (fp16) aligned


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class Slot:
    __slots__ = ("spec", "offset", "numel", "arena")

    def __init__(self, spec: ParamSpec, offset: int, arena: str):
        self.spec = spec
        self.offset = offset
        self.numel = spec.numel
        self.arena = arena


class ParameterStore:
    def __init__(self, graph: Graph, device, fp16_shadow: bool = True,
                 allocator: Optional[Callable[[int, torch.dtype], torch.Tensor]] = None):
        self.graph = graph
        self.device = torch.device(device)
        alloc = allocator or (lambda n, dt: torch.zeros(n, dtype=dt, device=self.device))
        self.slots: Dict[str, Slot] = {}
        trainable = [p for p in graph.params.values() if p.trainable]
        state = [p for p in graph.params.values() if not p.trainable]
        off = 0
        self.order: List[Slot] = []
        for p in reversed(trainable):  # backward-completion order
            s = Slot(p, off, "w")
            self.slots[p.name] = s
            self.order.append(s)
            off += _round_up(p.numel, ALIGN)
        self.num_trainable_padded = off
        off = 0
        for p in state:
            s = Slot(p, off, "s")
            self.slots[p.name] = s
            off += _round_up(p.numel, ALIGN)
        self.num_state_padded = max(off, ALIGN)
        n = max(self.num_trainable_padded, ALIGN)
        self.w32 = alloc(n, torch.float32)
        self.grad = alloc(n, torch.float32)
        self.w16 = alloc(n, torch.float16) if fp16_shadow else None
        self.state = torch.zeros(self.num_state_padded, dtype=torch.float32, device=self.device)
        self.opt_slots: Dict[str, torch.Tensor] = {}

    # -- views ----------------------------------------------------------------
    def _arena(self, slot: Slot, which: str) -> torch.Tensor:
        if slot.arena == "s":
            return self.state
        return {"w32": self.w32, "w16": self.w16, "grad": self.grad}[which]

    def view(self, name: str, which: str = "w32") -> torch.Tensor:
        s = self.slots[name]
        a = self._arena(s, which)
        return a[s.offset:s.offset + s.numel].view(s.spec.shape)

    def offset(self, name: str) -> int:
        return self.slots[name].offset

    def span(self, names) -> Tuple[int, int]:
        """[lo, hi) arena range (padded) covered by the given trainable variables."""
        lo, hi = None, None
        for n in names:
            s = self.slots[n]
            if s.arena != "w":
                continue
            a, b = s.offset, s.offset + _round_up(s.numel, ALIGN)
            lo = a if lo is None else min(lo, a)
            hi = b if hi is None else max(hi, b)
        return (0, 0) if lo is None else (lo, hi)

    def opt_slot(self, key: str) -> torch.Tensor:
        if key not in self.opt_slots:
            self.opt_slots[key] = torch.zeros_like(self.w32)
        return self.opt_slots[key]

    # -- initialisation ---------------------------------------------------------
    def initialize(self, seed: int = 1234) -> None:
        """Initialisers of T/convnet_builder_legacy.py:162-163,309-316,399-416 (Glorot-uniform
        conv kernels unless a stddev is given, truncated-normal FC weights, constant biases)."""
        gen = torch.Generator(device="cpu").manual_seed(int(seed))
        w = torch.zeros(self.w32.numel(), dtype=torch.float32)
        st = torch.zeros(self.state.numel(), dtype=torch.float32)
        for p in self.graph.params.values():  # creation order => reproducible stream
            s = self.slots[p.name]
            val = _init_value(p, gen)
            dst = st if s.arena == "s" else w
            dst[s.offset:s.offset + s.numel] = val.reshape(-1)
        self.w32.copy_(w)
        self.state.copy_(st)
        self.grad.zero_()
        self.sync_shadow()
        for t in self.opt_slots.values():
            t.zero_()

    def sync_shadow(self) -> None:
        if self.w16 is not None:
            self.w16.copy_(self.w32)

    # -- TF-layout import / export ------------------------------------------------
    def to_tf_arrays(self, prefix: str = "v0/", which: str = "w32") -> Dict[str, np.ndarray]:
        """{tf variable name: ndarray in the reference's layout}. ``which``
        selects the arena ("w32" weights, "grad" gradients of the trainable variables)."""
        out = {}
        for p in self.graph.params.values():
            if which != "w32" and not p.trainable:
                continue
            t = self.view(p.name, which).detach().float().cpu()
            out[prefix + p.name] = _native_to_tf(p, t).numpy()
        return out

    def from_tf_arrays(self, arrays: Dict[str, np.ndarray], prefix: str = "v0/",
                       strict: bool = True) -> List[str]:
        missing = []
        for p in self.graph.params.values():
            key = prefix + p.name
            if key not in arrays:
                missing.append(key)
                continue
            t = torch.from_numpy(np.asarray(arrays[key], dtype=np.float32))
            if tuple(t.shape) != tuple(p.tf_shape):
                raise ValueError(f"{key}: checkpoint shape {tuple(t.shape)} != {p.tf_shape}")
            self.view(p.name).copy_(_tf_to_native(p, t))
        if missing and strict:
            raise KeyError("variables missing from checkpoint: " + ", ".join(missing[:5]))
        self.sync_shadow()
        return missing

    def savable_names(self, prefix: str = "v0/") -> List[str]:
        return [prefix + n for n in self.graph.params]


def _trunc_normal(shape, std, gen) -> torch.Tensor:
    t = torch.empty(shape, dtype=torch.float32)
    torch.nn.init.trunc_normal_(t, mean=0.0, std=std, a=-2 * std, b=2 * std, generator=gen)
    return t


def _init_value(p: ParamSpec, gen) -> torch.Tensor:
    kind = p.init[0]
    tf_shape = p.tf_shape
    if kind == "zeros":
        v = torch.zeros(tf_shape)
    elif kind == "ones":
        v = torch.ones(tf_shape)
    elif kind == "const":
        v = torch.full(tf_shape, float(p.init[1]))
    elif kind == "trunc_normal":
        v = _trunc_normal(tf_shape, float(p.init[1]), gen)
    elif kind == "glorot_uniform":
        limit = math.sqrt(6.0 / (p.init[1] + p.init[2]))
        v = (torch.rand(tf_shape, generator=gen) * 2 - 1) * limit
    else:
        raise ValueError(f"unknown initializer {p.init}")
    return _tf_to_native(p, v)


def _tf_to_native(p: ParamSpec, t: torch.Tensor) -> torch.Tensor:
    if p.kind == "conv_kernel":            # HWIO -> OHWI
        v = t.permute(3, 0, 1, 2).contiguous()
        if p.in_map is None and p.out_pad is None:
            return v
        out = torch.zeros(p.shape, dtype=t.dtype)
        if p.in_map is None:
            out[:v.shape[0]] = v
        else:
            out[:v.shape[0], :, :, list(p.in_map)] = v
        return out
    if p.kind == "conv_kernel_flat":       # HWIO -> [O, pad(H*W*I)]
        o = t.shape[3]
        flat = t.permute(3, 0, 1, 2).reshape(o, -1)
        out = torch.zeros(p.shape, dtype=t.dtype)
        out[:o, :flat.shape[1]] = flat
        return out
    if p.kind == "dw_kernel":              # [H, W, C, 1] -> [H, W, Cphys]
        v = t[..., 0]
        if p.in_map is None:
            return v.contiguous()
        out = torch.zeros(p.shape, dtype=t.dtype)
        out[:, :, list(p.in_map)] = v
        return out
    if p.kind == "fc_weight":              # [in, out] -> [out, pad(in)]
        out = torch.zeros(p.shape, dtype=t.dtype)
        if p.in_map is None:
            out[:, :t.shape[0]] = t.t()
        else:
            out[:, list(p.in_map)] = t.t()
        return out
    if p.in_map is not None:               # per-channel vector on a padded layout
        out = torch.zeros(p.shape, dtype=t.dtype)
        out[list(p.in_map)] = t
        return out
    return t.contiguous()


def _native_to_tf(p: ParamSpec, t: torch.Tensor) -> torch.Tensor:
    if p.kind == "conv_kernel":
        k = p.tf_shape[3]
        if p.in_map is not None:
            t = t[:, :, :, list(p.in_map)]
        return t[:k].permute(1, 2, 3, 0).contiguous()
    if p.kind == "conv_kernel_flat":
        r, s, c = p.conv_rsc
        k = p.tf_shape[3]
        return t[:k, :r * s * c].reshape(k, r, s, c).permute(1, 2, 3, 0).contiguous()
    if p.kind == "dw_kernel":
        if p.in_map is not None:
            t = t[:, :, list(p.in_map)]
        return t.unsqueeze(-1).contiguous()
    if p.kind == "fc_weight":
        if p.in_map is not None:
            return t[:, list(p.in_map)].t().contiguous()
        return t[:, :p.tf_shape[0]].t().contiguous()
    if p.in_map is not None:
        return t[list(p.in_map)].contiguous()
    return t.contiguous()
