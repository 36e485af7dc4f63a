"""Reference executor: runs a ``Graph`` with plain PyTorch ops and autograd.

This is (a) the ``--device=cpu`` path of the benchmark (config #1 of BASELINE.json:
trivial model, NHWC, CPU replicas), and (b) the numerical oracle for the CUDA
executor: same graph, same ParameterStore layout, fp32 math. It is deliberately
simple; nothing here is on the B200 hot path.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch
import torch.nn.functional as F

from .graph import Graph, Node
from .params import ParameterStore


class _RoundFp16(torch.autograd.Function):
    """fp16 storage emulation: values and their gradients are rounded to half precision, the
    math in between stays fp32 - the numerics of the CUDA executor (fp16 activations and
    activation gradients, fp32 accumulation)."""

    @staticmethod
    def forward(ctx, x):
        return x.half().float()

    @staticmethod
    def backward(ctx, g):
        return g.half().float()


class RefExecutor:
    def __init__(self, graph: Graph, store: ParameterStore, nclass: int, aux_weight: float = 0.4,
                 dtype=torch.float32, dropout_seed: int = 4321, emulate_fp16: bool = False):
        self.graph = graph
        self.store = store
        self.nclass = nclass
        self.aux_weight = aux_weight
        self.dtype = dtype
        self.emulate_fp16 = emulate_fp16
        self.device = store.device
        self.images = None
        self.labels = None
        self._leaves: Dict[str, torch.Tensor] = {}
        self._vals: Dict[int, torch.Tensor] = {}
        self._loss = None
        self._gen = torch.Generator(device="cpu").manual_seed(dropout_seed)
        self.last = (float("nan"), 0, 0)
        self.step_index = 0

    # -- inputs ---------------------------------------------------------------
    def set_inputs(self, *tensors: torch.Tensor) -> None:
        """(primary input, *extra inputs, *labels) in the order the input source yields them."""
        g = self.graph
        ne = len(g.extra_inputs)

        def conv(t, dtype):
            if dtype == "i32":
                return t.to(self.device, torch.long)
            return t.to(self.device, self.dtype)
        self.images = conv(tensors[0], g.input.dtype)
        self.extras = [conv(t, ref.dtype) for t, ref in zip(tensors[1:1 + ne], g.extra_inputs)]
        labs = tensors[1 + ne:]
        self.label_tensors = [conv(t, spec[1]) for t, spec in zip(labs, g.label_specs)]
        self.labels = self.label_tensors[0]

    def set_primary_input(self, images: torch.Tensor) -> None:
        self.images = images.to(self.device, self.dtype if self.graph.input.dtype != "i32"
                                else torch.long)

    def _param(self, spec, train: bool) -> torch.Tensor:
        """Leaf tensor in the *TF* layout so gradients map back trivially."""
        v = self.store.view(spec.name).detach().to(self.dtype)
        if spec.kind == "conv_kernel":
            t = v.permute(0, 3, 1, 2)                       # OHWI -> OIHW view
        elif spec.kind == "conv_kernel_flat":
            r, s, c = spec.conv_rsc
            t = v[:, :r * s * c].reshape(v.shape[0], r, s, c).permute(0, 3, 1, 2)
        elif spec.kind == "fc_weight":
            t = v[:, :spec.logical_cols]
        elif spec.kind == "dw_kernel":
            t = v.permute(2, 0, 1).unsqueeze(1)             # [R,S,C] -> [C,1,R,S] view
        else:
            t = v
        t = t.clone()
        if train and spec.trainable:
            t.requires_grad_(True)
        self._leaves[spec.name] = t
        return t

    # -- forward ----------------------------------------------------------------
    def forward(self, train: bool = True) -> torch.Tensor:
        g = self.graph
        self._leaves.clear()
        vals = {g.input.id: self.images}
        for ref, t in zip(g.extra_inputs, getattr(self, "extras", [])):
            vals[ref.id] = t
        for node in g.nodes:
            y = self._run(node, [vals[t.id] for t in node.inputs], train)
            if self.emulate_fp16 and node.output is not g.logits and \
                    node.output is not g.aux_logits:
                y = _RoundFp16.apply(y)
            vals[node.output.id] = y
        self._vals = vals
        return vals[g.logits.id] if g.logits is not None else None

    def _run(self, node: Node, xs, train: bool) -> torch.Tensor:
        a = node.attrs
        op = node.op
        if op == "conv":
            x = xs[0].permute(0, 3, 1, 2)
            w = self._param(node.params["kernel"], train)
            ph, pw = a["pad_h"], a["pad_w"]
            x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
            b = self._param(node.params["biases"], train) if a["has_bias"] else None
            y = F.conv2d(x, w, b, stride=(a["sh"], a["sw"]))
            y = _act(y, a["activation"])
            return y.permute(0, 2, 3, 1)
        if op == "embedding":
            name = a["table"]
            if name not in self._leaves:
                self._param(self.graph.params[name], train)
            table = self._leaves[name]
            lo = a["col_lo"]
            return table[xs[0].clamp(0, a["num_rows"] - 1)][:, lo:lo + node.output.shape[-1]]
        if op == "mul":
            return xs[0] * xs[1]
        if op == "dwconv":
            x = xs[0].permute(0, 3, 1, 2)
            w = self._param(node.params["kernel"], train)
            ph, pw = a["pad_h"], a["pad_w"]
            x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]))
            y = F.conv2d(x, w, None, stride=(a["sh"], a["sw"]), groups=x.shape[1])
            return y.permute(0, 2, 3, 1)
        if op == "drop_path":
            if not train:
                return xs[0]
            kp = float(a["keep_prob_fn"](self.step_index))
            n = xs[0].shape[0]
            keep = (torch.rand((n,), generator=self._gen) < kp).to(xs[0])
            return xs[0] * (keep / kp).view(n, *([1] * (xs[0].dim() - 1)))
        if op == "bn":
            x = xs[0]
            beta = self._param(node.params["beta"], train)
            gamma = self._param(node.params["gamma"], train) if a["scale"] else None
            mm = self.store.view(node.params["moving_mean"].name)
            mv = self.store.view(node.params["moving_variance"].name)
            red = tuple(range(x.dim() - 1))
            if train:
                mean = x.mean(red)
                var = x.var(red, unbiased=False)
                cnt = x.numel() // x.shape[-1]
                with torch.no_grad():
                    d = a["decay"]
                    unb = var * (cnt / max(cnt - 1, 1))
                    mm.mul_(d).add_(mean.detach().to(mm.dtype) * (1 - d))
                    mv.mul_(d).add_(unb.detach().to(mv.dtype) * (1 - d))
            else:
                mean, var = mm.to(self.dtype), mv.to(self.dtype)
            y = (x - mean) * torch.rsqrt(var + a["epsilon"])
            if gamma is not None:
                y = y * gamma
            y = y + beta
            if len(xs) > 1:
                y = y + xs[1]
            return F.relu(y) if a["relu"] else y
        if op == "act":
            return _act(xs[0], a["kind"])
        if op == "add":
            y = xs[0] + xs[1]
            return F.relu(y) if a["relu"] else y
        if op == "pool":
            x = xs[0].permute(0, 3, 1, 2)
            ph, pw = a["pad_h"], a["pad_w"]
            k, s = (a["kh"], a["kw"]), (a["sh"], a["sw"])
            if a["kind"] == "max":
                x = F.pad(x, (pw[0], pw[1], ph[0], ph[1]), value=float("-inf"))
                y = F.max_pool2d(x, k, s)
            else:
                # TF 'SAME' average pooling divides by the number of in-bounds taps
                ones = torch.ones_like(x[:1, :1])
                xs_ = F.avg_pool2d(F.pad(x, (pw[0], pw[1], ph[0], ph[1])), k, s,
                                   divisor_override=1)
                cnt = F.avg_pool2d(F.pad(ones, (pw[0], pw[1], ph[0], ph[1])), k, s,
                                   divisor_override=1)
                y = xs_ / cnt.clamp(min=1.0)   # windows entirely in the padding give 0
            return y.permute(0, 2, 3, 1)
        if op == "spatial_mean":
            return xs[0].mean((1, 2))
        if op == "reshape":
            return xs[0].reshape(node.output.shape)
        if op == "lstm":
            return self._lstm(node, xs[0], train)
        if op == "affine":
            w = self._param(node.params["weights"], train)
            b = self._param(node.params["biases"], train)
            return _act(F.linear(xs[0], w, b), a["activation"])
        if op == "dropout":
            kp = a["keep_prob"]
            if not train or kp >= 1.0:
                return xs[0]
            mask = (torch.rand(xs[0].shape, generator=self._gen) < kp).to(xs[0])
            return xs[0] * mask / kp
        if op == "lrn":
            x = xs[0]
            sq = F.pad((x * x).unsqueeze(1), (a["radius"], a["radius"]))
            win = sq.unfold(-1, 2 * a["radius"] + 1, 1).sum(-1).squeeze(1)
            return x / (a["bias"] + a["alpha"] * win) ** a["beta"]
        if op == "concat":
            return torch.cat(xs, dim=-1)
        if op == "relayout":
            idx = torch.tensor(a["index"], dtype=torch.long, device=xs[0].device)
            y = xs[0][..., idx.clamp(min=0)]
            return y * (idx >= 0).to(y)
        if op == "pad_channels":
            return F.pad(xs[0], (a["before"], a["after"]))
        raise NotImplementedError(op)

    def _lstm(self, node: Node, x: torch.Tensor, train: bool) -> torch.Tensor:
        """BasicLSTMCell recurrences, mirroring the CUDA executor's storage precision when
        fp16 emulation is on (projections and hidden states are kept in half)."""
        a = node.attrs
        H = a["hidden"]
        rnd = _RoundFp16.apply if self.emulate_fp16 else (lambda v: v)
        n, T, _ = x.shape
        outs = []
        for d in a["dirs"]:
            wx = self._param(node.params[d + "_wx"], train)
            wh = self._param(node.params[d + "_wh"], train)
            b = self._param(node.params[d + "_b"], train)
            gx = rnd(x @ wx.t())
            h = torch.zeros(n, H, dtype=x.dtype, device=x.device)
            c = torch.zeros(n, H, dtype=x.dtype, device=x.device)
            hs = [None] * T
            for t in (range(T) if d == "fw" else range(T - 1, -1, -1)):
                g = gx[:, t] + rnd(h @ wh.t()) + b
                gi, gj, gf, go = g.split(H, dim=1)
                c = c * torch.sigmoid(gf + a["forget_bias"]) + torch.sigmoid(gi) * torch.tanh(gj)
                h = rnd(torch.tanh(c) * torch.sigmoid(go))
                hs[t] = h
            outs.append(torch.stack(hs, dim=1))
        return torch.cat(outs, dim=-1)

    # -- loss / backward ------------------------------------------------------------
    def loss_backward(self, loss_scale: float = 1.0, backward: bool = True,
                      on_grads_ready: Optional[Callable[[int, int], None]] = None,
                      dyn_scale=None) -> None:
        if dyn_scale is not None:       # device-resident loss scale (CUDA executor only)
            loss_scale = loss_scale * float(dyn_scale)
        g = self.graph
        logits = self._vals[g.logits.id].float() if g.logits is not None else None
        kind = g.loss["kind"]
        if kind == "softmax_xent":
            loss = F.cross_entropy(logits, self.labels)
            if g.aux_logits is not None:
                loss = loss + self.aux_weight * F.cross_entropy(
                    self._vals[g.aux_logits.id].float(), self.labels)
            with torch.no_grad():
                top5 = logits.topk(min(5, logits.shape[1]), dim=1).indices
                t1 = int((top5[:, 0] == self.labels).sum())
                t5 = int((top5 == self.labels[:, None]).any(1).sum())
        elif kind == "ncf":
            # softmax over [1, z]: T/models/experimental/official_ncf_model.py:85-97
            two = torch.cat([torch.ones_like(logits), logits], dim=1)
            loss = F.cross_entropy(two, (self.labels != 0).long())
            with torch.no_grad():
                t1 = int(((logits[:, 0] > 1.0) == (self.labels != 0)).sum())
                t5 = t1
        else:
            from . import losses_ref
            loss, t1, t5 = losses_ref.compute(kind, g, self._vals, self.label_tensors)
        self.last = (float(loss.detach()), t1, t5)
        if not backward:
            return
        (loss * loss_scale).backward()
        self.store.grad.zero_()
        for name, leaf in self._leaves.items():
            spec = self.graph.params[name]
            if not spec.trainable or leaf.grad is None:
                continue
            gv = self.store.view(name, "grad")
            gr = leaf.grad.to(gv.dtype)
            if spec.kind == "conv_kernel":
                gv.copy_(gr.permute(0, 2, 3, 1))
            elif spec.kind == "conv_kernel_flat":
                r, s, c = spec.conv_rsc
                gv[:, :r * s * c] = gr.permute(0, 2, 3, 1).reshape(gr.shape[0], -1)
            elif spec.kind == "fc_weight":
                gv[:, :spec.logical_cols] = gr
            elif spec.kind == "dw_kernel":
                gv.copy_(gr.squeeze(1).permute(1, 2, 0))
            else:
                gv.copy_(gr)
        if on_grads_ready is not None:
            on_grads_ready(0, self.store.num_trainable_padded)
        self.step_index += 1

    def read_loss(self):
        return self.last

    def logits(self) -> torch.Tensor:
        return self._vals[self.graph.logits.id]

    def value(self, tensor_id: int) -> torch.Tensor:
        return self._vals[tensor_id]


def _act(y, kind):
    if kind == "relu":
        return F.relu(y)
    if kind == "relu6":
        return F.relu6(y)
    if kind == "tanh":
        return torch.tanh(y)
    if kind == "sigmoid":
        return torch.sigmoid(y)
    return y
