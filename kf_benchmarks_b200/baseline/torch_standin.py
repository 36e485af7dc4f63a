"""Stand-in baseline: the same models and update rules in plain PyTorch.

This is NOT the product and NOT the reference. The reference (TF 2.5 + KungFu/Horovod) cannot
be installed on this image, so the same-box comparison point is a deliberately conventional
implementation: ``torch.nn.functional`` ops (cuDNN / cuBLAS) in
channels-last fp16 autocast with fp32 master weights, gradient exchange by NCCL
(``torch.distributed.all_reduce`` on a flat gradient buffer), a separate unscale pass and a
separate ``torch.optim.SGD(momentum, nesterov=True)`` step; KungFu SMA as an NCCL all-reduce of
the weights + torch blend; PairAveraging as an all-gather + torch blend. It is timed by the
same harness (CUDA events, max over ranks) so that "a path that only calls NCCL and cuDNN" has
a number next to the fused-kernel path.

The module is generated from the same ``nn.graph.Graph`` the product executes, so the
architecture is identical by construction.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..nn.graph import Graph


class GraphModule(nn.Module):
    """nn.Module executing a Graph with torch ops (NCHW logical shape, channels_last memory)."""

    def __init__(self, graph: Graph, nclass: int):
        super().__init__()
        self.graph = graph
        self.nclass = nclass
        self.params_by_node: Dict[int, Dict[str, nn.Parameter]] = {}
        plist = nn.ParameterList()
        self.buffers_by_node: Dict[int, Dict[str, torch.Tensor]] = {}
        for i, node in enumerate(graph.nodes):
            d = {}
            for key, spec in node.params.items():
                if not spec.trainable:
                    continue
                t = torch.empty(self._torch_shape(spec))
                self._init(spec, t)
                p = nn.Parameter(t)
                plist.append(p)
                d[key] = p
            self.params_by_node[i] = d
            if node.op == "bn":
                c = node.output.shape[-1]
                self.register_buffer("mm%d" % i, torch.zeros(c))
                self.register_buffer("mv%d" % i, torch.ones(c))
        self.plist = plist

    @staticmethod
    def _torch_shape(spec):
        if spec.kind in ("conv_kernel", "conv_kernel_flat"):
            r, s, c, k = spec.tf_shape
            return (k, c, r, s)
        if spec.kind == "fc_weight":
            i, o = spec.tf_shape
            return (o, i)
        return spec.tf_shape

    @staticmethod
    def _init(spec, t):
        kind = spec.init[0]
        with torch.no_grad():
            if kind == "zeros":
                t.zero_()
            elif kind == "ones":
                t.fill_(1.0)
            elif kind == "const":
                t.fill_(spec.init[1])
            elif kind == "trunc_normal":
                nn.init.trunc_normal_(t, std=spec.init[1], a=-2 * spec.init[1], b=2 * spec.init[1])
            else:
                limit = math.sqrt(6.0 / (spec.init[1] + spec.init[2]))
                t.uniform_(-limit, limit)

    def forward(self, images_nhwc: torch.Tensor):
        g = self.graph
        x = images_nhwc.permute(0, 3, 1, 2)  # NCHW view of NHWC memory == channels_last
        vals = {g.input.id: x}
        for i, node in enumerate(g.nodes):
            a = node.attrs
            xs = [vals[t.id] for t in node.inputs]
            p = self.params_by_node[i]
            op = node.op
            if op == "conv":
                ph, pw = a["pad_h"], a["pad_w"]
                inp = xs[0]
                if ph[0] != ph[1] or pw[0] != pw[1]:
                    inp = F.pad(inp, (pw[0], pw[1], ph[0], ph[1]))
                    pad = 0
                else:
                    pad = (ph[0], pw[0])
                y = F.conv2d(inp, p["kernel"], p.get("biases"), stride=(a["sh"], a["sw"]),
                             padding=pad)
                if a["activation"] == "relu":
                    y = F.relu(y)
            elif op == "bn":
                y = F.batch_norm(xs[0], getattr(self, "mm%d" % i), getattr(self, "mv%d" % i),
                                 p.get("gamma"), p["beta"], self.training, 1.0 - a["decay"],
                                 a["epsilon"])
                if len(xs) > 1:
                    y = y + xs[1]
                if a["relu"]:
                    y = F.relu(y)
            elif op == "act":
                y = F.relu(xs[0]) if a["kind"] == "relu" else torch.tanh(xs[0])
            elif op == "add":
                y = xs[0] + xs[1]
                if a["relu"]:
                    y = F.relu(y)
            elif op == "pool":
                ph, pw = a["pad_h"], a["pad_w"]
                k, s = (a["kh"], a["kw"]), (a["sh"], a["sw"])
                inp = xs[0]
                if a["kind"] == "max":
                    if any(ph) or any(pw):
                        inp = F.pad(inp, (pw[0], pw[1], ph[0], ph[1]), value=float("-inf"))
                    y = F.max_pool2d(inp, k, s)
                else:
                    if any(ph) or any(pw):
                        ones = torch.ones_like(inp[:1, :1])
                        num = F.avg_pool2d(F.pad(inp, (pw[0], pw[1], ph[0], ph[1])), k, s,
                                           divisor_override=1)
                        cnt = F.avg_pool2d(F.pad(ones, (pw[0], pw[1], ph[0], ph[1])), k, s,
                                           divisor_override=1)
                        y = num / cnt
                    else:
                        y = F.avg_pool2d(inp, k, s)
            elif op == "spatial_mean":
                y = xs[0].mean((2, 3))
            elif op == "reshape":
                y = xs[0].permute(0, 2, 3, 1).reshape(xs[0].shape[0], -1) if xs[0].dim() == 4 \
                    else xs[0].reshape(xs[0].shape[0], -1)
            elif op == "affine":
                y = F.linear(xs[0], p["weights"], p["biases"])
                if a["activation"] == "relu":
                    y = F.relu(y)
            elif op == "dropout":
                y = F.dropout(xs[0], 1.0 - a["keep_prob"], self.training)
            elif op == "lrn":
                y = F.local_response_norm(xs[0], 2 * a["radius"] + 1,
                                          alpha=a["alpha"] * (2 * a["radius"] + 1),
                                          beta=a["beta"], k=a["bias"])
            elif op == "concat":
                y = torch.cat(xs, dim=1)
            elif op == "pad_channels":
                y = F.pad(xs[0], (0, 0, 0, 0, a["before"], a["after"]))
            else:
                raise NotImplementedError(op)
            vals[node.output.id] = y
        aux = vals[g.aux_logits.id] if g.aux_logits is not None else None
        return vals[g.logits.id], aux


class StandinTrainer:
    """One rank of the stand-in: autocast fp16, static loss scale, NCCL all-reduce, SGD."""

    def __init__(self, params, model, graph, nclass, device, world, rank):
        import torch.distributed as dist
        self.p = params
        self.device = device
        self.world, self.rank = world, rank
        self.dist = dist if world > 1 else None
        torch.backends.cudnn.benchmark = True
        self.net = GraphModule(graph, nclass).to(device).to(memory_format=torch.channels_last)
        self.net.train()
        self.loss_scale = float(params.fp16_loss_scale or model.get_fp16_loss_scale())
        self.model = model
        mom = params.momentum if params.optimizer == "momentum" else 0.0
        self.opt = torch.optim.SGD(self.net.parameters(), lr=0.1, momentum=mom,
                                   nesterov=mom > 0, weight_decay=params.weight_decay or 0.0)
        self.aux_weight = model.aux_loss_weight
        from ..parallel.strategy import reduction_semantics
        self.algo, self.average = reduction_semantics(params)
        if world == 1:
            self.algo = "none"
        if self.dist is not None:
            flat = torch.cat([p.data.reshape(-1) for p in self.net.parameters()])
            self.dist.broadcast(flat, 0)
            off = 0
            for p in self.net.parameters():
                p.data.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        self.last_loss = None

    def step(self, images, labels, lr):
        for g in self.opt.param_groups:
            g["lr"] = lr
        self.opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.float16):
            logits, aux = self.net(images)
        loss = F.cross_entropy(logits.float(), labels.long())
        if aux is not None:
            loss = loss + self.aux_weight * F.cross_entropy(aux.float(), labels.long())
        (loss * self.loss_scale).backward()
        params = [p for p in self.net.parameters() if p.grad is not None]
        if self.algo == "allreduce":
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            self.dist.all_reduce(flat)
            scale = 1.0 / self.loss_scale / (self.world if self.average else 1)
            flat.mul_(scale)                       # separate unscale pass
            off = 0
            for p in params:
                p.grad.copy_(flat[off:off + p.numel()].view_as(p))
                off += p.numel()
        else:
            for p in params:
                p.grad.mul_(1.0 / self.loss_scale)
        if self.algo == "sma":
            flat = torch.cat([p.data.reshape(-1) for p in params])
            avg = flat.clone()
            self.dist.all_reduce(avg)
            avg.div_(self.world)
            off = 0
            for p in params:
                p.data.mul_(1 - self.p.sma_alpha).add_(avg[off:off + p.numel()].view_as(p),
                                                      alpha=self.p.sma_alpha)
                off += p.numel()
        elif self.algo == "pair":
            flat = torch.cat([p.data.reshape(-1) for p in params])
            models = [torch.empty_like(flat) for _ in range(self.world)]
            self.dist.all_gather(models, flat)
            peer = (self.rank + 1 + int(torch.randint(0, self.world - 1, (1,)))) % self.world
            off = 0
            for p in params:
                p.data.mul_(0.5).add_(models[peer][off:off + p.numel()].view_as(p), alpha=0.5)
                off += p.numel()
        self.opt.step()
        self.last_loss = loss.detach()
        return self.last_loss
