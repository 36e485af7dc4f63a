"""Stand-in baseline: the same models and update rules in plain PyTorch.

This is NOT the product and NOT the reference. The reference (TF 2.5 + KungFu/Horovod) cannot
be installed on this image, so the same-box comparison point is a deliberately conventional
implementation, tuned the way such a path usually is so that the
comparison is a fair one: ``torch.nn.functional`` ops (cuDNN / cuBLAS, ``cudnn.benchmark``) on a
channels-last **fp16 model** with fp32 master weights (no per-step autocast weight casts), fp32
batch-norm parameters, gradients written by autograd straight into one flat fp16 buffer that NCCL
all-reduces in place (the reference's compact gradient transfer), the update as a handful of
flat elementwise torch kernels, and the forward + backward pass replayed from a **CUDA graph**
(static shapes). SAME-padded pooling uses ``ceil_mode`` instead of a padded copy. KungFu SMA is an
NCCL all-reduce of the master weights + torch blend; PairAveraging an all-gather + torch blend.
It is timed by the same harness (CUDA events, max over ranks) so that "a path that only calls
NCCL and cuDNN" has a number next to the fused-kernel path.

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
                    if ph[0] == 0 and pw[0] == 0 and ph[1] < a["sh"] and pw[1] < a["sw"]:
                        # padding only after the last window (TF SAME): the clipped last window
                        # of ceil_mode is the same thing without a padded copy of the tensor
                        y = F.max_pool2d(inp, k, s, ceil_mode=bool(ph[1] or pw[1]))
                    else:
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
    """One rank of the stand-in: fp16 model + fp32 master, static loss scale, NCCL all-reduce of
    a flat fp16 gradient buffer, flat torch update kernels, CUDA-graphed forward + backward."""

    def __init__(self, params, model, graph, nclass, device, world, rank, cuda_graph=True,
                 group=None):
        import torch.distributed as dist
        self.p = params
        self.device = device
        self.world, self.rank = world, rank
        self.dist = dist if world > 1 else None
        self.group = group
        torch.backends.cudnn.benchmark = True
        net = GraphModule(graph, nclass).to(device)
        bn_ids = set()
        for i, node in enumerate(graph.nodes):
            if node.op == "bn":
                bn_ids.update(id(p) for p in net.params_by_node[i].values())
        net = net.to(memory_format=torch.channels_last)
        net.train()
        self.net = net
        plist = list(net.parameters())
        self.p16 = [p for p in plist if id(p) not in bn_ids]       # conv / FC: fp16 in the model
        self.p32 = [p for p in plist if id(p) in bn_ids]           # batch norm: fp32
        pad8 = lambda n: (n + 7) // 8 * 8              # noqa: E731  (16-byte aligned slices)
        n16 = sum(pad8(p.numel()) for p in self.p16)
        n32 = sum(pad8(p.numel()) for p in self.p32)
        self.n16 = n16
        # fp32 master of everything; the BN parameters ARE their slice of the master
        self.master = torch.zeros(n16 + n32, device=device)
        self._pad8 = pad8
        self.g16 = torch.zeros(n16, device=device, dtype=torch.float16)
        self.g32 = torch.zeros(max(n32, 1), device=device)
        off = 0
        for p in self.p16:
            n = p.numel()
            self.master[off:off + n].copy_(p.data.reshape(-1))
            p.data = p.data.half()                      # keeps the channels-last strides
            p.grad = torch.zeros_like(p.data)
            off += pad8(n)
        off32 = 0
        for p in self.p32:
            n = p.numel()
            self.master[n16 + off32:n16 + off32 + n].copy_(p.data.reshape(-1))
            p.data = self.master[n16 + off32:n16 + off32 + n].view(p.shape)
            p.grad = self.g32[off32:off32 + n].view(p.shape)
            off32 += pad8(n)
        self.mom = torch.zeros_like(self.master)
        self.loss_scale = float(params.fp16_loss_scale or model.get_fp16_loss_scale())
        self.model = model
        self.momentum = params.momentum if params.optimizer == "momentum" else 0.0
        self.wd = params.weight_decay or 0.0
        self.aux_weight = model.aux_loss_weight
        from ..parallel.strategy import reduction_semantics
        self.algo, self.average = reduction_semantics(params)
        if world == 1:
            self.algo = "none"
        if self.dist is not None:
            self.dist.broadcast(self.master, 0, group=self.group)
            self._emit_weights()
        self.last_loss = None
        self.use_graph = bool(cuda_graph)
        self._graph = None
        self._warm = 0

    # -- helpers --------------------------------------------------------------------------
    def _emit_weights(self):
        """fp16 model weights from the fp32 master."""
        off = 0
        for p in self.p16:
            n = p.numel()
            p.data.copy_(self.master[off:off + n].view(p.shape))
            off += self._pad8(n)

    def _gather_grads(self):
        """Flat fp16 gradient buffer (conv / FC) from the parameters' gradients."""
        off = 0
        for p in self.p16:
            n = p.numel()
            self.g16[off:off + n].copy_(p.grad.reshape(-1))
            off += self._pad8(n)

    def _fwd_bwd(self, images, labels):
        for p in self.p16:
            p.grad.zero_()
        self.g32.zero_()
        logits, aux = self.net(images)
        loss = F.cross_entropy(logits.float(), labels.long())
        if aux is not None:
            loss = loss + self.aux_weight * F.cross_entropy(aux.float(), labels.long())
        (loss * self.loss_scale).backward()
        self._gather_grads()
        return loss.detach()

    def step(self, images, labels, lr):
        images = images.half()
        if self.use_graph:
            if self._graph is None:
                self._static_in = (images.clone(), labels.clone())
                self._warm += 1
                if self._warm <= 3:                     # cudnn.benchmark + lazy init, eagerly
                    loss = self._fwd_bwd(*self._static_in)
                else:
                    torch.cuda.synchronize(self.device)
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._static_loss = self._fwd_bwd(*self._static_in)
                    self._graph = g
                    g.replay()
                    loss = self._static_loss
            else:
                self._static_in[0].copy_(images, non_blocking=True)
                self._static_in[1].copy_(labels, non_blocking=True)
                self._graph.replay()
                loss = self._static_loss
        else:
            loss = self._fwd_bwd(images, labels)
        scale = 1.0 / self.loss_scale
        if self.algo == "allreduce":
            self.dist.all_reduce(self.g16, group=self.group)   # fp16 on the wire, in place
            self.dist.all_reduce(self.g32, group=self.group)
            if self.average:
                scale /= self.world
        if self.algo == "sma":
            avg = self.master.clone()
            self.dist.all_reduce(avg, group=self.group)
            self.master.mul_(1 - self.p.sma_alpha).add_(avg, alpha=self.p.sma_alpha / self.world)
        elif self.algo == "pair":
            models = [torch.empty_like(self.master) for _ in range(self.world)]
            self.dist.all_gather(models, self.master, group=self.group)
            peer = (self.rank + 1 + int(torch.randint(0, self.world - 1, (1,)))) % self.world
            self.master.mul_(0.5).add_(models[peer], alpha=0.5)
        if not hasattr(self, "_lr_t"):
            self._lr_t = torch.zeros((), device=self.device)
            self._scale = scale
        self._lr_t.fill_(lr)
        if self.use_graph and self._graph is not None:
            if getattr(self, "_graph_upd", None) is None:
                self._update()                          # once eagerly (allocations)
                torch.cuda.synchronize(self.device)
                gu = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gu):
                    self._update()
                self._graph_upd = gu
            else:
                self._graph_upd.replay()
        else:
            self._update()
        self.last_loss = loss
        return self.last_loss

    def _update(self):
        """g = scale * grad + wd * w; Nesterov momentum as tf.train.MomentumOptimizer applies
        it; fp16 model weights re-emitted from the master. Flat torch kernels."""
        n16 = self.n16
        g = torch.empty_like(self.master)
        g[:n16] = self.g16
        g[n16:] = self.g32[:self.master.numel() - n16]
        g.mul_(self._scale)
        if self.wd:
            wd = self.wd * (self.world if (self.algo == "allreduce" and not self.average) else 1)
            g.add_(self.master, alpha=wd)
        if self.momentum > 0:
            self.mom.mul_(self.momentum).add_(g)
            g.add_(self.mom, alpha=self.momentum)
        self.master.sub_(g * self._lr_t)
        self._emit_weights()
