"""Channel legalisation: rewrite a ``Graph`` so every NHWC tensor has a channel count that is
a multiple of 8 (16 bytes of fp16 - the granularity of TMA boxes and of the vectorised
elementwise kernels).

The reference leaves odd channel counts to cuDNN (DenseNet's growth rate 12,
T/models/densenet_model.py:27-95; NASNet-mobile's 44 filters,
T/models/nasnet_model.py:39-111; SSD's 4*k / 81*k head widths,
T/models/ssd_model.py:138-221). Here the *physical* tensors get phantom channels that are
identically zero, and every parameter records where its logical channels live
(``ParamSpec.in_map`` / ``out_pad``):

* a convolution's output is padded to ``pad8(K)`` with all-zero filters; the input axis of its
  kernel follows the physical layout of the tensor it reads (zeros on phantom lanes);
* per-channel vectors (bias, gamma, beta, moving statistics) are scattered the same way;
* concat concatenates physical layouts, so a consumer's map is the concatenation of its
  producers' maps; flatten expands the map over the spatial positions.

Phantom lanes stay exactly zero through training: their activations are 0, every gradient
that reaches them is a product with a zero weight or a zero activation, and all the update
rules map (w=0, g=0) to 0. Checkpoints are written and read in the logical (reference)
layout, so nothing of this is visible outside the executor.

``legalize(graph)`` returns ``graph`` itself when nothing needs padding - the zoo's
mainstream models are untouched.
"""
from __future__ import annotations

from dataclasses import replace
from typing import Dict, Optional, Tuple

from .graph import Graph, ParamSpec, TensorRef

ChanMap = Optional[Tuple[int, ...]]  # physical index of each logical channel; None = identity


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def needs_legalization(graph: Graph) -> bool:
    for t in graph.tensors:
        if t is graph.input:
            continue
        if len(t.shape) == 4 and t.shape[-1] % 8 != 0:
            return True
    return False


def _ident(n: int) -> Tuple[int, ...]:
    return tuple(range(n))


def _full_map(m: ChanMap, n: int) -> Tuple[int, ...]:
    return _ident(n) if m is None else m


def _is_plain(m: ChanMap, logical: int, phys: int) -> bool:
    return phys == logical and (m is None or m == _ident(logical))


def _vec(spec: ParamSpec, cmap: ChanMap, phys: int) -> ParamSpec:
    """Per-channel vector living on a tensor with layout (cmap, phys)."""
    logical = spec.tf_shape[0]
    if _is_plain(cmap, logical, phys):
        return spec
    return replace(spec, shape=(phys,), in_map=_full_map(cmap, logical))


def legalize(graph: Graph) -> Graph:
    """Returns a padded copy of ``graph`` (or ``graph`` itself if it is already legal)."""
    if not needs_legalization(graph):
        return graph
    c = graph.input.shape[-1]
    out = Graph(graph.batch_size, input_shape=graph.input.shape, input_dtype=graph.input.dtype)
    out.logical = graph
    out.label_specs = list(graph.label_specs)
    out.loss = dict(graph.loss)
    # old tensor id -> (new tensor, channel map, physical channels)
    tmap: Dict[int, Tuple[TensorRef, ChanMap, int]] = {graph.input.id: (out.input, None, c)}
    for t in graph.extra_inputs:
        tmap[t.id] = (out.add_input(t.shape, t.dtype), None, t.shape[-1])

    def relayout(src, logical, want_map, want_phys):
        """Brings operand ``src`` = (tensor, map, phys) to the layout (want_map, want_phys)."""
        t, m, p = src
        if (m, p) == (want_map, want_phys):
            return src
        index = [-1] * want_phys
        for frm, to in zip(_full_map(m, logical), _full_map(want_map, logical)):
            index[to] = frm
        inverse = [-1] * p
        for dst, frm in enumerate(index):
            if frm >= 0:
                inverse[frm] = dst
        node = out.add_node("relayout", "relayout%d" % len(out.nodes), [t],
                            t.shape[:-1] + (want_phys,),
                            dict(index=tuple(index), inverse=tuple(inverse)))
        return (node.output, want_map, want_phys)

    for node in graph.nodes:
        ins = [tmap[t.id] for t in node.inputs]
        new_inputs = [i[0] for i in ins]
        a = dict(node.attrs)
        params = {}
        op = node.op
        t_out = node.output
        x, xm, xp = ins[0]
        logical_c = node.inputs[0].shape[-1]

        if op == "conv":
            k = t_out.shape[-1]
            kp = _pad8(k)
            spec = node.params["kernel"]
            r, s = a["kh"], a["kw"]
            if node.inputs[0] is graph.input and xp % 8 != 0:
                # tiny-C stem: explicit-im2col operand, only the filter count is padded
                if kp != k:
                    spec = replace(spec, shape=(kp, spec.shape[1]), out_pad=kp)
            else:
                plain_in = _is_plain(xm, logical_c, xp)
                if spec.kind != "conv_kernel" or not plain_in or kp != k:
                    spec = replace(spec, kind="conv_kernel", shape=(kp, r, s, xp),
                                   logical_cols=None, conv_rsc=None,
                                   in_map=None if plain_in else _full_map(xm, logical_c),
                                   out_pad=kp if kp != k else None)
            params["kernel"] = spec
            if "biases" in node.params:
                params["biases"] = _vec(node.params["biases"], None, kp)
            shape, omap, ophys = t_out.shape[:-1] + (kp,), None, kp
        elif op == "dwconv":
            params["kernel"] = replace(node.params["kernel"],
                                       shape=node.params["kernel"].shape[:-1] + (xp,),
                                       in_map=None if _is_plain(xm, logical_c, xp)
                                       else _full_map(xm, logical_c))
            shape, omap, ophys = t_out.shape[:-1] + (xp,), xm, xp
        elif op == "bn":
            for key, spec in node.params.items():
                params[key] = _vec(spec, xm, xp)
            for i in range(1, len(ins)):
                ins[i] = relayout(ins[i], logical_c, xm, xp)
            new_inputs = [i[0] for i in ins]
            shape, omap, ophys = t_out.shape[:-1] + (xp,), xm, xp
        elif op in ("act", "pool", "dropout", "spatial_mean", "drop_path"):
            shape, omap, ophys = t_out.shape[:-1] + (xp,), xm, xp
        elif op == "lrn":
            if not _is_plain(xm, logical_c, xp):
                raise NotImplementedError("lrn over a padded channel layout")
            shape, omap, ophys = t_out.shape, xm, xp
        elif op == "add":
            for i in range(1, len(ins)):
                ins[i] = relayout(ins[i], logical_c, xm, xp)
            new_inputs = [i[0] for i in ins]
            shape, omap, ophys = t_out.shape[:-1] + (xp,), xm, xp
        elif op == "embedding":
            shape, omap, ophys = t_out.shape, None, t_out.shape[-1]
        elif op == "mul":
            for i in range(1, len(ins)):
                ins[i] = relayout(ins[i], logical_c, xm, xp)
            new_inputs = [i[0] for i in ins]
            shape, omap, ophys = t_out.shape[:-1] + (xp,), xm, xp
        elif op == "concat":
            full, off = [], 0
            for (nt, m, p), t in zip(ins, node.inputs):
                full.extend(off + i for i in _full_map(m, t.shape[-1]))
                off += p
            ophys = off
            omap = tuple(full)
            if _is_plain(omap, t_out.shape[-1], ophys):
                omap = None
            shape = t_out.shape[:-1] + (ophys,)
        elif op == "pad_channels":
            bp, ap = _pad8(a["before"]), _pad8(a["after"])
            full = list(range(a["before"]))
            full += [bp + i for i in _full_map(xm, logical_c)]
            full += [bp + xp + i for i in range(a["after"])]
            ophys = bp + xp + ap
            omap = tuple(full)
            if _is_plain(omap, t_out.shape[-1], ophys):
                omap = None
            a["before"], a["after"] = bp, ap
            shape = t_out.shape[:-1] + (ophys,)
        elif op == "reshape":
            if len(node.inputs[0].shape) == 4:
                n_, hh, ww, _ = node.inputs[0].shape
                ophys = hh * ww * xp
                if _is_plain(xm, logical_c, xp):
                    omap = None
                else:
                    cm = _full_map(xm, logical_c)
                    omap = tuple(pos * xp + i for pos in range(hh * ww) for i in cm)
                shape = (n_, ophys)
            else:
                shape, omap, ophys = t_out.shape, xm, xp
        elif op == "affine":
            spec = node.params["weights"]
            n_out = t_out.shape[-1]
            if not _is_plain(xm, logical_c, xp):
                spec = replace(spec, shape=(n_out, _pad8(xp)), logical_cols=xp,
                               in_map=_full_map(xm, logical_c))
            params["weights"] = spec
            params["biases"] = node.params["biases"]
            shape, omap, ophys = t_out.shape, None, n_out
        else:
            raise NotImplementedError(f"legalize: unknown op {op}")

        for key, spec in node.params.items():
            params.setdefault(key, spec)
        new = out.add_node(op, node.name, new_inputs, shape, a,
                           {k: params[k] for k in node.params})
        tmap[t_out.id] = (new.output, omap, ophys)

    if graph.logits is not None:
        out.logits = tmap[graph.logits.id][0]
    if graph.aux_logits is not None:
        out.aux_logits = tmap[graph.aux_logits.id][0]
    for attr in ("extra",):
        if hasattr(graph, attr):
            setattr(out, attr, getattr(graph, attr))
    for key in ("loc", "conf"):     # loss heads that name tensors (SSD)
        if key in out.loss:
            out.loss[key] = [tmap[tid][0].id for tid in graph.loss[key]]
    out.tensor_layout = {tid: (v[0].id, v[1], v[2]) for tid, v in tmap.items()}
    return out
