"""NetBuilder: the layer-builder API the model zoo is written against.

It offers the same calls and bookkeeping as the reference's ``ConvNetBuilder``
(T/convnet_builder_legacy.py:32-468): a running ``top_layer`` / ``top_size``,
per-type name counters that are global to the network (``conv0, conv1, ...``,
``affine0``, ``batchnorm0`` ...), the ``SAME_RESNET`` padding mode, inception
columns with ``share``, and an auxiliary-head context. Instead of emitting TF ops
it records nodes of ``nn.graph.Graph`` and fuses as it goes:

* conv + bias + ReLU                -> one ``conv`` node (GEMM epilogue)
* batch-norm + ReLU                 -> one ``bn`` node
* batch-norm + residual add + ReLU  -> one ``bn`` node with a second input

Variable names reproduce the reference's scopes so checkpoints can be exported
in its layout.
"""
from __future__ import annotations

import contextlib
import math
from collections import defaultdict
from typing import List, Optional, Sequence, Tuple

from .graph import Graph, ParamSpec, TensorRef


def _same_pad(in_size: int, k: int, stride: int) -> Tuple[int, int]:
    out = -(-in_size // stride)
    total = max((out - 1) * stride + k - in_size, 0)
    return total // 2, total - total // 2


def resolve_padding(mode: str, in_size: int, k: int, stride: int) -> Tuple[int, int]:
    mode = mode.upper()
    if mode == "VALID":
        return 0, 0
    if mode == "SAME":
        return _same_pad(in_size, k, stride)
    if mode == "SAME_RESNET":
        if stride == 1:
            return _same_pad(in_size, k, 1)
        lo = (k - 1) // 2
        return lo, k - 1 - lo
    raise ValueError(f"unknown padding mode {mode!r}")


def out_size(in_size: int, k: int, stride: int, pad: Tuple[int, int]) -> int:
    return (in_size + pad[0] + pad[1] - k) // stride + 1


class NetBuilder:
    def __init__(self, graph: Graph, phase_train: bool = True):
        self.graph = graph
        self.phase_train = phase_train
        self.top_layer: TensorRef = graph.input
        self.top_size: int = graph.input.shape[-1]
        self.counts = defaultdict(int)
        self.use_batch_norm = False
        self.batch_norm_config = {}  # decay / epsilon / scale overrides
        self.aux_top_layer: Optional[TensorRef] = None
        self.aux_top_size = 0
        self._scopes: List[str] = ["cg"]
        self.data_format = "NHWC"
        self.channel_pos = "channels_last"

    # -- naming ---------------------------------------------------------------
    @contextlib.contextmanager
    def scope(self, name: str):
        self._scopes.append(name)
        try:
            yield
        finally:
            self._scopes.pop()

    def _full(self, leaf: str) -> str:
        return "/".join(self._scopes + [leaf])

    @contextlib.contextmanager
    def switch_to_aux_top_layer(self):
        """T/convnet_builder_legacy.py:97-110."""
        if self.aux_top_layer is None:
            raise RuntimeError("Empty auxiliary top layer in the network.")
        saved = (self.top_layer, self.top_size)
        self.top_layer, self.top_size = self.aux_top_layer, self.aux_top_size
        yield
        self.aux_top_layer, self.aux_top_size = self.top_layer, self.top_size
        self.top_layer, self.top_size = saved

    # -- layers ---------------------------------------------------------------
    def conv(self, num_out_channels, k_height, k_width, d_height=1, d_width=1, mode="SAME",
             input_layer=None, num_channels_in=None, use_batch_norm=None, stddev=None,
             activation="relu", bias=0.0, kernel_initializer=None, scope_name=None,
             kernel_leaf="conv2d/kernel", bias_leaf="biases", bn_scope_name=None,
             explicit_pad=None) -> TensorRef:
        """T/convnet_builder_legacy.py:143-226. ``scope_name`` / ``*_leaf`` override the
        variable names for the slim-style models (MobilenetV2/Conv/weights ...); activation
        "relu6" is accepted for those."""
        if input_layer is None:
            input_layer = self.top_layer
        if num_channels_in is None:
            num_channels_in = self.top_size
        if len(input_layer.shape) != 4:
            raise ValueError("conv expects an NHWC input")
        n, h, w, c = input_layer.shape
        if c != num_channels_in:
            raise ValueError(f"conv: input has {c} channels, num_channels_in={num_channels_in}")
        if scope_name is None:
            name = "conv" + str(self.counts["conv"])
            self.counts["conv"] += 1
        else:
            name = scope_name
        if use_batch_norm is None:
            use_batch_norm = self.use_batch_norm
        if explicit_pad is not None:     # tf.pad + VALID (DeepSpeech2's conv layers)
            pad_h = (int(explicit_pad[0]), int(explicit_pad[0]))
            pad_w = (int(explicit_pad[1]), int(explicit_pad[1]))
        else:
            pad_h = resolve_padding(mode, h, k_height, d_height)
            pad_w = resolve_padding(mode, w, k_width, d_width)
        p = out_size(h, k_height, d_height, pad_h)
        q = out_size(w, k_width, d_width, pad_w)
        fan_in = k_height * k_width * c
        fan_out = k_height * k_width * num_out_channels
        if kernel_initializer is not None:
            init = kernel_initializer
        elif stddev is not None:
            init = ("trunc_normal", float(stddev))
        else:
            init = ("glorot_uniform", fan_in, fan_out)
        with self.scope(name):
            params = {}
            tf_shape = (k_height, k_width, c, num_out_channels)
            if c % 8 == 0:
                params["kernel"] = ParamSpec(self._full(kernel_leaf),
                                             (num_out_channels, k_height, k_width, c), tf_shape,
                                             "conv_kernel", True, init)
            else:
                # tiny-C stem: filters stored as rows of an explicit-im2col GEMM, K padded to 8
                kpad = (fan_in + 7) // 8 * 8
                params["kernel"] = ParamSpec(self._full(kernel_leaf),
                                             (num_out_channels, kpad), tf_shape,
                                             "conv_kernel_flat", True, init, logical_cols=fan_in,
                                             conv_rsc=(k_height, k_width, c))
            has_bias = (not use_batch_norm) and bias is not None
            if has_bias:
                params["biases"] = ParamSpec(self._full(bias_leaf), (num_out_channels,),
                                             (num_out_channels,), "bias", True,
                                             ("const", float(bias)))
            if activation not in ("relu", "relu6", "linear", "tanh", None):
                raise KeyError("Invalid activation type '%s'" % activation)
            act = "linear" if activation is None else activation
            fused_act = act if act in ("relu", "linear") else "linear"
            node = self.graph.add_node(
                "conv", self._full(""), [input_layer], (n, p, q, num_out_channels),
                dict(kh=k_height, kw=k_width, sh=d_height, sw=d_width, pad_h=pad_h, pad_w=pad_w,
                     activation="linear" if use_batch_norm else fused_act,
                     has_bias=has_bias),
                params)
            self.top_layer, self.top_size = node.output, num_out_channels
            if use_batch_norm:
                self.batch_norm(scope_name=bn_scope_name, **self.batch_norm_config)
                if act == "relu":
                    self.relu()
            if act == "tanh":
                self.tanh()
            elif act == "relu6":
                self.relu6()
        return self.top_layer

    def depthwise_conv(self, k_height, k_width, d_height=1, d_width=1, mode="SAME",
                       input_layer=None, stddev=None, scope_name=None,
                       kernel_leaf="depthwise_weights", use_batch_norm=None, activation=None,
                       bn_scope_name=None) -> TensorRef:
        """Depthwise convolution with channel multiplier 1: the depthwise half of
        slim.separable_conv2d (T/models/mobilenet_conv_blocks.py:246-256,
        T/models/nasnet_utils.py:212-260). No bias; optional BN + activation after it."""
        x = self.top_layer if input_layer is None else input_layer
        n, h, w, c = x.shape
        if scope_name is None:
            scope_name = "depthwise" + str(self.counts["depthwise"])
            self.counts["depthwise"] += 1
        pad_h = resolve_padding(mode, h, k_height, d_height)
        pad_w = resolve_padding(mode, w, k_width, d_width)
        p = out_size(h, k_height, d_height, pad_h)
        q = out_size(w, k_width, d_width, pad_w)
        fan = k_height * k_width
        init = (("trunc_normal", float(stddev)) if stddev is not None
                else ("glorot_uniform", fan * c, fan * c))
        if use_batch_norm is None:
            use_batch_norm = self.use_batch_norm
        with self.scope(scope_name):
            spec = ParamSpec(self._full(kernel_leaf), (k_height, k_width, c),
                             (k_height, k_width, c, 1), "dw_kernel", True, init)
            node = self.graph.add_node(
                "dwconv", self._full(""), [x], (n, p, q, c),
                dict(kh=k_height, kw=k_width, sh=d_height, sw=d_width, pad_h=pad_h, pad_w=pad_w),
                {"kernel": spec})
            self.top_layer, self.top_size = node.output, c
            if use_batch_norm:
                self.batch_norm(scope_name=bn_scope_name, **self.batch_norm_config)
            if activation == "relu":
                self.relu()
            elif activation == "relu6":
                self.relu6()
        return self.top_layer

    def batch_norm(self, input_layer=None, decay=0.999, scale=False, epsilon=0.001,
                   scope_name=None) -> TensorRef:
        """T/convnet_builder_legacy.py:433-459 (defaults decay .999, eps 1e-3, no scale)."""
        if input_layer is None:
            input_layer = self.top_layer
        else:
            self.top_size = None
        if scope_name is None:
            name = "batchnorm" + str(self.counts["batchnorm"])
            self.counts["batchnorm"] += 1
        else:
            name = scope_name
        c = input_layer.shape[-1]
        with self.scope(name):
            params = {"beta": ParamSpec(self._full("beta"), (c,), (c,), "beta", True, ("zeros",))}
            if scale:
                params["gamma"] = ParamSpec(self._full("gamma"), (c,), (c,), "gamma", True,
                                            ("ones",))
            params["moving_mean"] = ParamSpec(self._full("moving_mean"), (c,), (c,),
                                              "moving_mean", False, ("zeros",))
            params["moving_variance"] = ParamSpec(self._full("moving_variance"), (c,), (c,),
                                                  "moving_variance", False, ("ones",))
            node = self.graph.add_node(
                "bn", self._full(""), [input_layer], input_layer.shape,
                dict(decay=float(decay), epsilon=float(epsilon), scale=bool(scale), relu=False),
                params)
        self.top_layer = node.output
        if self.top_size is None:
            self.top_size = c
        return self.top_layer

    def relu(self, input_layer=None, fold=True) -> TensorRef:
        """ReLU on the top layer; folded into a preceding batch-norm when possible. Pass
        ``fold=False`` when the pre-activation value has other (later) readers - NASNet's hidden
        states feed pooling branches raw and separable branches through a ReLU."""
        x = self.top_layer if input_layer is None else input_layer
        prod = x.producer
        if (fold and prod is not None and prod.op == "bn" and not prod.attrs["relu"] and
                not x.consumers and x is not self.graph.logits):
            prod.attrs["relu"] = True
            self.top_layer = x
            return x
        node = self.graph.add_node("act", "relu%d" % self.counts["relu"], [x], x.shape,
                                   dict(kind="relu"))
        self.counts["relu"] += 1
        self.top_layer = node.output
        return self.top_layer

    def relu6(self, input_layer=None) -> TensorRef:
        """min(max(x, 0), 6) - the activation of the MobileNet family."""
        x = self.top_layer if input_layer is None else input_layer
        node = self.graph.add_node("act", "relu6_%d" % self.counts["relu6"], [x], x.shape,
                                   dict(kind="relu6"))
        self.counts["relu6"] += 1
        self.top_layer = node.output
        return self.top_layer

    def drop_path(self, keep_prob_fn, input_layer=None) -> TensorRef:
        """Per-sample stochastic depth (T/models/nasnet_utils.py:113-138, 430-458).
        ``keep_prob_fn(step) -> float`` gives the scheduled keep probability; identity when
        not training."""
        x = self.top_layer if input_layer is None else input_layer
        if not self.phase_train:
            return x
        node = self.graph.add_node("drop_path", "drop_path%d" % self.counts["drop_path"], [x],
                                   x.shape, dict(keep_prob_fn=keep_prob_fn))
        self.counts["drop_path"] += 1
        self.top_layer = node.output
        return self.top_layer

    def tanh(self, input_layer=None) -> TensorRef:
        x = self.top_layer if input_layer is None else input_layer
        node = self.graph.add_node("act", "tanh%d" % self.counts["tanh"], [x], x.shape,
                                   dict(kind="tanh"))
        self.counts["tanh"] += 1
        self.top_layer = node.output
        return self.top_layer

    def add(self, shortcut: TensorRef, res: TensorRef, relu: bool = False,
            fold: bool = True) -> TensorRef:
        """``act(shortcut + res)``: the residual join of the ResNet blocks
        (T/models/resnet_model_legacy.py:77). Folded into the batch-norm that produced ``res``
        when that is legal."""
        if shortcut.shape != res.shape:
            raise ValueError(f"add: shapes differ {shortcut.shape} vs {res.shape}")
        prod = res.producer
        if (fold and prod is not None and prod.op == "bn" and not prod.attrs["relu"] and
                len(prod.inputs) == 1 and not res.consumers):
            prod.inputs.append(shortcut)
            shortcut.consumers.append(prod)
            prod.attrs["relu"] = bool(relu)
            self.top_layer = res
        else:
            node = self.graph.add_node("add", "add%d" % self.counts["add"], [shortcut, res],
                                       res.shape, dict(relu=bool(relu)))
            self.counts["add"] += 1
            self.top_layer = node.output
        self.top_size = res.shape[-1]
        return self.top_layer

    def _pool(self, kind, k_height, k_width, d_height, d_width, mode, input_layer,
              num_channels_in) -> TensorRef:
        """T/convnet_builder_legacy.py:228-286."""
        if input_layer is None:
            input_layer = self.top_layer
        else:
            self.top_size = num_channels_in
        name = kind + str(self.counts[kind])
        self.counts[kind] += 1
        n, h, w, c = input_layer.shape
        pad_h = resolve_padding(mode, h, k_height, d_height)
        pad_w = resolve_padding(mode, w, k_width, d_width)
        p = out_size(h, k_height, d_height, pad_h)
        q = out_size(w, k_width, d_width, pad_w)
        node = self.graph.add_node(
            "pool", name, [input_layer], (n, p, q, c),
            dict(kind="max" if kind == "mpool" else "avg", kh=k_height, kw=k_width, sh=d_height,
                 sw=d_width, pad_h=pad_h, pad_w=pad_w))
        self.top_layer = node.output
        if self.top_size is None:
            self.top_size = c
        return self.top_layer

    def mpool(self, k_height, k_width, d_height=2, d_width=2, mode="VALID", input_layer=None,
              num_channels_in=None) -> TensorRef:
        return self._pool("mpool", k_height, k_width, d_height, d_width, mode, input_layer,
                          num_channels_in)

    def apool(self, k_height, k_width, d_height=2, d_width=2, mode="VALID", input_layer=None,
              num_channels_in=None) -> TensorRef:
        return self._pool("apool", k_height, k_width, d_height, d_width, mode, input_layer,
                          num_channels_in)

    def subsample(self, stride: int, shift: int = 0, input_layer=None) -> TensorRef:
        """Keeps every ``stride``-th pixel starting at ``shift`` (zeros past the edge): the
        1x1 VALID average pool of NASNet's factorised reduction, whose second path reads the
        input shifted by one pixel (T/models/nasnet_utils.py:87-108)."""
        x = self.top_layer if input_layer is None else input_layer
        n, h, w, c = x.shape
        pad = (-shift, shift)
        p = out_size(h, 1, stride, pad)
        q = out_size(w, 1, stride, pad)
        node = self.graph.add_node(
            "pool", "subsample%d" % self.counts["subsample"], [x], (n, p, q, c),
            dict(kind="avg", kh=1, kw=1, sh=stride, sw=stride, pad_h=pad, pad_w=pad))
        self.counts["subsample"] += 1
        self.top_layer, self.top_size = node.output, c
        return self.top_layer

    def reshape(self, shape: Sequence[int], input_layer=None) -> TensorRef:
        """Only the flattening reshapes the zoo uses (``[-1, F]``) are supported."""
        if input_layer is None:
            input_layer = self.top_layer
        shape = list(shape)
        total = 1
        for s in input_layer.shape[1:]:
            total *= s
        if len(shape) != 2 or shape[0] != -1 or shape[1] != total:
            raise ValueError(f"reshape: unsupported target {shape} for input {input_layer.shape}")
        node = self.graph.add_node("reshape", "reshape%d" % self.counts["reshape"], [input_layer],
                                   (input_layer.shape[0], total))
        self.counts["reshape"] += 1
        self.top_layer, self.top_size = node.output, total
        return self.top_layer

    def affine(self, num_out_channels, input_layer=None, num_channels_in=None, bias=0.0,
               stddev=None, activation="relu", scope_name=None) -> TensorRef:
        """T/convnet_builder_legacy.py:288-325."""
        if input_layer is None:
            input_layer = self.top_layer
        if num_channels_in is None:
            num_channels_in = self.top_size
        if len(input_layer.shape) not in (2, 3) or input_layer.shape[-1] != num_channels_in:
            raise ValueError(f"affine: bad input {input_layer.shape} for in={num_channels_in}")
        if scope_name is None:
            name = "affine" + str(self.counts["affine"])
            self.counts["affine"] += 1
        else:
            name = scope_name
        if activation not in ("relu", "linear", None):
            raise KeyError("Invalid activation type '%s'" % activation)
        act = "linear" if activation is None else activation
        init_factor = 2.0 if act == "relu" else 1.0
        std = stddev or math.sqrt(init_factor / num_channels_in)
        with self.scope(name):
            params = {
                # native layout [out, in padded to 8]: K-major operand rows must be 16-byte
                # multiples for TMA; the pad columns stay zero
                "weights": ParamSpec(self._full("weights"),
                                     (num_out_channels, (num_channels_in + 7) // 8 * 8),
                                     (num_channels_in, num_out_channels), "fc_weight", True,
                                     ("trunc_normal", float(std)),
                                     logical_cols=num_channels_in),
                "biases": ParamSpec(self._full("biases"), (num_out_channels,),
                                    (num_out_channels,), "bias", True, ("const", float(bias))),
            }
            node = self.graph.add_node("affine", self._full(""), [input_layer],
                                       tuple(input_layer.shape[:-1]) + (num_out_channels,),
                                       dict(activation=act), params)
        self.top_layer, self.top_size = node.output, num_out_channels
        return self.top_layer

    def inception_module(self, name, cols, input_layer=None, in_size=None) -> TensorRef:
        """T/convnet_builder_legacy.py:327-365."""
        if input_layer is None:
            input_layer = self.top_layer
        if in_size is None:
            in_size = self.top_size
        # As in the reference, the counter that is bumped is the one of the *suffixed* name, so
        # every module of a kind ends up as "<kind>0" (conv counters keep the variables unique).
        # Reproduced on purpose: it defines the variable names of reference checkpoints.
        name += str(self.counts[name])
        self.counts[name] += 1
        with self.scope(name):
            col_layers, col_sizes = [], []
            for c, col in enumerate(cols):
                col_layers.append([])
                col_sizes.append([])
                for l, layer in enumerate(col):
                    ltype, args = layer[0], layer[1:]
                    kwargs = ({"input_layer": input_layer, "num_channels_in": in_size}
                              if l == 0 else {})
                    if ltype == "conv":
                        self.conv(*args, **kwargs)
                    elif ltype == "mpool":
                        self.mpool(*args, **kwargs)
                    elif ltype == "apool":
                        self.apool(*args, **kwargs)
                    elif ltype == "share":
                        self.top_layer = col_layers[c - 1][l]
                        self.top_size = col_sizes[c - 1][l]
                    else:
                        raise KeyError("Invalid layer type for inception module: '%s'" % ltype)
                    col_layers[c].append(self.top_layer)
                    col_sizes[c].append(self.top_size)
            self.concat([layers[-1] for layers in col_layers])
        return self.top_layer

    def concat(self, layers: Sequence[TensorRef]) -> TensorRef:
        base = layers[0].shape[:-1]
        for t in layers:
            if t.shape[:-1] != base:
                raise ValueError("concat: spatial shapes differ")
        total = sum(t.shape[-1] for t in layers)
        node = self.graph.add_node("concat", "concat%d" % self.counts["concat"], list(layers),
                                   tuple(base) + (total,))
        self.counts["concat"] += 1
        self.top_layer, self.top_size = node.output, total
        return self.top_layer

    def merge_last_dims(self, input_layer=None) -> TensorRef:
        """[N, T, F, C] -> [N, T, F*C]: the reshape between DeepSpeech2's convolutions and its
        recurrent stack (T/models/experimental/deepspeech.py:336-340)."""
        x = self.top_layer if input_layer is None else input_layer
        n, t, f, c = x.shape
        node = self.graph.add_node("reshape", "reshape%d" % self.counts["reshape"], [x],
                                   (n, t, f * c))
        self.counts["reshape"] += 1
        self.top_layer, self.top_size = node.output, f * c
        return self.top_layer

    def lstm(self, hidden: int, bidirectional: bool = True, forget_bias: float = 1.0,
             scope_name: str = "rnn", input_layer=None) -> TensorRef:
        """(Bi)LSTM layer over [N, T, I] with BasicLSTMCell semantics (gate order i, j, f, o,
        forget_bias added to f; zero initial state; outputs of the two directions concatenated)
        - T/models/experimental/deepspeech.py:232-268. TF keeps one ``kernel`` of shape
        [I + H, 4H] per cell; here it is split into the input part (``kernel_x``, applied to all
        time steps by one GEMM) and the recurrent part (``kernel_h``)."""
        x = self.top_layer if input_layer is None else input_layer
        n, t, i = x.shape
        dirs = ("fw", "bw") if bidirectional else ("fw",)
        params = {}
        with self.scope(scope_name):
            for d in dirs:
                init = ("glorot_uniform", i + hidden, 4 * hidden)
                params[d + "_wx"] = ParamSpec(self._full(d + "/kernel_x"),
                                              (4 * hidden, (i + 7) // 8 * 8), (i, 4 * hidden),
                                              "fc_weight", True, init, logical_cols=i)
                params[d + "_wh"] = ParamSpec(self._full(d + "/kernel_h"),
                                              (4 * hidden, (hidden + 7) // 8 * 8),
                                              (hidden, 4 * hidden), "fc_weight", True, init,
                                              logical_cols=hidden)
                params[d + "_b"] = ParamSpec(self._full(d + "/bias"), (4 * hidden,),
                                             (4 * hidden,), "bias", True, ("zeros",))
            node = self.graph.add_node("lstm", self._full(""), [x],
                                       (n, t, hidden * len(dirs)),
                                       dict(hidden=hidden, dirs=dirs,
                                            forget_bias=float(forget_bias)), params)
        self.top_layer, self.top_size = node.output, hidden * len(dirs)
        return self.top_layer

    def embedding(self, ids: TensorRef, num_rows: int, dim: int, col_lo: int = 0,
                  cols: Optional[int] = None, scope_name: str = "embedding",
                  share_with: Optional[str] = None) -> TensorRef:
        """Embedding lookup ``table[ids][col_lo:col_lo+cols]`` (Keras ``Embedding`` of the
        official NeuMF model, T/models/experimental/official_ncf_model.py:62-83). The first
        lookup of a table creates it (Glorot-uniform); later ones pass ``share_with`` (the
        table's variable name) to read another column range of the same table."""
        cols = dim if cols is None else cols
        n = ids.shape[0]
        params = {}
        if share_with is None:
            with self.scope(scope_name):
                spec = ParamSpec(self._full("embeddings"), (num_rows, dim), (num_rows, dim),
                                 "embedding", True, ("glorot_uniform", num_rows, dim))
            params["table"] = spec
            table = spec.name
        else:
            table = share_with
        node = self.graph.add_node("embedding", "%s_%d" % (scope_name, self.counts["embedding"]),
                                   [ids], (n, cols),
                                   dict(table=table, num_rows=num_rows, dim=dim, col_lo=col_lo),
                                   params)
        self.counts["embedding"] += 1
        self.top_layer, self.top_size = node.output, cols
        return self.top_layer

    def mul(self, a: TensorRef, b: TensorRef) -> TensorRef:
        if a.shape != b.shape:
            raise ValueError("mul: shapes differ")
        node = self.graph.add_node("mul", "mul%d" % self.counts["mul"], [a, b], a.shape)
        self.counts["mul"] += 1
        self.top_layer, self.top_size = node.output, a.shape[-1]
        return self.top_layer

    def pad_channels(self, before: int, after: int, input_layer=None) -> TensorRef:
        """Zero-pad the channel axis ("plan A" shortcut, T/models/resnet_model_legacy.py:239-247)."""
        x = self.top_layer if input_layer is None else input_layer
        node = self.graph.add_node("pad_channels", "pad%d" % self.counts["pad"], [x],
                                   x.shape[:-1] + (x.shape[-1] + before + after,),
                                   dict(before=before, after=after))
        self.counts["pad"] += 1
        self.top_layer, self.top_size = node.output, node.output.shape[-1]
        return self.top_layer

    def spatial_mean(self, keep_dims=False) -> TensorRef:
        """T/convnet_builder_legacy.py:367-373."""
        name = "spatial_mean" + str(self.counts["spatial_mean"])
        self.counts["spatial_mean"] += 1
        x = self.top_layer
        n, h, w, c = x.shape
        node = self.graph.add_node("spatial_mean", name, [x], (n, c))
        self.top_layer = node.output
        return self.top_layer

    def dropout(self, keep_prob=0.5, input_layer=None) -> TensorRef:
        """T/convnet_builder_legacy.py:375-390 (the counter is not advanced there either)."""
        if input_layer is None:
            input_layer = self.top_layer
        else:
            self.top_size = None
        if not self.phase_train:
            keep_prob = 1.0
        node = self.graph.add_node("dropout", "dropout%d_%d" % (self.counts["dropout"],
                                                               len(self.graph.nodes)),
                                   [input_layer], input_layer.shape,
                                   dict(keep_prob=float(keep_prob)))
        self.top_layer = node.output
        if self.top_size is None:
            self.top_size = input_layer.shape[-1]
        return self.top_layer

    def lrn(self, depth_radius, bias, alpha, beta) -> TensorRef:
        """T/convnet_builder_legacy.py:462-468."""
        name = "lrn" + str(self.counts["lrn"])
        self.counts["lrn"] += 1
        node = self.graph.add_node("lrn", name, [self.top_layer], self.top_layer.shape,
                                   dict(radius=int(depth_radius), bias=float(bias),
                                        alpha=float(alpha), beta=float(beta)))
        self.top_layer = node.output
        return self.top_layer
