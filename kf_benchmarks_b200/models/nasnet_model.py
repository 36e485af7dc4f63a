"""NASNet-A: ``nasnet`` (mobile, ImageNet), ``nasnetlarge`` and ``nasnet`` for CIFAR-10.

Reference: T/models/nasnet_model.py:39-111 (hyper-parameters), :262-297 (stems), :452-540
(network assembly), :543-579 (CNNModel wrappers) and T/models/nasnet_utils.py:44-491 (cells,
factorised reduction, stacked separable convs, scheduled drop-path).

What the reference actually runs - and what is restated here:

* the wrappers call ``build_nasnet_*`` with ``num_classes=None``, so the network ends at the
  global average pool (no auxiliary head, no dense dropout) and the CNNModel's own final
  affine layer produces the logits;
* the wrappers do **not** enter ``nasnet_*_arg_scope``, so every slim layer runs with slim's
  defaults: convolutions (1x1 and the pointwise half of the separable ones) have a bias and
  a ReLU, batch norm has no scale, decay .999, eps 1e-3, weights are Glorot-uniform.
  ``NasnetConfig.as_run=False`` builds the arg-scoped variant instead (no bias / activation
  on convs, BN with scale, decay .9997 / .9) for users who want the published network.

Cells (NASNet-A): five "combinations" per cell, each adding two branches picked from
{stacked separable k x k, 3x3 avg/max pool, identity}; normal cells concatenate the previous
cell's (reduced) output with the five new states, reduction cells the last four. Odd filter
counts (11 / 22 in the mobile stem cells, 44-multiples elsewhere) are fine on the CUDA
executor thanks to nn/legalize.py.
"""
from dataclasses import dataclass

from . import model


@dataclass
class NasnetConfig:
    stem_multiplier: float
    num_cells: int
    num_conv_filters: int
    drop_path_keep_prob: float
    total_training_steps: int
    skip_reduction_layer_input: bool
    stem_type: str
    filter_scaling_rate: float = 2.0
    num_reduction_layers: int = 2
    as_run: bool = True
    arg_scope_bn_decay: float = 0.9997
    arg_scope_bn_eps: float = 1e-3


def cifar_config(phase_train=True, total_steps=None):
    return NasnetConfig(3.0, 18, 32, 0.6 if phase_train else 1.0, total_steps or 937500, False,
                        "cifar", arg_scope_bn_decay=0.9, arg_scope_bn_eps=1e-5)


def large_imagenet_config(phase_train=True, total_steps=None):
    return NasnetConfig(3.0, 18, 168, 0.7 if phase_train else 1.0, total_steps or 250000, True,
                        "imagenet")


def mobile_imagenet_config(total_steps=None):
    return NasnetConfig(1.0, 12, 44, 1.0, total_steps or 250000, False, "imagenet")


def calc_reduction_layers(num_cells, num_reduction_layers):
    """T/models/nasnet_utils.py:44-52."""
    return [int(float(i) / (num_reduction_layers + 1) * num_cells)
            for i in range(1, num_reduction_layers + 1)]


NORMAL_CELL = dict(
    operations=["separable_5x5_2", "separable_3x3_2", "separable_5x5_2", "separable_3x3_2",
                "avg_pool_3x3", "none", "avg_pool_3x3", "avg_pool_3x3", "separable_3x3_2", "none"],
    used=[1, 0, 0, 0, 0, 0, 0],
    indices=[0, 1, 1, 1, 0, 1, 1, 1, 0, 0])
REDUCTION_CELL = dict(
    operations=["separable_5x5_2", "separable_7x7_2", "max_pool_3x3", "separable_7x7_2",
                "avg_pool_3x3", "separable_5x5_2", "none", "avg_pool_3x3", "separable_3x3_2",
                "max_pool_3x3"],
    used=[1, 1, 1, 0, 0, 0, 0],
    indices=[0, 1, 0, 1, 0, 1, 3, 2, 2, 0])


class _Net:
    """slim-flavoured layer helpers on top of the NetBuilder."""

    def __init__(self, cnn, cfg: NasnetConfig, total_num_cells: int):
        self.cnn = cnn
        self.cfg = cfg
        self.total_num_cells = total_num_cells
        self._relu_cache = {}

    def relu(self, x):
        # several branches read relu(x) of the same hidden state: compute it once
        if x.id not in self._relu_cache:
            self._relu_cache[x.id] = self.cnn.relu(x, fold=False)
        return self._relu_cache[x.id]

    def conv(self, x, filters, k, stride, scope, mode="SAME", leaf="weights"):
        as_run = self.cfg.as_run
        return self.cnn.conv(filters, k, k, stride, stride, mode=mode, input_layer=x,
                             num_channels_in=x.shape[-1], use_batch_norm=False,
                             activation="relu" if as_run else None,
                             bias=0.0 if as_run else None, scope_name=scope, kernel_leaf=leaf)

    def bn(self, x, scope):
        if self.cfg.as_run:
            return self.cnn.batch_norm(x, decay=0.999, epsilon=0.001, scale=False,
                                       scope_name=scope)
        return self.cnn.batch_norm(x, decay=self.cfg.arg_scope_bn_decay,
                                   epsilon=self.cfg.arg_scope_bn_eps, scale=True,
                                   scope_name=scope)

    def separable(self, x, filters, k, stride, scope):
        x = self.cnn.depthwise_conv(k, k, stride, stride, mode="SAME", input_layer=x,
                                    scope_name=scope, use_batch_norm=False)
        return self.conv(x, filters, 1, 1, scope, leaf="pointwise_weights")

    def factorized_reduction(self, x, filters, stride):
        """T/models/nasnet_utils.py:87-108."""
        if stride == 1:
            return self.bn(self.conv(x, filters, 1, 1, "path_conv"), "path_bn")
        if filters % 2:
            raise ValueError("Need even number of filters when using this factorized reduction.")
        p1 = self.conv(self.cnn.subsample(stride, 0, input_layer=x), filters // 2, 1, 1,
                       "path1_conv")
        p2 = self.conv(self.cnn.subsample(stride, 1, input_layer=x), filters // 2, 1, 1,
                       "path2_conv")
        return self.bn(self.cnn.concat([p1, p2]), "final_path_bn")

    def stacked_separable(self, x, stride, operation, filters):
        parts = operation.split("_")
        k, layers = int(parts[1][0]), int(parts[2])
        for i in range(layers):
            x = self.relu(x)
            x = self.separable(x, filters, k, stride, "separable_%dx%d_%d" % (k, k, i + 1))
            x = self.bn(x, "bn_sep_%dx%d_%d" % (k, k, i + 1))
            stride = 1
        return x

    def drop_path(self, x, cell_num):
        base = self.cfg.drop_path_keep_prob
        if base >= 1.0 or not self.cnn.phase_train:
            return x
        layer_ratio = (cell_num + 1) / float(self.total_num_cells)
        keep_layer = 1.0 - layer_ratio * (1.0 - base)
        total = float(self.cfg.total_training_steps)

        def keep_prob(step, keep_layer=keep_layer, total=total):
            return 1.0 - min(1.0, step / total) * (1.0 - keep_layer)
        return self.cnn.drop_path(keep_prob, input_layer=x)


def _apply_operation(net: _Net, x, operation, stride, from_original_input, filters, cell_num):
    """T/models/nasnet_utils.py:336-361. Returns (tensor, fresh) - ``fresh`` says the tensor
    was produced here by a batch norm nobody else reads, so an add may be folded into it."""
    cnn = net.cnn
    if stride > 1 and not from_original_input:
        stride = 1
    in_filters = x.shape[-1]
    fresh = False
    if "separable" in operation:
        x = net.stacked_separable(x, stride, operation, filters)
        fresh = True
    elif operation == "none":
        if stride > 1 or in_filters != filters:
            x = net.bn(net.conv(net.relu(x), filters, 1, stride, "1x1"), "bn_1")
            fresh = True
    elif "pool" in operation:
        kind, _, shape = operation.split("_")
        k = int(shape[0])
        pool = cnn.apool if kind == "avg" else cnn.mpool
        x = pool(k, k, stride, stride, mode="SAME", input_layer=x, num_channels_in=in_filters)
        if in_filters != filters:
            x = net.bn(net.conv(x, filters, 1, 1, "1x1"), "bn_1")
            fresh = True
    else:
        raise ValueError("Unimplemented operation %s" % operation)
    if operation != "none":
        y = net.drop_path(x, cell_num)
        fresh = fresh and y is x
        x = y
    return x, fresh


def build_cell(net: _Net, spec, x, prev, scope, filter_scaling, stride, cell_num):
    """One NASNet-A cell (T/models/nasnet_utils.py:263-334, 363-388)."""
    cnn = net.cnn
    filters = int(net.cfg.num_conv_filters * filter_scaling)
    with cnn.scope(scope):
        # bring the previous cell's output to this cell's resolution / width
        if prev is None:
            prev = x
        elif prev.shape[2] != x.shape[2]:
            prev = net.factorized_reduction(net.relu(prev), filters, 2)
        elif prev.shape[-1] != filters:
            prev = net.bn(net.conv(net.relu(prev), filters, 1, 1, "prev_1x1"), "prev_bn")
        h = net.bn(net.conv(net.relu(x), filters, 1, 1, "1x1"), "beginning_bn")
        states = [h, prev]
        ops, idx = spec["operations"], spec["indices"]
        for it in range(5):
            with cnn.scope("comb_iter_%d" % it):
                li, ri = idx[2 * it], idx[2 * it + 1]
                with cnn.scope("left"):
                    h1, f1 = _apply_operation(net, states[li], ops[2 * it], stride, li < 2,
                                              filters, cell_num)
                with cnn.scope("right"):
                    h2, f2 = _apply_operation(net, states[ri], ops[2 * it + 1], stride, ri < 2,
                                              filters, cell_num)
                # fold the add into the right branch's batch norm only: it is the later node,
                # so its second input (the left branch) already exists in execution order
                states.append(cnn.add(h1, h2, relu=False, fold=f2))
        with cnn.scope("cell_output"):
            final = states[-1]
            for i, used in enumerate(spec["used"]):
                s = states[i]
                differs = s.shape[2] != final.shape[2] or s.shape[-1] != final.shape[-1]
                if differs and not used:
                    with cnn.scope("reduction_%d" % i):
                        states[i] = net.factorized_reduction(
                            s, final.shape[-1], 2 if s.shape[2] != final.shape[2] else 1)
            out = cnn.concat([s for s, used in zip(states, spec["used"]) if not used])
    return out


def build_nasnet(cnn, cfg: NasnetConfig):
    """T/models/nasnet_model.py:452-540 with num_classes=None (ends at the global pool)."""
    total_num_cells = cfg.num_cells + 2 + (2 if cfg.stem_type == "imagenet" else 0)
    net = _Net(cnn, cfg, total_num_cells)
    images = cnn.top_layer
    if cfg.stem_type == "imagenet":
        x = net.conv(images, int(32 * cfg.stem_multiplier), 3, 2, "conv0", mode="VALID")
        x = net.bn(x, "conv0_bn")
        cell_outputs = [None, x]
        scaling = 1.0 / (cfg.filter_scaling_rate ** 2)
        for i in range(2):
            x = build_cell(net, REDUCTION_CELL, x, cell_outputs[-2], "cell_stem_%d" % i, scaling,
                           2, i)
            cell_outputs.append(x)
            scaling *= cfg.filter_scaling_rate
        true_cell_num = 2
    else:
        x = net.conv(images, int(cfg.num_conv_filters * cfg.stem_multiplier), 3, 1,
                     "l1_stem_3x3")
        x = net.bn(x, "l1_stem_bn")
        cell_outputs = [None, x]
        true_cell_num = 0
    reduction_indices = calc_reduction_layers(cfg.num_cells, cfg.num_reduction_layers)
    scaling = 1.0
    prev = None
    for cell_num in range(cfg.num_cells):
        if cfg.skip_reduction_layer_input:
            prev = cell_outputs[-2]
        if cell_num in reduction_indices:
            scaling *= cfg.filter_scaling_rate
            x = build_cell(net, REDUCTION_CELL, x, cell_outputs[-2],
                           "reduction_cell_%d" % reduction_indices.index(cell_num), scaling, 2,
                           true_cell_num)
            true_cell_num += 1
            cell_outputs.append(x)
        if not cfg.skip_reduction_layer_input:
            prev = cell_outputs[-2]
        x = build_cell(net, NORMAL_CELL, x, prev, "cell_%d" % cell_num, scaling, 1,
                       true_cell_num)
        true_cell_num += 1
        cell_outputs.append(x)
    with cnn.scope("final_layer"):
        cnn.relu(x, fold=False)
        cnn.spatial_mean()
    cnn.top_size = x.shape[-1]


class _NasnetBase(model.CNNModel):
    # field overrides for NasnetConfig (e.g. {"num_cells": 6, "as_run": False})
    config_overrides = None

    def _config(self, cnn) -> NasnetConfig:
        raise NotImplementedError

    def add_inference(self, cnn):
        import dataclasses
        cnn.use_batch_norm = False
        cfg = self._config(cnn)
        if self.config_overrides:
            cfg = dataclasses.replace(cfg, **self.config_overrides)
        build_nasnet(cnn, cfg)


class NasnetModel(_NasnetBase):
    def __init__(self, params=None):
        super().__init__("nasnet", 224, 32, 0.005, params=params)

    def _config(self, cnn):
        return mobile_imagenet_config()


class NasnetLargeModel(_NasnetBase):
    def __init__(self, params=None):
        super().__init__("nasnet", 331, 16, 0.005, params=params)

    def _config(self, cnn):
        return large_imagenet_config(cnn.phase_train)


class NasnetCifarModel(_NasnetBase):
    def __init__(self, params=None):
        super().__init__("nasnet", 32, 32, 0.025, params=params)

    def _config(self, cnn):
        return cifar_config(cnn.phase_train)
