"""ResNet v1 / v1.5 / v2 for ImageNet (50/101/152) and the CIFAR-10 ResNets (20..110).

Canonical definition: T/models/resnet_model_legacy.py:42-329 (the live fork's
Keras-layer rewrite is numerically broken, SURVEY.md 0.5, and is not reproduced).
v1 puts the stride on the first 1x1 of a bottleneck, v1.5 on the 3x3, v2 is
pre-activation. Batch norm: decay .9, eps 1e-5, with scale.
"""
from . import model
from .. import datasets

_BN = {"decay": 0.9, "epsilon": 1e-5, "scale": True}


def _shortcut(cnn, x, in_size, depth, stride, source=None, use_bn=True):
    """identity | 1x1 average-pool subsample | 1x1 projection (+BN)."""
    if depth == in_size:
        if stride == 1:
            return x
        return cnn.apool(1, 1, stride, stride, input_layer=x, num_channels_in=in_size)
    return cnn.conv(depth, 1, 1, stride, stride, activation=None, use_batch_norm=use_bn,
                    input_layer=x if source is None else source, num_channels_in=in_size,
                    bias=None)


def bottleneck_block_v1(cnn, depth, depth_bottleneck, stride, first_stride_on_3x3=False):
    x, in_size = cnn.top_layer, cnn.top_size
    key = "resnet_v1.5" if first_stride_on_3x3 else "resnet_v1"
    name = key + str(cnn.counts[key])
    cnn.counts[key] += 1
    s1, s3 = (1, stride) if first_stride_on_3x3 else (stride, 1)
    with cnn.scope(name):
        shortcut = _shortcut(cnn, x, in_size, depth, stride)
        cnn.conv(depth_bottleneck, 1, 1, s1, s1, input_layer=x, num_channels_in=in_size,
                 use_batch_norm=True, bias=None)
        cnn.conv(depth_bottleneck, 3, 3, s3, s3, mode="SAME_RESNET", use_batch_norm=True,
                 bias=None)
        res = cnn.conv(depth, 1, 1, 1, 1, activation=None, use_batch_norm=True, bias=None)
        cnn.add(shortcut, res, relu=True)


def bottleneck_block_v2(cnn, depth, depth_bottleneck, stride):
    x, in_size = cnn.top_layer, cnn.top_size
    name = "resnet_v2" + str(cnn.counts["resnet_v2"])
    cnn.counts["resnet_v2"] += 1
    cnn.batch_norm()
    preact = cnn.relu()
    with cnn.scope(name):
        shortcut = _shortcut(cnn, x, in_size, depth, stride, source=preact, use_bn=False)
        cnn.conv(depth_bottleneck, 1, 1, stride, stride, input_layer=preact,
                 num_channels_in=in_size, use_batch_norm=True, bias=None)
        cnn.conv(depth_bottleneck, 3, 3, 1, 1, mode="SAME_RESNET", use_batch_norm=True,
                 bias=None)
        res = cnn.conv(depth, 1, 1, 1, 1, activation=None, use_batch_norm=False, bias=None)
        cnn.add(shortcut, res, relu=False)


def bottleneck_block(cnn, depth, depth_bottleneck, stride, version):
    if version == "v2":
        bottleneck_block_v2(cnn, depth, depth_bottleneck, stride)
    elif version == "v1.5":
        bottleneck_block_v1(cnn, depth, depth_bottleneck, stride, first_stride_on_3x3=True)
    else:
        bottleneck_block_v1(cnn, depth, depth_bottleneck, stride)


def residual_block(cnn, depth, stride, version, projection_shortcut=False):
    """Basic 3x3+3x3 block of the CIFAR ResNets with the zero-padded "plan A" shortcut
    (T/models/resnet_model_legacy.py:199-257)."""
    pre_activation = version == "v2"
    x, in_size = cnn.top_layer, cnn.top_size
    if projection_shortcut:
        shortcut = cnn.conv(depth, 1, 1, stride, stride, activation=None, use_batch_norm=True,
                            input_layer=x, num_channels_in=in_size, bias=None)
    elif in_size != depth:
        shortcut = cnn.apool(1, 1, stride, stride, input_layer=x, num_channels_in=in_size)
        padding = (depth - in_size) // 2
        shortcut = cnn.pad_channels(padding, padding, input_layer=shortcut)
    else:
        shortcut = x
    if pre_activation:
        cnn.batch_norm(x)
        res = cnn.relu()
    else:
        res = x
    cnn.conv(depth, 3, 3, stride, stride, input_layer=res, num_channels_in=in_size,
             use_batch_norm=True, bias=None)
    if pre_activation:
        res = cnn.conv(depth, 3, 3, 1, 1, activation=None, use_batch_norm=False, bias=None)
        cnn.add(shortcut, res, relu=False)
    else:
        res = cnn.conv(depth, 3, 3, 1, 1, activation=None, use_batch_norm=True, bias=None)
        cnn.add(shortcut, res, relu=True)


class ResnetModel(model.CNNModel):
    _DEFAULT_BATCH = {"resnet50": 64, "resnet50_v1.5": 64, "resnet50_v2": 64}

    def __init__(self, name, layer_counts, params=None):
        self.base_lr_batch_size = 256
        super().__init__(name, 224, self._DEFAULT_BATCH.get(name, 32), 0.128, layer_counts,
                         params=params)
        self.version = "v2" if "v2" in name else ("v1.5" if "v1.5" in name else "v1")

    def add_inference(self, cnn):
        if self.layer_counts is None:
            raise ValueError("Layer counts not specified for %s" % self.get_model_name())
        cnn.use_batch_norm = True
        cnn.batch_norm_config = dict(_BN)
        cnn.conv(64, 7, 7, 2, 2, mode="SAME_RESNET", use_batch_norm=True)
        cnn.mpool(3, 3, 2, 2, mode="SAME")
        for width, count, first_stride in zip((64, 128, 256, 512), self.layer_counts,
                                              (1, 2, 2, 2)):
            for i in range(count):
                bottleneck_block(cnn, width * 4, width, first_stride if i == 0 else 1,
                                 self.version)
        if self.version == "v2":
            cnn.batch_norm()
            cnn.relu()
        cnn.spatial_mean()

    def get_scaled_base_learning_rate(self, batch_size):
        """Replicated mode sums gradients, so the base LR is divided by num_gpus
        (T/models/resnet_model_legacy.py:311-328)."""
        base_lr = self.learning_rate
        if self.params is not None and self.params.variable_update == "replicated":
            base_lr = self.learning_rate / self.params.num_gpus
        return base_lr * (batch_size / self.base_lr_batch_size)

    def get_learning_rate(self, global_step, batch_size):
        """5 warm-up epochs, then x{1, .1, .01, .001, .0001} at epochs {30, 60, 80, 90}."""
        lr0 = self.get_scaled_base_learning_rate(batch_size)
        per_epoch = float(datasets.IMAGENET_NUM_TRAIN_IMAGES) / batch_size
        boundaries = [int(per_epoch * x) for x in (30, 60, 80, 90)]
        values = [lr0 * v for v in (1, 0.1, 0.01, 0.001, 0.0001)]
        sched = model.LRSchedule(values, boundaries, warmup_steps=int(per_epoch * 5),
                                 warmup_target=lr0)
        return sched(global_step)


class ResnetCifar10Model(model.CNNModel):
    def __init__(self, name, layer_counts, params=None):
        self.version = "v2" if "v2" in name else "v1"
        super().__init__(name, 32, 128, 0.1, layer_counts, params=params)

    def add_inference(self, cnn):
        cnn.use_batch_norm = True
        cnn.batch_norm_config = dict(_BN)
        if self.version == "v2":
            cnn.conv(16, 3, 3, 1, 1, use_batch_norm=True)
        else:
            cnn.conv(16, 3, 3, 1, 1, activation=None, use_batch_norm=True)
        for width, count, first_stride in zip((16, 32, 64), self.layer_counts, (1, 2, 2)):
            for i in range(count):
                residual_block(cnn, width, first_stride if i == 0 else 1, self.version)
        if self.version == "v2":
            cnn.batch_norm()
            cnn.relu()
        cnn.spatial_mean()

    def get_learning_rate(self, global_step, batch_size):
        per_epoch = int(50000 / batch_size)
        sched = model.LRSchedule([0.1, 0.01, 0.001, 0.0002],
                                 [per_epoch * e for e in (82, 123, 300)])
        return sched(global_step)


_IMAGENET = {"resnet50": (3, 4, 6, 3), "resnet101": (3, 4, 23, 3), "resnet152": (3, 8, 36, 3)}
_CIFAR = {"resnet20": 3, "resnet32": 5, "resnet44": 7, "resnet56": 9, "resnet110": 18}


def imagenet_factories():
    out = {}
    for base, counts in _IMAGENET.items():
        for suffix in ("", "_v1.5", "_v2"):
            name = base + suffix
            out[name] = (lambda n, c: (lambda params=None: ResnetModel(n, c, params)))(name, counts)
    return out


def cifar_factories():
    out = {}
    for base, n in _CIFAR.items():
        for suffix in ("", "_v2"):
            name = base + suffix
            out[name] = (lambda nm, c: (lambda params=None: ResnetCifar10Model(nm, c, params)))(
                name, (n, n, n))
    return out
