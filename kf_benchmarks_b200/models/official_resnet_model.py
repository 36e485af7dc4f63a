""""Official" ImageNet ResNets: official_resnet{18,34,50,101,152,200}{,_v2}.

The reference wraps ``official.resnet.imagenet_main.ImagenetModel`` of the external
tensorflow/models repository (T/models/official_resnet_model.py:26-77) and only adds the batch
size / learning-rate policy. That package is not part of the reference tree, so the network
is restated here from its published definition, on the same builder the rest of the zoo uses:

* 7x7/2 stem with "fixed padding" (explicit (k-1)/2 pads for strided convs = ``SAME_RESNET``),
  3x3/2 SAME max-pool; v1 applies BN+ReLU to the stem, v2 leaves that to the blocks;
* sizes 18/34 use the basic 3x3+3x3 block, 50..200 the bottleneck with the stride on the 3x3;
* the first block of every stage has a 1x1 projection shortcut (also when the shape does not
  change), v1 normalises it, v2 takes it from the pre-activation;
* batch norm momentum .997, eps 1e-5, with scale; no convolution biases;
* v2 ends with BN+ReLU; global mean; dense 1001.

Learning rate: 0.0125 * default_batch / 32, rescaled by batch/default, x{1,.1,.01,.001,.0001}
at epochs {30,60,80,90} (T/models/official_resnet_model.py:49-58).
"""
from . import model
from .. import datasets

_BN = {"decay": 0.997, "epsilon": 1e-5, "scale": True}
_BLOCKS = {18: (2, 2, 2, 2), 34: (3, 4, 6, 3), 50: (3, 4, 6, 3), 101: (3, 4, 23, 3),
           152: (3, 8, 36, 3), 200: (3, 24, 36, 3)}


def _block(cnn, filters, stride, bottleneck, version, project):
    x, in_size = cnn.top_layer, cnn.top_size
    out = filters * 4 if bottleneck else filters
    key = "block_v%d" % version
    name = key + "_" + str(cnn.counts[key])
    cnn.counts[key] += 1
    if version == 2:
        cnn.batch_norm(**_BN)
        src = cnn.relu()
    else:
        src = x
    with cnn.scope(name):
        if project:
            shortcut = cnn.conv(out, 1, 1, stride, stride, mode="SAME_RESNET", activation=None,
                                use_batch_norm=(version == 1), input_layer=src,
                                num_channels_in=in_size, bias=None)
        else:
            shortcut = x
        last_bn = version == 1
        if bottleneck:
            cnn.conv(filters, 1, 1, 1, 1, input_layer=src, num_channels_in=in_size,
                     use_batch_norm=True, bias=None)
            cnn.conv(filters, 3, 3, stride, stride, mode="SAME_RESNET", use_batch_norm=True,
                     bias=None)
            res = cnn.conv(out, 1, 1, 1, 1, activation=None, use_batch_norm=last_bn, bias=None)
        else:
            cnn.conv(filters, 3, 3, stride, stride, mode="SAME_RESNET", input_layer=src,
                     num_channels_in=in_size, use_batch_norm=True, bias=None)
            res = cnn.conv(out, 3, 3, 1, 1, activation=None, use_batch_norm=last_bn, bias=None)
        cnn.add(shortcut, res, relu=(version == 1))


class ImagenetResnetModel(model.CNNModel):
    def __init__(self, resnet_size, version=2, params=None):
        if resnet_size not in _BLOCKS:
            raise ValueError("official_resnet: unsupported size %s" % resnet_size)
        batch_size = {50: 128, 101: 32, 152: 32}.get(resnet_size, 32)
        lr = 0.0125 * batch_size / 32
        super().__init__("official_resnet_%d_v%d" % (resnet_size, version), 224, batch_size, lr,
                         params=params)
        self.resnet_size = resnet_size
        self.version = version

    def add_inference(self, cnn):
        cnn.use_batch_norm = False
        cnn.batch_norm_config = dict(_BN)
        v1 = self.version == 1
        cnn.conv(64, 7, 7, 2, 2, mode="SAME_RESNET", use_batch_norm=v1,
                 activation="relu" if v1 else None, bias=None)
        cnn.mpool(3, 3, 2, 2, mode="SAME")
        bottleneck = self.resnet_size >= 50
        for filters, count, stride in zip((64, 128, 256, 512), _BLOCKS[self.resnet_size],
                                          (1, 2, 2, 2)):
            for i in range(count):
                _block(cnn, filters, stride if i == 0 else 1, bottleneck, self.version,
                       project=(i == 0))
        if not v1:
            cnn.batch_norm(**_BN)
            cnn.relu()
        cnn.spatial_mean()

    def get_learning_rate(self, global_step, batch_size):
        per_epoch = float(datasets.IMAGENET_NUM_TRAIN_IMAGES) / batch_size
        boundaries = [int(per_epoch * x) for x in (30, 60, 80, 90)]
        lr0 = self.learning_rate / self.default_batch_size * batch_size
        sched = model.LRSchedule([lr0 * v for v in (1, 0.1, 0.01, 0.001, 0.0001)], boundaries)
        return sched(global_step)


def factories():
    out = {}
    for size in _BLOCKS:
        for version, suffix in ((1, ""), (2, "_v2")):
            out["official_resnet%d%s" % (size, suffix)] = (
                lambda s, v: (lambda params=None: ImagenetResnetModel(s, v, params)))(size, version)
    return out
