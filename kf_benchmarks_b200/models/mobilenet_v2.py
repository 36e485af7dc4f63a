"""MobileNet-v2 (``--model=mobilenet``).

Reference: T/models/mobilenet_v2.py:42-198 (conv defs and the CNNModel wrapper),
T/models/mobilenet_conv_blocks.py:163-313 (``expanded_conv``), T/models/mobilenet.py:363-382
(the logits head) and :414-466 (training scope: truncated-normal(.09) weights, BN decay .997,
dropout keep .8). The reference builds it through tf.contrib.slim; here the same layers are
recorded on the NetBuilder with slim's variable names:

    conv 3x3/2 x32 -> 17 x expanded_conv -> conv 1x1 x1280 -> global pool -> dropout ->
    1x1 conv (with bias) to nclass

``expanded_conv`` = 1x1 expand (x6, skipped for the first block) -> 3x3 depthwise -> 1x1
linear projection, every conv followed by BN (center+scale) and ReLU6 except the projection;
a residual joins input and output when the stride is 1 and the depth does not change.

As in the reference, the CNNModel wrapper then stacks its usual final affine layer on top of
MobileNet's own logits (the wrapper does not override ``skip_final_affine_layer``).
"""
from . import model

_BN = {"decay": 0.997, "epsilon": 0.001, "scale": True}
_STDDEV = 0.09
# (stride, num_outputs) of the 17 expanded_conv blocks; the first one has expansion 1
_BLOCKS = [(1, 16), (2, 24), (1, 24), (2, 32), (1, 32), (1, 32), (2, 64), (1, 64), (1, 64),
           (1, 64), (1, 96), (1, 96), (1, 96), (2, 160), (1, 160), (1, 160), (1, 320)]


def _make_divisible(v, divisor=8, min_value=None):
    """T/models/mobilenet_conv_blocks.py:33-40."""
    if min_value is None:
        min_value = divisor
    new_v = max(min_value, int(v + divisor / 2) // divisor * divisor)
    if new_v < 0.9 * v:
        new_v += divisor
    return new_v


def expanded_conv(cnn, num_outputs, stride, expansion, scope):
    x, in_size = cnn.top_layer, cnn.top_size
    inner = in_size * expansion if expansion == 1 else _make_divisible(in_size * expansion, 8)
    with cnn.scope(scope):
        if inner > in_size:
            cnn.conv(inner, 1, 1, 1, 1, input_layer=x, num_channels_in=in_size,
                     use_batch_norm=True, activation="relu6", stddev=_STDDEV, bias=None,
                     scope_name="expand", kernel_leaf="weights", bn_scope_name="BatchNorm")
        cnn.depthwise_conv(3, 3, stride, stride, mode="SAME", stddev=_STDDEV,
                           scope_name="depthwise", use_batch_norm=True, activation="relu6",
                           bn_scope_name="BatchNorm")
        out = cnn.conv(num_outputs, 1, 1, 1, 1, use_batch_norm=True, activation=None,
                       stddev=_STDDEV, bias=None, scope_name="project", kernel_leaf="weights",
                       bn_scope_name="BatchNorm")
        if stride == 1 and num_outputs == in_size:
            cnn.add(x, out, relu=False)


class MobilenetModel(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("mobilenet", 224, 32, 0.005, params=params)

    def add_inference(self, cnn):
        cnn.use_batch_norm = True
        cnn.batch_norm_config = dict(_BN)
        with cnn.scope("MobilenetV2"):
            cnn.conv(32, 3, 3, 2, 2, mode="SAME", activation="relu6", stddev=_STDDEV, bias=None,
                     scope_name="Conv", kernel_leaf="weights", bn_scope_name="BatchNorm")
            for i, (stride, depth) in enumerate(_BLOCKS):
                expanded_conv(cnn, depth, stride, 1 if i == 0 else 6,
                              "expanded_conv" if i == 0 else "expanded_conv_%d" % i)
            cnn.conv(1280, 1, 1, 1, 1, activation="relu6", stddev=_STDDEV, bias=None,
                     scope_name="Conv_1", kernel_leaf="weights", bn_scope_name="BatchNorm")
            with cnn.scope("Logits"):
                cnn.spatial_mean()
                cnn.top_size = 1280
                cnn.dropout(0.8)
                # slim's 1x1 conv on the pooled [N,1,1,1280] map == a dense layer
                cnn.affine(1001, stddev=_STDDEV, activation="linear",
                           scope_name="Conv2d_1c_1x1")
        cnn.use_batch_norm = False
