"""VGG 11/16/19 (configurations A, D, E of arXiv:1409.1556), no batch norm.
Layer plan of T/models/vgg_model.py:30-79."""
from . import model

_PLANS = {"vgg11": [1, 1, 2, 2, 2], "vgg16": [2, 2, 3, 3, 3], "vgg19": [2, 2, 4, 4, 4]}


class VggModel(model.CNNModel):
    def __init__(self, name, params=None):
        super().__init__(name, 224, 64, 0.005, params=params)
        self.plan = _PLANS[name]

    def add_inference(self, cnn):
        for width, reps in zip((64, 128, 256, 512, 512), self.plan):
            for _ in range(reps):
                cnn.conv(width, 3, 3)
            cnn.mpool(2, 2)
        cnn.reshape([-1, 512 * 7 * 7])
        cnn.affine(4096)
        cnn.dropout()
        cnn.affine(4096)
        cnn.dropout()


def Vgg11Model(params=None):
    return VggModel("vgg11", params)


def Vgg16Model(params=None):
    return VggModel("vgg16", params)


def Vgg19Model(params=None):
    return VggModel("vgg19", params)
