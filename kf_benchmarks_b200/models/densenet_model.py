"""DenseNet-BC-less CIFAR-10 variants (arXiv:1608.06993), plan of T/models/densenet_model.py:27-95."""
import math

from . import model

_BN = {"decay": 0.9, "epsilon": 1e-5, "scale": True}


class DensenetCifar10Model(model.CNNModel):
    def __init__(self, name, layer_counts, growth_rate, params=None):
        self.growth_rate = growth_rate
        super().__init__(name, 32, 64, 0.1, layer_counts=layer_counts, params=params)

    def dense_block(self, cnn, growth_rate):
        x, in_size = cnn.top_layer, cnn.top_size
        cnn.batch_norm(x, **_BN)
        c = cnn.relu()
        c = cnn.conv(growth_rate, 3, 3, 1, 1, stddev=math.sqrt(2.0 / 9 / growth_rate),
                     activation=None, input_layer=c, num_channels_in=in_size)
        cnn.concat([x, c])

    def transition_layer(self, cnn):
        in_size = cnn.top_size
        cnn.batch_norm(**_BN)
        cnn.relu()
        cnn.conv(in_size, 1, 1, 1, 1, stddev=math.sqrt(2.0 / 9 / in_size))
        cnn.apool(2, 2, 2, 2)

    def add_inference(self, cnn):
        cnn.conv(16, 3, 3, 1, 1, activation=None)
        for stage, count in enumerate(self.layer_counts):
            for _ in range(count):
                self.dense_block(cnn, self.growth_rate)
            if stage < 2:
                self.transition_layer(cnn)
        cnn.batch_norm(**_BN)
        cnn.relu()
        cnn.spatial_mean()

    def get_learning_rate(self, global_step, batch_size):
        per_epoch = int(50000 / batch_size)
        sched = model.LRSchedule([0.1, 0.01, 0.001, 0.0001],
                                 [per_epoch * e for e in (150, 225, 300)])
        return sched(global_step)


def create_densenet40_k12_model(params=None):
    return DensenetCifar10Model("densenet40_k12", (12, 12, 12), 12, params)


def create_densenet100_k12_model(params=None):
    return DensenetCifar10Model("densenet100_k12", (32, 32, 32), 12, params)


def create_densenet100_k24_model(params=None):
    return DensenetCifar10Model("densenet100_k24", (32, 32, 32), 24, params)
