"""AlexNet for ImageNet (227x227, batch 512) and the CIFAR-10 tutorial variant with LRN and an
exponentially decayed learning rate (T/models/alexnet_model.py:27-89)."""
from . import model
from .plain_stack import build_stack

_IMAGENET_PLAN = [("conv", 64, 11, 4, "VALID"), ("maxpool", 3, 2),
                  ("conv", 192, 5, 1, "SAME"), ("maxpool", 3, 2),
                  ("conv", 384, 3, 1, "SAME"), ("conv", 384, 3, 1, "SAME"),
                  ("conv", 256, 3, 1, "SAME"), ("maxpool", 3, 2), ("flatten",),
                  ("fc", 4096), ("dropout",), ("fc", 4096), ("dropout",)]

_LRN = ("lrn", 4, 1.0, 0.001 / 9.0, 0.75)
_CIFAR_PLAN = [("conv", 64, 5, 1, "SAME", {"stddev": 5e-2}), ("maxpool", 3, 2, "SAME"), _LRN,
               ("conv", 64, 5, 1, "SAME", {"stddev": 5e-2, "bias": 0.1}), _LRN,
               ("maxpool", 3, 2, "SAME"), ("flatten",),
               ("fc", 384, {"stddev": 0.04, "bias": 0.1}),
               ("fc", 192, {"stddev": 0.04, "bias": 0.1})]


class AlexnetModel(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("alexnet", 227, 512, 0.005, params=params)

    def add_inference(self, cnn):
        build_stack(cnn, _IMAGENET_PLAN)


class AlexnetCifar10Model(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("alexnet", 32, 128, 0.1, params=params)

    def add_inference(self, cnn):
        build_stack(cnn, _CIFAR_PLAN)

    def get_learning_rate(self, global_step, batch_size):
        """x0.1 (staircase) every 100 epochs of the 50 000 training images."""
        steps_per_decay = int(100 * 50000 / batch_size)
        return self.learning_rate * 0.1 ** (global_step // steps_per_decay)
