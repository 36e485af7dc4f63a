"""LeNet-5 on 28x28 inputs (T/models/lenet_model.py:27-40): two 5x5 SAME conv + 2x2 max-pool
stages (32 and 64 filters) and a 512-unit FC layer."""
from . import model
from .plain_stack import build_stack

_PLAN = [("conv", 32, 5, 1, "SAME"), ("maxpool", 2, 2),
         ("conv", 64, 5, 1, "SAME"), ("maxpool", 2, 2),
         ("flatten",), ("fc", 512)]


class Lenet5Model(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("lenet5", 28, 32, 0.005, params=params)

    def add_inference(self, cnn):
        build_stack(cnn, _PLAN)
