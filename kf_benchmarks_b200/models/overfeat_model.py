"""OverFeat (arXiv:1312.6229) on 231x231 inputs, layer plan of T/models/overfeat_model.py:28-49:
11x11/4 and 5x5 VALID convs with 2x2 pools, three 3x3 SAME convs, pool, FC 3072 and 4096 with
dropout."""
from . import model
from .plain_stack import build_stack

_PLAN = [("conv", 96, 11, 4, "VALID"), ("maxpool", 2, 2),
         ("conv", 256, 5, 1, "VALID"), ("maxpool", 2, 2),
         ("conv", 512, 3, 1, "SAME"), ("conv", 1024, 3, 1, "SAME"), ("conv", 1024, 3, 1, "SAME"),
         ("maxpool", 2, 2), ("flatten",),
         ("fc", 3072), ("dropout",), ("fc", 4096), ("dropout",)]


class OverfeatModel(model.CNNModel):
    def __init__(self, params=None):
        super().__init__("overfeat", 231, 32, 0.005, params=params)

    def add_inference(self, cnn):
        build_stack(cnn, _PLAN)
