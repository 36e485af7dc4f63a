"""Trivial models: a flatten and two tiny FC layers in front of the classifier
(T/models/trivial_model.py:20-43). Config #1 of BASELINE.json (CPU plumbing) and the fastest
smoke model."""
from . import model
from .plain_stack import build_stack

_PLAN = [("flatten",), ("fc", 1), ("fc", 4096)]


class _Trivial(model.CNNModel):
    image = 0

    def __init__(self, params=None):
        super().__init__("trivial", self.image, 32, 0.005, params=params)

    def add_inference(self, cnn):
        build_stack(cnn, _PLAN)


class TrivialModel(_Trivial):
    image = 227


class TrivialCifar10Model(_Trivial):
    image = 32
