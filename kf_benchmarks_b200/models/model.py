"""Model base classes: batch size / learning-rate defaults, graph construction, loss rules.

Mirrors the contract of the reference's ``Model`` / ``CNNModel``
(T/models/model.py:31-312): a model knows its default per-device batch size, image
size, base learning rate and fp16 loss scale, adds its layers through a builder in
``add_inference`` and gets a final linear ``affine(nclass)`` appended (the canonical
legacy behaviour, T/models/model_legacy.py; the live fork commented it out, which is
one of the bugs we do not reproduce - SURVEY.md appendix A).
"""
from __future__ import annotations

from typing import Callable, List, Optional

from ..nn.builder import NetBuilder
from ..nn.graph import Graph


class LRSchedule:
    """Piecewise-constant schedule with optional linear warm-up, evaluated on the host.

    ``boundaries`` are global steps; ``values`` has one more entry. TF's
    ``piecewise_constant`` semantics: ``values[i]`` applies while ``step <= boundaries[i]``.
    """

    def __init__(self, values: List[float], boundaries: Optional[List[int]] = None,
                 warmup_steps: int = 0, warmup_target: Optional[float] = None,
                 fn: Optional[Callable[[int], float]] = None):
        self.values = list(values)
        self.boundaries = list(boundaries or [])
        self.warmup_steps = int(warmup_steps)
        self.warmup_target = warmup_target
        self.fn = fn

    def __call__(self, step: int) -> float:
        if self.warmup_steps > 0 and step < self.warmup_steps:
            return float(self.warmup_target) * float(step) / float(self.warmup_steps)
        if self.fn is not None:
            return float(self.fn(step))
        for b, v in zip(self.boundaries, self.values):
            if step <= b:
                return float(v)
        return float(self.values[-1])


class Model:
    def __init__(self, model_name: str, batch_size: int, learning_rate: float,
                 fp16_loss_scale: float, params=None):
        self.model_name = model_name
        self.batch_size = batch_size
        self.default_batch_size = batch_size
        self.learning_rate = learning_rate
        self.fp16_loss_scale = fp16_loss_scale
        self.params = params
        self.use_fp16 = bool(params.use_fp16) if params else False
        self.fp16_vars = bool(params.fp16_vars) if params else False

    def get_model_name(self):
        return self.model_name

    def get_batch_size(self):
        return self.batch_size

    def set_batch_size(self, batch_size):
        self.batch_size = batch_size

    def get_default_batch_size(self):
        return self.default_batch_size

    def get_fp16_loss_scale(self):
        return self.fp16_loss_scale

    def l2_decay_filter(self, var_name: str) -> bool:
        """Whether variable ``var_name`` takes part in the L2 (weight decay) term. The default
        decays every trainable variable (T/benchmark_cnn.py:3068-3099); models override it the
        way the reference overrides ``custom_l2_loss`` (T/models/model.py:115-131)."""
        del var_name
        return True

    def get_learning_rate(self, global_step: int, batch_size: int) -> float:
        del global_step, batch_size
        return self.learning_rate

    def postprocess(self, results):
        return results


class CNNModel(Model):
    """T/models/model.py:134-312."""

    def __init__(self, model, image_size, batch_size, learning_rate, layer_counts=None,
                 fp16_loss_scale=128, params=None):
        super().__init__(model, batch_size, learning_rate, fp16_loss_scale, params=params)
        self.image_size = image_size
        self.layer_counts = layer_counts
        self.depth = 3
        self.data_format = params.data_format if params else "NHWC"

    def get_layer_counts(self):
        return self.layer_counts

    def skip_final_affine_layer(self) -> bool:
        return False

    def add_inference(self, cnn: NetBuilder) -> None:
        raise NotImplementedError("Must be implemented in derived classes")

    def get_input_shapes(self, subset="train"):
        del subset
        return [[self.batch_size, self.image_size, self.image_size, self.depth],
                [self.batch_size]]

    def build_graph(self, nclass: int = 1001, phase_train: bool = True,
                    batch_size: Optional[int] = None) -> Graph:
        """Records the network (T/models/model.py:239-283) and returns the graph."""
        n = batch_size or self.batch_size
        g = Graph(n, (self.image_size, self.image_size, self.depth))
        cnn = NetBuilder(g, phase_train=phase_train)
        self.add_inference(cnn)
        if not self.skip_final_affine_layer():
            cnn.affine(nclass, activation="linear")
        g.logits = cnn.top_layer
        if cnn.aux_top_layer is not None:
            with cnn.switch_to_aux_top_layer():
                cnn.affine(nclass, activation="linear", stddev=0.001)
            g.aux_logits = cnn.aux_top_layer
        return g

    # aux loss weight of T/models/model.py:297-302
    aux_loss_weight = 0.4
