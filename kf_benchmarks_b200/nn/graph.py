"""Static layer graph: the intermediate representation between the model zoo and the
executors.

The reference builds a TensorFlow graph through ``ConvNetBuilder``
(T/convnet_builder_legacy.py:32-468); here the builder records a small static DAG
of coarse, fusion-friendly nodes that two executors run:

* ``nn.executor_cuda.CudaExecutor`` - hand-written sm_100a kernels, fp16 NHWC
  activations, parameters in flat arenas;
* ``nn.executor_ref.RefExecutor``   - plain PyTorch fp32 autograd on any device
  (the ``--device=cpu`` path and the numerical oracle of the GPU tests).

Shapes are NHWC ``(N, H, W, C)`` or flat ``(N, F)``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple


@dataclass
class ParamSpec:
    """One variable. ``name`` follows the reference's TF layout below the tower scope
    (``cg/conv0/conv2d/kernel`` ...), see SURVEY.md 2.7."""
    name: str
    shape: Tuple[int, ...]          # native layout (what the kernels read)
    tf_shape: Tuple[int, ...]       # layout in a reference checkpoint
    kind: str                       # conv_kernel | conv_kernel_flat | fc_weight | bias | gamma |
                                    # beta | moving_mean | moving_variance
    trainable: bool = True
    init: Tuple[Any, ...] = ("zeros",)  # ("zeros",) ("ones",) ("const", v) ("trunc_normal", std)
                                        # ("glorot_uniform", fan_in, fan_out)
    logical_cols: Optional[int] = None  # conv_kernel_flat: R*S*C before padding to a multiple of 8
    conv_rsc: Optional[Tuple[int, int, int]] = None
    # channel legalisation (nn/legalize.py): physical position of each logical input channel
    # (or of each element, for per-channel vectors) and the padded output-channel count
    in_map: Optional[Tuple[int, ...]] = None
    out_pad: Optional[int] = None

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass(eq=False)
class TensorRef:
    id: int
    shape: Tuple[int, ...]
    producer: Optional["Node"] = None
    consumers: List["Node"] = field(default_factory=list)
    dtype: str = "f16"              # f16 activations; i32 ids / labels; f32 targets

    @property
    def channels(self) -> int:
        return self.shape[-1]

    @property
    def rows(self) -> int:
        n = 1
        for s in self.shape[:-1]:
            n *= s
        return n


@dataclass(eq=False)
class Node:
    op: str
    name: str
    inputs: List[TensorRef]
    output: TensorRef
    attrs: Dict[str, Any] = field(default_factory=dict)
    params: Dict[str, ParamSpec] = field(default_factory=dict)


class Graph:
    def __init__(self, batch_size: int, image_shape: Optional[Tuple[int, int, int]] = None,
                 input_shape: Optional[Tuple[int, ...]] = None, input_dtype: str = "f16"):
        """``image_shape`` (h, w, c) gives the usual NHWC image input; models with other
        inputs (NCF ids, spectrograms ...) pass the full ``input_shape`` / ``input_dtype``."""
        self.batch_size = batch_size
        self.nodes: List[Node] = []
        self.tensors: List[TensorRef] = []
        self.params: Dict[str, ParamSpec] = {}
        if input_shape is None:
            h, w, c = image_shape
            input_shape = (batch_size, h, w, c)
        self.input = self.new_tensor(input_shape)
        self.input.dtype = input_dtype
        # further per-step inputs (NCF item ids ...), in the order the input source yields them
        # after the primary input and before the labels
        self.extra_inputs: List[TensorRef] = []
        # label tensors: (shape, dtype); classification has one int32 label per example
        self.label_specs: List[Tuple[Tuple[int, ...], str]] = [((batch_size,), "i32")]
        # loss head run on ``logits`` (see the executors): softmax_xent | ncf | ssd | ctc
        self.loss: Dict[str, Any] = {"kind": "softmax_xent"}
        self.logits: Optional[TensorRef] = None
        self.aux_logits: Optional[TensorRef] = None

    def add_input(self, shape, dtype: str = "f16") -> TensorRef:
        t = self.new_tensor(shape)
        t.dtype = dtype
        self.extra_inputs.append(t)
        return t

    def new_tensor(self, shape) -> TensorRef:
        t = TensorRef(len(self.tensors), tuple(int(s) for s in shape))
        self.tensors.append(t)
        return t

    def add_node(self, op, name, inputs, out_shape, attrs=None, params=None) -> Node:
        out = self.new_tensor(out_shape)
        node = Node(op, name, list(inputs), out, dict(attrs or {}), dict(params or {}))
        out.producer = node
        for t in inputs:
            t.consumers.append(node)
        for p in node.params.values():
            if p.name in self.params:
                raise ValueError(f"duplicate parameter name {p.name}")
            self.params[p.name] = p
        self.nodes.append(node)
        return node

    # -- statistics ---------------------------------------------------------
    def num_trainable(self) -> int:
        def logical(p):
            n = 1
            for d in p.tf_shape:
                n *= d
            return n
        return sum(logical(p) for p in self.params.values() if p.trainable)

    def trainable_params(self) -> List[ParamSpec]:
        return [p for p in self.params.values() if p.trainable]

    def describe(self) -> str:
        lines = []
        for n in self.nodes:
            ins = ",".join(str(t.id) for t in n.inputs)
            lines.append(f"{n.op:14s} {n.name:40s} [{ins}] -> {n.output.id} {n.output.shape}")
        return "\n".join(lines)
