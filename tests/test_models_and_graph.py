"""Model zoo: parameter counts, variable names and shapes (reference names pinned by
T/benchmark_cnn_test.py:262-311; sizes from SURVEY.md 2.5)."""
import pytest

from kf_benchmarks_b200 import datasets
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.nn.params import ParameterStore


def _graph(name, data="imagenet", batch=2):
    ds = datasets.create_dataset(None, data)
    m = model_config.get_model_config(name, ds, None)
    return m, m.build_graph(ds.num_classes, batch_size=batch)


@pytest.mark.parametrize("name,params,tensors", [
    ("trivial", 4263877, 6), ("vgg16", 138361641, 32), ("alexnet", 61842345, 16),
    ("resnet50", 25559081, 161), ("resnet50_v1.5", 25559081, 161), ("inception3", 23819401, 190),
    ("googlenet", 6999577, 116), ("vgg11", 132867433, 22), ("vgg19", 143671337, 38),
    ("resnet101", 44551209, 314), ("resnet152", 60194857, 467), ("overfeat", 145924969, 16),
])
def test_imagenet_model_sizes(name, params, tensors):
    m, g = _graph(name)
    assert g.num_trainable() == params
    assert sum(1 for p in g.params.values() if p.trainable) == tensors
    assert g.logits.shape == (2, 1001)


def test_default_batch_sizes_and_image_sizes():
    expect = {"trivial": (32, 227), "vgg16": (64, 224), "alexnet": (512, 227),
              "resnet50": (64, 224), "resnet101": (32, 224), "inception3": (32, 299),
              "inception4": (32, 299), "googlenet": (32, 224), "overfeat": (32, 231),
              "lenet": (32, 28)}
    for name, (bs, size) in expect.items():
        m, _ = _graph(name)
        assert (m.get_batch_size(), m.image_size) == (bs, size), name
    m, _ = _graph("resnet20", "cifar10")
    assert (m.get_batch_size(), m.image_size, m.get_fp16_loss_scale()) == (128, 32, 128)


def test_resnet50_variable_names_follow_reference_layout():
    _, g = _graph("resnet50")
    names = list(g.params)
    assert names[0] == "cg/conv0/conv2d/kernel"
    assert "cg/conv0/batchnorm0/gamma" in names and "cg/conv0/batchnorm0/moving_variance" in names
    assert "cg/resnet_v10/conv1/conv2d/kernel" in names        # projection shortcut comes first
    assert "cg/resnet_v10/conv4/batchnorm4/beta" in names
    assert "cg/resnet_v115/conv52/conv2d/kernel" in names
    assert names[-2:] == ["cg/affine0/weights", "cg/affine0/biases"]
    k = g.params["cg/resnet_v10/conv3/conv2d/kernel"]
    assert k.tf_shape == (3, 3, 64, 64) and k.shape == (64, 3, 3, 64)   # HWIO vs OHWI
    fc = g.params["cg/affine0/weights"]
    assert fc.tf_shape == (2048, 1001) and fc.shape == (1001, 2048)
    # no biases on batch-normed convs, BN params are per channel
    assert "cg/conv0/biases" not in names
    assert g.params["cg/conv0/batchnorm0/gamma"].shape == (64,)


def test_other_name_scopes():
    _, g = _graph("resnet50_v2")
    assert "cg/batchnorm0/beta" not in g.params            # v2 pre-activation BN sits outside
    assert any(n.startswith("cg/resnet_v20/") for n in g.params)
    assert not any(n.endswith("gamma") and "/batchnorm1/" in n and n.count("/") == 2
                   for n in g.params)
    _, g = _graph("inception3")
    assert "cg/incept_v3_a0/conv5/conv2d/kernel" in g.params
    # the reference bumps the counter of the *suffixed* name, so every module is "...0"
    assert "cg/incept_v3_e0/conv93/batchnorm93/beta" in g.params         # 94 convs, all BN
    assert not any(n.endswith("gamma") for n in g.params)    # builder default: no scale
    _, g = _graph("vgg16")
    assert "cg/conv0/biases" in g.params and "cg/affine2/weights" in g.params


def test_fusion_rules():
    _, g = _graph("resnet50")
    ops = [n.op for n in g.nodes]
    assert "add" not in ops and "act" not in ops             # folded into the bn nodes
    joins = [n for n in g.nodes if n.op == "bn" and len(n.inputs) == 2]
    assert len(joins) == 16 and all(n.attrs["relu"] for n in joins)
    _, g = _graph("resnet50_v2")
    assert sum(1 for n in g.nodes if n.op == "add") == 16    # pre-activation: plain adds


def test_cifar_and_registry():
    _, g = _graph("resnet20", "cifar10")
    assert g.logits.shape == (2, 11) and g.num_trainable() == 269787
    _, g = _graph("densenet40_k12", "cifar10")
    assert g.num_trainable() == 1021083
    with pytest.raises(ValueError):
        _graph("resnet20")                                   # cifar model on imagenet
    with pytest.raises(ValueError):
        _graph("ssd300")
    model_config.register_model("my_trivial", "imagenet",
                                model_config._model_name_to_imagenet_model["trivial"])
    assert _graph("my_trivial")[1].num_trainable() == 4263877
    with pytest.raises(ValueError):
        model_config.register_model("my_trivial", "imagenet", None)


def test_arena_layout_is_backward_completion_order():
    _, g = _graph("resnet20", "cifar10")
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    order = [s.spec.name for s in st.order]
    assert order[0] == "cg/affine0/biases" and order[-1] == "cg/conv0/conv2d/kernel"
    offs = [st.slots[n].offset for n in order]
    assert offs == sorted(offs) and all(o % 128 == 0 for o in offs)
    lo, hi = st.span(["cg/affine0/biases", "cg/affine0/weights"])
    assert lo == 0 and hi == st.slots[order[2]].offset


def test_channel_legalization_is_transparent():
    """DenseNet's growth rate 12 gets zero phantom channels for the CUDA executor; the padded
    graph computes the same loss and the same logical gradients, and the phantom lanes of every
    gradient are exactly zero."""
    import numpy as np
    import torch
    from kf_benchmarks_b200.nn.executor_ref import RefExecutor
    from kf_benchmarks_b200.nn.legalize import legalize, needs_legalization

    _, g = _graph("densenet40_k12", "cifar10", batch=2)
    assert needs_legalization(g)
    pg = legalize(g)
    assert not needs_legalization(pg) and pg.num_trainable() == g.num_trainable()
    _, r50 = _graph("resnet50")
    assert legalize(r50) is r50                       # mainstream models are untouched
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(7)
    pst = ParameterStore(pg, "cpu", fp16_shadow=False)
    pst.from_tf_arrays(st.to_tf_arrays())
    gen = torch.Generator().manual_seed(0)
    img = torch.randn(2, 32, 32, 3, generator=gen)
    lab = torch.randint(0, 10, (2,), generator=gen)
    res = []
    for gg, ss in ((g, st), (pg, pst)):
        ex = RefExecutor(gg, ss, 11)
        ex.set_inputs(img, lab)
        ex.forward(True)
        ex.loss_backward(1.0)
        res.append((ex.read_loss()[0], ss.to_tf_arrays(which="grad")))
    assert abs(res[0][0] - res[1][0]) < 1e-5
    for k, a in res[0][1].items():
        assert np.abs(a - res[1][1][k]).max() <= 1e-4 * np.abs(a).max() + 1e-6, k
    logical = sum(float(np.abs(v).sum()) for v in res[1][1].values())
    assert abs(float(pst.grad.abs().sum()) - logical) <= 1e-4 * logical
