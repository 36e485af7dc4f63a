"""Model families beyond the classic CNN zoo: graph sizes / names and one CPU
training step through the PyTorch executor for each input / loss type."""
import numpy as np
import pytest
import torch

from kf_benchmarks_b200 import datasets
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.nn.executor_ref import RefExecutor
from kf_benchmarks_b200.nn.params import ParameterStore


def _model(name, data):
    ds = datasets.create_dataset(None, data)
    return ds, model_config.get_model_config(name, ds, None)


@pytest.mark.parametrize("name,data,params", [
    ("mobilenet", "imagenet", 4509155),          # 3 506 153 (MobileNet-v2, 1001) + final affine
    ("official_resnet18", "imagenet", 11694249),
    ("official_resnet34", "imagenet", 21802409),
    ("official_resnet50", "imagenet", 25559081),
    ("official_resnet50_v2", "imagenet", 25551401),
    ("nasnet", "imagenet", 5291035),
    ("nasnet", "cifar10", 3347307),
    ("ssd300", "coco", 19972534),
])
def test_model_sizes(name, data, params):
    ds, m = _model(name, data)
    g = m.build_graph(ds.num_classes, batch_size=2)
    assert g.num_trainable() == params


def test_every_reference_model_name_resolves():
    for data, names in (("imagenet", ["mobilenet", "nasnet", "nasnetlarge", "ncf"] +
                         ["official_resnet%d%s" % (d, v) for d in (18, 34, 50, 101, 152, 200)
                          for v in ("", "_v2")]),
                        ("cifar10", ["nasnet", "densenet40_k12"]), ("coco", ["ssd300"]),
                        ("librispeech", ["deepspeech2"])):
        ds = datasets.create_dataset(None, data)
        for n in names:
            assert model_config.get_model_config(n, ds, None) is not None


def test_mobilenet_names_and_defaults():
    ds, m = _model("mobilenet", "imagenet")
    g = m.build_graph(1001, batch_size=2)
    names = list(g.params)
    assert names[0] == "cg/MobilenetV2/Conv/weights"
    assert "cg/MobilenetV2/expanded_conv/depthwise/depthwise_weights" in names
    assert "cg/MobilenetV2/expanded_conv_16/project/BatchNorm/gamma" in names
    assert "cg/MobilenetV2/Logits/Conv2d_1c_1x1/biases" in names
    assert g.params["cg/MobilenetV2/expanded_conv_1/depthwise/depthwise_weights"].tf_shape == \
        (3, 3, 96, 1)
    assert (m.get_batch_size(), m.image_size) == (32, 224)
    assert any(n.op == "act" and n.attrs["kind"] == "relu6" for n in g.nodes)


def test_nasnet_structure():
    ds, m = _model("nasnet", "imagenet")
    g = m.build_graph(1001, batch_size=2)
    assert g.nodes[0].output.shape == (2, 111, 111, 32)           # 3x3/2 VALID stem
    widths = {t.shape[-1] for t in g.tensors if len(t.shape) == 4}
    assert {11, 22, 44, 88, 176, 264, 528, 1056} <= widths
    assert "cg/cell_stem_0/comb_iter_0/left/separable_5x5_1/depthwise_weights" in g.params
    assert "cg/cell_11/comb_iter_4/left/separable_3x3_2/pointwise_weights" in g.params
    # as run by the reference (no arg scope): biases on convolutions, batch norm without scale
    assert "cg/conv0/biases" in g.params and "cg/conv0_bn/gamma" not in g.params
    m.config_overrides = {"as_run": False}
    g2 = m.build_graph(1001, batch_size=2)
    assert "cg/conv0/biases" not in g2.params and "cg/conv0_bn/gamma" in g2.params
    ds, large = _model("nasnetlarge", "imagenet")
    assert (large.image_size, large.get_batch_size()) == (331, 16)
    gl = large.build_graph(1001, batch_size=1)
    assert any(n.op == "drop_path" for n in gl.nodes)
    kp = [n.attrs["keep_prob_fn"] for n in gl.nodes if n.op == "drop_path"]
    assert kp[0](0) == 1.0 and kp[-1](250000) == pytest.approx(0.7, abs=1e-6)
    assert kp[0](250000) > kp[-1](250000)                          # deeper cells drop more


def test_ncf_trains_on_cpu():
    ds, m = _model("ncf", "imagenet")
    small = type(m)(num_users=300, num_items=200)
    g = small.build_graph(None, batch_size=64)
    assert g.num_trainable() == 300 * 192 + 200 * 192 + 256 * 256 + 256 + 256 * 128 + 128 + \
        128 * 64 + 64 + 128 + 1
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(1)
    src = small.make_input_source(64, "cpu", 5)
    ex = RefExecutor(g, st, 2)
    losses = []
    for _ in range(25):
        ex.set_inputs(*src.next_batch())
        ex.forward(True)
        ex.loss_backward(1.0)
        losses.append(ex.read_loss()[0])
        st.w32 -= 0.5 * st.grad
    assert losses[-1] < losses[0] * 0.7
    # only the looked-up rows of an embedding table receive gradient
    users = src.next_batch()[0].long().unique()
    gtab = st.view("cg/embedding_user/embeddings", "grad")
    untouched = torch.ones(300, dtype=torch.bool)
    untouched[users] = False
    assert float(gtab[untouched].abs().sum()) == 0.0 and float(gtab[users].abs().sum()) > 0.0


def test_ssd_loss_matches_hand_computation():
    from kf_benchmarks_b200.nn.losses_ref import ssd_loss
    torch.manual_seed(0)
    n, a, L = 2, 12, 5
    loc, conf = torch.randn(n, a, 4), torch.randn(n, a, L)
    gt_loc = torch.randn(n, a, 4)
    gt = torch.zeros(n, a)
    gt[0, 3], gt[1, 7], gt[1, 8] = 2, 1, 4
    nm = torch.tensor([1.0, 2.0])
    total, cls, box = ssd_loss(loc, conf, gt_loc, gt, nm, negs_per_positive=3)
    want_box, want_cls = [], []
    for i in range(n):
        pos = gt[i] > 0
        d = (loc[i] - gt_loc[i]).abs()
        h = torch.where(d <= 1, 0.5 * d * d, d - 0.5).sum(1)
        want_box.append(float(h[pos].sum() / nm[i]))
        xe = torch.nn.functional.cross_entropy(conf[i], gt[i].long(), reduction="none")
        neg = xe.clone()
        neg[pos] = 0
        k = int(nm[i]) * 3
        top = torch.topk(neg, k).values.sum()
        want_cls.append(float((xe[pos].sum() + top) / nm[i]))
    assert float(box) == pytest.approx(np.mean(want_box), rel=1e-5)
    assert float(cls) == pytest.approx(np.mean(want_cls), rel=1e-5)
    assert float(total) == pytest.approx(np.mean(want_box) + np.mean(want_cls), rel=1e-5)


def test_ssd_graph_heads_and_lr():
    ds, m = _model("ssd300", "coco")
    g = m.build_graph(81, batch_size=1)
    sizes = [g.tensors[t].shape[1] for t in g.loss["loc"]]
    assert sizes == [38, 19, 10, 5, 3, 1]
    assert [g.tensors[t].shape[-1] for t in g.loss["conf"]] == [324, 486, 486, 486, 324, 324]
    assert sum(s * s * d for s, d in zip(sizes, g.loss["num_dboxes"])) == 8732
    assert g.label_specs[0][0] == (1, 8732, 4)
    assert m.get_learning_rate(0, 32) == 0.0
    assert m.get_learning_rate(10 ** 5, 32) == pytest.approx(1e-3)
    assert m.get_learning_rate(170000, 32) == pytest.approx(1e-4)
    assert m.get_learning_rate(50000, 64) == pytest.approx(2e-3)


def test_deepspeech2_tiny_trains_and_decoder():
    ds, m = _model("deepspeech2", "librispeech")
    full = m.build_graph(29, batch_size=1)
    assert full.logits.shape == (1, 874, 29)
    assert [n.op for n in full.nodes].count("lstm") == 5
    assert full.params["cg/rnn_1/fw/kernel_x"].tf_shape == (2592, 3200)
    assert full.params["cg/rnn_2/bw/kernel_h"].tf_shape == (800, 3200)
    m.max_time_steps, m.num_feature_bins, m.max_label_length = 48, 21, 4
    m.rnn_hidden_size, m.num_rnn_layers = 16, 2
    g = m.build_graph(29, batch_size=2)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(2)
    src = m.make_input_source(2, "cpu", 3)
    ex = RefExecutor(g, st, 29)
    losses = []
    for _ in range(8):
        ex.set_inputs(*src.next_batch())
        ex.forward(True)
        ex.loss_backward(1.0)
        losses.append(ex.read_loss()[0])
        st.w32 -= 0.02 * st.grad
    assert losses[-1] < losses[0]
    from kf_benchmarks_b200.models.experimental.deepspeech import DeepSpeechDecoder
    dec = DeepSpeechDecoder(" abcdefghijklmnopqrstuvwxyz'-")
    probs = np.zeros((6, 29))
    for t, k in enumerate([8, 8, 28, 9, 9, 28]):       # "h h _ i i _"
        probs[t, k] = 1.0
    assert dec.decode_logits(probs) == "hi"
    assert dec.cer("hallo", "hello") == 1 and dec.wer("a b c", "a x c") == 1
    out = m.postprocess({"probs": [probs], "labels": [[8, 9]]})
    assert out["cer"] == 0.0


def test_ctc_reference_ignores_infeasible_samples():
    from kf_benchmarks_b200.nn.ctc_ref import ctc_input_lengths, ctc_loss
    z = torch.randn(2, 6, 5)
    labels = torch.tensor([[0, 1, 2, 3], [0, 0, 0, 0]])
    loss = ctc_loss(z, labels, torch.tensor([6, 6]), torch.tensor([4, 4]))
    # sample 1 needs 4 symbols + 3 separating blanks = 7 > 6 frames -> contributes 0
    only0 = ctc_loss(z[:1], labels[:1], torch.tensor([6]), torch.tensor([4]))
    assert float(loss) == pytest.approx(float(only0) / 2, rel=1e-5)
    assert ctc_input_lengths(torch.tensor([[3494], [1747]]), 874, 3494).tolist() == [874, 437]


def test_librispeech_tfrecords_and_eval_hooks(tmp_path):
    from kf_benchmarks_b200 import params as P
    from kf_benchmarks_b200.data import tfrecord
    rng = np.random.RandomState(0)
    recs = []
    for n_frames, labels in ((30, [8, 9]), (48, [1, 2, 3, 4]), (20, [5])):
        feats = rng.rand(n_frames, 21).astype(np.float32)
        recs.append(tfrecord.encode_sequence_example(
            {"labels": labels, "input_length": [n_frames], "label_length": [len(labels)]},
            {"features": [[float(v) for v in row] for row in feats]}))
        ctx, lists = tfrecord.parse_sequence_example(recs[-1])
        assert ctx["labels"] == labels and ctx["input_length"] == [n_frames]
        assert np.allclose(np.asarray(lists["features"], np.float32), feats)
    tfrecord.write_records(str(tmp_path / "train-clean-100.tfrecords"), recs)
    tfrecord.write_records(str(tmp_path / "test-clean.tfrecords"), recs[:2])
    ds = datasets.create_dataset(str(tmp_path), "librispeech")
    p = P.make_params(model="deepspeech2", data_name="librispeech", data_dir=str(tmp_path),
                      device="cpu", data_format="NHWC")
    m = model_config.get_model_config("deepspeech2", ds, p)
    m.max_time_steps, m.num_feature_bins, m.max_label_length = 48, 21, 4
    m.rnn_hidden_size, m.num_rnn_layers = 16, 1
    src = m.make_input_source(2, "cpu", seed=1, train=False)
    x, lab, ilen, llen = src.next_batch()
    assert x.shape == (2, 48, 21, 1) and lab.shape == (2, 4)
    assert ilen.view(-1).tolist() == [30, 48] and llen.view(-1).tolist() == [2, 4]
    assert float(x[0, 30:].abs().sum()) == 0.0          # zero padding past the utterance
    g = m.build_graph(29, phase_train=False, batch_size=2)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(1)
    ex = RefExecutor(g, st, 29)
    loss = m.eval_step(ex, (x, lab, ilen, llen), src)
    assert loss > 0
    out = m.eval_finish(p)
    assert "cer" in out and "wer" in out and out["cer"] >= 0
