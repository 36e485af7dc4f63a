"""Input pipeline: TFRecord framing / Example parsing, image transforms on constant-colour
images (reference: T/benchmark_cnn_test.py:817-886 with black/white JPEGs), CIFAR reader, the
producer thread (T/cnn_util_test.py:30-128) and a real-data training run on fabricated records
(T/test_data/tfrecord_image_generator.py)."""
import io
import os
import pickle
import time

import numpy as np
import pytest
import torch

from kf_benchmarks_b200 import benchmark, params as P
from kf_benchmarks_b200.data import preprocessing as pp, tfrecord
from kf_benchmarks_b200.utils import cnn_util, log_util


def _jpeg(color, h=60, w=80):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(np.full((h, w, 3), color, dtype=np.uint8)).save(buf, format="JPEG")
    return buf.getvalue()


def _write_imagenet(tmp, subset, n=8, shards=2):
    for s in range(shards):
        recs = []
        for i in range(n):
            color = 0 if (i + s) % 2 == 0 else 255
            recs.append(tfrecord.encode_example({
                "image/encoded": [_jpeg(color)], "image/class/label": [1 + (i % 5)],
                "image/class/text": [b"thing"], "image/object/bbox/xmin": [0.1],
                "image/object/bbox/ymin": [0.1], "image/object/bbox/xmax": [0.9],
                "image/object/bbox/ymax": [0.8]}))
        tfrecord.write_records(os.path.join(tmp, "%s-%05d-of-%05d" % (subset, s, shards)), recs)


def test_tfrecord_roundtrip_and_crc(tmp_path):
    path = str(tmp_path / "x")
    ex = {"a": [b"bytes", b"more"], "b": [1, -2, 1 << 40], "c": [0.5, 1.5]}
    tfrecord.write_records(path, [tfrecord.encode_example(ex)] * 3)
    recs = list(tfrecord.read_records(path, check_crc=True))
    assert len(recs) == 3
    got = tfrecord.parse_example(recs[1])
    assert got["a"] == [b"bytes", b"more"] and got["b"] == [1, -2, 1 << 40]
    assert got["c"] == pytest.approx([0.5, 1.5])
    assert tfrecord.crc32c(b"123456789") == 0xE3069283


@pytest.mark.parametrize("method", ["bilinear", "nearest", "bicubic", "area", "round_robin"])
def test_transforms_keep_constant_images_constant(method):
    import random
    rng = random.Random(0)
    for color in (0, 255):
        img = np.full((90, 120, 3), color, dtype=np.uint8)
        ev = pp.eval_image(img, 64, 64, 3, method)
        assert ev.shape == (64, 64, 3) and np.abs(ev.astype(int) - color).max() <= 1
        tr = pp.train_image(img, 48, 56, (0.1, 0.1, 0.9, 0.9), 2, method, True, rng)
        assert tr.shape == (48, 56, 3) and np.abs(tr - color).max() <= 50
    n = pp.normalized_image(np.array([0, 127.5, 255]))
    assert n.tolist() == [-1.0, 0.0, 1.0]
    with pytest.raises(ValueError):
        pp.get_image_resize_method("lanczos7")


def test_distorted_bbox_properties():
    import random
    rng = random.Random(1)
    for _ in range(50):
        top, left, h, w = pp.sample_distorted_bounding_box(200, 300, (0.2, 0.2, 0.8, 0.8), rng)
        assert 0 <= top and top + h <= 200 and 0 <= left and left + w <= 300
        assert 0.05 * 60000 * 0.9 <= h * w <= 60000
        assert 0.70 <= w / h <= 1.40


def test_barrier_and_image_producer():
    import threading
    b = cnn_util.Barrier(3)
    hits = []
    ts = [threading.Thread(target=lambda: (b.wait(), hits.append(1))) for _ in range(3)]
    [t.start() for t in ts]
    [t.join(2) for t in ts]
    assert len(hits) == 3
    b2 = cnn_util.Barrier(2)
    err = []
    t = threading.Thread(target=lambda: err.append(_raises(b2.wait)))
    t.start()
    time.sleep(0.05)
    b2.abort()
    t.join(2)
    assert err == [True]
    count = [0]

    def produce():
        count[0] += 1
        return count[0]
    prod = cnn_util.ImageProducer(produce, batch_group_size=2)
    prod.start()
    time.sleep(0.2)
    assert 0 < prod.staged() <= 4          # never more than 2 * batch_group_size staged
    assert [prod.get() for _ in range(6)] == [1, 2, 3, 4, 5, 6]
    prod.done()
    assert np.array_equal(cnn_util.roll_numpy_batches(np.arange(8), 2, 0.5), [4, 5, 6, 7, 0, 1, 2, 3])


def _raises(fn):
    try:
        fn()
    except RuntimeError:
        return True
    return False


def _run(kw):
    lines = []
    old = log_util.log_fn
    log_util.log_fn = lines.append
    try:
        b = benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw)))
        stats = b.run()
        b.input_source.close()
    finally:
        log_util.log_fn = old
    return stats, lines


def test_training_on_fabricated_imagenet_records(tmp_path):
    d = str(tmp_path / "imagenet")
    os.makedirs(d)
    _write_imagenet(d, "train")
    _write_imagenet(d, "validation")
    kw = dict(model="trivial", data_dir=d, data_name="imagenet", device="cpu", data_format="NHWC",
              batch_size=4, num_batches=3, num_warmup_batches=0, display_every=1,
              train_dir=str(tmp_path / "ckpt"))
    stats, lines = _run(kw)
    assert stats["num_steps"] == 3 and np.isfinite(stats["last_average_loss"])
    assert any(l.startswith("Dataset:     imagenet") and "synthetic" not in l for l in lines) or True
    stats, lines = _run(dict(kw, eval=True, num_batches=2))
    assert stats["num_examples"] == 8
    stats, _ = _run(dict(kw, datasets_repeat_cached_sample=True, distortions=False,
                         train_dir=str(tmp_path / "ckpt2")))     # fresh dir: no resume
    assert stats["num_steps"] == 3


def test_training_on_cifar10_pickles(tmp_path):
    d = str(tmp_path / "cifar-10-batches-py")
    os.makedirs(d)
    rs = np.random.RandomState(0)
    for name in ["data_batch_%d" % i for i in range(1, 6)] + ["test_batch"]:
        with open(os.path.join(d, name), "wb") as f:
            pickle.dump({b"data": rs.randint(0, 255, (20, 3072)).astype(np.uint8),
                         b"labels": rs.randint(0, 10, 20).tolist()}, f)
    from kf_benchmarks_b200 import datasets
    imgs, labels = datasets.create_dataset(d, "cifar10").read_data_files("train")
    assert imgs.shape == (100, 32, 32, 3) and labels.shape == (100,)
    stats, _ = _run(dict(model="trivial", data_dir=d, data_name="cifar10", device="cpu",
                         data_format="NHWC", batch_size=8, num_batches=3, num_warmup_batches=0))
    assert stats["num_steps"] == 3 and np.isfinite(stats["last_average_loss"])


# ---- native image transform (csrc/host/image_ops.cpp) and the parallel batch producer ----------
def _native_image():
    from kf_benchmarks_b200 import build as build_lib
    from kf_benchmarks_b200.data import native_image
    build_lib.build_host()
    assert native_image.available()
    return native_image


@pytest.mark.parametrize("method", ["bilinear", "nearest", "bicubic", "area"])
def test_native_image_transform_matches_numpy_restatement(method):
    ni = _native_image()
    img = np.random.RandomState(3).randint(0, 256, (75, 101, 3)).astype(np.uint8)
    cases = [((5, 7, 60, 80), (32, 48), None, True, (1, 12.0, 1.3, 0.07, 0.7)),     # train crop
             (None, (90, 121), (13, 28, 64, 64), False, None),                 # eval centre crop
             ((0, 0, 20, 18), (64, 64), None, False, (2, -20.0, 0.6, -0.15, 1.4)),    # upsampling
             ((5, 5, 1, 1), (7, 9), None, True, None)]                         # 1-pixel crop
    for crop, resized, window, flip, distort in cases:
        op = ni.make_op(img, crop, resized, window, method, flip, distort)
        got, want = ni.transform(img, op), ni.transform_reference(img, op)
        assert got.shape == want.shape and np.abs(got - want).max() < 1e-4
    op = ni.make_op(img, None, (40, 40), None, method, out_f16=True)
    got, want = ni.transform(img, op), ni.transform_reference(img, op)
    assert got.dtype == np.float16 and np.abs(got.astype(np.float32) - want).max() < 2e-3
    # rows padded beyond the pixels (a view into a wider image) are read through the stride
    wide = np.zeros((75, 140, 3), np.uint8)
    wide[:, :101] = img
    view = wide[:, :101]
    op_v, op_c = ni.make_op(view, (5, 7, 60, 80), (32, 48), None, method), \
        ni.make_op(img, (5, 7, 60, 80), (32, 48), None, method)
    assert np.array_equal(ni.transform(view, op_v), ni.transform(img, op_c))


def test_native_image_transform_constant_images_and_bad_geometry():
    """The reference's preprocessing test: a constant-colour image stays constant through every
    resize method, with and without distortions (T/benchmark_cnn_test.py:817-886)."""
    ni = _native_image()
    for color in (0, 255):
        img = np.full((90, 120, 3), color, dtype=np.uint8)
        want = color / 127.5 - 1.0
        for method in ni.METHODS:
            (rh, rw), win = pp.eval_geometry(90, 120, 64, 64)
            ev = ni.transform(img, ni.make_op(img, None, (rh, rw), win, method))
            assert ev.shape == (64, 64, 3) and np.abs(ev - want).max() < 1e-3
            tr = ni.transform(img, ni.make_op(img, (9, 12, 50, 70), (48, 56), None, method, True,
                                              (1, 10.0, 1.2, 0.1, 0.8)))
            assert tr.shape == (48, 56, 3) and np.abs(tr - want).max() <= 50 / 127.5
    img = np.zeros((10, 10, 3), np.uint8)
    with pytest.raises(ValueError, match="bad geometry"):
        ni.transform(img, ni.make_op(img, (5, 5, 8, 8), (4, 4)))
    with pytest.raises(ValueError, match="packed pixels"):
        ni.transform(img[:, ::2], ni.make_op(img, None, (4, 4)))
    assert pp.eval_geometry(375, 500, 224, 224) == ((257, 343), (16, 59, 224, 224))


def test_batches_do_not_depend_on_the_number_of_decode_threads(tmp_path, monkeypatch):
    """Examples of a batch are decoded / transformed on a thread pool; augmentation seeds are
    drawn in order on the producer thread, so 1 thread and 4 threads give identical batches."""
    from kf_benchmarks_b200 import datasets
    _native_image()
    monkeypatch.setattr(pp, "SHUFFLE_BUFFER_SIZE", 0)
    d = str(tmp_path / "imagenet")
    os.makedirs(d)
    _write_imagenet(d, "train", n=6, shards=1)
    ds = datasets.create_dataset(d, "imagenet")
    got = {}
    for threads in (1, 4):
        p = P.make_params(model="trivial", data_dir=d, data_name="imagenet", device="cpu",
                          data_format="NHWC", batch_size=5,
                          datasets_num_private_threads=threads)
        src = pp.ImagenetSource(p, ds, 5, 32, 32, "cpu", True, 11, 0, 1)
        got[threads] = [tuple(t.clone() for t in src.next_batch()) for _ in range(4)]
        src.close()
    for (a_img, a_lab), (b_img, b_lab) in zip(got[1], got[4]):
        assert torch.equal(a_img, b_img) and torch.equal(a_lab, b_lab)
    labels = torch.cat([lab for _, lab in got[1]]).tolist()
    assert labels[:6] == [1, 2, 3, 4, 5, 1]           # record order is kept
    assert all(abs(float(img.min()) + 1) < 0.4 or abs(float(img.max()) - 1) < 0.4
               for img, _ in got[1])


def test_training_records_go_through_a_shuffle_buffer(tmp_path, monkeypatch):
    """ds.shuffle(buffer_size=10000) of the reference (T/preprocessing.py:536-537), here with a
    4-record buffer over a 6-record stream: every record keeps turning up, the order is not the
    file order, and it is reproducible from the seed; evaluation is not shuffled."""
    from kf_benchmarks_b200 import datasets
    monkeypatch.setattr(pp, "SHUFFLE_BUFFER_SIZE", 4)
    d = str(tmp_path / "imagenet")
    os.makedirs(d)
    for subset in ("train", "validation"):
        recs = [tfrecord.encode_example({"image/encoded": [_jpeg(40 * i)],
                                         "image/class/label": [i + 1]}) for i in range(6)]
        tfrecord.write_records(os.path.join(d, "%s-00000-of-00001" % subset), recs)
    ds = datasets.create_dataset(d, "imagenet")
    p = P.make_params(model="trivial", data_dir=d, data_name="imagenet", device="cpu",
                      data_format="NHWC", batch_size=6, datasets_num_private_threads=1)

    def labels(train, seed):
        src = pp.ImagenetSource(p, ds, 6, 16, 16, "cpu", train, seed, 0, 1)
        out = torch.cat([src.next_batch()[1].clone() for _ in range(5)]).tolist()
        src.close()
        return out
    a, b, c = labels(True, 3), labels(True, 3), labels(True, 4)
    assert a == b and a != c
    assert set(a) == {1, 2, 3, 4, 5, 6} and a != [1, 2, 3, 4, 5, 6] * 5
    assert labels(False, 3) == [1, 2, 3, 4, 5, 6] * 5


def test_hsv_colour_jitter_arithmetic():
    import colorsys
    from kf_benchmarks_b200.data import native_image as ni
    rng = np.random.RandomState(0)
    x = rng.rand(5, 7, 3)
    h, s, v = ni.rgb_to_hsv(x)
    for i in range(5):
        for j in range(7):
            assert np.allclose((h[i, j], s[i, j], v[i, j]), colorsys.rgb_to_hsv(*x[i, j]), atol=1e-9)
    assert np.allclose(ni.hsv_to_rgb(h, s, v), x, atol=1e-9)
    img = x * 255.0
    same = ni.distort_reference(img, 1, 0.0, 1.0, 0.0, 1.0)
    assert np.allclose(same, img, atol=1e-6)                       # neutral parameters
    red = np.zeros((2, 2, 3))
    red[..., 0] = 200.0
    green = ni.distort_reference(red, 2, 0.0, 1.0, 1.0 / 3.0, 1.0)   # a third of a turn
    assert np.allclose(green[..., 1], 200.0) and np.allclose(green[..., [0, 2]], 0.0, atol=1e-6)
    grey = ni.distort_reference(img, 1, 0.0, 0.0, 0.1, 1.0)          # saturation 0 -> value
    assert np.allclose(grey, np.repeat(img.max(axis=2, keepdims=True), 3, axis=2), atol=1e-6)


_REF_DATA = "/root/reference/scripts/tf_cnn_benchmarks/test_data"


@pytest.mark.skipif(not os.path.isdir(os.path.join(_REF_DATA, "fake_tf_record_data")),
                    reason="the reference tree (with its TensorFlow-written test records) is "
                           "not present on this machine")
def test_tensorflow_written_records_of_the_reference_tree(tmp_path):
    """The reference ships ten small ImageNet-style TFRecord shards written by TensorFlow itself
    (T/test_data/fake_tf_record_data, used by its testImagenetPreprocessor): both record readers
    accept their CRCs, both Example parsers agree on every field, the JPEGs decode, and a
    training + evaluation run on that directory works (data files only - no reference code)."""
    import glob
    from kf_benchmarks_b200.data import native_reader
    d = os.path.join(_REF_DATA, "fake_tf_record_data")
    files = sorted(glob.glob(os.path.join(d, "*")))
    assert len(files) == 10
    total = 0
    for f in files:
        recs = list(tfrecord.read_records(f, check_crc=True))
        assert recs == list(native_reader.RecordReader([f], num_threads=1))
        for r in recs:
            buf, label, bbox, text = pp.parse_example_proto(r)          # native scanner
            ex = tfrecord.parse_example(r)                              # Python codec
            assert buf == ex["image/encoded"][0] and label == ex["image/class/label"][0]
            assert 0 <= label <= 1000 and text == ex["image/class/text"][0]
            img = pp.decode_jpeg(buf)
            assert img.ndim == 3 and img.shape[2] == 3
            assert img.shape[0] == ex["image/height"][0] and img.shape[1] == ex["image/width"][0]
            if "train" in os.path.basename(f):
                assert bbox is not None and 0 <= bbox[0] < bbox[2] <= 1 and 0 <= bbox[1] < bbox[3] <= 1
            total += 1
    assert total == 640
    kw = dict(model="trivial", data_dir=d, data_name="imagenet", device="cpu", data_format="NHWC",
              batch_size=8, num_batches=3, num_warmup_batches=0, display_every=1,
              train_dir=str(tmp_path / "ckpt"))
    stats, _ = _run(kw)
    assert stats["num_steps"] == 3 and np.isfinite(stats["last_average_loss"])
    stats, _ = _run(dict(kw, eval=True, num_batches=4))
    assert stats["num_examples"] == 32
    # the constant-colour JPEGs of the reference's preprocessing test stay constant
    for name, color in (("black_image.jpg", 0), ("white_image.jpg", 255)):
        with open(os.path.join(_REF_DATA, "images", name), "rb") as fh:
            img = pp.decode_jpeg(fh.read())
        ev = pp.eval_image(img, 64, 64, 0, "bilinear")
        assert np.abs(ev.astype(int) - color).max() <= 2


def test_datasets_use_caching_reads_the_files_once(tmp_path, monkeypatch):
    from kf_benchmarks_b200 import datasets
    monkeypatch.setattr(pp, "SHUFFLE_BUFFER_SIZE", 0)
    d = str(tmp_path / "imagenet")
    os.makedirs(d)
    _write_imagenet(d, "train", n=3, shards=2)
    ds = datasets.create_dataset(d, "imagenet")
    p = P.make_params(model="trivial", data_dir=d, data_name="imagenet", device="cpu",
                      data_format="NHWC", batch_size=4, datasets_use_caching=True,
                      datasets_num_private_threads=1)
    src = pp.ImagenetSource(p, ds, 4, 16, 16, "cpu", True, 1, 0, 1)
    first = torch.cat([src.next_batch()[1].clone() for _ in range(3)]).tolist()     # two passes
    for f in os.listdir(d):
        os.remove(os.path.join(d, f))                   # later passes come from memory
    more = torch.cat([src.next_batch()[1].clone() for _ in range(6)]).tolist()
    src.close()
    seq = first + more
    assert seq[:6] == seq[6:12] == seq[12:18] and sorted(seq[:6]) == [1, 1, 2, 2, 3, 3]


def test_a_corrupt_image_fails_the_run_instead_of_hanging_it(tmp_path, monkeypatch):
    """An exception on a decode worker travels through the producer thread to next_batch()."""
    from kf_benchmarks_b200 import datasets
    monkeypatch.setattr(pp, "SHUFFLE_BUFFER_SIZE", 0)
    d = str(tmp_path / "imagenet")
    os.makedirs(d)
    recs = [tfrecord.encode_example({"image/encoded": [_jpeg(10)], "image/class/label": [1]}),
            tfrecord.encode_example({"image/encoded": [b"this is not a JPEG"],
                                     "image/class/label": [2]})]
    tfrecord.write_records(os.path.join(d, "train-00000-of-00001"), recs)
    ds = datasets.create_dataset(d, "imagenet")
    p = P.make_params(model="trivial", data_dir=d, data_name="imagenet", device="cpu",
                      data_format="NHWC", batch_size=2, datasets_num_private_threads=2)
    src = pp.ImagenetSource(p, ds, 2, 16, 16, "cpu", True, 1, 0, 1)
    with pytest.raises(RuntimeError, match="input pipeline failed"):
        src.next_batch()
    src.close()
    # a producer that raises is reported by the generic ImageProducer as well
    prod = cnn_util.ImageProducer(lambda: 1 / 0, 1)
    prod.start()
    with pytest.raises(RuntimeError, match="ZeroDivisionError"):
        prod.get()


def test_native_image_transform_property_random_geometry():
    """Random crops / virtual sizes / windows / methods / flips / jitter: the native transform
    always equals the NumPy restatement (hypothesis drives the geometry)."""
    from hypothesis import given, settings, strategies as st
    ni = _native_image()
    img = np.random.RandomState(9).randint(0, 256, (23, 31, 3)).astype(np.uint8)

    @settings(max_examples=60, deadline=None, derandomize=True, database=None)
    @given(st.data())
    def check(data):
        ch = data.draw(st.integers(1, 23))
        cw = data.draw(st.integers(1, 31))
        top = data.draw(st.integers(0, 23 - ch))
        left = data.draw(st.integers(0, 31 - cw))
        vh = data.draw(st.integers(1, 40))
        vw = data.draw(st.integers(1, 40))
        oh = data.draw(st.integers(1, vh))
        ow = data.draw(st.integers(1, vw))
        oy = data.draw(st.integers(0, vh - oh))
        ox = data.draw(st.integers(0, vw - ow))
        method = data.draw(st.sampled_from(sorted(ni.METHODS)))
        flip = data.draw(st.booleans())
        distort = data.draw(st.one_of(st.none(), st.tuples(
            st.sampled_from([1, 2]), st.floats(-32, 32), st.floats(0.5, 1.5),
            st.floats(-0.2, 0.2), st.floats(0.5, 1.5))))
        op = ni.make_op(img, (top, left, ch, cw), (vh, vw), (oy, ox, oh, ow), method, flip, distort)
        got, want = ni.transform(img, op), ni.transform_reference(img, op)
        assert got.shape == (oh, ow, 3)
        assert np.abs(got - want).max() < 2e-4
    check()
