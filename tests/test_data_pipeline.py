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
