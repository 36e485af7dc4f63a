"""SSD data path (default boxes, matching / encoding, crop, jitter, COCO TFRecords) and the NumPy
COCO metric (reference: T/ssd_dataloader.py, T/coco_metric.py)."""
import io
import json

import numpy as np
import pytest
import torch

from kf_benchmarks_b200 import coco_metric
from kf_benchmarks_b200.data import ssd_dataloader as sd
from kf_benchmarks_b200.data import tfrecord


def test_default_boxes_layout():
    db = sd.DefaultBoxes()
    assert db("xywh").shape == (8732, 4) and db("ltrb").shape == (8732, 4)
    xywh = db("xywh")
    # first map: 38x38, first default box is the square of scale 21/300, row-major centres
    assert np.allclose(xywh[0], [0.5 / 37.5, 0.5 / 37.5, 0.07, 0.07], atol=1e-6)
    assert np.allclose(xywh[1], [0.5 / 37.5, 1.5 / 37.5, 0.07, 0.07], atol=1e-6)
    assert np.allclose(xywh[38], [1.5 / 37.5, 0.5 / 37.5, 0.07, 0.07], atol=1e-6)
    # second default box of the first map: sqrt(21*45)/300
    assert np.allclose(xywh[38 * 38][2:], [np.sqrt(21 * 45) / 300] * 2, atol=1e-6)
    # the last map has one position and 4 boxes; sizes are clipped to [0, 1]
    assert np.allclose(xywh[-4][:2], [0.5, 0.5]) and xywh.max() <= 1.0
    counts = [s * s * d for s, d in zip(sd.FEATURE_SIZES, sd.NUM_DEFAULTS)]
    assert sum(counts) == 8732


def test_iou_and_box_coder_roundtrip():
    a = np.array([[0.1, 0.1, 0.5, 0.5], [0.0, 0.0, 1.0, 1.0]], np.float32)
    b = np.array([[0.1, 0.1, 0.5, 0.5], [0.3, 0.3, 0.7, 0.7]], np.float32)
    iou = sd.calc_iou(a, b)
    assert iou[0, 0] == pytest.approx(1.0)
    assert iou[0, 1] == pytest.approx(0.04 / (0.16 + 0.16 - 0.04))
    assert iou[1, 1] == pytest.approx(0.16)
    anchors = sd.default_boxes()("ltrb")[1000:1010]
    gt = np.clip(anchors + 0.03, 0, 1)
    rel = sd.encode_boxes(gt, anchors)
    assert np.allclose(sd.decode_boxes(rel, anchors), gt, atol=1e-5)
    # identical box -> zero target; scale factors (10, 10, 5, 5)
    assert np.allclose(sd.encode_boxes(anchors, anchors), 0, atol=1e-5)
    shifted = anchors.copy()
    shifted[:, [0, 2]] += 0.01
    h = anchors[:, 2] - anchors[:, 0]
    assert np.allclose(sd.encode_boxes(shifted, anchors)[:, 0], 0.01 / h * 10, rtol=1e-4)


def test_encode_labels_matches_and_forces_a_match():
    anchors = sd.default_boxes()("ltrb")
    big = anchors[8732 - 40].copy()          # an anchor of a coarse map: exact match
    tiny = np.array([0.50, 0.50, 0.51, 0.51], np.float32)   # below every IoU threshold
    classes, boxes, n = sd.encode_labels(np.stack([big, tiny]), np.array([7, 3], np.float32))
    assert classes.shape == (8732, 1) and boxes.shape == (8732, 4)
    assert classes[8732 - 40, 0] == 7 and np.allclose(boxes[8732 - 40], 0, atol=1e-5)
    forced = np.flatnonzero(classes[:, 0] == 3)
    assert len(forced) == 1                              # only the forced best anchor
    assert n == float((classes[:, 0] > 0).sum()) and n >= 2
    iou = sd.calc_iou(big[None], anchors)[0]
    matched7 = set(np.flatnonzero(classes[:, 0] == 7))
    assert matched7 == set(np.flatnonzero(iou >= 0.5)) - set(forced)
    c0, b0, n0 = sd.encode_labels(np.zeros((0, 4), np.float32), np.zeros((0,), np.float32))
    assert n0 == 0 and not c0.any() and not b0.any()


def test_crop_flip_jitter_keep_boxes_consistent():
    rng = np.random.RandomState(0)
    image = rng.rand(120, 160, 3).astype(np.float32)
    boxes = np.array([[0.2, 0.2, 0.6, 0.7], [0.5, 0.1, 0.9, 0.4]], np.float32)
    classes = np.array([1.0, 2.0], np.float32)
    for _ in range(20):
        img, bx, cl = sd.ssd_crop(image, boxes, classes, rng)
        assert img.shape == (300, 300, 3) and len(bx) == len(cl) >= 1
        assert bx.min() >= -1e-6 and bx.max() <= 1 + 1e-6
        assert (bx[:, 2] > bx[:, 0]).all() and (bx[:, 3] > bx[:, 1]).all()
    flipped, fb = sd.random_horizontal_flip(image, boxes, np.random.RandomState(1))
    if flipped is not image:
        assert np.allclose(fb[:, 1], 1 - boxes[:, 3]) and np.allclose(fb[:, 3], 1 - boxes[:, 1])
        assert np.array_equal(flipped, image[:, ::-1])
    j = sd.color_jitter(image, np.random.RandomState(2))
    assert j.shape == image.shape and np.isfinite(j).all()
    # hue / saturation round trip: zero jitter is the identity
    same = sd.color_jitter(image, np.random.RandomState(3), brightness=0, contrast=0,
                           saturation=1e-9, hue=0)
    assert np.allclose(same, image, atol=1e-4)
    norm = sd.normalize_image(np.full((2, 2, 3), 0.5, np.float32))
    assert np.allclose(norm[0, 0], (0.5 - np.array(sd.NORMALIZATION_MEAN)) /
                       np.array(sd.NORMALIZATION_STD), atol=1e-6)


def _coco_record(image, boxes, labels, source_id):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(image).save(buf, format="JPEG")
    feats = {"image/encoded": [buf.getvalue()], "image/source_id": [str(source_id).encode()],
             "image/object/class/label": [int(l) for l in labels]}
    for k, col in zip(("ymin", "xmin", "ymax", "xmax"), np.asarray(boxes).T):
        feats["image/object/bbox/" + k] = [float(v) for v in col]
    return tfrecord.encode_example(feats)


def test_coco_input_source_reads_tfrecords(tmp_path):
    rng = np.random.RandomState(4)
    recs = []
    for i in range(6):
        img = (rng.rand(90, 120, 3) * 255).astype(np.uint8)
        boxes = [[0.1, 0.1, 0.8, 0.7], [0.3, 0.4, 0.9, 0.95]]
        recs.append(_coco_record(img, boxes, [1, 18], 100 + i))    # COCO ids: person, dog
    tfrecord.write_records(str(tmp_path / "train-00000-of-00001"), recs)
    tfrecord.write_records(str(tmp_path / "val-00000-of-00001"), recs[:3])
    src = sd.CocoInputSource(str(tmp_path), 4, "cpu", seed=1)
    images, boxes, classes, nm = src.next_batch()
    assert images.shape == (4, 300, 300, 3) and boxes.shape == (4, 8732, 4)
    assert classes.shape == (4, 8732, 1) and nm.shape == (4,) and float(nm.min()) >= 1
    present = set(np.unique(classes.numpy()).astype(int)) - {0}
    assert present <= {sd.CLASS_MAP[1], sd.CLASS_MAP[18]} and present
    ev = sd.CocoInputSource(str(tmp_path), 2, "cpu", seed=1, train=False)
    im, b, c = ev.next_batch()
    assert im.shape == (2, 300, 300, 3) and b.shape == (2, 200, 4) and c.shape == (2, 200, 1)
    assert [m[0] for m in ev.last_eval_meta] == [100, 101]
    assert ev.last_eval_meta[0][1] == (90, 120, 3)


def test_nms_decode():
    boxes = np.array([[0.1, 0.1, 0.5, 0.5], [0.11, 0.1, 0.5, 0.5], [0.6, 0.6, 0.9, 0.9]], np.float32)
    scores = np.zeros((3, 3), np.float32)
    scores[:, 1] = [0.9, 0.8, 0.7]            # two overlapping + one separate box of class 1
    scores[2, 2] = 0.04                       # below the score threshold
    b, l, s = coco_metric.decode_single(boxes, scores)
    assert len(b) == 2 and set(l) == {1} and sorted(s.tolist()) == pytest.approx([0.7, 0.9])
    eb, el, es = coco_metric.decode_single(boxes, np.zeros((3, 3), np.float32))
    assert es[0] == coco_metric.DUMMY_SCORE


def _ann(i, img, cat, bbox, crowd=0):
    return dict(id=i, image_id=img, category_id=cat, bbox=bbox, area=bbox[2] * bbox[3],
                iscrowd=crowd)


def test_coco_eval_corner_cases():
    gts = [_ann(1, 1, 1, [10, 10, 50, 50]), _ann(2, 1, 1, [100, 100, 40, 40]),
           _ann(3, 2, 2, [20, 20, 200, 100])]
    perfect = [[1, 10, 10, 50, 50, 0.9, 1], [1, 100, 100, 40, 40, 0.8, 1],
               [2, 20, 20, 200, 100, 0.7, 2]]
    r = coco_metric.coco_eval(gts, perfect)
    assert r["COCO/AP"] == pytest.approx(1.0) and r["COCO/AP50"] == pytest.approx(1.0)
    assert r["COCO/ARmax100"] == pytest.approx(1.0)
    # a detection shifted to IoU 0.6 counts at thresholds .50 and .55 and .60 only
    shifted = [[1, 10, 10, 50, 50, 0.9, 1], [1, 100, 100, 40, 40, 0.8, 1],
               [2, 20, 20, 200 * 0.6, 100, 0.7, 2]]
    r2 = coco_metric.coco_eval(gts, shifted)
    assert r2["COCO/AP50"] == pytest.approx(1.0) and r2["COCO/AP75"] == pytest.approx(0.5)
    assert r2["COCO/AP"] == pytest.approx((1.0 + 0.3) / 2)
    # a duplicate detection is a false positive ranked below the true positive: AP unchanged;
    # ranked above a true positive of the same class it lowers precision
    dup_low = perfect + [[1, 10, 10, 50, 50, 0.1, 1]]
    assert coco_metric.coco_eval(gts, dup_low)["COCO/AP"] == pytest.approx(1.0)
    fp_high = perfect + [[1, 300, 300, 30, 30, 0.99, 1]]
    r3 = coco_metric.coco_eval(gts, fp_high)
    assert r3["COCO/AP50"] < 1.0 and r3["COCO/AP50"] > 0.5
    # detections on a crowd region are ignored, not penalised
    gts_c = gts + [_ann(4, 1, 1, [300, 300, 100, 100], crowd=1)]
    on_crowd = perfect + [[1, 310, 310, 30, 30, 0.95, 1]]
    assert coco_metric.coco_eval(gts_c, on_crowd)["COCO/AP"] == pytest.approx(1.0)
    # missing half of the objects of a class: recall capped
    r4 = coco_metric.coco_eval(gts, perfect[1:])
    assert r4["COCO/ARmax100"] == pytest.approx((0.5 + 1.0) / 2)
    # area ranges: 50x50 and 40x40 are "medium", 200x100 is "large", nothing is "small"
    assert r["COCO/APm"] == pytest.approx(1.0) and r["COCO/APl"] == pytest.approx(1.0)
    assert r["COCO/APs"] == -1.0


def test_compute_map_end_to_end(tmp_path):
    anchors = sd.default_boxes()("ltrb")
    k = 8732 - 30
    box = [float(v) for v in anchors[k]]
    ann = {"annotations": [_ann(1, 5, 18, [box[1] * 640, box[0] * 480, (box[3] - box[1]) * 640,
                                           (box[2] - box[0]) * 480])]}
    path = tmp_path / "instances_val2017.json"
    path.write_text(json.dumps(ann))
    scores = np.zeros((8732, 81), np.float32)
    scores[:, 0] = 1.0
    scores[k] = 0
    scores[k, sd.CLASS_MAP[18]] = 0.9
    rel = np.zeros((8732, 4), np.float32)                 # zero offsets decode to the anchors
    preds = [{"pred_boxes": sd.decode_boxes(rel, anchors), "pred_scores": scores,
              "source_id": 5, "raw_shape": (480, 640, 3)}]
    out = coco_metric.compute_map(preds, str(path))
    assert out["COCO/AP"] == pytest.approx(1.0)


def test_ssd_eval_hooks_on_the_cpu_executor(tmp_path):
    """eval_step decodes the heads of a forward pass into per-image detections; eval_finish
    turns what was collected into COCO metrics (here against an empty annotation set)."""
    import torch
    from kf_benchmarks_b200 import datasets, params as P
    from kf_benchmarks_b200.models import model_config
    from kf_benchmarks_b200.nn.executor_ref import RefExecutor
    from kf_benchmarks_b200.nn.params import ParameterStore
    ds = datasets.create_dataset(None, "coco")
    (tmp_path / "annotations").mkdir()
    (tmp_path / "annotations" / "instances_val2017.json").write_text(json.dumps(
        {"annotations": [_ann(1, 0, 1, [10, 10, 100, 100])]}))
    p = P.make_params(model="ssd300", data_name="coco", device="cpu", data_format="NHWC",
                      data_dir=str(tmp_path))
    m = model_config.get_model_config("ssd300", ds, p)
    g = m.build_graph(81, phase_train=False, batch_size=1)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(0)
    ex = RefExecutor(g, st, 81)
    image = torch.rand(1, 300, 300, 3)
    assert m.eval_step(ex, (image, None, None), src=None) == 0.0
    boxes, scores = m.predict(ex)
    assert boxes.shape == (1, 8732, 4) and scores.shape == (1, 8732, 81)
    assert np.allclose(scores.sum(-1), 1.0, atol=1e-4)
    out = m.eval_finish(p)
    assert 0.0 <= out["simple_value:COCO/AP"] <= 1.0 and m.predictions == {}
    from kf_benchmarks_b200.benchmark import _eval_summary_record
    rec = _eval_summary_record(7, out)
    assert rec["step"] == 7 and "eval/COCO/AP" in rec and rec["eval/Accuracy@1"] == 0.0


def test_backbone_loading_and_graph_dumps(tmp_path):
    import torch
    from kf_benchmarks_b200 import benchmark, datasets, params as P
    from kf_benchmarks_b200.models import model_config
    from kf_benchmarks_b200.nn.params import ParameterStore
    ds = datasets.create_dataset(None, "coco")
    m = model_config.get_model_config("ssd300", ds, None)
    g = m.build_graph(81, batch_size=1)
    src_store = ParameterStore(g, "cpu", fp16_shadow=False)
    src_store.initialize(11)
    arrays = src_store.to_tf_arrays()
    # (a) a checkpoint of this framework
    np.savez(str(tmp_path / "model.ckpt-5.npz"), **arrays)
    dst = ParameterStore(g, "cpu", fp16_shadow=False)
    dst.initialize(99)
    n = m.load_backbone_model(dst, str(tmp_path / "model.ckpt-5"))
    bb = [k for k in g.params if "resnet34_backbone" in k]
    assert n == len(bb) and n > 100
    for k in bb:
        assert torch.equal(dst.view(k), src_store.view(k))
    head = [k for k in g.params if "resnet34_backbone" not in k][0]
    assert not torch.equal(dst.view(head), src_store.view(head))       # heads keep their init
    # (b) an npz with the TF official-ResNet names (conv2d, conv2d_1, batch_normalization_3/gamma)
    official = {}
    import re
    for k in bb:
        mm = re.search(r"conv(\d+)/conv2d/(.+)$", k)
        if mm:
            official["conv2d" + ("_" + mm.group(1) if int(mm.group(1)) else "") + "/" +
                     mm.group(2)] = arrays["v0/" + k]
            continue
        mm = re.search(r"batchnorm(\d+)/(.+)$", k)
        official["batch_normalization" + ("_" + mm.group(1) if int(mm.group(1)) else "") + "/" +
                 mm.group(2)] = arrays["v0/" + k]
    np.savez(str(tmp_path / "official.npz"), **official)
    dst2 = ParameterStore(g, "cpu", fp16_shadow=False)
    dst2.initialize(5)
    assert m.load_backbone_model(dst2, str(tmp_path / "official")) == len(bb)
    assert torch.equal(dst2.view(bb[0]), src_store.view(bb[0]))
    # graph dumps
    p = P.make_params(model="trivial", device="cpu", data_format="NHWC", batch_size=2,
                      num_batches=1, num_warmup_batches=0, graph_file=str(tmp_path / "g.json"),
                      partitioned_graph_file_prefix=str(tmp_path / "part_"))
    benchmark.BenchmarkCNN(benchmark.setup(p)).run()
    doc = json.loads((tmp_path / "g.json").read_text())
    assert [n["op"] for n in doc["nodes"]][:2] == ["reshape", "affine"]
    assert "cg/affine0/weights" in doc["variables"]
    part = json.loads((tmp_path / "part_0.json").read_text())
    assert part["rank"] == 0 and part["buckets"]


def test_native_crop_and_resize_and_ssd_train_transform(tmp_path):
    """The SSD training image goes through one native call (tf.image.crop_and_resize arithmetic,
    mirror, un-clamped colour jitter, per-channel normalisation): checked against the NumPy
    restatement, and the pooled COCO source yields the same batches with 1 and 4 threads."""
    from kf_benchmarks_b200 import build as build_lib
    from kf_benchmarks_b200.data import native_image as ni
    build_lib.build_host()
    assert ni.available()
    img = np.random.RandomState(1).randint(0, 256, (48, 64, 3)).astype(np.uint8)
    cases = [((0.1, 0.2, 0.8, 0.9), False, None, True),
             ((0.0, 0.0, 1.0, 1.0), True, (2, 10.0, 1.2, 0.03, 0.8), False),
             ((-0.2, 0.3, 0.7, 1.3), True, (2, -8.0, 0.7, -0.04, 1.3), True)]   # partly outside
    for box, flip, distort, clamp in cases:
        op = ni.make_crop_and_resize_op(img, box, (30, 31), flip, distort, (1 / 255 / 0.2,) * 3,
                                        (-2.0, -1.9, -1.8), clamp)
        got, want = ni.transform(img, op), ni.transform_reference(img, op)
        assert got.shape == (30, 31, 3) and np.abs(got - want).max() < 1e-4
    # the identity window reproduces the image (then normalised)
    op = ni.make_crop_and_resize_op(img, (0, 0, 1, 1), (48, 64), scale3=(1, 1, 1))
    assert np.array_equal(ni.transform(img, op), img.astype(np.float32))

    rng = np.random.RandomState(0)
    recs = []
    for i in range(6):
        im = rng.randint(0, 255, (90, 120, 3)).astype(np.uint8)
        boxes = np.array([[0.1, 0.1, 0.6, 0.5], [0.3, 0.4, 0.9, 0.95]], np.float32)
        recs.append(_coco_record(im, boxes, [1, 18], 100 + i))
    tfrecord.write_records(str(tmp_path / "train-00000-of-00001"), recs)
    got = {}
    for threads in (1, 4):
        src = sd.CocoInputSource(str(tmp_path), 4, "cpu", seed=5, num_threads=threads)
        got[threads] = [tuple(t.clone() for t in src.next_batch()) for _ in range(3)]
        src.close()
    for a, b in zip(got[1], got[4]):
        assert all(torch.equal(x, y) for x, y in zip(a, b))
    images, boxes, classes, nm = got[1][0]
    assert images.dtype == torch.float16 and images.shape == (4, 300, 300, 3)
    assert float(nm.min()) >= 1 and torch.isfinite(images.float()).all()
    # normalised pixels: (v - mean) / std of values in [0, 1] plus jitter stay within a few units
    assert float(images.float().abs().max()) < 6
