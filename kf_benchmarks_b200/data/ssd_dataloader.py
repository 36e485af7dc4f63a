"""SSD-300 data path: default boxes, IoU, IoU-biased random crop, colour jitter, normalisation,
target assignment (anchor matching + box encoding) and the COCO TFRecord input source.

Reference: T/ssd_dataloader.py:35-290 (which leans on the external ``object_detection``
package for the matcher / box coder / flip) and ``COCOPreprocessor`` (T/preprocessing.py:742-
900). Everything here is plain NumPy + PIL, so the COCO path works without that package.
Boxes are ``[ymin, xmin, ymax, xmax]`` normalised to [0, 1] (the TF object-detection
convention); anchors are ordered feature map -> default box -> row -> column, which is the
order ``csrc/loss_kernels.cu`` walks the head convolutions in.
"""
from __future__ import annotations

import glob
import itertools
import math
import os
from typing import Optional

import numpy as np
import torch

IMAGE_SIZE = 300
FEATURE_SIZES = (38, 19, 10, 5, 3, 1)
STEPS = (8, 16, 32, 64, 100, 300)
SCALES = (21, 45, 99, 153, 207, 261, 315)
ASPECT_RATIOS = ((2,), (2, 3), (2, 3), (2, 3), (2,), (2,))
NUM_DEFAULTS = (4, 6, 6, 6, 4, 4)
NUM_SSD_BOXES = 8732
SCALE_XY, SCALE_HW = 0.1, 0.2
BOX_CODER_SCALES = (1 / SCALE_XY, 1 / SCALE_XY, 1 / SCALE_HW, 1 / SCALE_HW)
MATCH_THRESHOLD = 0.5
NORMALIZATION_MEAN = (0.485, 0.456, 0.406)
NORMALIZATION_STD = (0.229, 0.224, 0.225)
NUM_CROP_PASSES = 50
CROP_MIN_IOU_CHOICES = (0, 0.1, 0.3, 0.5, 0.7, 0.9)
P_NO_CROP_PER_PASS = 1 / (len(CROP_MIN_IOU_CHOICES) + 1)
MAX_NUM_EVAL_BOXES = 200
# COCO category ids actually used (80 classes + background) and the inverse map
CLASS_INV_MAP = (
    0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 27,
    28, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 41, 42, 43, 44, 46, 47, 48, 49, 50, 51, 52, 53, 54,
    55, 56, 57, 58, 59, 60, 61, 62, 63, 64, 65, 67, 70, 72, 73, 74, 75, 76, 77, 78, 79, 80, 81, 82,
    84, 85, 86, 87, 88, 89, 90)
_MAP = {j: i for i, j in enumerate(CLASS_INV_MAP)}
CLASS_MAP = tuple(_MAP.get(i, -1) for i in range(max(CLASS_INV_MAP) + 1))


class DefaultBoxes:
    """The 8732 default boxes of SSD-300 (T/ssd_dataloader.py:35-79): per feature map two
    square sizes (s_k and sqrt(s_k s_k+1)) plus two boxes per extra aspect ratio, laid out
    default-box-major, then row, then column. ``'xywh'`` is (cy, cx, h, w), ``'ltrb'`` is
    (ymin, xmin, ymax, xmax)."""

    def __init__(self):
        fk = IMAGE_SIZE / np.array(STEPS, dtype=np.float64)
        boxes = []
        for idx, fsize in enumerate(FEATURE_SIZES):
            sk1 = SCALES[idx] / IMAGE_SIZE
            sk2 = SCALES[idx + 1] / IMAGE_SIZE
            sk3 = math.sqrt(sk1 * sk2)
            sizes = [(sk1, sk1), (sk3, sk3)]
            for alpha in ASPECT_RATIOS[idx]:
                w, h = sk1 * math.sqrt(alpha), sk1 / math.sqrt(alpha)
                sizes.append((w, h))
                sizes.append((h, w))
            assert len(sizes) == NUM_DEFAULTS[idx]
            for w, h in sizes:
                for i, j in itertools.product(range(fsize), repeat=2):
                    cx, cy = (j + 0.5) / fk[idx], (i + 0.5) / fk[idx]
                    boxes.append([min(max(v, 0.0), 1.0) for v in (cy, cx, h, w)])
        self.xywh = np.asarray(boxes, dtype=np.float32)
        assert self.xywh.shape == (NUM_SSD_BOXES, 4)
        cy, cx, h, w = self.xywh.T
        self.ltrb = np.stack([cy - h / 2, cx - w / 2, cy + h / 2, cx + w / 2], 1)

    def __call__(self, order="ltrb") -> np.ndarray:
        return self.ltrb if order == "ltrb" else self.xywh


_DEFAULT_BOXES: Optional[DefaultBoxes] = None


def default_boxes() -> DefaultBoxes:
    global _DEFAULT_BOXES
    if _DEFAULT_BOXES is None:
        _DEFAULT_BOXES = DefaultBoxes()
    return _DEFAULT_BOXES


def calc_iou(box1: np.ndarray, box2: np.ndarray) -> np.ndarray:
    """IoU matrix [N, M] of two box sets (T/ssd_dataloader.py:82-110)."""
    b1, b2 = box1[:, None, :], box2[None, :, :]
    lt = np.maximum(b1[..., :2], b2[..., :2])
    rb = np.minimum(b1[..., 2:], b2[..., 2:])
    delta = np.maximum(rb - lt, 0)
    inter = delta[..., 0] * delta[..., 1]
    a1 = (b1[..., 2] - b1[..., 0]) * (b1[..., 3] - b1[..., 1])
    a2 = (b2[..., 2] - b2[..., 0]) * (b2[..., 3] - b2[..., 1])
    return inter / np.maximum(a1 + a2 - inter, 1e-12)


def _crop_and_resize(image: np.ndarray, bounds, size: int) -> np.ndarray:
    """Bilinear crop of the normalised window ``bounds`` = (y0, x0, y1, x1) to size x size."""
    from PIL import Image
    h, w = image.shape[:2]
    y0, x0, y1, x1 = bounds
    box = (x0 * (w - 1), y0 * (h - 1), x1 * (w - 1) + 1, y1 * (h - 1) + 1)
    chans = [np.asarray(Image.fromarray(image[..., c].astype(np.float32), mode="F").resize(
        (size, size), Image.BILINEAR, box=box)) for c in range(image.shape[2])]
    return np.stack(chans, -1)


def ssd_crop(image, boxes, classes, rng: np.random.RandomState):
    """IoU-biased random crop: geometry from ``ssd_crop_geometry``, pixels through the PIL
    stand-in of crop_and_resize (the native path resamples with the exact TF arithmetic)."""
    bounds, kept, cls = ssd_crop_geometry(boxes, classes, rng)
    return _crop_and_resize(image, bounds, IMAGE_SIZE), kept, cls


def ssd_crop_geometry(boxes, classes, rng: np.random.RandomState):
    """-> (window (y0, x0, y1, x1), boxes kept / clipped / renormalised, their classes).
    IoU-biased random crop (T/ssd_dataloader.py:113-227): every pass either keeps the whole
    image (p = 1/7) or draws 50 windows (sides in [0.3, 1]) and a minimum IoU from
    {0, .1, .3, .5, .7, .9}; a window is valid if height/width < 2, every box overlaps it by
    more than the minimum IoU and at least one box centre lies inside; the last valid window
    wins, boxes whose centre is outside are dropped, the rest are clipped and renormalised."""
    n = boxes.shape[0]
    while True:
        if n == 0 or rng.uniform() < P_NO_CROP_PER_PASS:
            bounds, mask = np.array([0, 0, 1, 1], np.float32), np.ones(n, bool)
            break
        width = rng.uniform(0.3, 1, (NUM_CROP_PASSES, 1))
        height = rng.uniform(0.3, 1, (NUM_CROP_PASSES, 1))
        left = rng.uniform(0, 1, (NUM_CROP_PASSES, 1)) * (1 - width)
        top = rng.uniform(0, 1, (NUM_CROP_PASSES, 1)) * (1 - height)
        ltrb = np.concatenate([left, top, left + width, top + height], 1)
        min_iou = CROP_MIN_IOU_CHOICES[rng.randint(len(CROP_MIN_IOU_CHOICES))]
        ious = calc_iou(ltrb, boxes)
        c0 = 0.5 * (boxes[:, 0] + boxes[:, 2])[None, :]
        c1 = 0.5 * (boxes[:, 1] + boxes[:, 3])[None, :]
        masks = (c0 > ltrb[:, 0:1]) & (c0 < ltrb[:, 2:3]) & (c1 > ltrb[:, 1:2]) & \
            (c1 < ltrb[:, 3:4])
        valid = ((height / width < 2)[:, 0] & (ious > min_iou).all(1) & masks.any(1))
        if valid.any():
            k = int(np.flatnonzero(valid)[-1])
            bounds, mask = ltrb[k].astype(np.float32), masks[k]
            break
    kept = boxes[mask]
    kept = np.stack([np.maximum(kept[:, 0], bounds[0]), np.maximum(kept[:, 1], bounds[1]),
                     np.minimum(kept[:, 2], bounds[2]), np.minimum(kept[:, 3], bounds[3])], 1)
    span = np.array([bounds[2] - bounds[0], bounds[3] - bounds[1]] * 2, np.float32)
    off = np.array([bounds[0], bounds[1]] * 2, np.float32)
    kept = (kept - off) / span
    return bounds, kept.astype(np.float32), classes[mask]


def random_horizontal_flip(image, boxes, rng: np.random.RandomState):
    if rng.uniform() < 0.5:
        image = image[:, ::-1]
        boxes = np.stack([boxes[:, 0], 1 - boxes[:, 3], boxes[:, 2], 1 - boxes[:, 1]], 1)
    return image, boxes


def _rgb_to_hsv(rgb):
    r, g, b = rgb[..., 0], rgb[..., 1], rgb[..., 2]
    mx, mn = rgb.max(-1), rgb.min(-1)
    d = mx - mn
    s = np.where(mx > 0, d / np.maximum(mx, 1e-12), 0)
    safe = np.maximum(d, 1e-12)
    h = np.where(mx == r, ((g - b) / safe) % 6, np.where(mx == g, (b - r) / safe + 2,
                                                         (r - g) / safe + 4))
    h = np.where(d > 0, h / 6.0, 0)
    return np.stack([h, s, mx], -1)


def _hsv_to_rgb(hsv):
    h, s, v = hsv[..., 0], hsv[..., 1], hsv[..., 2]
    i = np.floor(h * 6).astype(np.int32) % 6
    f = h * 6 - np.floor(h * 6)
    p, q, t = v * (1 - s), v * (1 - f * s), v * (1 - (1 - f) * s)
    choices = [np.stack(c, -1) for c in ((v, t, p), (q, v, p), (p, v, t), (p, q, v), (t, p, v),
                                         (v, p, q))]
    out = np.zeros_like(hsv)
    for k, c in enumerate(choices):
        out = np.where((i == k)[..., None], c, out)
    return out


def color_jitter(image, rng: np.random.RandomState, brightness=0.125, contrast=0.5,
                 saturation=0.5, hue=0.05):
    """tf.image.random_{brightness,contrast,saturation,hue} in that order
    (T/ssd_dataloader.py:230-244) on a float image in [0, 1]."""
    if brightness > 0:
        image = image + rng.uniform(-brightness, brightness)
    if contrast > 0:
        f = rng.uniform(1 - contrast, 1 + contrast)
        mean = image.mean((0, 1), keepdims=True)
        image = (image - mean) * f + mean
    if saturation > 0 or hue > 0:
        hsv = _rgb_to_hsv(np.clip(image, 0, 1))
        if saturation > 0:
            hsv[..., 1] = np.clip(hsv[..., 1] * rng.uniform(1 - saturation, 1 + saturation), 0, 1)
        if hue > 0:
            hsv[..., 0] = (hsv[..., 0] + rng.uniform(-hue, hue)) % 1.0
        image = _hsv_to_rgb(hsv)
    return image.astype(np.float32)


def normalize_image(image):
    return ((image - np.asarray(NORMALIZATION_MEAN, np.float32)) /
            np.asarray(NORMALIZATION_STD, np.float32)).astype(np.float32)


def encode_boxes(gt: np.ndarray, anchors_ltrb: np.ndarray) -> np.ndarray:
    """Faster-RCNN box coder with scale factors (10, 10, 5, 5): (ty, tx, th, tw)."""
    eps = 1e-8
    ha = anchors_ltrb[:, 2] - anchors_ltrb[:, 0]
    wa = anchors_ltrb[:, 3] - anchors_ltrb[:, 1]
    ya = anchors_ltrb[:, 0] + 0.5 * ha
    xa = anchors_ltrb[:, 1] + 0.5 * wa
    h = gt[:, 2] - gt[:, 0]
    w = gt[:, 3] - gt[:, 1]
    y = gt[:, 0] + 0.5 * h
    x = gt[:, 1] + 0.5 * w
    ha, wa, h, w = (np.maximum(v, eps) for v in (ha, wa, h, w))
    s = BOX_CODER_SCALES
    return np.stack([(y - ya) / ha * s[0], (x - xa) / wa * s[1], np.log(h / ha) * s[2],
                     np.log(w / wa) * s[3]], 1).astype(np.float32)


def decode_boxes(rel: np.ndarray, anchors_ltrb: np.ndarray) -> np.ndarray:
    """Inverse of ``encode_boxes``: predicted offsets -> (ymin, xmin, ymax, xmax)."""
    ha = anchors_ltrb[:, 2] - anchors_ltrb[:, 0]
    wa = anchors_ltrb[:, 3] - anchors_ltrb[:, 1]
    ya = anchors_ltrb[:, 0] + 0.5 * ha
    xa = anchors_ltrb[:, 1] + 0.5 * wa
    s = BOX_CODER_SCALES
    y = rel[..., 0] / s[0] * ha + ya
    x = rel[..., 1] / s[1] * wa + xa
    h = np.exp(rel[..., 2] / s[2]) * ha
    w = np.exp(rel[..., 3] / s[3]) * wa
    return np.stack([y - h / 2, x - w / 2, y + h / 2, x + w / 2], -1).astype(np.float32)


def encode_labels(gt_boxes: np.ndarray, gt_labels: np.ndarray):
    """Anchor labelling (T/ssd_dataloader.py:257-290): arg-max matcher with matched ==
    unmatched threshold .5 and a forced match for every ground-truth box (its best anchor),
    Faster-RCNN encoding of the matched box, zeros elsewhere. Returns
    (classes [8732, 1], boxes [8732, 4], number of matched anchors)."""
    anchors = default_boxes()("ltrb")
    classes = np.zeros((NUM_SSD_BOXES, 1), np.float32)
    boxes = np.zeros((NUM_SSD_BOXES, 4), np.float32)
    if gt_boxes.shape[0] == 0:
        return classes, boxes, np.float32(0.0)
    iou = calc_iou(gt_boxes.astype(np.float32), anchors)            # [n_gt, A]
    match = iou.argmax(0)
    match[iou.max(0) < MATCH_THRESHOLD] = -1
    forced = iou.argmax(1)                                          # best anchor of each gt
    match[forced] = np.arange(gt_boxes.shape[0])
    pos = match >= 0
    classes[pos, 0] = np.asarray(gt_labels, np.float32).reshape(-1)[match[pos]]
    boxes[pos] = encode_boxes(gt_boxes[match[pos]], anchors[pos])
    return classes, boxes, np.float32(pos.sum())


def preprocess_train(image_u8: np.ndarray, boxes: np.ndarray, labels: np.ndarray,
                     rng: np.random.RandomState):
    """uint8 HWC image + COCO category ids -> (image [300,300,3] fp32 normalised, encoded boxes,
    encoded classes, number of matched anchors), T/preprocessing.py:781-809."""
    image = image_u8.astype(np.float32) / 255.0
    classes = np.asarray([CLASS_MAP[int(c)] if 0 <= int(c) < len(CLASS_MAP) else -1
                          for c in labels], np.float32)
    image, boxes, classes = ssd_crop(image, boxes.astype(np.float32), classes, rng)
    image, boxes = random_horizontal_flip(image, boxes, rng)
    image = normalize_image(color_jitter(image, rng))
    enc_classes, enc_boxes, num_matched = encode_labels(boxes, classes)
    return image, enc_boxes, enc_classes, num_matched


def preprocess_train_native(image_u8: np.ndarray, boxes: np.ndarray, labels: np.ndarray,
                            rng: np.random.RandomState, out_ptr: int, out_f16: bool):
    """``preprocess_train`` with the pixel work (crop_and_resize, flip, colour jitter,
    normalisation, cast) in one native call that writes the image at ``out_ptr``
    (csrc/host/image_ops.cpp). Draws from ``rng`` in the same order as ``preprocess_train``.
    Returns (encoded boxes, encoded classes, matched anchors)."""
    from . import native_image as ni
    classes = np.asarray([CLASS_MAP[int(c)] if 0 <= int(c) < len(CLASS_MAP) else -1
                          for c in labels], np.float32)
    bounds, boxes, classes = ssd_crop_geometry(boxes.astype(np.float32), classes, rng)
    flip = rng.uniform() < 0.5
    if flip:
        boxes = np.stack([boxes[:, 0], 1 - boxes[:, 3], boxes[:, 2], 1 - boxes[:, 1]], 1)
    # colour jitter of color_jitter(): brightness, contrast, saturation, hue - no clamp
    brightness = 255.0 * rng.uniform(-0.125, 0.125)
    contrast = rng.uniform(0.5, 1.5)
    saturation = rng.uniform(0.5, 1.5)
    hue = rng.uniform(-0.05, 0.05)
    distort = (2, brightness, saturation, hue, contrast)
    mean, std = np.asarray(NORMALIZATION_MEAN, np.float64), np.asarray(NORMALIZATION_STD, np.float64)
    op = ni.make_crop_and_resize_op(image_u8, bounds, (IMAGE_SIZE, IMAGE_SIZE), flip, distort,
                                    scale3=1.0 / (255.0 * std), bias3=-mean / std, clamp=False,
                                    out_f16=out_f16)
    ni.transform_into(image_u8, op, out_ptr)
    enc_classes, enc_boxes, num_matched = encode_labels(boxes, classes)
    return enc_boxes, enc_classes, num_matched


def preprocess_eval(image_u8: np.ndarray, boxes: np.ndarray, labels: np.ndarray):
    """Resize to 300x300, normalise, pad / trim the ground truth to 200 boxes
    (T/preprocessing.py:811-834)."""
    image = _crop_and_resize(image_u8.astype(np.float32) / 255.0, (0, 0, 1, 1), IMAGE_SIZE)
    image = normalize_image(image)
    classes = np.asarray([CLASS_MAP[int(c)] if 0 <= int(c) < len(CLASS_MAP) else -1
                          for c in labels], np.float32).reshape(-1, 1)
    b = np.zeros((MAX_NUM_EVAL_BOXES, 4), np.float32)
    c = np.zeros((MAX_NUM_EVAL_BOXES, 1), np.float32)
    k = min(len(boxes), MAX_NUM_EVAL_BOXES)
    b[:k], c[:k] = boxes[:k], classes[:k]
    return image, b, c


# ---------------------------------------------------------------------------------------------
# COCO TFRecords (tensorflow/models object_detection create_coco_tf_record layout)
# ---------------------------------------------------------------------------------------------
def parse_coco_example(record: bytes):
    """-> (image uint8 HWC, boxes [n,4] normalised ymin/xmin/ymax/xmax, category ids [n],
    source id, raw (h, w, 3))."""
    from . import tfrecord
    from .preprocessing import decode_jpeg
    ex = tfrecord.parse_example(record)
    img = decode_jpeg(ex["image/encoded"][0])
    cols = [np.asarray(ex.get("image/object/bbox/" + k, []), np.float32)
            for k in ("ymin", "xmin", "ymax", "xmax")]
    boxes = np.stack(cols, 1) if len(cols[0]) else np.zeros((0, 4), np.float32)
    labels = np.asarray(ex.get("image/object/class/label", []), np.int64)
    sid = ex.get("image/source_id", [b"0"])[0]
    sid = int(sid.decode() if isinstance(sid, bytes) else sid)
    return img, boxes, labels, sid, (img.shape[0], img.shape[1], 3)


def coco_files(data_dir: str, subset: str):
    pats = ["%s-*" % subset, "*%s*record*" % subset, "*%s*.tfrecord*" % subset]
    files = []
    for p in pats:
        files = sorted(glob.glob(os.path.join(data_dir, p)))
        if files:
            break
    if not files:
        raise FileNotFoundError("no COCO TFRecords for subset %r under %s" % (subset, data_dir))
    return files


class CocoInputSource:
    """Real-data input source for SSD-300: decodes, augments and labels on the host, yields
    pinned batches (image fp16 NHWC, encoded boxes, encoded classes, matched counts).

    Training batches are produced ahead of the consumer by a background thread; the examples of
    a batch are spread over a thread pool (JPEG decode, the native image transform and the
    NumPy anchor matching all release the GIL for most of their time). Every example gets its
    own seed, drawn in record order, and rejected examples (no box left after the crop) are
    dropped in that order, so batches do not depend on the number of threads. Evaluation walks
    the files once, in order, on the caller's thread."""
    host_resident = True

    def __init__(self, data_dir: str, batch: int, device, seed: int, rank: int = 0, world: int = 1,
                 train: bool = True, num_threads: Optional[int] = None):
        from concurrent.futures import ThreadPoolExecutor
        from . import native_image, tfrecord
        from ..utils.cnn_util import ImageProducer
        self.batch, self.train = batch, train
        self.rng = np.random.RandomState(seed * 9973 + rank)
        files = coco_files(data_dir, "train" if train else "val")
        self.files = files[rank::world] or files
        self._tf = tfrecord
        self._it = self._records()
        self.pin = torch.device(device).type == "cuda" and torch.cuda.is_available()
        self.bytes_per_step = batch * (IMAGE_SIZE * IMAGE_SIZE * 3 * 2 + NUM_SSD_BOXES * 5 * 4 + 4)
        self._native = native_image if native_image.available() else None
        if num_threads is None:
            local = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")) or 1)
            num_threads = max(1, min(16, (os.cpu_count() or 2) // max(local, 1)))
        self._pool = ThreadPoolExecutor(num_threads, thread_name_prefix="kfb-coco") \
            if (train and num_threads > 1) else None
        self._producer = None
        if train:
            self._ring = [self._alloc_slot() for _ in range(5)]
            self._ring_pos = 0
            self._producer = ImageProducer(self._make_train_batch, 1)
            self._producer.start()

    def _alloc_slot(self):
        b, pin = self.batch, self.pin
        return (torch.empty((b, IMAGE_SIZE, IMAGE_SIZE, 3), dtype=torch.float16, pin_memory=pin),
                torch.empty((b, NUM_SSD_BOXES, 4), dtype=torch.float32, pin_memory=pin),
                torch.empty((b, NUM_SSD_BOXES, 1), dtype=torch.float32, pin_memory=pin),
                torch.empty((b,), dtype=torch.float32, pin_memory=pin))

    def _records(self):
        while True:
            order = list(self.files)
            if self.train:
                self.rng.shuffle(order)
            for f in order:
                for rec in self._tf.read_records(f):
                    yield rec
            if not self.train:
                return

    # ---- training: pooled, produced ahead ---------------------------------------------------
    def _train_example(self, job):
        """Worker: (record, seed) -> (image fp16 [300,300,3], boxes, classes, matched) or None."""
        rec, seed = job
        rng = np.random.RandomState(seed)
        img, boxes, labels, _, _ = parse_coco_example(rec)
        if len(boxes) == 0:
            return None
        if self._native is not None:
            image = torch.empty((IMAGE_SIZE, IMAGE_SIZE, 3), dtype=torch.float16)
            eb, ec, nm = preprocess_train_native(np.ascontiguousarray(img), boxes, labels, rng,
                                                 image.data_ptr(), True)
        else:
            im, eb, ec, nm = preprocess_train(img, boxes, labels, rng)
            image = torch.from_numpy(im).half()
        if nm <= 0:
            return None
        return image, eb, ec, nm

    def _make_train_batch(self):
        slot = self._ring[self._ring_pos % len(self._ring)]
        self._ring_pos += 1
        filled = 0
        while filled < self.batch:
            jobs = [(next(self._it), int(self.rng.randint(0, 2 ** 31 - 1)))
                    for _ in range(self.batch - filled)]
            results = list(self._pool.map(self._train_example, jobs)) if self._pool is not None \
                else [self._train_example(j) for j in jobs]
            for r in results:
                if r is None:
                    continue
                image, eb, ec, nm = r
                slot[0][filled] = image
                slot[1][filled] = torch.from_numpy(eb)
                slot[2][filled] = torch.from_numpy(ec)
                slot[3][filled] = float(nm)
                filled += 1
        return slot

    def close(self):
        if self._producer is not None:
            self._producer.done()
        if self._pool is not None:
            self._pool.shutdown(wait=False)

    # ---- evaluation: one ordered pass ---------------------------------------------------------
    def next_batch(self):
        if self.train:
            return self._producer.get()
        imgs, bxs, cls, extra = [], [], [], []
        while len(imgs) < self.batch:
            try:
                rec = next(self._it)
            except StopIteration:
                if not imgs:
                    raise
                break
            img, boxes, labels, sid, raw = parse_coco_example(rec)
            image, eb, ec = preprocess_eval(img, boxes, labels)
            extra.append((sid, raw))
            imgs.append(image)
            bxs.append(eb)
            cls.append(ec)
        out = [torch.from_numpy(np.stack(imgs)).half(), torch.from_numpy(np.stack(bxs)),
               torch.from_numpy(np.stack(cls))]
        self.last_eval_meta = extra
        if self.pin:
            out = [t.pin_memory() for t in out]
        return tuple(out)
