#!/usr/bin/env python
"""Writes a small ImageNet-style TFRecord data set of synthetic photographs (counterpart of the
reference's T/test_data/tfrecord_image_generator.py:204-224): `train-xxxxx-of-yyyyy` /
`validation-...` shards of tf.Example records with `image/encoded` (JPEG), `image/class/label`,
`image/class/text`, `image/height|width` and one bounding box.

    python tools/make_fake_imagenet.py /tmp/imnet --images 64 --shards 4 --size 375x500
"""
import argparse
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from kf_benchmarks_b200.data import tfrecord  # noqa: E402


def natural_image(rng, h, w):
    """Smooth colour field plus fine noise: compresses like a photograph (~85 KB at 375x500)."""
    from PIL import Image
    small = rng.randint(0, 255, (h // 16 + 1, w // 16 + 1, 3)).astype(np.uint8)
    img = np.asarray(Image.fromarray(small).resize((w, h), Image.BICUBIC)).astype(np.int16)
    return np.clip(img + rng.randint(-20, 20, (h, w, 3)), 0, 255).astype(np.uint8)


def example(rng, h, w, label, quality=90):
    from PIL import Image
    buf = io.BytesIO()
    Image.fromarray(natural_image(rng, h, w)).save(buf, format="JPEG", quality=quality)
    y0, x0 = rng.uniform(0.0, 0.3, 2)
    y1, x1 = rng.uniform(0.6, 1.0, 2)
    return tfrecord.encode_example({
        "image/encoded": [buf.getvalue()], "image/format": [b"JPEG"],
        "image/class/label": [int(label)], "image/class/text": [b"class %d" % label],
        "image/height": [h], "image/width": [w], "image/channels": [3],
        "image/object/bbox/ymin": [float(y0)], "image/object/bbox/xmin": [float(x0)],
        "image/object/bbox/ymax": [float(y1)], "image/object/bbox/xmax": [float(x1)]})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out_dir")
    ap.add_argument("--images", type=int, default=64, help="images per subset")
    ap.add_argument("--shards", type=int, default=4)
    ap.add_argument("--size", default="375x500", help="HxW; every other image is transposed")
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    h, w = (int(v) for v in a.size.split("x"))
    os.makedirs(a.out_dir, exist_ok=True)
    rng = np.random.RandomState(a.seed)
    total = 0
    for subset in ("train", "validation"):
        per = (a.images + a.shards - 1) // a.shards
        for s in range(a.shards):
            recs = [example(rng, *((h, w) if i % 2 == 0 else (w, h)), label=1 + (s * per + i) % 1000)
                    for i in range(per)]
            total += sum(len(r) for r in recs)
            tfrecord.write_records(os.path.join(a.out_dir, "%s-%05d-of-%05d" % (subset, s, a.shards)),
                                   recs)
    print("wrote 2 x %d shards, %.1f KB per record on average" %
          (a.shards, total / 1024.0 / (2 * a.shards * per)))


if __name__ == "__main__":
    main()
