"""CIFAR-10 python-pickle reader (T/datasets.py:155-181): data_batch_1..5 / test_batch, each a
dict with ``data`` uint8 [N, 3072] (CHW planes) and ``labels``."""
from __future__ import annotations

import os
import pickle

import numpy as np


def read_data_files(data_dir: str, subset: str = "train"):
    if subset == "train":
        names = ["data_batch_%d" % i for i in range(1, 6)]
    elif subset == "validation":
        names = ["test_batch"]
    else:
        raise ValueError('Invalid data subset "%s"' % subset)
    images, labels = [], []
    for n in names:
        with open(os.path.join(data_dir, n), "rb") as f:
            d = pickle.load(f, encoding="bytes")
        images.append(np.asarray(d[b"data"], dtype=np.uint8))
        labels.append(np.asarray(d[b"labels"], dtype=np.int32))
    imgs = np.concatenate(images).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)  # NHWC
    return np.ascontiguousarray(imgs), np.concatenate(labels)
