#!/usr/bin/env python
"""Convert checkpoints between this framework's ``.npz`` (keys = the reference's TF variable
names, HWIO / [in,out] layouts) and TensorFlow V2 checkpoint bundles (``.index`` +
``.data-00000-of-00001``) that the reference's ``tf.train.Saver`` / ``load_checkpoint``
(T/benchmark_cnn.py:905-950) reads and writes. No TensorFlow needed
(kf_benchmarks_b200/tf_bundle.py); training can also write bundles directly with
``--checkpoint_format=tf``.

    python tools/export_tf_checkpoint.py train_dir/model.ckpt-100.npz out_dir/model.ckpt-100
    python tools/export_tf_checkpoint.py --to_npz ref_dir/model.ckpt-100 out.npz
    python tools/export_tf_checkpoint.py --list train_dir/model.ckpt-100[.npz]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from kf_benchmarks_b200 import ckpt, tf_bundle  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("src", help="model.ckpt-N.npz, or the prefix of a TF bundle")
    ap.add_argument("out", nargs="?")
    ap.add_argument("--list", action="store_true", help="print variable names / shapes and exit")
    ap.add_argument("--to_npz", action="store_true", help="TF bundle -> .npz")
    a = ap.parse_args()
    arrays = ckpt.load_arrays(a.src)
    if a.list or not a.out:
        for k in sorted(arrays):
            print("%-70s %-20s %s" % (k, arrays[k].shape, arrays[k].dtype))
        return 0
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    if a.to_npz:
        np.savez(a.out, **arrays)
        print("wrote %s (%d tensors)" % (a.out, len(arrays)))
    else:
        tf_bundle.write_bundle(a.out, arrays)
        print("wrote %s.index / .data-00000-of-00001 (%d tensors)" % (a.out, len(arrays)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
