#!/usr/bin/env python
"""Turn a kf-benchmarks-b200 checkpoint (``model.ckpt-N.npz``, keys = the reference's TF variable
names, HWIO / [in,out] layouts) into a TensorFlow V2-format checkpoint bundle that the reference's
``tf.train.Saver`` / ``load_checkpoint`` (T/benchmark_cnn.py:905-950) can restore.

Needs TensorFlow, which is NOT available on the B200 image: run it wherever TF is installed.

    python tools/export_tf_checkpoint.py train_dir/model.ckpt-100.npz out_dir/model.ckpt-100
    python tools/export_tf_checkpoint.py --list train_dir/model.ckpt-100.npz     # no TF needed
"""
import argparse
import sys

import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("npz")
    ap.add_argument("out", nargs="?")
    ap.add_argument("--list", action="store_true", help="print variable names / shapes and exit")
    a = ap.parse_args()
    with np.load(a.npz) as z:
        arrays = {k: z[k] for k in z.files}
    if a.list or not a.out:
        for k in sorted(arrays):
            print("%-70s %-20s %s" % (k, arrays[k].shape, arrays[k].dtype))
        return 0
    try:
        import tensorflow as tf
    except ImportError:
        print("TensorFlow is required to write a TF checkpoint bundle; this image has none. "
              "Copy the .npz to a machine with TensorFlow and run this script there.",
              file=sys.stderr)
        return 1
    tf1 = tf.compat.v1
    tf1.disable_eager_execution()
    with tf1.Graph().as_default():
        variables = []
        for name, value in arrays.items():
            dtype = tf.int64 if value.dtype.kind == "i" else tf.float32
            variables.append(tf1.get_variable(name, shape=value.shape, dtype=dtype,
                                              initializer=tf1.constant_initializer(value)))
        saver = tf1.train.Saver(variables, save_relative_paths=True)
        with tf1.Session() as sess:
            sess.run(tf1.global_variables_initializer())
            path = saver.save(sess, a.out, write_meta_graph=False)
    print("wrote", path)
    return 0


if __name__ == "__main__":
    sys.exit(main())
