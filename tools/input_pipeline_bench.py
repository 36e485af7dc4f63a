#!/usr/bin/env python
"""Host-side throughput of the real-data input pipeline (no GPU needed): images/sec delivered by
the ImageNet source (data/preprocessing.py) from TFRecords into batch buffers.

    python tools/make_fake_imagenet.py /tmp/imnet
    python tools/input_pipeline_bench.py /tmp/imnet --threads 8 [--eval] [--no_distortions]
"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("data_dir")
    ap.add_argument("--batch_size", type=int, default=64)
    ap.add_argument("--image_size", type=int, default=224)
    ap.add_argument("--threads", type=int, default=None)
    ap.add_argument("--seconds", type=float, default=8.0)
    ap.add_argument("--eval", action="store_true")
    ap.add_argument("--no_distortions", action="store_true")
    ap.add_argument("--shuffle_buffer", type=int, default=None)
    ap.add_argument("--gpu", action="store_true",
                    help="device pipeline: --gpu_preprocess transform kernel, nvJPEG decode "
                         "unless KFB_NVJPEG=0")
    a = ap.parse_args()
    from kf_benchmarks_b200 import datasets, params as P
    from kf_benchmarks_b200.data import native_image, preprocessing
    if a.shuffle_buffer is not None:
        preprocessing.SHUFFLE_BUFFER_SIZE = a.shuffle_buffer
    p = P.make_params(model="resnet50", batch_size=a.batch_size, data_dir=a.data_dir,
                      data_name="imagenet", device="gpu" if a.gpu else "cpu",
                      data_format="NHWC", distortions=not a.no_distortions,
                      datasets_num_private_threads=a.threads, gpu_preprocess=a.gpu,
                      use_fp16=a.gpu)
    ds = datasets.create_dataset(a.data_dir, "imagenet")
    src = preprocessing.ImagenetSource(p, ds, a.batch_size, a.image_size, a.image_size,
                                       "cuda" if a.gpu else "cpu", not a.eval, 1, 0, 1)
    src.next_batch()
    t0, n = time.time(), 0
    while time.time() - t0 < a.seconds:
        src.next_batch()
        n += a.batch_size
    if a.gpu:
        import torch
        torch.cuda.synchronize()
    dt = time.time() - t0
    decode = "Pillow"
    if a.gpu and getattr(src, "_jpeg", None) is not None:
        decode = "nvJPEG " + src._jpeg.backend
    print("%s: %.1f images/sec (%d decode threads, native transform %s, %d cores, decode: %s, "
          "transform on %s)" %
          ("eval" if a.eval else "train", n / dt, preprocessing._decode_workers(p),
           "on" if native_image.available() else "off", os.cpu_count(), decode,
           "the GPU" if a.gpu else "host threads"))
    src.close()


if __name__ == "__main__":
    main()
