"""Logging and throughput statistics in the reference's text format.

``log_fn`` is the single sink (monkey-patchable, like T/cnn_util.py:37-38); the per-step
line, header and footer reproduce T/benchmark_cnn.py:834-846,887-902,2276-2281,2350-2354
byte for byte because the reference's tests and users' scripts parse them.
"""
from __future__ import annotations

import numpy as np

LOSS_AND_ACCURACY_DIGITS_TO_SHOW = 3


def log_fn(log: str) -> None:
    print(log, flush=True)


def get_perf_timing(batch_size, step_train_times, scale=1):
    """mean speed, its standard error and a robust jitter (1.4826 * MAD)."""
    times = np.array(step_train_times)
    speeds = batch_size / times
    speed_mean = scale * batch_size / np.mean(times)
    speed_uncertainty = np.std(speeds) / np.sqrt(float(len(speeds)))
    speed_jitter = 1.4826 * np.median(np.abs(speeds - np.median(speeds)))
    return speed_mean, speed_uncertainty, speed_jitter


def get_perf_timing_str(speed_mean, speed_uncertainty, speed_jitter, scale=1):
    if scale == 1:
        return ("images/sec: %.1f +/- %.1f (jitter = %.1f)" %
                (speed_mean, speed_uncertainty, speed_jitter))
    return "images/sec: %.1f" % speed_mean


def step_line(step, batch_size, step_train_times, lossval, top1=None, top5=None) -> str:
    mean, unc, jit = get_perf_timing(batch_size, step_train_times)
    s = "%i\t%s\t%.*f" % (step, get_perf_timing_str(mean, unc, jit),
                          LOSS_AND_ACCURACY_DIGITS_TO_SHOW, lossval)
    if top1 is not None:
        s += "\t%.*f\t%.*f" % (LOSS_AND_ACCURACY_DIGITS_TO_SHOW, top1,
                               LOSS_AND_ACCURACY_DIGITS_TO_SHOW, top5)
    return s


def header_line(with_accuracy: bool, loss_type: str = "total_loss") -> str:
    h = "Step\tImg/sec\t" + loss_type
    if with_accuracy:
        h += "\ttop_1_accuracy\ttop_5_accuracy"
    return h
