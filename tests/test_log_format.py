"""The text log format is parsed by users' scripts and by the reference's own tests
(T/test_util.py:101-162): keep it byte compatible."""
import re

import numpy as np

from kf_benchmarks_b200.utils import log_util


def test_perf_timing_statistics():
    times = [0.1, 0.2, 0.1, 0.4]
    mean, unc, jit = log_util.get_perf_timing(32, times)
    speeds = 32 / np.array(times)
    assert mean == 32 / np.mean(times)
    assert unc == np.std(speeds) / 2.0
    assert jit == 1.4826 * np.median(np.abs(speeds - np.median(speeds)))


def test_step_line_matches_reference_regex():
    line = log_util.step_line(10, 64, [0.1, 0.11, 0.09], 7.12345)
    assert re.match(r"^10\timages/sec: \d+\.\d \+/- \d+\.\d \(jitter = \d+\.\d\)\t7\.123$", line)
    line = log_util.step_line(1, 64, [0.1], 2.0, 0.25, 0.5)
    assert line.endswith("\t2.000\t0.250\t0.500")
    assert log_util.header_line(False) == "Step\tImg/sec\ttotal_loss"
    assert log_util.header_line(True, "base_loss").endswith("\ttop_1_accuracy\ttop_5_accuracy")
