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


def test_lines_have_the_shape_of_a_committed_reference_log():
    """The reference tree keeps the output of real runs (T/slurm-2810608-200.out:93-110): the
    header, step and footer lines produced here must be indistinguishable in shape."""
    import os

    import pytest
    path = "/root/reference/scripts/tf_cnn_benchmarks/slurm-2810608-200.out"
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    lines = open(path, errors="replace").read().splitlines()
    start = lines.index("Step\tImg/sec\ttotal_loss")
    assert log_util.header_line(False) == lines[start]
    real_steps = [l for l in lines[start + 1:] if re.match(r"^\d+\timages/sec: ", l)]
    assert len(real_steps) == 11 and real_steps[0].startswith("1\t") and \
        real_steps[-1].startswith("100\t")
    # a pattern that accepts every real step line (losses there are 'nan') ...
    pat = re.compile(r"^\d+\timages/sec: \d+\.\d \+/- \d+\.\d \(jitter = \d+\.\d\)\t(nan|\d+\.\d{3})$")
    assert all(pat.match(l) for l in real_steps)
    # ... also accepts ours, for a finite and for a nan loss
    assert pat.match(log_util.step_line(10, 200, [0.48, 0.479, 0.481], 6.90775))
    assert pat.match(log_util.step_line(1, 200, [0.48], float("nan")))
    # real first line: one timing sample gives '+/- 0.0 (jitter = 0.0)'
    ours = log_util.step_line(1, 200, [200 / 416.4], float("nan"))
    assert ours == real_steps[0]
    total = [l for l in lines if l.startswith("total images/sec: ")]
    assert total and re.match(r"^total images/sec: \d+\.\d{2}$", total[0])
    assert "-" * 64 in lines
