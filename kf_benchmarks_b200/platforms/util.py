"""Site-specific hooks (T/platforms/util.py, T/platforms/default/util.py:28-77): where tests put
their outputs, where test data lives, how to run a python module, and a process-level
``initialize``. Override by replacing this module's functions."""
from __future__ import annotations

import os
import sys
import tempfile


def define_platform_params():
    """No extra flags on the default platform."""


def get_test_output_dir():
    base = os.environ.get("TEST_OUTPUTS_DIR", os.path.join(tempfile.gettempdir(),
                                                          "kf_benchmarks_b200_test_outputs"))
    os.makedirs(base, exist_ok=True)
    return base


def get_test_data_dir():
    return os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__)))), "tests", "test_data")


def get_command_to_run_python_module(module):
    return [sys.executable, "-m", "kf_benchmarks_b200." + module]


def get_cluster_manager(params, config_proto=None):
    """T/platforms/default/util.py:52-66. Returns a description of the cluster; workers are
    ranks of the process group (benchmark._apply_cluster_role)."""
    del config_proto
    from ..utils.cnn_util import GrpcClusterManager
    return GrpcClusterManager(params)


def initialize(params, config_proto=None):
    """Process-level initialisation hook; nothing to do by default."""
