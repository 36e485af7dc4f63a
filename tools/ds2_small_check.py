"""Reduced-geometry DeepSpeech2 executor check (see tests/test_gpu_executor.py)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import executor_check  # noqa: E402


def small(m):
    m.max_time_steps, m.num_feature_bins, m.max_label_length = 96, 33, 8
    m.rnn_hidden_size, m.num_rnn_layers = 64, 2


print("RESULT " + json.dumps(executor_check.compare("deepspeech2", "librispeech", 4, tweak=small,
                                                    verbose="-v" in sys.argv)))
