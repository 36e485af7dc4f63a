"""Shared constants (T/constants.py:7-49)."""


class BenchmarkMode:
    TRAIN = "training"
    EVAL = "evaluation"
    TRAIN_AND_EVAL = "training + evaluation"
    FORWARD_ONLY = "forward only"


class NetworkTopology:
    """Kept for flag compatibility. On an NVSwitch box every GPU reaches every peer at full
    NVLink 5 bandwidth, so the reference's DGX-1 / GCP_V100 P2P matrices have no role."""
    DGX1 = "dgx1"
    GCP_V100 = "gcp_v100"


TRAIN_ACCURACY_TOP_1_PREFIX = "train_top_1_accuracy"
EVAL_ACCURACY_TOP_1_PREFIX = "eval_top_1_accuracy"
SIMPLE_VALUE_RESULT_PREFIX = "simple_value:"
EVAL_LOSS_PREFIX = "eval_average_loss"
