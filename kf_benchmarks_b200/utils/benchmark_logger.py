"""``--benchmark_log_dir``: JSON run info + metrics, in the spirit of the model-garden
BenchmarkFileLogger the reference writes through (T/benchmark_cnn.py:1594-1608,1694-1724):
``benchmark_run.log`` (one JSON object: model, dataset, params, machine) and ``metric.log``
(one JSON object per line: name, value, unit, global_step, timestamp)."""
from __future__ import annotations

import datetime
import json
import os
import platform

import torch


def _now():
    return datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.%fZ")


def write(params, stats, bench) -> None:
    d = params.benchmark_log_dir
    os.makedirs(d, exist_ok=True)
    info = {
        "model_name": params.model, "dataset": {"name": bench.dataset.name},
        "test_id": params.benchmark_test_id, "run_date": _now(),
        "run_parameters": [{"name": k, "string_value": str(v)}
                           for k, v in sorted(params._asdict().items())],
        "machine_config": {
            "platform": platform.platform(), "torch": torch.__version__,
            "gpu": torch.cuda.get_device_name(0) if torch.cuda.is_available() else None,
            "gpu_count": torch.cuda.device_count() if torch.cuda.is_available() else 0},
    }
    with open(os.path.join(d, "benchmark_run.log"), "w") as f:
        json.dump(info, f)
        f.write("\n")
    with open(os.path.join(d, "metric.log"), "a") as f:
        for name, unit in (("average_examples_per_sec", "examples/sec"),
                           ("last_average_loss", None), ("num_steps", None)):
            key = {"average_examples_per_sec": "images_per_sec"}.get(name, name)
            f.write(json.dumps({"name": name, "value": float(stats[key]), "unit": unit,
                                "global_step": bench.global_step, "timestamp": _now()}) + "\n")
