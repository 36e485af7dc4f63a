"""``--benchmark_log_dir``: JSON run info + metrics, in the spirit of the model-garden
BenchmarkFileLogger the reference writes through (T/benchmark_cnn.py:1594-1608,1694-1724):
``benchmark_run.log`` (one JSON object: model, dataset, params, machine) and ``metric.log``
(one JSON object per line: name, value, unit, global_step, timestamp). Logged like the
reference: ``current_examples_per_sec`` (+ ``top_1_accuracy`` / ``top_5_accuracy`` when shown) at
every displayed step (:847-854), ``average_examples_per_sec`` at the end (:2368-2371) and the
evaluation results ``eval_top_1_accuracy`` / ``eval_top_5_accuracy`` /
``eval_average_examples_per_sec`` (:1915-1922)."""
from __future__ import annotations

import datetime
import json
import os
import platform

import torch


def _now():
    return datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.%fZ")


def log_metric(params, name, value, global_step, unit=None) -> None:
    """Append one metric record to ``<benchmark_log_dir>/metric.log``."""
    d = params.benchmark_log_dir
    if not d:
        return
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "metric.log"), "a") as f:
        f.write(json.dumps({"name": name, "value": float(value), "unit": unit,
                            "global_step": int(global_step), "timestamp": _now()}) + "\n")


def log_evaluation_result(params, result, global_step) -> None:
    log_metric(params, "eval_top_1_accuracy", result["top_1_accuracy"], global_step)
    log_metric(params, "eval_top_5_accuracy", result["top_5_accuracy"], global_step)
    log_metric(params, "eval_average_examples_per_sec", result["images_per_sec"], global_step,
               "examples/sec")


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
