"""Tracing / profiling outputs (reference: --trace_file at warm-up step -2 via TF Timeline,
--tfprof_file for the first 10 timed steps, T/benchmark_cnn.py:801-884,1208-1228).

* ``--trace_file``: a Chrome-trace JSON (or a plain-text kernel table when
  ``--use_chrome_trace_format=false``) of the second-to-last warm-up step, recorded with
  ``torch.profiler`` (CUPTI).
* ``--tfprof_file``: per-kernel table (calls, total/avg device microseconds, share) over the
  first 10 timed steps, printed (top 20) and written to the file, followed - when the model
  graph is known - by a per-kernel-class table against the analytic roofline of the step
  (``utils/roofline.py``: FLOPs, bytes, roof, measured, achieved fraction).

Nothing here runs unless one of the flags is given; numbers taken while a profiler is active
are not benchmark results.
"""
from __future__ import annotations

import os

import torch

from . import log_util

_NUM_STEPS_TO_PROFILE = 10


class StepProfiler:
    def __init__(self, params, device, graph=None):
        self.p = params
        self.device = device
        self.graph = graph
        self.trace_prof = None
        self.table_prof = None

    def _activities(self):
        acts = [torch.profiler.ProfilerActivity.CPU]
        if self.device.type == "cuda":
            acts.append(torch.profiler.ProfilerActivity.CUDA)
        return acts

    def before_step(self, step: int) -> None:
        if self.p.trace_file and step == -2:
            self.trace_prof = torch.profiler.profile(activities=self._activities())
            self.trace_prof.__enter__()
        if self.p.tfprof_file and step == 0:
            self.table_prof = torch.profiler.profile(activities=self._activities())
            self.table_prof.__enter__()

    def after_step(self, step: int) -> None:
        if self.trace_prof is not None and step == -2:
            self._sync()
            self.trace_prof.__exit__(None, None, None)
            path = self.p.trace_file
            log_util.log_fn("Dumping trace to %s" % path)
            d = os.path.dirname(path)
            if d:
                os.makedirs(d, exist_ok=True)
            if self.p.use_chrome_trace_format:
                self.trace_prof.export_chrome_trace(path)
            else:
                with open(path, "w") as f:
                    f.write(self._table(self.trace_prof, 1000))
            self.trace_prof = None
        if self.table_prof is not None and step == _NUM_STEPS_TO_PROFILE - 1:
            self._finish_table()

    def _sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def _table(self, prof, rows):
        key = "cuda_time_total" if self.device.type == "cuda" else "cpu_time_total"
        return prof.key_averages().table(sort_by=key, row_limit=rows)

    def _finish_table(self):
        self._sync()
        self.table_prof.__exit__(None, None, None)
        path = self.p.tfprof_file
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        roof = self._roofline_table(self.table_prof)
        with open(path, "w") as f:
            f.write(self._table(self.table_prof, 1000))
            if roof:
                f.write("\n\n" + roof + "\n")
        log_util.log_fn("Top kernels over the first %d steps:" % _NUM_STEPS_TO_PROFILE)
        log_util.log_fn(self._table(self.table_prof, 20))
        if roof:
            log_util.log_fn(roof)
        self.table_prof = None

    def _roofline_table(self, prof) -> str:
        """Device time per kernel class and step against the analytic roof of that class."""
        if self.graph is None:
            return ""
        try:
            from collections import defaultdict

            from . import roofline
            bw, fl = roofline.peaks()
            costs = roofline.model_costs(self.graph, bw, fl)
            per_class = defaultdict(float)
            other = 0.0
            for ev in prof.key_averages():
                us = float(getattr(ev, "device_time_total", 0.0) or
                           getattr(ev, "cuda_time_total", 0.0) or 0.0)
                if us <= 0.0:
                    continue
                cls = roofline.class_of(ev.key)
                if cls is None:
                    other += us
                else:
                    per_class[cls] += us
            steps = float(_NUM_STEPS_TO_PROFILE)
            lines = ["Kernel classes against the analytic roofline (per step; HBM %.0f GB/s, "
                     "16-bit GEMM %.0f TFLOP/s):" % (bw / 1e9, fl / 1e12),
                     "%-44s %9s %9s %9s %11s %9s" % ("class", "GFLOP", "GB", "roof ms",
                                                     "measured ms", "achieved")]
            for cls, (flops, nbytes, roof_s) in costs.items():
                ms = per_class.get(cls, 0.0) / steps / 1e3
                lines.append("%-44s %9.1f %9.3f %9.3f %11s %9s" %
                             (cls, flops / 1e9, nbytes / 1e9, roof_s * 1e3,
                              "%.3f" % ms if ms > 0 else "-",
                              "%.0f %%" % (100 * roof_s * 1e3 / ms) if ms > 0 else "-"))
            lines.append("(device time of kernels outside these classes: %.3f ms per step)" %
                         (other / steps / 1e3))
            return "\n".join(lines)
        except Exception as e:  # a reporting extra must never break the run
            return "(roofline table not available: %r)" % (e,)

    def finish(self) -> None:
        if self.table_prof is not None:
            self._finish_table()
