"""Training summaries (T/benchmark_cnn.py:2811-2846): scalars at ``--summary_verbosity`` >= 1
(learning rate, loss, loss scale), a gradient-magnitude histogram at >= 2 and per-variable /
per-gradient histograms at >= 3, written every ``--save_summaries_steps`` steps by rank 0 as
JSON lines in ``train_dir/summaries.jsonl`` and - like the reference, which goes through
``tf.summary.FileWriter`` - as a TensorBoard event file ``train_dir/events.out.tfevents.*``
(written by ``utils/tfevents.py`` without TensorFlow) with the reference's tags:
``learning_rate``, the reported loss type, ``loss_scale``, ``log_gradients``, ``<variable>`` and
``<variable>/gradients``."""
from __future__ import annotations

import json
import os

import torch


def _hist(t: torch.Tensor, bins: int = 20):
    t = t.detach().float().flatten()
    if t.numel() == 0:
        return {"min": 0.0, "max": 0.0, "counts": []}
    lo, hi = float(t.min()), float(t.max())
    counts = torch.histc(t.cpu(), bins=bins, min=lo, max=hi if hi > lo else lo + 1.0)
    return {"min": lo, "max": hi, "counts": [int(c) for c in counts]}


class SummaryWriter:
    def __init__(self, params, directory):
        self.verbosity = params.summary_verbosity
        self.every = params.save_summaries_steps
        self.path = None
        if self.verbosity > 0 and self.every > 0 and directory:
            os.makedirs(directory, exist_ok=True)
            self.path = os.path.join(directory, "summaries.jsonl")
        self.loss_tag = getattr(params, "loss_type_to_report", None) or "total_loss"
        self.use_loss_scale = bool(getattr(params, "use_fp16", False))
        self._events = None

    def maybe_write(self, step, lr, loss, engine, store) -> None:
        if self.path is None or step % self.every != 0:
            return
        rec = {"step": step, "learning_rate": lr, "loss": loss, "loss_scale": engine.loss_scale}
        if self.verbosity >= 2:
            g = store.grad.detach().float().abs() / max(engine.loss_scale, 1e-30)
            rec["log_gradients"] = _hist(torch.log10(g[g > 0]))
        if self.verbosity >= 3:
            rec["variables"] = {}
            rec["gradients"] = {}
            for name, slot in store.slots.items():
                if slot.arena != "w":
                    continue
                rec["variables"][name] = _hist(store.view(name))
                rec["gradients"][name] = _hist(store.view(name, "grad"))
        with open(self.path, "a") as f:
            f.write(json.dumps(rec) + "\n")
        self._write_event(step, lr, loss, engine, store)

    def _write_event(self, step, lr, loss, engine, store) -> None:
        from . import tfevents
        if self._events is None:
            self._events = tfevents.EventFileWriter(os.path.dirname(self.path))
        scalars = {"learning_rate": lr, self.loss_tag: loss}
        if self.use_loss_scale:
            scalars["loss_scale"] = engine.loss_scale
        hists = {}
        if self.verbosity >= 2:
            g = store.grad.detach().float().abs() / max(engine.loss_scale, 1e-30)
            hists["log_gradients"] = tfevents.histogram_of(torch.log10(g[g > 0]).cpu().numpy())
        if self.verbosity >= 3:
            for name, slot in store.slots.items():
                if slot.arena != "w":
                    continue
                hists[name] = tfevents.histogram_of(store.view(name).detach().float().cpu().numpy())
                hists[name + "/gradients"] = tfevents.histogram_of(
                    store.view(name, "grad").detach().float().cpu().numpy())
        self._events.add(step, scalars, hists)

    def close(self) -> None:
        if self._events is not None:
            self._events.close()
