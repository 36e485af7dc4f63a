"""Checkpoints: save / resume and export in the reference's variable-name layout.

The reference saves ``train_dir/model.ckpt-<global_step>`` through ``tf.train.Saver`` plus a
``checkpoint`` state file (T/benchmark_cnn.py:2076-2082,2373-2378) and loads the newest or a
named one (``load_checkpoint`` :905-950). The tensors are keyed by *the reference's TF variable
names* (``v0/cg/conv0/conv2d/kernel`` ..., HWIO / [in,out] layouts, fp32, plus ``global_step``,
``loss_scale``, ``loss_scale_normal_steps`` and ``<var>/Momentum`` style optimizer slots) and
stored, by default (``--checkpoint_format=tf``), as a TensorFlow V2 bundle (``tf_bundle.py``,
no TensorFlow needed: ``model.ckpt-N.index`` + ``.data-00000-of-00001``, CRC32C through the
host library) like the reference's own train_dir, or as one ``.npz`` per checkpoint
(``--checkpoint_format=npz``); either format - including bundles written by the reference - is
restored transparently. ``tools/export_tf_checkpoint.py`` converts between the two. The
``checkpoint`` state file uses TF's text format with relative paths so a moved ``train_dir``
still resolves.
"""
from __future__ import annotations

import os
import re
from typing import Dict, Optional, Tuple

import numpy as np

_SUFFIX = ".npz"


class CheckpointNotFoundException(Exception):
    pass


def _state_path(train_dir: str) -> str:
    return os.path.join(train_dir, "checkpoint")


def _write_state(train_dir: str, names) -> None:
    with open(_state_path(train_dir), "w") as f:
        f.write('model_checkpoint_path: "%s"\n' % names[-1])
        for n in names:
            f.write('all_model_checkpoint_paths: "%s"\n' % n)


def _read_state(train_dir: str):
    path = _state_path(train_dir)
    if not os.path.exists(path):
        return []
    names = []
    with open(path) as f:
        for line in f:
            m = re.match(r'all_model_checkpoint_paths: "(.*)"', line.strip())
            if m:
                names.append(m.group(1))
    return names


def _exists(base: str) -> bool:
    return os.path.exists(base + _SUFFIX) or os.path.exists(base + ".index")


def _remove(base: str) -> None:
    import glob
    for path in [base + _SUFFIX, base + ".index"] + glob.glob(base + ".data-*"):
        try:
            os.remove(path)
        except OSError:
            pass


def save_checkpoint(train_dir: str, global_step: int, store, extra: Optional[Dict] = None,
                    max_to_keep: int = 5, opt_slots: Optional[Dict] = None,
                    fmt: str = "npz", prefix: str = "v0/") -> str:
    """``fmt="npz"`` (this function's default; one ``.npz`` file) or ``"tf"``: a TensorFlow V2
    bundle (``model.ckpt-N.index`` + ``.data-00000-of-00001``, see tf_bundle.py) the reference's
    Saver can restore. The benchmark driver passes ``--checkpoint_format``, whose default is
    ``tf`` - the files the reference's train_dir holds."""
    os.makedirs(train_dir, exist_ok=True)
    arrays = store.to_tf_arrays(prefix)
    if opt_slots:
        from .nn.params import _native_to_tf
        for suffix, arena in opt_slots.items():
            for p in store.graph.params.values():
                if not p.trainable:
                    continue
                s = store.slots[p.name]
                t = arena[s.offset:s.offset + s.numel].view(p.shape).detach().float().cpu()
                arrays["%s%s/%s" % (prefix, p.name, suffix)] = _native_to_tf(p, t).numpy()
    for k, v in (extra or {}).items():
        arrays[k] = np.asarray(v)
    name = "model.ckpt-%d" % global_step
    if fmt == "tf":
        from . import tf_bundle
        tf_bundle.write_bundle(os.path.join(train_dir, name), arrays)
    else:
        tmp = os.path.join(train_dir, name + ".tmp" + _SUFFIX)
        np.savez(tmp, **arrays)
        os.replace(tmp, os.path.join(train_dir, name + _SUFFIX))
    names = [n for n in _read_state(train_dir) if n != name] + [name]
    while max_to_keep and len(names) > max_to_keep:
        _remove(os.path.join(train_dir, names.pop(0)))
    _write_state(train_dir, names)
    return os.path.join(train_dir, name)


def latest_checkpoint(path: str) -> Tuple[str, int]:
    """Resolves a directory (newest checkpoint) or an explicit ``.../model.ckpt-N`` path and
    returns (path without suffix, global step) - the rules of T/benchmark_cnn.py:905-950."""
    if path is None:
        raise CheckpointNotFoundException("No checkpoint directory given")
    m = re.match(r".*ckpt-(\d+)$", path[:-len(_SUFFIX)] if path.endswith(_SUFFIX) else path)
    if m:
        base = path[:-len(_SUFFIX)] if path.endswith(_SUFFIX) else path
        if not _exists(base):
            raise CheckpointNotFoundException("No checkpoint file found at %s" % base)
        return base, int(m.group(1))
    names = _read_state(path)
    if not names:
        raise CheckpointNotFoundException("No checkpoint file found in dir:`%s`" % path)
    name = os.path.basename(names[-1])
    base = os.path.join(path, name)
    if not _exists(base):
        raise CheckpointNotFoundException("No checkpoint file found at %s" % base)
    return base, int(name.split("-")[-1])


def variable_prefix(variable_update: str) -> str:
    """Outer variable scope of the reference: ``v/`` with a parameter server (one shared copy,
    T/variable_mgr.py:212-214), ``v0/`` (tower 0 of the replicas) otherwise (:175-177,
    307-309, 358-365)."""
    return "v/" if variable_update == "parameter_server" else "v0/"


def restore_into(store, arrays: Dict[str, np.ndarray], prefix: str = "v0/"):
    """Loads ``arrays`` into ``store``; a checkpoint written under the other outer scope
    (``v/`` vs ``v0/``) is accepted too, so runs can switch variable_update modes."""
    other = "v/" if prefix == "v0/" else "v0/"
    have = sum(1 for k in arrays if k.startswith(prefix))
    alt = sum(1 for k in arrays if k.startswith(other))
    use = prefix if have >= alt else other
    return store.from_tf_arrays(arrays, prefix=use, strict=False)


def load_arrays(base_path: str) -> Dict[str, np.ndarray]:
    """Tensors of a checkpoint: the framework's ``.npz`` or a TensorFlow V2 bundle (for
    example one written by the reference's Saver)."""
    if base_path.endswith(_SUFFIX):
        base_path = base_path[:-len(_SUFFIX)]
    if not os.path.exists(base_path + _SUFFIX) and os.path.exists(base_path + ".index"):
        from . import tf_bundle
        return tf_bundle.read_bundle(base_path)
    with np.load(base_path + _SUFFIX) as z:
        return {k: z[k] for k in z.files}
