"""Checkpoint save / load / export (T/benchmark_cnn.py:905-950,2076-2082,2373-2378; tests
T/benchmark_cnn_test.py:74-142,697-728)."""
import os
import shutil

import numpy as np
import pytest
import torch

from kf_benchmarks_b200 import ckpt, datasets
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.nn.params import ParameterStore


def _store(seed=3):
    ds = datasets.create_dataset(None, "cifar10")
    g = model_config.get_model_config("resnet20", ds, None).build_graph(11, batch_size=2)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(seed)
    return st


def test_export_uses_tf_names_and_layouts(tmp_path):
    st = _store()
    st.view("cg/conv0/batchnorm0/moving_mean").fill_(0.25)
    arrays = st.to_tf_arrays("v0/")
    k = arrays["v0/cg/conv1/conv2d/kernel"]
    assert k.shape == (3, 3, 16, 16) and k.dtype == np.float32          # HWIO
    native = st.view("cg/conv1/conv2d/kernel")
    assert np.allclose(k[1, 2, 5, 7], float(native[7, 1, 2, 5]))
    assert arrays["v0/cg/conv0/conv2d/kernel"].shape == (3, 3, 3, 16)    # flat stem kernel
    assert arrays["v0/cg/affine0/weights"].shape == (64, 11)            # [in, out]
    assert np.all(arrays["v0/cg/conv0/batchnorm0/moving_mean"] == 0.25)
    # round trip into a differently initialised store
    other = _store(seed=99)
    assert not torch.equal(other.w32, st.w32)
    other.from_tf_arrays(arrays)
    assert torch.equal(other.w32, st.w32) and torch.equal(other.state, st.state)


def test_save_latest_max_to_keep_and_moved_dir(tmp_path):
    st = _store()
    d = str(tmp_path / "train")
    for step in (10, 20, 30, 40):
        ckpt.save_checkpoint(d, step, st, {"global_step": step}, max_to_keep=2)
    files = sorted(f for f in os.listdir(d) if f.startswith("model.ckpt"))
    assert files == ["model.ckpt-30.npz", "model.ckpt-40.npz"]
    path, step = ckpt.latest_checkpoint(d)
    assert step == 40 and path.endswith("model.ckpt-40")
    assert int(ckpt.load_arrays(path)["global_step"]) == 40
    path, step = ckpt.latest_checkpoint(os.path.join(d, "model.ckpt-30"))
    assert step == 30
    moved = str(tmp_path / "moved")
    shutil.move(d, moved)                                  # relative paths in the state file
    assert ckpt.latest_checkpoint(moved)[1] == 40
    with pytest.raises(ckpt.CheckpointNotFoundException):
        ckpt.latest_checkpoint(str(tmp_path / "empty"))
    with pytest.raises(ckpt.CheckpointNotFoundException):
        ckpt.latest_checkpoint(os.path.join(moved, "model.ckpt-1243"))


def test_optimizer_slots_are_saved(tmp_path):
    st = _store()
    m = st.opt_slot("m")
    m.fill_(0.5)
    path = ckpt.save_checkpoint(str(tmp_path), 7, st, opt_slots={"Momentum": m})
    arrays = ckpt.load_arrays(path)
    mom = arrays["v0/cg/conv1/conv2d/kernel/Momentum"]
    assert mom.shape == (3, 3, 16, 16) and np.all(mom == 0.5)


def test_tf_bundle_checkpoints_save_resume_and_interchange(tmp_path):
    """--checkpoint_format=tf writes TensorFlow V2 bundles; both formats restore through the
    same entry points, and the two are interchangeable tensor for tensor."""
    from kf_benchmarks_b200 import tf_bundle
    st = _store()
    g = st.graph
    d = str(tmp_path / "tfck")
    ckpt.save_checkpoint(d, 7, st, {"global_step": np.asarray(7, np.int64)}, fmt="tf")
    assert os.path.exists(os.path.join(d, "model.ckpt-7.index"))
    assert os.path.exists(os.path.join(d, "model.ckpt-7.data-00000-of-00001"))
    assert not os.path.exists(os.path.join(d, "model.ckpt-7.npz"))
    path, step = ckpt.latest_checkpoint(d)
    assert step == 7
    arrays = ckpt.load_arrays(path)
    assert int(arrays["global_step"]) == 7
    want = st.to_tf_arrays("v0/")
    for k, v in want.items():
        assert np.array_equal(arrays[k], v), k
    assert arrays["v0/cg/conv0/conv2d/kernel"].shape[:2] == (3, 3)       # HWIO, as TF stores it
    st2 = ParameterStore(g, "cpu", fp16_shadow=False)
    st2.from_tf_arrays(arrays)
    assert np.array_equal(st2.w32.numpy(), st.w32.numpy())
    # max_to_keep removes both files of old bundles
    for s in (8, 9, 10):
        ckpt.save_checkpoint(d, s, st, fmt="tf", max_to_keep=2)
    left = sorted(f for f in os.listdir(d) if f.startswith("model.ckpt-"))
    assert left == ["model.ckpt-10.data-00000-of-00001", "model.ckpt-10.index",
                    "model.ckpt-9.data-00000-of-00001", "model.ckpt-9.index"]
    # explicit path to a bundle
    assert ckpt.latest_checkpoint(os.path.join(d, "model.ckpt-9"))[1] == 9
    assert "v0/cg/affine0/biases" in tf_bundle.list_bundle(os.path.join(d, "model.ckpt-9"))
