"""TensorFlow V2 checkpoint bundle codec (kf_benchmarks_b200/tf_bundle.py)."""
import struct

import numpy as np
import pytest

from kf_benchmarks_b200 import tf_bundle
from kf_benchmarks_b200.data.tfrecord import masked_crc


def _arrays(n_extra=0):
    rng = np.random.RandomState(0)
    a = {"v0/cg/conv0/conv2d/kernel": rng.randn(7, 7, 3, 64).astype(np.float32),
         "v0/cg/conv0/batchnorm0/gamma": np.ones(64, np.float32),
         "v0/cg/affine0/weights": rng.randn(2048, 1001).astype(np.float32),
         "global_step": np.asarray(1243, np.int64),
         "loss_scale": np.asarray(128.0, np.float32),
         "half": rng.randn(5).astype(np.float16)}
    for i in range(n_extra):
        a["v0/cg/extra/var_%04d" % i] = rng.randn(3).astype(np.float32)
    return a


def test_roundtrip_and_listing(tmp_path):
    prefix = str(tmp_path / "model.ckpt-1243")
    arrays = _arrays()
    tf_bundle.write_bundle(prefix, arrays)
    assert tf_bundle.is_bundle(prefix)
    back = tf_bundle.read_bundle(prefix)
    assert set(back) == set(arrays)
    for k, v in arrays.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape
        assert np.array_equal(back[k], v)
    listing = tf_bundle.list_bundle(prefix)
    assert listing["v0/cg/conv0/conv2d/kernel"] == (np.dtype("float32"), (7, 7, 3, 64))
    assert listing["global_step"] == (np.dtype("int64"), ())


def test_format_invariants(tmp_path):
    prefix = str(tmp_path / "m")
    arrays = _arrays()
    tf_bundle.write_bundle(prefix, arrays)
    idx = open(prefix + ".index", "rb").read()
    # footer: 48 bytes ending in the table magic
    assert struct.unpack("<Q", idx[-8:])[0] == 0xDB4775248B80FB57
    # first data block starts with the header entry: shared 0, key length 0, BundleHeaderProto
    assert idx[0] == 0 and idx[1] == 0
    hdr_len = idx[2]
    assert idx[3:3 + hdr_len] == bytes([0x08, 0x01, 0x1A, 0x02, 0x08, 0x01])
    # the data file is the tensors in key order, back to back
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    order = sorted(arrays, key=lambda n: n.encode())
    assert len(data) == sum(arrays[k].nbytes for k in order)
    assert data[:arrays[order[0]].nbytes] == arrays[order[0]].tobytes()
    # entry protos carry the masked CRC32C of the tensor bytes
    entries = dict(tf_bundle._read_table(prefix + ".index"))
    _, shape, shard, offset, size, crc = tf_bundle._parse_entry(entries[b"loss_scale"])
    assert shape == () and shard == 0 and size == 4
    assert crc == masked_crc(arrays["loss_scale"].tobytes())
    assert data[offset:offset + 4] == struct.pack("<f", 128.0)


def test_many_keys_span_blocks_and_restarts(tmp_path, monkeypatch):
    prefix = str(tmp_path / "big")
    arrays = _arrays(n_extra=600)
    monkeypatch.setattr(tf_bundle, "BLOCK_SIZE", 4096)   # force many data blocks
    tf_bundle.write_bundle(prefix, arrays)
    items = tf_bundle._read_table(prefix + ".index")
    keys = [k for k, _ in items]
    assert keys == sorted(keys) and len(keys) == len(arrays) + 1
    back = tf_bundle.read_bundle(prefix)
    assert np.array_equal(back["v0/cg/extra/var_0599"], arrays["v0/cg/extra/var_0599"])
    buf = open(prefix + ".index", "rb").read()
    footer = buf[-48:]
    _, pos = tf_bundle._get_varint(footer, 0)
    _, pos = tf_bundle._get_varint(footer, pos)
    idx_off, pos = tf_bundle._get_varint(footer, pos)
    idx_size, _ = tf_bundle._get_varint(footer, pos)
    assert len(tf_bundle._read_block(buf, idx_off, idx_size)) >= 3     # several data blocks


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "c")
    tf_bundle.write_bundle(prefix, _arrays())
    data = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    data[100] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(data))
    with pytest.raises(ValueError):
        tf_bundle.read_bundle(prefix)
    assert tf_bundle.read_bundle(prefix, check_crc=False)
    idx = bytearray(open(prefix + ".index", "rb").read())
    idx[10] ^= 1
    open(prefix + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError):
        tf_bundle.read_bundle(prefix)
    open(prefix + ".index", "wb").write(b"not a table")
    with pytest.raises(ValueError):
        tf_bundle.read_bundle(prefix)
