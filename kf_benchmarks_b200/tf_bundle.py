"""TensorFlow V2 checkpoint bundles ("TensorBundle": ``<prefix>.index`` +
``<prefix>.data-00000-of-00001``) read and written without TensorFlow.

The reference saves and restores through ``tf.train.Saver`` (T/benchmark_cnn.py:2076-2082,
905-950), i.e. in this format. It is restated here from its on-disk definition so checkpoints
can move between the two code bases in both directions:

* data file: the raw little-endian bytes of every tensor, concatenated in key order;
* index file: an immutable sorted string table (the LevelDB table format TensorFlow embeds):
  data blocks of prefix-compressed ``(key, value)`` entries with a restart point every 16
  entries and a ``[restart offsets..., count]`` tail, each block followed by a 5-byte trailer
  (compression type 0 + masked CRC32C), then a metaindex block, an index block mapping a
  separator key to each data block's ``(offset, size)`` handle, and a 48-byte footer ending
  in the magic number ``0xdb4775248b80fb57``;
* key ``""`` holds a ``BundleHeaderProto`` (num_shards, endianness, version), every other key
  is a variable name whose value is a ``BundleEntryProto`` (dtype, shape, shard, offset, size,
  masked CRC32C of the bytes).

TensorFlow is not installable on this image, so the writer is validated by the reader, by
format invariants (tests/test_tf_bundle.py) and by hand-assembled golden bytes - not against a
live TensorFlow.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Tuple

import numpy as np

from .data.tfrecord import masked_crc

MAGIC = 0xDB4775248B80FB57
BLOCK_SIZE = 256 << 10
RESTART_INTERVAL = 16

# tensorflow DataType enum values
_DTYPES = {np.dtype("float32"): 1, np.dtype("float64"): 2, np.dtype("int32"): 3,
           np.dtype("uint8"): 4, np.dtype("int16"): 5, np.dtype("int8"): 6,
           np.dtype("int64"): 9, np.dtype("bool"): 10, np.dtype("float16"): 19}
_DTYPES_INV = {v: k for k, v in _DTYPES.items()}


# ---- varints / tiny protobuf helpers --------------------------------------------------------
def _put_varint(x: int) -> bytes:
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _get_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    result = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _field_varint(num: int, value: int) -> bytes:
    return _put_varint(num << 3) + _put_varint(value)


def _field_bytes(num: int, payload: bytes) -> bytes:
    return _put_varint((num << 3) | 2) + _put_varint(len(payload)) + payload


def _fields(buf: bytes):
    pos = 0
    while pos < len(buf):
        key, pos = _get_varint(buf, pos)
        num, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = _get_varint(buf, pos)
        elif wire == 2:
            n, pos = _get_varint(buf, pos)
            val, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            val, pos = struct.unpack_from("<I", buf, pos)[0], pos + 4
        elif wire == 1:
            val, pos = struct.unpack_from("<Q", buf, pos)[0], pos + 8
        else:
            raise ValueError("unsupported wire type %d" % wire)
        yield num, wire, val


def _header_proto() -> bytes:
    # BundleHeaderProto{num_shards=1, endianness=LITTLE(0, default), version{producer=1}}
    return _field_varint(1, 1) + _field_bytes(3, _field_varint(1, 1))


def _entry_proto(arr: np.ndarray, offset: int, size: int, crc: int) -> bytes:
    shape = b"".join(_field_bytes(2, _field_varint(1, int(d))) for d in arr.shape)
    out = _field_varint(1, _DTYPES[arr.dtype])
    out += _field_bytes(2, shape)
    if offset:
        out += _field_varint(4, offset)
    out += _field_varint(5, size)
    out += _put_varint((6 << 3) | 5) + struct.pack("<I", crc)
    return out


def _parse_entry(buf: bytes):
    dtype, shape, shard, offset, size, crc = 0, [], 0, 0, 0, None
    for num, wire, val in _fields(buf):
        if num == 1:
            dtype = val
        elif num == 2:
            for n2, _, dim in _fields(val):
                if n2 == 2:
                    d = 0
                    for n3, _, v3 in _fields(dim):
                        if n3 == 1:
                            d = v3 - (1 << 64) if v3 >> 63 else v3
                    shape.append(d)
        elif num == 3:
            shard = val
        elif num == 4:
            offset = val
        elif num == 5:
            size = val
        elif num == 6:
            crc = val
    return dtype, tuple(shape), shard, offset, size, crc


# ---- sorted string table --------------------------------------------------------------------
class _BlockBuilder:
    def __init__(self):
        self.buf = bytearray()
        self.restarts = [0]
        self.count = 0
        self.last_key = b""

    def add(self, key: bytes, value: bytes) -> None:
        shared = 0
        if self.count % RESTART_INTERVAL == 0 and self.count:
            self.restarts.append(len(self.buf))
        elif self.count:
            limit = min(len(key), len(self.last_key))
            while shared < limit and key[shared] == self.last_key[shared]:
                shared += 1
        self.buf += _put_varint(shared) + _put_varint(len(key) - shared) + _put_varint(len(value))
        self.buf += key[shared:] + value
        self.last_key = key
        self.count += 1

    def size_estimate(self) -> int:
        return len(self.buf) + 4 * len(self.restarts) + 4

    def finish(self) -> bytes:
        tail = b"".join(struct.pack("<I", r) for r in self.restarts)
        return bytes(self.buf) + tail + struct.pack("<I", len(self.restarts))


def _block_with_trailer(contents: bytes) -> bytes:
    return contents + b"\x00" + struct.pack("<I", masked_crc(contents + b"\x00"))


def _handle(offset: int, size: int) -> bytes:
    return _put_varint(offset) + _put_varint(size)


def _write_table(path: str, items: List[Tuple[bytes, bytes]]) -> None:
    """``items`` sorted by key."""
    out = bytearray()
    index = _BlockBuilder()
    block = _BlockBuilder()

    def flush():
        nonlocal block
        if not block.count:
            return
        contents = block.finish()
        index.add(block.last_key, _handle(len(out), len(contents)))
        out.extend(_block_with_trailer(contents))
        block = _BlockBuilder()
    for key, value in items:
        block.add(key, value)
        if block.size_estimate() >= BLOCK_SIZE:
            flush()
    flush()
    meta = _BlockBuilder().finish()
    meta_handle = _handle(len(out), len(meta))
    out.extend(_block_with_trailer(meta))
    idx = index.finish()
    idx_handle = _handle(len(out), len(idx))
    out.extend(_block_with_trailer(idx))
    footer = meta_handle + idx_handle
    footer += b"\x00" * (40 - len(footer)) + struct.pack("<Q", MAGIC)
    out.extend(footer)
    with open(path, "wb") as f:
        f.write(bytes(out))


def _read_block(buf: bytes, offset: int, size: int, check: bool = True) -> List[Tuple[bytes, bytes]]:
    contents = buf[offset:offset + size]
    ctype = buf[offset + size]
    crc = struct.unpack_from("<I", buf, offset + size + 1)[0]
    if ctype != 0:
        raise ValueError("compressed table blocks are not supported (type %d)" % ctype)
    if check and masked_crc(contents + bytes([ctype])) != crc:
        raise ValueError("corrupt table block at offset %d" % offset)
    num_restarts = struct.unpack_from("<I", contents, len(contents) - 4)[0]
    end = len(contents) - 4 - 4 * num_restarts
    items, pos, key = [], 0, b""
    while pos < end:
        shared, pos = _get_varint(contents, pos)
        non_shared, pos = _get_varint(contents, pos)
        vlen, pos = _get_varint(contents, pos)
        key = key[:shared] + contents[pos:pos + non_shared]
        pos += non_shared
        items.append((key, contents[pos:pos + vlen]))
        pos += vlen
    return items


def _read_table(path: str) -> List[Tuple[bytes, bytes]]:
    with open(path, "rb") as f:
        buf = f.read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != MAGIC:
        raise ValueError("%s is not a TensorFlow checkpoint index (bad magic)" % path)
    footer = buf[len(buf) - 48:]
    _, pos = _get_varint(footer, 0)
    _, pos = _get_varint(footer, pos)
    idx_off, pos = _get_varint(footer, pos)
    idx_size, pos = _get_varint(footer, pos)
    items = []
    for _, handle in _read_block(buf, idx_off, idx_size):
        off, p = _get_varint(handle, 0)
        size, _ = _get_varint(handle, p)
        items.extend(_read_block(buf, off, size))
    return items


# ---- public API -----------------------------------------------------------------------------
def write_bundle(prefix: str, arrays: Dict[str, np.ndarray]) -> None:
    """Writes ``prefix.index`` and ``prefix.data-00000-of-00001``."""
    items = [(b"", _header_proto())]
    offset = 0
    with open(prefix + ".data-00000-of-00001", "wb") as data:
        for name in sorted(arrays, key=lambda n: n.encode("utf-8")):
            arr = np.asarray(arrays[name])
            if arr.dtype not in _DTYPES:
                raise TypeError("%s: dtype %s cannot be stored in a bundle" % (name, arr.dtype))
            raw = np.ascontiguousarray(arr).astype(arr.dtype.newbyteorder("<"), copy=False)\
                .tobytes()
            data.write(raw)
            items.append((name.encode("utf-8"), _entry_proto(arr, offset, len(raw),
                                                             masked_crc(raw))))
            offset += len(raw)
    _write_table(prefix + ".index", items)


def list_bundle(prefix: str) -> Dict[str, Tuple[np.dtype, Tuple[int, ...]]]:
    out = {}
    for key, value in _read_table(prefix + ".index"):
        if key == b"":
            continue
        dtype, shape, _, _, _, _ = _parse_entry(value)
        out[key.decode("utf-8")] = (_DTYPES_INV.get(dtype), shape)
    return out


def read_bundle(prefix: str, check_crc: bool = True) -> Dict[str, np.ndarray]:
    """All tensors of a V2 checkpoint (single- or multi-shard, uncompressed, unsliced)."""
    entries = _read_table(prefix + ".index")
    num_shards = 1
    for key, value in entries:
        if key == b"":
            for num, _, val in _fields(value):
                if num == 1:
                    num_shards = val
                elif num == 2 and val != 0:
                    raise ValueError("big-endian bundles are not supported")
    shards = {}
    out: Dict[str, np.ndarray] = {}
    for key, value in entries:
        if key == b"":
            continue
        dtype, shape, shard, offset, size, crc = _parse_entry(value)
        if dtype not in _DTYPES_INV:
            continue                      # strings / resources / variants: not model variables
        if shard not in shards:
            with open("%s.data-%05d-of-%05d" % (prefix, shard, num_shards), "rb") as f:
                shards[shard] = f.read()
        raw = shards[shard][offset:offset + size]
        if check_crc and crc is not None and masked_crc(raw) != crc:
            raise ValueError("checksum mismatch for tensor %r" % key.decode("utf-8", "replace"))
        out[key.decode("utf-8")] = np.frombuffer(raw, dtype=_DTYPES_INV[dtype].newbyteorder("<"))\
            .reshape(shape).copy()
    return out


def is_bundle(prefix: str) -> bool:
    import os
    return os.path.exists(prefix + ".index")
