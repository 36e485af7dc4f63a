"""Minimal TFRecord + tf.Example reader (no TensorFlow needed).

Record framing: uint64 length | uint32 masked-crc(length) | data | uint32 masked-crc(data).
``tf.Example`` is decoded with a tiny protobuf wire-format parser: Example{features=1} ->
Features{feature map=1} -> Feature{bytes_list=1 | float_list=2 | int64_list=3}. Keys used by
the ImageNet records are those of the reference's parser (T/preprocessing.py:27-72):
``image/encoded``, ``image/class/label``, ``image/class/text``, ``image/object/bbox/{xmin,
ymin,xmax,ymax}``. A writer is included so tests can fabricate records
(reference: T/test_data/tfrecord_image_generator.py).
"""
from __future__ import annotations

import struct
from typing import Dict, Iterator


def _crc32c_table():
    poly = 0x82F63B78
    table = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ poly if c & 1 else c >> 1
        table.append(c)
    return table


_TABLE = _crc32c_table()


def _crc32c_python(data: bytes) -> int:
    c = 0xFFFFFFFF
    for b in data:
        c = _TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


_native_crc = None          # None: not looked up yet; False: host library not built


def crc32c(data: bytes) -> int:
    """CRC32C (Castagnoli). Uses the host library's hardware implementation
    (csrc/host/record_io.cpp, several GB/s) when it is built - checkpoint bundles put hundreds
    of megabytes through this - and the table-driven Python loop otherwise."""
    global _native_crc
    if _native_crc is None:
        try:
            from . import native_reader
            _native_crc = native_reader.crc32c if native_reader.available() else False
        except Exception:  # pragma: no cover - a broken build must not break the codec
            _native_crc = False
    if _native_crc and len(data) >= 64:
        return _native_crc(bytes(data) if not isinstance(data, bytes) else data)
    return _crc32c_python(data)


def masked_crc(data: bytes) -> int:
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def read_records(path: str, check_crc: bool = False) -> Iterator[bytes]:
    with open(path, "rb") as f:
        while True:
            head = f.read(12)
            if len(head) < 12:
                return
            (length,) = struct.unpack("<Q", head[:8])
            data = f.read(length)
            tail = f.read(4)
            if len(data) < length or len(tail) < 4:
                raise IOError("truncated TFRecord file %s" % path)
            if check_crc and struct.unpack("<I", tail)[0] != masked_crc(data):
                raise IOError("corrupt record in %s" % path)
            yield data


def write_records(path: str, records) -> None:
    with open(path, "wb") as f:
        for data in records:
            head = struct.pack("<Q", len(data))
            f.write(head + struct.pack("<I", masked_crc(head)) + data +
                    struct.pack("<I", masked_crc(data)))


# ---- protobuf wire format ------------------------------------------------------------
def _varint(buf: bytes, pos: int):
    result = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        result |= (b & 0x7F) << shift
        if not b & 0x80:
            return result, pos
        shift += 7


def _fields(buf: bytes):
    pos = 0
    while pos < len(buf):
        key, pos = _varint(buf, pos)
        field, wire = key >> 3, key & 7
        if wire == 0:
            val, pos = _varint(buf, pos)
        elif wire == 1:
            val, pos = buf[pos:pos + 8], pos + 8
        elif wire == 2:
            n, pos = _varint(buf, pos)
            val, pos = buf[pos:pos + n], pos + n
        elif wire == 5:
            val, pos = buf[pos:pos + 4], pos + 4
        else:
            raise ValueError("unsupported wire type %d" % wire)
        yield field, wire, val


def _parse_feature(feature: bytes) -> list:
    """tf.train.Feature -> python list (bytes / floats / ints)."""
    values: list = []
    for kind, _, payload in _fields(feature):
        if kind == 1:      # BytesList
            values = [v for f, _, v in _fields(payload) if f == 1]
        elif kind == 2:    # FloatList (packed or not)
            for f, w, v in _fields(payload):
                if f == 1 and w == 2:
                    values.extend(struct.unpack("<%df" % (len(v) // 4), v))
                elif f == 1:
                    values.append(struct.unpack("<f", v)[0])
        elif kind == 3:    # Int64List
            for f, w, v in _fields(payload):
                if f == 1 and w == 2:
                    p = 0
                    while p < len(v):
                        x, p = _varint(v, p)
                        values.append(x - (1 << 64) if x >> 63 else x)
                elif f == 1:
                    values.append(v - (1 << 64) if v >> 63 else v)
    return values


def _parse_features(features: bytes) -> Dict[str, list]:
    """tf.train.Features (map<string, Feature>) -> dict."""
    out: Dict[str, list] = {}
    for f2, _, entry in _fields(features):
        if f2 != 1:
            continue
        key, feature = None, b""
        for f3, _, v in _fields(entry):
            if f3 == 1:
                key = v.decode("utf-8")
            elif f3 == 2:
                feature = v
        if key is not None:
            out[key] = _parse_feature(feature)
    return out


def parse_example(data: bytes) -> Dict[str, list]:
    out: Dict[str, list] = {}
    for f1, _, features in _fields(data):
        if f1 == 1:
            out.update(_parse_features(features))
    return out


def parse_sequence_example(data: bytes):
    """tf.train.SequenceExample -> (context dict, {name: [list per step]}): the record format
    of the preprocessed LibriSpeech set (T/preprocessing.py:1079-1112)."""
    context: Dict[str, list] = {}
    lists: Dict[str, list] = {}
    for f1, _, payload in _fields(data):
        if f1 == 1:
            context.update(_parse_features(payload))
        elif f1 == 2:
            for f2, _, entry in _fields(payload):       # map<string, FeatureList>
                if f2 != 1:
                    continue
                key, flist = None, b""
                for f3, _, v in _fields(entry):
                    if f3 == 1:
                        key = v.decode("utf-8")
                    elif f3 == 2:
                        flist = v
                if key is not None:
                    lists[key] = [_parse_feature(v) for f, _, v in _fields(flist) if f == 1]
    return context, lists


# ---- encoder (tests / tools) --------------------------------------------------------------
def _enc_varint(x: int) -> bytes:
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _ld(field: int, payload: bytes) -> bytes:
    return _enc_varint((field << 3) | 2) + _enc_varint(len(payload)) + payload


def _enc_feature(values: list) -> bytes:
    if values and isinstance(values[0], (bytes, bytearray)):
        return _ld(1, b"".join(_ld(1, bytes(v)) for v in values))
    if values and isinstance(values[0], float):
        return _ld(2, _ld(1, struct.pack("<%df" % len(values), *values)))
    return _ld(3, _ld(1, b"".join(_enc_varint(int(v)) for v in values)))


def _enc_features(features: Dict[str, list]) -> bytes:
    entries = b""
    for key, values in features.items():
        entries += _ld(1, _ld(1, key.encode("utf-8")) + _ld(2, _enc_feature(values)))
    return entries


def encode_example(features: Dict[str, list]) -> bytes:
    return _ld(1, _enc_features(features))


def encode_sequence_example(context: Dict[str, list], feature_lists: Dict[str, list]) -> bytes:
    lists = b""
    for key, steps in feature_lists.items():
        flist = b"".join(_ld(1, _enc_feature(list(step))) for step in steps)
        lists += _ld(1, _ld(1, key.encode("utf-8")) + _ld(2, flist))
    return _ld(1, _enc_features(context)) + _ld(2, lists)
