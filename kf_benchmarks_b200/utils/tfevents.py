"""TensorBoard event files without TensorFlow.

The reference writes its summaries with ``tf.summary.FileWriter`` (T/benchmark_cnn.py:2052-2056,
2286-2302, 1771): a file ``events.out.tfevents.<time>.<host>`` of TFRecord-framed ``Event``
protocol buffers. This module writes (and, for the tests, reads back) that format with the
record framing of ``data/tfrecord.py`` and a hand-rolled encoder for the four messages involved:

    Event          { double wall_time = 1; int64 step = 2; string file_version = 3;
                     Summary summary = 5; }
    Summary        { repeated Value value = 1; }
    Summary.Value  { string tag = 1; float simple_value = 2; HistogramProto histo = 5; }
    HistogramProto { double min = 1, max = 2, num = 3, sum = 4, sum_squares = 5;
                     repeated double bucket_limit = 6 [packed]; repeated double bucket = 7 [packed]; }

so ``tensorboard --logdir <train_dir>`` shows the curves of a run of this framework.
"""
from __future__ import annotations

import os
import socket
import struct
import time
from typing import Dict, Iterator, List, Optional, Tuple

from ..data import tfrecord
from ..data.tfrecord import _enc_varint, _fields, _ld


def _double(field: int, x: float) -> bytes:
    return _enc_varint((field << 3) | 1) + struct.pack("<d", float(x))


def _float(field: int, x: float) -> bytes:
    return _enc_varint((field << 3) | 5) + struct.pack("<f", float(x))


def _int(field: int, x: int) -> bytes:
    return _enc_varint(field << 3) + _enc_varint(int(x) & 0xFFFFFFFFFFFFFFFF)


def encode_histogram(h: Dict) -> bytes:
    """``h``: {'min', 'max', 'num', 'sum', 'sum_squares', 'bucket_limit': [...], 'bucket': [...]}"""
    out = b"".join(_double(i + 1, h[k]) for i, k in
                   enumerate(("min", "max", "num", "sum", "sum_squares")))
    out += _ld(6, b"".join(struct.pack("<d", float(v)) for v in h["bucket_limit"]))
    out += _ld(7, b"".join(struct.pack("<d", float(v)) for v in h["bucket"]))
    return out


def encode_event(step: int, wall_time: float, scalars: Optional[Dict[str, float]] = None,
                 histograms: Optional[Dict[str, Dict]] = None,
                 file_version: Optional[str] = None) -> bytes:
    ev = _double(1, wall_time) + _int(2, step)
    if file_version is not None:
        return ev + _ld(3, file_version.encode())
    values = b""
    for tag, v in (scalars or {}).items():
        values += _ld(1, _ld(1, tag.encode()) + _float(2, v))
    for tag, h in (histograms or {}).items():
        values += _ld(1, _ld(1, tag.encode()) + _ld(5, encode_histogram(h)))
    return ev + _ld(5, values)


def histogram_of(values, bins: int = 30) -> Dict:
    """HistogramProto fields of a 1-D array-like (numpy or torch)."""
    import numpy as np
    a = np.asarray(values, dtype=np.float64).ravel()
    if a.size == 0:
        return {"min": 0.0, "max": 0.0, "num": 0.0, "sum": 0.0, "sum_squares": 0.0,
                "bucket_limit": [0.0], "bucket": [0.0]}
    lo, hi = float(a.min()), float(a.max())
    counts, edges = np.histogram(a, bins=bins, range=(lo, hi if hi > lo else lo + 1.0))
    return {"min": lo, "max": hi, "num": float(a.size), "sum": float(a.sum()),
            "sum_squares": float((a * a).sum()), "bucket_limit": [float(e) for e in edges[1:]],
            "bucket": [float(c) for c in counts]}


class EventFileWriter:
    """Appends events to ``<logdir>/events.out.tfevents.<time>.<host>`` (one file per writer,
    starting with the ``brain.Event:2`` version record like TensorFlow's)."""

    def __init__(self, logdir: str, filename_suffix: str = ""):
        os.makedirs(logdir, exist_ok=True)
        self.path = os.path.join(logdir, "events.out.tfevents.%010d.%s%s" %
                                 (int(time.time()), socket.gethostname(), filename_suffix))
        self._f = open(self.path, "ab")
        self._write(encode_event(0, time.time(), file_version="brain.Event:2"))

    def _write(self, payload: bytes) -> None:
        head = struct.pack("<Q", len(payload))
        self._f.write(head + struct.pack("<I", tfrecord.masked_crc(head)) + payload +
                      struct.pack("<I", tfrecord.masked_crc(payload)))
        self._f.flush()

    def add(self, step: int, scalars: Optional[Dict[str, float]] = None,
            histograms: Optional[Dict[str, Dict]] = None) -> None:
        self._write(encode_event(step, time.time(), scalars, histograms))

    def close(self) -> None:
        if not self._f.closed:
            self._f.close()


# ---------------------------------------------------------------------------------------------
# reader (tests, and `tools` that want the curves back)
# ---------------------------------------------------------------------------------------------
def _packed_doubles(b: bytes) -> List[float]:
    return list(struct.unpack("<%dd" % (len(b) // 8), b))


def parse_event(data: bytes) -> Dict:
    ev: Dict = {"scalars": {}, "histograms": {}}
    for num, wire, val in _fields(data):
        if num == 1 and wire == 1:
            ev["wall_time"] = struct.unpack("<d", val)[0]
        elif num == 2 and wire == 0:
            ev["step"] = val if val < (1 << 63) else val - (1 << 64)
        elif num == 3 and wire == 2:
            ev["file_version"] = val.decode()
        elif num == 5 and wire == 2:
            for n2, w2, value in _fields(val):
                if n2 != 1 or w2 != 2:
                    continue
                tag, simple, histo = "", None, None
                for n3, w3, v3 in _fields(value):
                    if n3 == 1 and w3 == 2:
                        tag = v3.decode()
                    elif n3 == 2 and w3 == 5:
                        simple = struct.unpack("<f", v3)[0]
                    elif n3 == 5 and w3 == 2:
                        histo = {}
                        names = {1: "min", 2: "max", 3: "num", 4: "sum", 5: "sum_squares"}
                        for n4, w4, v4 in _fields(v3):
                            if n4 in names and w4 == 1:
                                histo[names[n4]] = struct.unpack("<d", v4)[0]
                            elif n4 == 6 and w4 == 2:
                                histo["bucket_limit"] = _packed_doubles(v4)
                            elif n4 == 7 and w4 == 2:
                                histo["bucket"] = _packed_doubles(v4)
                if simple is not None:
                    ev["scalars"][tag] = simple
                if histo is not None:
                    ev["histograms"][tag] = histo
    return ev


def read_events(path: str) -> Iterator[Dict]:
    for rec in tfrecord.read_records(path, check_crc=True):
        yield parse_event(rec)


def latest_event_file(logdir: str) -> Optional[str]:
    files = sorted(f for f in os.listdir(logdir) if f.startswith("events.out.tfevents."))
    return os.path.join(logdir, files[-1]) if files else None


__all__: Tuple[str, ...] = ("EventFileWriter", "encode_event", "histogram_of", "parse_event",
                            "read_events", "latest_event_file")
