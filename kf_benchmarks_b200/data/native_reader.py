"""ctypes binding of the host-side input runtime (``_lib/libkfb200_host.so``, sources in
``csrc/host/``): multi-threaded TFRecord streaming with CRC32C validation and an in-place
tf.Example field scanner. ``available()`` is False when the library has not been built; callers
(``data/preprocessing.py``) then use the pure-Python codec of ``data/tfrecord.py``.

Reference counterpart: TensorFlow's C++ RecordInput / tf.data readers that
T/preprocessing.py:358-399, 601-609 drives.
"""
from __future__ import annotations

import ctypes
import os
from typing import Iterator, List, Optional, Sequence

_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib",
                     "libkfb200_host.so")
_lib = None


def available() -> bool:
    return os.path.exists(_PATH)


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(_PATH)
        lib.kfb_host_last_error.restype = ctypes.c_char_p
        lib.kfb_crc32c.restype = ctypes.c_uint32
        lib.kfb_crc32c.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        lib.kfb_masked_crc32c.restype = ctypes.c_uint32
        lib.kfb_masked_crc32c.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        lib.kfb_rr_open.restype = ctypes.c_void_p
        lib.kfb_rr_open.argtypes = [ctypes.POINTER(ctypes.c_char_p), ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_uint64]
        lib.kfb_rr_open2.restype = ctypes.c_void_p
        lib.kfb_rr_open2.argtypes = lib.kfb_rr_open.argtypes + [ctypes.c_int]
        lib.kfb_rr_next.restype = ctypes.c_int
        lib.kfb_rr_next.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p),
                                    ctypes.POINTER(ctypes.c_uint64)]
        lib.kfb_rr_close.argtypes = [ctypes.c_void_p]
        lib.kfb_example_get_bytes.restype = ctypes.c_int
        lib.kfb_example_get_bytes.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p,
                                              ctypes.POINTER(ctypes.c_void_p),
                                              ctypes.POINTER(ctypes.c_uint64)]
        lib.kfb_example_get_numbers.restype = ctypes.c_int
        lib.kfb_example_get_numbers.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p,
                                                ctypes.POINTER(ctypes.c_double), ctypes.c_int]
        _lib = lib
    return _lib


def crc32c(data: bytes) -> int:
    return int(_load().kfb_crc32c(data, len(data)))


def masked_crc32c(data: bytes) -> int:
    return int(_load().kfb_masked_crc32c(data, len(data)))


class RecordReader:
    """Iterates the records of TFRecord files read by ``num_threads`` background threads
    (file i belongs to thread i % num_threads). ``repeat`` cycles forever, ``shuffle_files``
    reshuffles each thread's file list every pass. ``deterministic`` takes one record from each
    reader in turn (tf.data interleave, block length 1): the order then depends on the file
    list only; otherwise records come in whatever order the threads deliver them (sloppy)."""

    def __init__(self, paths: Sequence[str], num_threads: int = 1, capacity: int = 256,
                 check_crc: bool = True, repeat: bool = False, shuffle_files: bool = False,
                 seed: int = 0, deterministic: bool = False):
        lib = _load()
        arr = (ctypes.c_char_p * len(paths))(*[os.fsencode(p) for p in paths])
        self._h = lib.kfb_rr_open2(arr, len(paths), int(num_threads), int(capacity),
                                   int(check_crc), int(repeat), int(shuffle_files), int(seed),
                                   int(deterministic))
        if not self._h:
            raise IOError(lib.kfb_host_last_error().decode())

    def __iter__(self) -> Iterator[bytes]:
        return self

    def __next__(self) -> bytes:
        if self._h is None:
            raise StopIteration
        lib = _load()
        ptr, n = ctypes.c_void_p(), ctypes.c_uint64()
        rc = lib.kfb_rr_next(self._h, ctypes.byref(ptr), ctypes.byref(n))
        if rc == 1:
            return ctypes.string_at(ptr, n.value)
        msg = lib.kfb_host_last_error().decode() if rc < 0 else ""
        self.close()
        if rc < 0:
            raise IOError(msg)
        raise StopIteration

    def close(self) -> None:
        if self._h is not None:
            _load().kfb_rr_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def example_bytes(record: bytes, key: str) -> Optional[bytes]:
    """First bytes value of feature ``key`` (None when missing)."""
    ptr, n = ctypes.c_void_p(), ctypes.c_uint64()
    if _load().kfb_example_get_bytes(record, len(record), key.encode(), ctypes.byref(ptr),
                                     ctypes.byref(n)) != 1:
        return None
    return ctypes.string_at(ptr, n.value)


def example_numbers(record: bytes, key: str, cap: int = 1024) -> Optional[List[float]]:
    """Float / int64 values of feature ``key`` (None when missing)."""
    buf = (ctypes.c_double * cap)()
    n = _load().kfb_example_get_numbers(record, len(record), key.encode(), buf, cap)
    if n < 0:
        return None
    if n > cap:
        return example_numbers(record, key, n)
    return list(buf[:n])
