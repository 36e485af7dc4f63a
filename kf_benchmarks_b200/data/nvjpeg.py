"""nvJPEG batch decoder (csrc/jpeg_decode.cu): JPEG bitstreams in host memory -> interleaved RGB
images in device memory, ready for the batch transform kernel (data/gpu_image.py).

The reference decodes on host threads inside tf.data (tf.image.decode_jpeg,
T/preprocessing.py:95-104). Here the worker threads only cut the bytes out of the tf.Example and
read the frame header; the decode runs on the GPU (hardware JPEG engine where there is one).
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence, Tuple

import torch

from ..ops import native

BACKENDS = {0: "default", 1: "hybrid", 2: "gpu_hybrid", 3: "hardware", 4: "gpu_hybrid_device",
            5: "hardware_device", 6: "lossless"}


def jpeg_header(data: bytes):
    """(height, width, components, sof) of a JPEG bitstream, or None if it is not one. ``sof`` is
    the frame marker (0xC0 baseline, 0xC1 extended sequential, 0xC2 progressive, ...). A pure
    marker scan: safe to call from any number of worker threads."""
    n = len(data)
    if n < 4 or data[0] != 0xFF or data[1] != 0xD8:
        return None
    i = 2
    while i + 3 < n:
        if data[i] != 0xFF:
            i += 1
            continue
        m = data[i + 1]
        if m == 0xFF:                      # fill byte
            i += 1
            continue
        if m in (0x01, 0xD8) or 0xD0 <= m <= 0xD7:
            i += 2
            continue
        seg = (data[i + 2] << 8) | data[i + 3]
        if 0xC0 <= m <= 0xCF and m not in (0xC4, 0xC8, 0xCC):
            if i + 9 >= n:
                return None
            h = (data[i + 5] << 8) | data[i + 6]
            w = (data[i + 7] << 8) | data[i + 8]
            return h, w, data[i + 9], m
        if m == 0xDA or seg < 2:           # start of scan before any frame header
            return None
        i += 2 + seg
    return None


def device_decodable(data: bytes):
    """(height, width) if the bitstream is a baseline 1- or 3-component JPEG (what every nvJPEG
    backend, the hardware engine included, decodes), else None (host decode)."""
    hd = jpeg_header(data)
    if hd is None:
        return None
    h, w, comps, sof = hd
    if sof != 0xC0 or comps not in (1, 3) or h <= 0 or w <= 0:
        return None
    return h, w


def available() -> bool:
    try:
        return bool(native.load().kfb_jpeg_available())
    except Exception:  # library not built
        return False


class NvJpegDecoder:
    """One decoder per producer thread. ``decode`` returns (buffer, pointers, pitches): the images
    sit back to back (256-byte aligned) in ``buffer``, which stays valid until the decoder has
    been used ``slots`` more times."""

    def __init__(self, device, backend: int = 0, cpu_threads: int = 4, slots: int = 2):
        lib = native.load()
        self._lib = lib
        self.device = torch.device(device)
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            native.check(lib.kfb_jpeg_create(ctypes.c_int(backend), ctypes.c_int(cpu_threads),
                                             ctypes.byref(h)), "jpeg_create")
        self._h = h
        self.backend = BACKENDS.get(int(lib.kfb_jpeg_backend(h)), "?")
        self.slots = slots
        self._slot = 0
        self._buf = [None] * slots
        self._keep = [None] * slots

    def info(self, data: bytes) -> Tuple[int, int, int]:
        """(height, width, components) from the frame header."""
        w, h, c = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
        native.check(self._lib.kfb_jpeg_image_info(
            self._h, ctypes.c_char_p(data), ctypes.c_ulonglong(len(data)), ctypes.byref(w),
            ctypes.byref(h), ctypes.byref(c)), "jpeg_image_info")
        return h.value, w.value, c.value

    def decode(self, datas: Sequence[bytes], dims: Sequence[Tuple[int, int]], stream=None):
        n = len(datas)
        offsets: List[int] = []
        total = 0
        for (h, w) in dims:
            offsets.append(total)
            total += (h * w * 3 + 255) // 256 * 256
        slot = self._slot
        self._slot = (slot + 1) % self.slots
        buf = self._buf[slot]
        if buf is None or buf.numel() < total:
            buf = torch.empty(int(total * 1.25) + 4096, dtype=torch.uint8, device=self.device)
            self._buf[slot] = buf
        self._keep[slot] = list(datas)          # the bitstreams must outlive the async decode
        base = buf.data_ptr()
        ptrs = [base + off for off in offsets]
        pitches = [w * 3 for (_, w) in dims]
        arr_data = (ctypes.c_char_p * n)(*datas)
        arr_len = (ctypes.c_ulonglong * n)(*[len(d) for d in datas])
        arr_dst = (ctypes.c_void_p * n)(*ptrs)
        arr_pitch = (ctypes.c_int * n)(*pitches)
        s = stream or torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            native.check(self._lib.kfb_jpeg_decode_batch(
                self._h, arr_data, arr_len, ctypes.c_int(n), arr_dst, arr_pitch,
                ctypes.c_void_p(s.cuda_stream)), "jpeg_decode_batch")
        return buf, ptrs, pitches

    def close(self):
        if self._h is not None:
            self._lib.kfb_jpeg_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
