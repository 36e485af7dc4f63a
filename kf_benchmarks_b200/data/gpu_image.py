"""Device-side image preprocessing: raw decoded RGB images go to the GPU as they are and one
kernel (csrc/image_ops.cu) crops, resizes, mirrors, colour-jitters, normalises and casts the
whole batch to fp16 NHWC - the stage the reference runs in tf.data's C++ runtime
(T/preprocessing.py:192-307) and that the host library (csrc/host/image_ops.cpp) runs per image
on the decode threads. The descriptors are the same ``ImageOp`` structures ``native_image``
builds, so the two paths are interchangeable (and tested against each other to one fp16 ulp).

Supported on the device: bilinear, nearest, crop_and_resize (the defaults of the ImageNet and
SSD pipelines); bicubic and area resizing stay on the host path.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence

import numpy as np
import torch

from ..ops import native
from ..ops.native import i32, i64, ptr
from .native_image import ImageOp

DEVICE_METHODS = (0, 1, 4)


def supports(op: ImageOp) -> bool:
    return int(op.method) in DEVICE_METHODS


class GpuImageBatcher:
    """Packs the raw images of a batch into a pinned staging buffer, uploads bytes + descriptors
    with one copy each and launches the batch transform. Buffers grow on demand and are reused;
    ``run`` returns a device tensor [n, oh, ow, 3] that stays valid until the next call with the
    same slot."""

    def __init__(self, device, batch: int, out_h: int, out_w: int, out_dtype=torch.float16,
                 slots: int = 2):
        lib = native.load()
        if int(lib.kfb_image_op_bytes_device()) != ctypes.sizeof(ImageOp):
            raise native.KfbError("ImageOp mirror does not match the device library")
        self.device = torch.device(device)
        self.batch, self.oh, self.ow = batch, out_h, out_w
        self.out_dtype = out_dtype
        self.slots = slots
        self._slot = 0
        self._raw_host = [None] * slots      # pinned uint8
        self._raw_dev = [None] * slots
        self._ops_host = [torch.empty(batch * ctypes.sizeof(ImageOp), dtype=torch.uint8,
                                      pin_memory=True) for _ in range(slots)]
        self._ptr_host = [torch.empty(batch, dtype=torch.int64, pin_memory=True)
                          for _ in range(slots)]
        self._ops_dev = [torch.empty(batch * ctypes.sizeof(ImageOp), dtype=torch.uint8,
                                     device=self.device) for _ in range(slots)]
        self._ptr_dev = [torch.empty(batch, dtype=torch.int64, device=self.device)
                         for _ in range(slots)]
        self._means = torch.zeros(batch * 3, dtype=torch.float32, device=self.device)
        self._out = [torch.empty((batch, out_h, out_w, 3), dtype=out_dtype, device=self.device)
                     for _ in range(slots)]
        self.h2d_bytes_last = 0

    def _ensure_raw(self, slot: int, nbytes: int) -> None:
        cur = self._raw_host[slot]
        if cur is None or cur.numel() < nbytes:
            cap = int(nbytes * 1.25) + 4096
            self._raw_host[slot] = torch.empty(cap, dtype=torch.uint8, pin_memory=True)
            self._raw_dev[slot] = torch.empty(cap, dtype=torch.uint8, device=self.device)

    def run(self, images: Sequence[np.ndarray], ops: Sequence[ImageOp], stream=None) -> torch.Tensor:
        n = len(images)
        if n != len(ops) or n > self.batch:
            raise ValueError("gpu_image: %d images, %d descriptors, batch %d"
                             % (n, len(ops), self.batch))
        slot = self._slot
        self._slot = (slot + 1) % self.slots
        offsets: List[int] = []
        total = 0
        for img in images:
            offsets.append(total)
            total += (img.shape[0] * img.shape[1] * 3 + 15) // 16 * 16
        self._ensure_raw(slot, total)
        raw = self._raw_host[slot].numpy()
        ops_bytes = self._ops_host[slot].numpy()
        ptrs = self._ptr_host[slot].numpy()
        base = self._raw_dev[slot].data_ptr()
        size = ctypes.sizeof(ImageOp)
        for k, (img, op, off) in enumerate(zip(images, ops, offsets)):
            if not supports(op):
                raise ValueError("gpu_image: resize method %d runs on the host path" % op.method)
            if not (img.flags["C_CONTIGUOUS"] and img.dtype == np.uint8):
                img = np.ascontiguousarray(img, dtype=np.uint8)
            nb = img.shape[0] * img.shape[1] * 3
            raw[off:off + nb] = img.reshape(-1)
            dop = ImageOp.from_buffer_copy(op)          # the staged copy is densely packed
            dop.src_stride = img.shape[1] * 3
            dop.out_f16 = int(self.out_dtype == torch.float16)
            ops_bytes[k * size:(k + 1) * size] = np.frombuffer(bytes(dop), dtype=np.uint8)
            ptrs[k] = base + off
        s = stream or torch.cuda.current_stream(self.device)
        with torch.cuda.stream(s):
            self._raw_dev[slot][:total].copy_(self._raw_host[slot][:total], non_blocking=True)
            self._ops_dev[slot].copy_(self._ops_host[slot], non_blocking=True)
            self._ptr_dev[slot].copy_(self._ptr_host[slot], non_blocking=True)
            out = self._out[slot]
            native.call("image_transform_batch", ptr(self._ptr_dev[slot]), ptr(self._ops_dev[slot]),
                        i32(n), i32(self.oh * self.ow), ptr(self._means), ptr(out),
                        i64(self.oh * self.ow * 3), ctypes.c_void_p(s.cuda_stream))
        self.h2d_bytes_last = total + n * (size + 8)
        return out[:n]

    def run_device(self, dev_ptrs: Sequence[int], pitches: Sequence[int], ops: Sequence[ImageOp],
                   stream=None) -> torch.Tensor:
        """The same transform on images that are already on the device (decoded there by
        data/nvjpeg.py): ``dev_ptrs[k]`` is interleaved RGB with row pitch ``pitches[k]`` bytes.
        Only the descriptors and pointers are uploaded."""
        n = len(dev_ptrs)
        if n != len(ops) or n > self.batch:
            raise ValueError("gpu_image: %d images, %d descriptors, batch %d"
                             % (n, len(ops), self.batch))
        slot = self._slot
        self._slot = (slot + 1) % self.slots
        ops_bytes = self._ops_host[slot].numpy()
        ptrs = self._ptr_host[slot].numpy()
        size = ctypes.sizeof(ImageOp)
        for k, (dp, pitch, op) in enumerate(zip(dev_ptrs, pitches, ops)):
            if not supports(op):
                raise ValueError("gpu_image: resize method %d runs on the host path" % op.method)
            dop = ImageOp.from_buffer_copy(op)
            dop.src_stride = int(pitch)
            dop.out_f16 = int(self.out_dtype == torch.float16)
            ops_bytes[k * size:(k + 1) * size] = np.frombuffer(bytes(dop), dtype=np.uint8)
            ptrs[k] = dp
        s = stream or torch.cuda.current_stream(self.device)
        with torch.cuda.stream(s):
            self._ops_dev[slot].copy_(self._ops_host[slot], non_blocking=True)
            self._ptr_dev[slot].copy_(self._ptr_host[slot], non_blocking=True)
            out = self._out[slot]
            native.call("image_transform_batch", ptr(self._ptr_dev[slot]), ptr(self._ops_dev[slot]),
                        i32(n), i32(self.oh * self.ow), ptr(self._means), ptr(out),
                        i64(self.oh * self.ow * 3), ctypes.c_void_p(s.cuda_stream))
        self.h2d_bytes_last = n * (size + 8)
        return out[:n]
