"""ctypes binding of the native image transform (``csrc/host/image_ops.cpp`` in
``_lib/libkfb200_host.so``) and a NumPy restatement of the same arithmetic.

One call takes a decoded uint8 RGB image and writes the cropped, resized, optionally flipped /
colour-jittered, normalised fp32 or fp16 result into a caller-provided buffer (a slot of the
pinned batch). ctypes releases the GIL for the call, so ``data/preprocessing.py`` runs it from a
thread pool. Resize semantics are those of the TF1 ``tf.image.resize_*`` kernels with
``align_corners=False`` that the reference's preprocessing uses (T/preprocessing.py:130-307):
source coordinate = destination index * in/out, no antialiasing.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import numpy as np

from . import native_reader

METHODS = {"bilinear": 0, "nearest": 1, "bicubic": 2, "area": 3}
CROP_AND_RESIZE = 4     # tf.image.crop_and_resize of a normalised float window (SSD)


class ImageOp(ctypes.Structure):
    """Mirror of ``kfb_image_op`` (csrc/host/image_ops.cpp)."""
    _fields_ = [(n, ctypes.c_int) for n in ("ih", "iw", "src_stride", "top", "left", "ch", "cw",
                                            "vh", "vw", "oy", "ox", "oh", "ow", "method", "flip",
                                            "distort_order")] + \
               [(n, ctypes.c_float) for n in ("brightness", "saturation", "hue", "contrast",
                                              "scale", "bias")] + \
               [("out_f16", ctypes.c_int), ("box", ctypes.c_float * 4),
                ("per_channel", ctypes.c_int), ("scale3", ctypes.c_float * 3),
                ("bias3", ctypes.c_float * 3), ("no_clamp", ctypes.c_int)]


_fn = None


def available() -> bool:
    return native_reader.available() and _function() is not None


def _function():
    global _fn
    if _fn is None:
        lib = native_reader._load()
        try:
            fn = lib.kfb_image_transform
        except AttributeError:          # a host library built before image_ops.cpp existed
            return None
        fn.restype = ctypes.c_int
        fn.argtypes = [ctypes.c_void_p, ctypes.POINTER(ImageOp), ctypes.c_void_p]
        lib.kfb_image_last_error.restype = ctypes.c_char_p
        try:
            size = int(lib.kfb_image_op_bytes())
        except AttributeError:
            size = ctypes.sizeof(ImageOp)
        if size != ctypes.sizeof(ImageOp):
            raise RuntimeError("kfb_image_op is %d bytes in the library but %d in the binding: "
                               "rebuild the host library" % (size, ctypes.sizeof(ImageOp)))
        _fn = fn
    return _fn


def make_op(img: np.ndarray, crop: Optional[Tuple[int, int, int, int]], resized: Tuple[int, int],
            window: Optional[Tuple[int, int, int, int]] = None, method: str = "bilinear",
            flip: bool = False, distort: Optional[Tuple[int, float, float, float, float]] = None,
            scale: float = 1.0 / 127.5, bias: float = -1.0, out_f16: bool = False) -> ImageOp:
    """crop = (top, left, h, w) or None for the whole image; resized = (h, w) the crop is scaled
    to; window = (top, left, h, w) of the resized image to produce (None: all of it);
    distort = (order 1|2, brightness delta, saturation factor, hue shift in turns, contrast
    factor) - see ``distort_color`` in data/preprocessing.py."""
    ih, iw = img.shape[:2]
    top, left, ch, cw = crop if crop is not None else (0, 0, ih, iw)
    vh, vw = resized
    oy, ox, oh, ow = window if window is not None else (0, 0, vh, vw)
    op = ImageOp()
    op.ih, op.iw, op.src_stride = ih, iw, img.strides[0]
    op.top, op.left, op.ch, op.cw = int(top), int(left), int(ch), int(cw)
    op.vh, op.vw, op.oy, op.ox, op.oh, op.ow = int(vh), int(vw), int(oy), int(ox), int(oh), int(ow)
    op.method, op.flip = METHODS[method], int(bool(flip))
    if distort is not None:
        op.distort_order, op.brightness, op.saturation, op.hue, op.contrast = distort
    op.scale, op.bias, op.out_f16 = scale, bias, int(bool(out_f16))
    return op


def make_crop_and_resize_op(img: np.ndarray, box, size: Tuple[int, int], flip: bool = False,
                            distort: Optional[Tuple[int, float, float, float, float]] = None,
                            scale3=(1.0, 1.0, 1.0), bias3=(0.0, 0.0, 0.0), clamp: bool = True,
                            out_f16: bool = False) -> ImageOp:
    """tf.image.crop_and_resize of the normalised window ``box`` = (y0, x0, y1, x1) to ``size``,
    then mirror, colour jitter (on the 0..255 scale) and per-channel ``v * scale3 + bias3``:
    the SSD training transform (T/ssd_dataloader.py:218-254, T/preprocessing.py:792-800)."""
    op = ImageOp()
    op.ih, op.iw, op.src_stride = img.shape[0], img.shape[1], img.strides[0]
    op.oh, op.ow = int(size[0]), int(size[1])
    op.method, op.flip = CROP_AND_RESIZE, int(bool(flip))
    op.box = (ctypes.c_float * 4)(*[float(v) for v in box])
    if distort is not None:
        op.distort_order, op.brightness, op.saturation, op.hue, op.contrast = distort
    op.per_channel = 1
    op.scale3 = (ctypes.c_float * 3)(*[float(v) for v in scale3])
    op.bias3 = (ctypes.c_float * 3)(*[float(v) for v in bias3])
    op.no_clamp = int(not clamp)
    op.out_f16 = int(bool(out_f16))
    return op


def transform_into(img: np.ndarray, op: ImageOp, out_ptr: int) -> None:
    """Run the native transform; ``img`` is uint8 HxWx3 with contiguous pixels, ``out_ptr`` the
    address of oh*ow*3 fp32 (or fp16) elements."""
    if img.dtype != np.uint8 or img.ndim != 3 or img.shape[2] != 3 or img.strides[1] != 3 \
            or img.strides[2] != 1:
        raise ValueError("native image transform needs a uint8 HxWx3 image with packed pixels")
    rc = _function()(img.ctypes.data, ctypes.byref(op), out_ptr)
    if rc != 0:
        raise ValueError(native_reader._load().kfb_image_last_error().decode())


def transform(img: np.ndarray, op: ImageOp) -> np.ndarray:
    out = np.empty((op.oh, op.ow, 3), dtype=np.float16 if op.out_f16 else np.float32)
    transform_into(img, op, out.ctypes.data)
    return out


# ---------------------------------------------------------------------------------------------
# NumPy restatement (oracle of the tests, and the fallback when the library is not built)
# ---------------------------------------------------------------------------------------------
def _cubic_weights(frac: np.ndarray) -> np.ndarray:
    a = -0.75
    x = np.rint(frac * 1024.0) / 1024.0

    def near(t):
        return ((a + 2.0) * t - (a + 3.0)) * t * t + 1.0

    def far(t):
        return ((a * t - 5.0 * a) * t + 8.0 * a) * t - 4.0 * a
    return np.stack([far(x + 1.0), near(x), near(1.0 - x), far(2.0 - x)], axis=-1)


def _axis_matrix(n_in: int, n_virtual: int, off: int, n_out: int, method: int) -> np.ndarray:
    """[n_out, n_in] interpolation matrix of one axis."""
    s = np.float32(n_in) / np.float32(n_virtual)
    m = np.zeros((n_out, n_in), dtype=np.float64)
    o = np.arange(off, off + n_out)
    pos = (o * s).astype(np.float32)
    base = np.floor(pos).astype(int)
    rows = np.arange(n_out)
    if method == 1:
        m[rows, np.minimum(base, n_in - 1)] = 1.0
    elif method == 0:
        lo = np.minimum(base, n_in - 1)
        hi = np.minimum(lo + 1, n_in - 1)
        w = (pos - np.floor(pos)).astype(np.float64)
        np.add.at(m, (rows, lo), 1.0 - w)
        np.add.at(m, (rows, hi), w)
    elif method == 2:
        w = _cubic_weights((pos - base).astype(np.float32)).astype(np.float64)
        for k in range(4):
            np.add.at(m, (rows, np.clip(base - 1 + k, 0, n_in - 1)), w[:, k])
    else:
        for r, oo in enumerate(o):
            a, b = np.float32(oo * s), np.float32((oo + 1) * s)
            for i in range(int(np.floor(a)), int(np.ceil(b))):
                if i < a:
                    wt = (b - a) if i + 1 > b else (i + 1 - a)
                else:
                    wt = (b - i) if i + 1 > b else 1.0
                m[r, min(i, n_in - 1)] += float(wt) / float(s)
    return m


def rgb_to_hsv(x: np.ndarray):
    """tf.image.rgb_to_hsv arithmetic on arbitrary floats (last axis RGB)."""
    r, g, b = x[..., 0], x[..., 1], x[..., 2]
    v = x.max(axis=-1)
    rng = v - x.min(axis=-1)
    s = np.where(v > 0, rng / np.where(v > 0, v, 1.0), 0.0)
    norm = 1.0 / (6.0 * np.where(rng > 0, rng, 1.0))
    h = np.where(r == v, norm * (g - b),
                 np.where(g == v, norm * (b - r) + 2.0 / 6.0, norm * (r - g) + 4.0 / 6.0))
    h = np.where(rng > 0, h, 0.0)
    return np.where(h < 0, h + 1.0, h), s, v


def hsv_to_rgb(h, s, v) -> np.ndarray:
    c = s * v
    m = v - c
    dh = h * 6.0
    x = c * (1.0 - np.abs(dh - 2.0 * np.floor(dh * 0.5) - 1.0))
    cat = dh.astype(np.int64)
    z = np.zeros_like(c)
    r = np.select([cat == 0, cat == 1, cat == 2, cat == 3, cat == 4], [c, x, z, z, x], c)
    g = np.select([cat == 0, cat == 1, cat == 2, cat == 3, cat == 4], [x, c, c, x, z], z)
    b = np.select([cat == 0, cat == 1, cat == 2, cat == 3, cat == 4], [z, z, x, c, c], x)
    return np.stack([r + m, g + m, b + m], axis=-1)


def distort_reference(x: np.ndarray, order: int, brightness: float, saturation: float,
                      hue: float, contrast: float, clamp: bool = True) -> np.ndarray:
    """Colour jitter on 0..255 floats with the arithmetic of tf.image.adjust_brightness /
    adjust_saturation / adjust_hue / adjust_contrast; order 1 = brightness, saturation + hue,
    contrast; order 2 = brightness, contrast, saturation + hue; clipped at the end
    (T/preprocessing.py:268-307)."""
    x = x.astype(np.float64) + brightness

    def sat_hue(v):
        h, s, val = rgb_to_hsv(v)
        s = np.clip(s * saturation, 0.0, 1.0)
        h = h + hue
        return hsv_to_rgb(h - np.floor(h), s, val)

    def con(v):
        m = v.mean(axis=(0, 1), keepdims=True)
        return m + (v - m) * contrast
    x = con(sat_hue(x)) if order == 1 else sat_hue(con(x))
    return np.clip(x, 0.0, 255.0) if clamp else x


def _crop_and_resize_reference(img: np.ndarray, op: ImageOp) -> np.ndarray:
    H, W = img.shape[:2]
    y0, x0, y1, x1 = [np.float32(v) for v in op.box]
    out = np.zeros((op.oh, op.ow, 3), np.float64)
    f32 = np.float32
    for i in range(op.oh):
        iy = y0 * f32(H - 1) + f32(i) * ((y1 - y0) * f32(H - 1) / f32(op.oh - 1)) if op.oh > 1 \
            else f32(0.5) * (y0 + y1) * f32(H - 1)
        if iy < 0 or iy > H - 1:
            continue
        t, b, wy = int(np.floor(iy)), int(np.ceil(iy)), float(iy - np.floor(iy))
        for j in range(op.ow):
            jj = op.ow - 1 - j if op.flip else j
            ix = x0 * f32(W - 1) + f32(jj) * ((x1 - x0) * f32(W - 1) / f32(op.ow - 1)) \
                if op.ow > 1 else f32(0.5) * (x0 + x1) * f32(W - 1)
            if ix < 0 or ix > W - 1:
                continue
            lo, hi, wx = int(np.floor(ix)), int(np.ceil(ix)), float(ix - np.floor(ix))
            top = img[t, lo] + (img[t, hi].astype(np.float64) - img[t, lo]) * wx
            bot = img[b, lo] + (img[b, hi].astype(np.float64) - img[b, lo]) * wx
            out[i, j] = top + (bot - top) * wy
    return out


def transform_reference(img: np.ndarray, op: ImageOp) -> np.ndarray:
    if op.method == CROP_AND_RESIZE:
        x = _crop_and_resize_reference(img, op)
        if op.distort_order:
            x = distort_reference(x, op.distort_order, op.brightness, op.saturation, op.hue,
                                  op.contrast, clamp=not op.no_clamp)
        x = x * np.asarray(list(op.scale3)) + np.asarray(list(op.bias3))
        return x.astype(np.float16 if op.out_f16 else np.float32)
    crop = img[op.top:op.top + op.ch, op.left:op.left + op.cw].astype(np.float64)
    if op.flip:                       # the reference flips the crop, then resizes it
        crop = crop[:, ::-1]
    my = _axis_matrix(op.ch, op.vh, op.oy, op.oh, op.method)
    mx = _axis_matrix(op.cw, op.vw, op.ox, op.ow, op.method)
    x = np.einsum("yi,ijc,xj->yxc", my, crop, mx)
    if op.distort_order:
        x = distort_reference(x, op.distort_order, op.brightness, op.saturation, op.hue,
                              op.contrast)
    x = x * op.scale + op.bias
    return x.astype(np.float16 if op.out_f16 else np.float32)
