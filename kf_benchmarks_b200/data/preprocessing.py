"""Real-data input pipelines: ImageNet TFRecords and CIFAR-10, on host threads, delivered to the
device through a pinned double buffer (the executor prefetches the next batch on a copy stream).

Image transforms follow the reference (T/preprocessing.py:130-307,653-737):

* normalisation ``x / 127.5 - 1``;
* ImageNet eval: resize so the short side covers 1.15x the target, centre crop;
* ImageNet train: sample a distorted bounding box (object coverage >= 0.1, aspect ratio in
  [3/4, 4/3], area in [0.05, 1], up to 100 attempts), crop, random horizontal flip, resize
  (bilinear / nearest / bicubic / area, ``round_robin`` cycles through them by batch position);
  with ``--distortions`` also the colour jitter (brightness 32/255, saturation [0.5, 1.5] and
  hue +-0.2 in HSV, contrast [0.5, 1.5], order alternating with batch position);
* training records pass through a 10 000-record shuffle buffer like the reference's
  ``ds.shuffle(buffer_size=10000)``;
* CIFAR-10 train: pad 4, random 32x32 crop, random flip; eval: centre crop / pad.

A background ``ImageProducer`` keeps ``2 * batch_group_size`` batches staged
(T/cnn_util.py:118-198); ``--datasets_repeat_cached_sample`` repeats one decoded sample to take
IO out of a benchmark; workers start at different file offsets (``shift_ratio`` =
rank / world, T/benchmark_cnn.py:3217-3221).
"""
from __future__ import annotations

import glob
import io
import os
import random
from typing import Optional, Tuple

import numpy as np
import torch

from ..utils.cnn_util import ImageProducer
from . import tfrecord
from .inputs import InputSource

# records held by the training shuffle buffer (ds.shuffle(buffer_size=10000),
# T/preprocessing.py:536-537)
SHUFFLE_BUFFER_SIZE = 10000

_RESIZE = {"nearest": 0, "bilinear": 2, "bicubic": 3, "area": 4}  # PIL resample codes
_ROUND_ROBIN = ["bilinear", "nearest", "bicubic", "area"]


def normalized_image(img: np.ndarray) -> np.ndarray:
    return img.astype(np.float32) * (1.0 / 127.5) - 1.0


def get_image_resize_method(resize_method: str, batch_position: int = 0) -> str:
    if resize_method == "round_robin":
        return _ROUND_ROBIN[batch_position % len(_ROUND_ROBIN)]
    if resize_method not in _RESIZE:
        raise ValueError("Invalid resize_method %s" % resize_method)
    return resize_method


def _resize(img: np.ndarray, h: int, w: int, method: str) -> np.ndarray:
    from PIL import Image
    pil = Image.fromarray(img.astype(np.uint8))
    return np.asarray(pil.resize((w, h), resample=_RESIZE[method]))


def eval_geometry(ih: int, iw: int, height: int, width: int):
    """((resized h, w), (top, left, h, w) of the centre crop): the short side is resized to cover
    1.15x the target (T/preprocessing.py:216-259)."""
    ratio = max(height / float(ih), width / float(iw))
    rh, rw = max(int(ih * ratio * 1.15), height), max(int(iw * ratio * 1.15), width)
    return (rh, rw), ((rh - height) // 2, (rw - width) // 2, height, width)


def eval_image(img: np.ndarray, height: int, width: int, batch_position: int = 0,
               resize_method: str = "bilinear") -> np.ndarray:
    (rh, rw), (top, left, _, _) = eval_geometry(img.shape[0], img.shape[1], height, width)
    out = _resize(img, rh, rw, get_image_resize_method(resize_method, batch_position))
    return out[top:top + height, left:left + width]


def sample_distorted_bounding_box(ih: int, iw: int, bbox: Optional[Tuple[float, ...]], rng,
                                  min_object_covered=0.1, aspect=(0.75, 1.33), area=(0.05, 1.0),
                                  max_attempts=100) -> Tuple[int, int, int, int]:
    """(top, left, h, w) of a random crop; falls back to the whole image."""
    if bbox is None:
        bbox = (0.0, 0.0, 1.0, 1.0)  # ymin, xmin, ymax, xmax
    by0, bx0, by1, bx1 = bbox
    barea = max((by1 - by0) * (bx1 - bx0), 1e-6)
    for _ in range(max_attempts):
        a = rng.uniform(area[0], area[1]) * ih * iw
        r = rng.uniform(aspect[0], aspect[1])
        w = int(round((a * r) ** 0.5))
        h = int(round((a / r) ** 0.5))
        if w < 1 or h < 1 or w > iw or h > ih:
            continue
        top, left = rng.randint(0, ih - h), rng.randint(0, iw - w)
        y0, x0, y1, x1 = top / ih, left / iw, (top + h) / ih, (left + w) / iw
        inter = max(0.0, min(y1, by1) - max(y0, by0)) * max(0.0, min(x1, bx1) - max(x0, bx0))
        if inter / barea >= min_object_covered:
            return top, left, h, w
    return 0, 0, ih, iw


def draw_color_distortion(batch_position: int, rng):
    """(order, brightness delta, saturation factor, hue shift, contrast factor) of one example:
    the random parameters of the reference's distort_color (T/preprocessing.py:268-307) -
    brightness +-32/255, saturation and contrast in [0.5, 1.5], hue +-0.2 turns; even batch
    positions apply saturation / hue before contrast, odd ones after."""
    return (1 if batch_position % 2 == 0 else 2, rng.uniform(-32.0, 32.0), rng.uniform(0.5, 1.5),
            rng.uniform(-0.2, 0.2), rng.uniform(0.5, 1.5))


def distort_color(img: np.ndarray, batch_position: int, rng) -> np.ndarray:
    """Brightness / saturation / hue / contrast jitter on [0, 255] floats (NumPy path; the
    native transform applies the same arithmetic)."""
    from .native_image import distort_reference
    return distort_reference(img.astype(np.float32), *draw_color_distortion(batch_position, rng)) \
        .astype(np.float32)


def train_image(img: np.ndarray, height: int, width: int, bbox, batch_position: int,
                resize_method: str, distortions: bool, rng, distort_colors: bool = True):
    """Training transform of the reference (T/preprocessing.py:192-265): sample a distorted
    bounding box, crop, random horizontal flip, resize - always - and, with --distortions, the
    colour jitter. (NumPy / PIL path; the native transform is used when the host library is
    built.)"""
    top, left, h, w = sample_distorted_bounding_box(img.shape[0], img.shape[1], bbox, rng)
    img = img[top:top + h, left:left + w]
    if rng.random() < 0.5:
        img = img[:, ::-1]
    img = _resize(np.ascontiguousarray(img), height, width,
                  get_image_resize_method(resize_method, batch_position))
    if distortions and distort_colors:
        img = distort_color(img, batch_position, rng)
    return np.ascontiguousarray(img)


class _ImageDims:
    """Stand-in for a decoded image where only its geometry is needed (the pixels are decoded
    on the device): what ``native_image.make_op`` and the crop samplers read."""

    def __init__(self, h: int, w: int):
        self.shape = (h, w, 3)
        self.strides = (w * 3, 3, 1)


def decode_jpeg(buf: bytes) -> np.ndarray:
    from PIL import Image
    return np.asarray(Image.open(io.BytesIO(buf)).convert("RGB"))


def parse_example_proto(record: bytes):
    """(image bytes, label, bbox or None, text) - T/preprocessing.py:27-72. Uses the native
    tf.Example scanner (csrc/host/example_scan.cpp) when the host library is built."""
    from . import native_reader
    if native_reader.available():
        buf = native_reader.example_bytes(record, "image/encoded")
        lab = native_reader.example_numbers(record, "image/class/label", cap=4)
        if buf is not None and lab:
            bbox = None
            xmin = native_reader.example_numbers(record, "image/object/bbox/xmin", cap=64)
            if xmin:
                cols = [native_reader.example_numbers(record, "image/object/bbox/" + k, cap=64)
                        for k in ("ymin", "xmin", "ymax", "xmax")]
                bbox = tuple(float(c[0]) for c in cols)
            text = native_reader.example_bytes(record, "image/class/text")
            return buf, int(lab[0]), bbox, text or b""
    ex = tfrecord.parse_example(record)
    label = int(ex["image/class/label"][0])
    bbox = None
    if ex.get("image/object/bbox/xmin"):
        bbox = (float(ex["image/object/bbox/ymin"][0]), float(ex["image/object/bbox/xmin"][0]),
                float(ex["image/object/bbox/ymax"][0]), float(ex["image/object/bbox/xmax"][0]))
    text = ex.get("image/class/text", [b""])
    return ex["image/encoded"][0], label, bbox, text[0] if text else b""


_logged = set()


def _log_once(msg: str) -> None:
    if msg not in _logged:
        _logged.add(msg)
        from ..utils import log_util
        log_util.log_fn(msg)


def _decode_workers(params) -> int:
    """Threads that decode / transform examples of one batch in parallel
    (--datasets_num_private_threads; default: this rank's share of the cores, at most 16)."""
    n = getattr(params, "datasets_num_private_threads", None)
    if n:
        return max(int(n), 1)
    local = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")) or 1)
    return max(1, min(16, (os.cpu_count() or 2) // max(local, 1)))


class _BatchSource(InputSource):
    """Common machinery: a producer thread filling a ring of (pinned) batch buffers, the
    examples of one batch spread over a pool of worker threads. JPEG decode (Pillow) and the
    native transform (csrc/host/image_ops.cpp) both run without the GIL, so the pool scales
    with the cores like the reference's tf.data map(num_parallel_calls) does
    (T/preprocessing.py:566-651). Augmentation randomness is drawn per example from a seed
    taken in order on the producer thread: the batches do not depend on the thread count."""
    host_resident = True

    def __init__(self, params, batch, height, width, device, train, seed):
        from concurrent.futures import ThreadPoolExecutor
        self.p = params
        self.batch, self.h, self.w = batch, height, width
        self.device = torch.device(device)
        self.train = train
        self.rng = random.Random(seed)
        self.out_dtype = torch.float16 if self.device.type == "cuda" else torch.float32
        group = max(int(params.batch_group_size or 1), 1)
        pin = self.device.type == "cuda" and torch.cuda.is_available()
        # 2*group batches may be staged, one is being filled, one or two are still being read
        # by the consumer (host->device copy in flight): a slot comes round again after that
        self._ring = [(torch.empty((batch, height, width, 3), dtype=self.out_dtype,
                                   pin_memory=pin),
                       torch.empty((batch,), dtype=torch.int32, pin_memory=pin))
                      for _ in range(2 * group + 3)]
        self._ring_pos = 0
        nthreads = _decode_workers(params)
        self._pool = ThreadPoolExecutor(nthreads, thread_name_prefix="kfb-decode") \
            if nthreads > 1 else None
        # --gpu_preprocess: workers hand over (decoded image, descriptor) pairs and the batch is
        # transformed by one CUDA kernel (data/gpu_image.py); the source then yields device
        # tensors, one output buffer per ring slot
        self._gpu = None
        if getattr(params, "gpu_preprocess", False) and pin:
            from . import gpu_image, native_image
            if native_image.available():
                self._gpu = gpu_image.GpuImageBatcher(self.device, batch, height, width,
                                                      self.out_dtype, slots=len(self._ring))
                self._gpu_stream = torch.cuda.Stream(device=self.device)
                self._labels_dev = [torch.empty((batch,), dtype=torch.int32, device=self.device)
                                    for _ in range(len(self._ring))]
                self.host_resident = False
                self._raw_jobs = [None] * batch
                # JPEG decode on the device as well (data/nvjpeg.py): workers then hand over
                # (bitstream, descriptor, dims) for baseline JPEGs; KFB_NVJPEG=0 keeps Pillow
                self._jpeg = None
                self._final = {}
                if os.environ.get("KFB_NVJPEG", "1") != "0":
                    from . import nvjpeg
                    if nvjpeg.available():
                        try:
                            self._jpeg = nvjpeg.NvJpegDecoder(
                                self.device, int(os.environ.get("KFB_NVJPEG_BACKEND", "0")),
                                cpu_threads=max(nthreads, 1), slots=len(self._ring))
                            _log_once("JPEG decode on the GPU: nvJPEG backend '%s'"
                                      % self._jpeg.backend)
                        except Exception as e:  # no usable backend on this GPU
                            _log_once("nvJPEG unavailable (%s): decoding with Pillow" % e)
        self.producer = ImageProducer(self._make_batch, group)
        self.producer.start()
        self.bytes_per_step = batch * height * width * 3 * (2 if self.out_dtype ==
                                                            torch.float16 else 4) + batch * 4

    def _draw(self, position: int):
        """Producer thread, in batch order: everything sequential about example `position`
        (next record, its augmentation seed). Returns the job handed to `_fill`."""
        raise NotImplementedError

    def _fill(self, images: torch.Tensor, labels: np.ndarray, position: int, job) -> None:
        """Worker thread: decode / transform the example and write slot `position`."""
        raise NotImplementedError

    def _make_batch(self):
        images, labels_t = self._ring[self._ring_pos % len(self._ring)]
        self._ring_pos += 1
        labels = labels_t.numpy()
        jobs = [self._draw(i) for i in range(self.batch)]
        if self._pool is not None:
            list(self._pool.map(lambda ij: self._fill(images, labels, ij[0], ij[1]),
                                enumerate(jobs)))
        else:
            for i, job in enumerate(jobs):
                self._fill(images, labels, i, job)
        if self._gpu is not None:
            return self._finish_on_gpu(images, labels_t)
        return images, labels_t

    def _finish_on_gpu(self, images, labels_t):
        """Producer thread: positions whose descriptor the device kernel supports were left as
        (image, descriptor) pairs by the workers; they are uploaded raw and transformed in one
        launch on the source's own stream. Positions the host had to transform itself (bicubic /
        area) are copied over from the pinned batch."""
        slot = (self._ring_pos - 1) % len(self._ring)
        jobs = self._raw_jobs
        jidx = [i for i, j in enumerate(jobs) if j is not None and len(j) == 3]   # bitstreams
        if jidx:
            try:
                _, ptrs, pitches = self._jpeg.decode([jobs[i][0] for i in jidx],
                                                     [jobs[i][2] for i in jidx],
                                                     stream=self._gpu_stream)
            except Exception as e:
                # a bitstream nvJPEG rejects: this batch is decoded on the host instead
                _log_once("nvJPEG decode failed (%s): batch decoded with Pillow" % e)
                for i in jidx:
                    jobs[i] = (np.ascontiguousarray(decode_jpeg(jobs[i][0])), jobs[i][1])
                jidx = []
        idx = [i for i, j in enumerate(jobs) if j is not None and len(j) == 2]    # decoded
        with torch.cuda.stream(self._gpu_stream):
            out_j = self._gpu.run_device(ptrs, pitches, [jobs[i][1] for i in jidx],
                                         stream=self._gpu_stream) if jidx else None
            out = self._gpu.run([jobs[i][0] for i in idx], [jobs[i][1] for i in idx],
                                stream=self._gpu_stream) if idx else None
            if len(jidx) == self.batch:
                dev = out_j
            elif len(idx) == self.batch:
                dev = out
            else:
                # mixed batch (device-decoded, host-decoded, host-transformed positions): the
                # pieces are gathered into a batch buffer of this ring slot
                dev = self._final.get(slot)
                if dev is None:
                    dev = self._final[slot] = torch.empty(
                        (self.batch, self.h, self.w, 3), dtype=self.out_dtype, device=self.device)
                if len(jidx) + len(idx) != self.batch:
                    dev.copy_(images, non_blocking=True)      # host-transformed positions
                if jidx:
                    dev.index_copy_(0, torch.tensor(jidx, device=self.device), out_j)
                if idx:
                    dev.index_copy_(0, torch.tensor(idx, device=self.device), out)
            labels_dev = self._labels_dev[slot]
            labels_dev.copy_(labels_t, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._gpu_stream)
        self.bytes_per_step = self._gpu.h2d_bytes_last + self.batch * 4
        self._raw_jobs = [None] * self.batch
        return dev, labels_dev, ev

    def next_batch(self):
        out = self.producer.get()
        if self._gpu is not None:
            torch.cuda.current_stream(self.device).wait_event(out[2])
            return out[0], out[1]
        return out

    def close(self):
        self.producer.done()
        if self._pool is not None:
            self._pool.shutdown(wait=False)


class ImagenetSource(_BatchSource):
    def __init__(self, params, dataset, batch, height, width, device, train, seed, rank, world):
        subset = "train" if train else "validation"
        files = sorted(glob.glob(dataset.tf_record_pattern(subset)))
        if not files:
            raise ValueError("Found no files in --data_dir matching: %s" %
                             dataset.tf_record_pattern(subset))
        shift = int(len(files) * (rank / float(max(world, 1))))
        self.files = files[shift:] + files[:shift]       # shift_ratio = rank / world
        self.p = params
        self.records = self._record_iter()
        self.cached = None
        self._shuffle_n = SHUFFLE_BUFFER_SIZE if train else 0
        self._shuffle_buf = []
        from . import native_image
        self._native = native_image if native_image.available() else None
        super().__init__(params, batch, height, width, device, train, seed * 7919 + rank)

    def _record_iter(self):
        """Endless record stream. With the host library built, files are read and CRC-checked
        by background C++ threads: --datasets_parallel_interleave_cycle_length of them (default
        10 like the reference's interleave, T/preprocessing.py:516-519), file i on thread
        i % n, merged one record per file in turn - or first come first served with
        --datasets_sloppy_parallel_interleave. Otherwise the Python codec, file by file."""
        from . import native_reader
        if native_reader.available():
            cycle = int(getattr(self.p, "datasets_parallel_interleave_cycle_length", None) or 10)
            sloppy = bool(getattr(self.p, "datasets_sloppy_parallel_interleave", False))
            caching = bool(getattr(self.p, "datasets_use_caching", False))
            reader = native_reader.RecordReader(
                self.files, num_threads=max(1, min(cycle, 16, len(self.files))), capacity=512,
                check_crc=True, repeat=not caching, deterministic=not sloppy)
            return self._cached(reader) if caching else iter(reader)
        if getattr(self.p, "datasets_use_caching", False):
            return self._cached(rec for path in self.files for rec in tfrecord.read_records(path))
        return self._python_record_iter()

    @staticmethod
    def _cached(one_pass):
        """--datasets_use_caching (ds.cache(), T/preprocessing.py:533-534): the first pass over
        the files keeps every record in host memory, later passes never touch the disk."""
        cache = []
        for rec in one_pass:
            cache.append(rec)
            yield rec
        while cache:
            for rec in cache:
                yield rec

    def _python_record_iter(self):
        while True:
            order = list(self.files)
            for path in order:
                for rec in tfrecord.read_records(path):
                    yield rec

    def _draw(self, position: int):
        if self.p.datasets_repeat_cached_sample and self.cached is not None:
            record = self.cached
        elif self.p.datasets_repeat_cached_sample:
            # ds.take(1).cache().repeat() sits in front of the shuffle (T/preprocessing.py:
            # 520-522): one record, no buffer to fill
            record = self.cached = next(self.records)
        else:
            record = self._next_record()
            self.cached = record
        return record, self.rng.getrandbits(62)

    def _next_record(self):
        """Next record of the stream; in training through a shuffle buffer: a uniformly drawn
        buffered record leaves, the next record of the stream takes its place."""
        if self._shuffle_n <= 0:
            return next(self.records)
        buf = self._shuffle_buf
        if len(buf) < self._shuffle_n:
            _log_once("Filling up shuffle buffer (this may take a while): %d records"
                      % self._shuffle_n)
            while len(buf) < self._shuffle_n:
                buf.append(next(self.records))
        i = self.rng.randrange(len(buf))
        record, buf[i] = buf[i], next(self.records)
        return record

    def _fill(self, images, labels, position, job) -> None:
        record, seed = job
        p = self.p
        buf, label, bbox, _ = parse_example_proto(record)
        dims = None
        if self._gpu is not None and self._jpeg is not None and self._native is not None:
            from . import nvjpeg
            dims = nvjpeg.device_decodable(buf)
        # with a device decode only the geometry of the image is needed on the host
        img = _ImageDims(dims[0], dims[1]) if dims is not None else decode_jpeg(buf)
        rng = random.Random(seed)
        labels[position] = label
        if self._native is None:
            if self.train:
                out = train_image(img, self.h, self.w, bbox, position, p.resize_method,
                                  p.distortions, rng)
            else:
                out = eval_image(img, self.h, self.w, position, p.resize_method)
            images[position] = torch.from_numpy(normalized_image(out)).to(images.dtype)
            return
        ni = self._native
        method = get_image_resize_method(p.resize_method, position)
        crop, window, flip, distort = None, None, False, None
        resized = (self.h, self.w)
        if self.train:
            crop = sample_distorted_bounding_box(img.shape[0], img.shape[1], bbox, rng)
            flip = rng.random() < 0.5
            if p.distortions:
                distort = draw_color_distortion(position, rng)
        else:
            resized, window = eval_geometry(img.shape[0], img.shape[1], self.h, self.w)
        op = ni.make_op(img, crop, resized, window, method, flip, distort,
                        out_f16=images.dtype == torch.float16)
        if self._gpu is not None:
            from . import gpu_image
            if gpu_image.supports(op):
                self._raw_jobs[position] = (buf, op, dims) if dims is not None else \
                    (np.ascontiguousarray(img), op)
                return
        if dims is not None:                      # (resize method the device kernel lacks)
            img = decode_jpeg(buf)
        ni.transform_into(np.ascontiguousarray(img), op,
                          images.data_ptr() + position * images.stride(0) * images.element_size())


class Cifar10Source(_BatchSource):
    def __init__(self, params, dataset, batch, device, train, seed, rank, world):
        self.images, self.labels = dataset.read_data_files("train" if train else "validation")
        self.order = np.arange(len(self.labels))
        self.pos = 0
        self.np_rng = np.random.RandomState(seed * 7919 + rank)
        if train:
            self.np_rng.shuffle(self.order)
        super().__init__(params, batch, 32, 32, device, train, seed * 7919 + rank)

    def _draw(self, position: int):
        if self.pos >= len(self.order):
            self.pos = 0
            if self.train:
                self.np_rng.shuffle(self.order)
        idx = self.order[self.pos]
        self.pos += 1
        return idx, self.rng.getrandbits(62)

    def _fill(self, images, labels, position, job) -> None:
        idx, seed = job
        rng = random.Random(seed)
        img = self.images[idx]
        if self.train and self.p.distortions:
            padded = np.zeros((40, 40, 3), dtype=img.dtype)
            padded[4:36, 4:36] = img
            top, left = rng.randint(0, 8), rng.randint(0, 8)
            img = padded[top:top + 32, left:left + 32]
            if rng.random() < 0.5:
                img = img[:, ::-1]
        images[position] = torch.from_numpy(normalized_image(np.ascontiguousarray(img))) \
            .to(images.dtype)
        labels[position] = int(self.labels[idx])


_PREPROCESSORS = {"imagenet": {"default": ImagenetSource}, "cifar10": {"default": Cifar10Source}}


def get_preprocessor(dataset_name: str, name: str = "default"):
    try:
        return _PREPROCESSORS[dataset_name][name]
    except KeyError:
        raise ValueError("no input preprocessor %r for dataset %r" % (name, dataset_name))


def create_real_data_source(params, dataset, model, batch, device, nclass, seed, rank, world,
                            train) -> InputSource:
    if dataset.name == "imagenet":
        return ImagenetSource(params, dataset, batch, model.image_size, model.image_size, device,
                              train, seed, rank, world)
    if dataset.name == "cifar10":
        return Cifar10Source(params, dataset, batch, device, train, seed, rank, world)
    raise ValueError("no real-data pipeline for dataset %s" % dataset.name)
