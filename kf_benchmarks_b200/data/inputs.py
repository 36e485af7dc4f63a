"""Input sources: where a step's (images, labels) come from.

* ``SyntheticDeviceSource`` - the headline path. Images ~ truncated normal(127, 60) and labels
  uniform in [0, nclass-1) are generated once on the device by the Philox kernels
  (``kfb_synth_images`` / ``kfb_synth_labels``) and reused every step, like the reference's
  cached ``gpu_cached_inputs`` (T/models/model.py:220-237, T/benchmark_cnn.py:1945).
* ``SyntheticHostSource`` - the same distribution produced on the host into pinned memory and
  copied host->device every step (used for end-to-end timing and the CPU path).
* real data (ImageNet TFRecords / CIFAR-10 pickles) through ``data.preprocessing``.
"""
from __future__ import annotations

import ctypes

import torch


class InputSource:
    def next_batch(self):
        raise NotImplementedError


class SyntheticDeviceSource(InputSource):
    def __init__(self, batch, image_size, depth, nclass, device, seed):
        from ..ops import native
        from ..ops.native import f32, i32, i64, ptr
        self.images = torch.empty((batch, image_size, image_size, depth), dtype=torch.float16,
                                  device=device)
        self.labels = torch.empty((batch,), dtype=torch.int32, device=device)
        s = native.stream_ptr()
        native.call("synth_images", ptr(self.images), i64(self.images.numel()), f32(127.0),
                    f32(60.0), ctypes.c_ulonglong(seed), s)
        native.call("synth_labels", ptr(self.labels), i32(batch), i32(0), i32(max(nclass - 1, 1)),
                    ctypes.c_ulonglong(seed + 1), s)

    def next_batch(self):
        return self.images, self.labels


class SyntheticHostSource(InputSource):
    """A small ring of pinned host batches; every step pays the H2D copy."""
    host_resident = True

    def __init__(self, batch, image_size, depth, nclass, device, seed, ring=2,
                 dtype=torch.float16):
        gen = torch.Generator().manual_seed(seed)
        self.device = torch.device(device)
        self.ring = []
        pin = self.device.type == "cuda"
        for _ in range(ring):
            img = torch.empty((batch, image_size, image_size, depth), dtype=torch.float32)
            torch.nn.init.trunc_normal_(img, mean=127.0, std=60.0, a=127.0 - 120.0,
                                        b=127.0 + 120.0, generator=gen)
            img = img.to(dtype)
            lab = torch.randint(0, max(nclass - 1, 1), (batch,), generator=gen, dtype=torch.int32)
            if pin:
                img, lab = img.pin_memory(), lab.pin_memory()
            self.ring.append((img, lab))
        self.i = 0
        self.bytes_per_step = (self.ring[0][0].numel() * self.ring[0][0].element_size() +
                               self.ring[0][1].numel() * 4)

    def next_batch(self):
        b = self.ring[self.i % len(self.ring)]
        self.i += 1
        return b


def create_input_source(params, dataset, model, batch, device, nclass, seed, rank, world,
                        train=True) -> InputSource:
    device = torch.device(device)
    if hasattr(model, "make_input_source"):
        # models whose inputs are not images (NCF ids, spectrograms, detection targets)
        import inspect
        kw = {}
        accepted = inspect.signature(model.make_input_source).parameters
        if "world" in accepted:
            kw["world"] = world
        if "train" in accepted:
            kw["train"] = train
        return model.make_input_source(batch, device, seed, rank, **kw)
    image_size, depth = model.image_size, model.depth
    if dataset.use_synthetic_gpu_inputs():
        if device.type == "cuda":
            return SyntheticDeviceSource(batch, image_size, depth, nclass, device,
                                         seed * 1000003 + rank)
        return SyntheticHostSource(batch, image_size, depth, nclass, device,
                                   seed * 1000003 + rank, dtype=torch.float32)
    from . import preprocessing
    return preprocessing.create_real_data_source(params, dataset, model, batch, device, nclass,
                                                 seed, rank, world, train)
