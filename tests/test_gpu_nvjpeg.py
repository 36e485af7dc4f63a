"""nvJPEG batch decode (csrc/jpeg_decode.cu, data/nvjpeg.py) against Pillow, and the device input
pipeline (--gpu_preprocess with the decode on the GPU) against the same pipeline decoding with
Pillow. Reference stage: tf.image.decode_jpeg inside tf.data, T/preprocessing.py:95-104."""
import io
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _jpeg(rng, h, w, gray=False, quality=90, subsampling=0):
    from PIL import Image
    small = rng.randint(0, 255, (h // 16 + 1, w // 16 + 1, 3)).astype(np.uint8)
    img = np.asarray(Image.fromarray(small).resize((w, h), Image.BICUBIC)).astype(np.int16)
    img = np.clip(img + rng.randint(-12, 12, (h, w, 3)), 0, 255).astype(np.uint8)
    pil = Image.fromarray(img)
    if gray:
        pil = pil.convert("L")
    buf = io.BytesIO()
    if gray:
        pil.save(buf, format="JPEG", quality=quality)
    else:
        pil.save(buf, format="JPEG", quality=quality, subsampling=subsampling)
    return buf.getvalue()


def test_header_scan_and_decode_match_pillow():
    from PIL import Image
    from kf_benchmarks_b200.data import nvjpeg
    if not nvjpeg.available():
        pytest.skip("libnvjpeg not loadable")
    rng = np.random.RandomState(0)
    shapes = [(375, 500, False), (500, 375, False), (64, 48, False), (301, 203, True),
              (224, 224, False), (33, 1000, False)]
    # 4:4:4 for the first images (only IDCT rounding differs between decoders), 4:2:0 for the
    # last two (chroma upsampling filters differ as well)
    sub = [0, 0, 0, 0, 2, 2]
    datas = [_jpeg(rng, h, w, g, subsampling=ss) for (h, w, g), ss in zip(shapes, sub)]
    dims = []
    for d, (h, w, _) in zip(datas, shapes):
        got = nvjpeg.device_decodable(d)
        assert got == (h, w)
        dims.append(got)
    dec = nvjpeg.NvJpegDecoder("cuda:0", cpu_threads=2)
    for h, w, c in (dec.info(d) for d in datas[:2]):
        assert (h, w) in ((375, 500), (500, 375)) and c == 3
    buf, ptrs, pitches = dec.decode(datas, dims)
    torch.cuda.synchronize()
    base = buf.data_ptr()
    for d, (h, w), p, pitch, ss in zip(datas, dims, ptrs, pitches, sub):
        assert pitch == 3 * w
        got = buf[p - base:p - base + h * w * 3].reshape(h, w, 3).cpu().numpy().astype(np.int16)
        want = np.asarray(Image.open(io.BytesIO(d)).convert("RGB")).astype(np.int16)
        diff = np.abs(got - want)
        if ss == 0:
            assert diff.mean() < 1.0 and diff.max() <= 6, (diff.mean(), diff.max())
        else:
            assert diff.mean() < 5.0, (diff.mean(), diff.max())
    # a second batch of another size re-initialises the batched decoder
    buf2, ptrs2, _ = dec.decode(datas[:3], dims[:3])
    torch.cuda.synchronize()
    assert len(ptrs2) == 3
    dec.close()


_PIPE = """
import os, sys, torch
from kf_benchmarks_b200 import datasets, params as P
from kf_benchmarks_b200.data import preprocessing
p = P.make_params(model='resnet50', batch_size=8, data_dir=sys.argv[1], data_name='imagenet',
                  device='gpu', data_format='NHWC', use_fp16=True, gpu_preprocess=True,
                  datasets_num_private_threads=2)
ds = datasets.create_dataset(sys.argv[1], 'imagenet')
src = preprocessing.ImagenetSource(p, ds, 8, 224, 224, 'cuda', False, 1, 0, 1)
x, y = src.next_batch()
torch.cuda.synchronize()
print('DECODER', 'nvjpeg' if getattr(src, '_jpeg', None) is not None else 'pillow')
torch.save({'x': x.float().cpu(), 'y': y.cpu()}, sys.argv[2])
src.close()
"""


def test_device_pipeline_with_nvjpeg_matches_the_pillow_decode(tmp_path):
    from kf_benchmarks_b200.data import nvjpeg
    if not nvjpeg.available():
        pytest.skip("libnvjpeg not loadable")
    data = str(tmp_path / "imnet")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_fake_imagenet.py"), data,
                    "--images", "16", "--shards", "2", "--size", "120x160"], check=True,
                   cwd=ROOT, capture_output=True)
    outs = {}
    for mode in ("1", "0"):
        env = dict(os.environ, KFB_NVJPEG=mode)
        out = str(tmp_path / ("batch_%s.pt" % mode))
        res = subprocess.run([sys.executable, "-c", _PIPE, data, out], cwd=ROOT, env=env,
                             capture_output=True, text=True, timeout=300)
        assert res.returncode == 0, res.stdout[-1500:] + res.stderr[-1500:]
        assert ("DECODER nvjpeg" if mode == "1" else "DECODER pillow") in res.stdout
        outs[mode] = torch.load(out)
    a, b = outs["1"], outs["0"]
    assert torch.equal(a["y"], b["y"])
    assert a["x"].shape == (8, 224, 224, 3) and torch.isfinite(a["x"]).all()
    # normalised to [-1, 1]: one LSB of the decoded image is 1 / 127.5
    assert float((a["x"] - b["x"]).abs().mean()) < 5.0 / 127.5
