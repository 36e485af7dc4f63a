"""The JPEG frame-header scan that routes records to the device decoder (data/nvjpeg.py):
baseline 1- / 3-component bitstreams go to nvJPEG, everything else (progressive, CMYK, PNG in a
.JPEG record - all present in ImageNet, T/preprocessing.py:95-104 decodes them with TF) stays on
the Pillow path."""
import io

import numpy as np
import pytest

from kf_benchmarks_b200.data import nvjpeg

Image = pytest.importorskip("PIL.Image")


def _img(h=37, w=53):
    rng = np.random.RandomState(0)
    return Image.fromarray(rng.randint(0, 255, (h, w, 3)).astype(np.uint8))


def _bytes(img, **kw):
    b = io.BytesIO()
    img.save(b, **kw)
    return b.getvalue()


def test_baseline_colour_and_grey_go_to_the_device():
    d = _bytes(_img(), format="JPEG", quality=90)
    assert nvjpeg.jpeg_header(d) == (37, 53, 3, 0xC0)
    assert nvjpeg.device_decodable(d) == (37, 53)
    g = _bytes(_img().convert("L"), format="JPEG")
    assert nvjpeg.jpeg_header(g) == (37, 53, 1, 0xC0)
    assert nvjpeg.device_decodable(g) == (37, 53)


def test_progressive_cmyk_and_non_jpeg_stay_on_the_host():
    p = _bytes(_img(), format="JPEG", progressive=True)
    assert nvjpeg.jpeg_header(p)[3] == 0xC2 and nvjpeg.device_decodable(p) is None
    c = _bytes(_img().convert("CMYK"), format="JPEG")
    assert nvjpeg.jpeg_header(c)[2] == 4 and nvjpeg.device_decodable(c) is None
    png = _bytes(_img(), format="PNG")
    assert nvjpeg.jpeg_header(png) is None and nvjpeg.device_decodable(png) is None
    assert nvjpeg.jpeg_header(b"") is None and nvjpeg.jpeg_header(b"\xff\xd8\xff") is None


def test_markers_before_the_frame_header_are_skipped():
    # EXIF / comment segments of arbitrary length precede SOF0 in camera files
    d = _bytes(_img(64, 48), format="JPEG", quality=80)
    comment = b"\xff\xfe" + (1000 + 2).to_bytes(2, "big") + b"x" * 1000
    app1 = b"\xff\xe1" + (300 + 2).to_bytes(2, "big") + b"\xff\xc0" * 150   # decoy bytes inside
    patched = d[:2] + comment + app1 + d[2:]
    assert nvjpeg.jpeg_header(patched) == (64, 48, 3, 0xC0)
    assert np.asarray(Image.open(io.BytesIO(patched))).shape == (64, 48, 3)
    # truncated in the middle of the frame header
    sof = patched.index(b"\xff\xc0", len(comment) + len(app1))
    assert nvjpeg.jpeg_header(patched[:sof + 6]) is None
