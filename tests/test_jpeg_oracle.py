"""CPU test: oracle/jpeg_oracle.py against the REAL Pillow call the reference makes
(main.py:100-103 `image.save(img_io, 'JPEG')`, src/utils.py:12-16): byte-identical files.
This is what pins the JPEG oracle; the GPU tests compare the CUDA bytes with Pillow too."""
import io

import numpy as np
import pytest
from PIL import Image

from oracle import blur_oracle as B
from oracle import jpeg_oracle as J


def pillow_jpeg(arr: np.ndarray, **kw) -> bytes:
    buf = io.BytesIO()
    Image.fromarray(arr, "RGB").save(buf, "JPEG", **kw)          # the reference passes no options
    return buf.getvalue()


def make_image(kind: str, h: int, w: int, rng) -> np.ndarray:
    if kind == "rand":
        return rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    if kind == "flat":
        return np.full((h, w, 3), int(rng.integers(0, 256)), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.stack([(xx * 3 + yy) % 256, (xx + yy * 2) % 256, (xx * yy) % 256], -1).astype(np.uint8)
    if kind == "photo":                                          # smooth + mild noise, blurred like the served image
        img = np.clip(img.astype(int) + rng.integers(-12, 12, img.shape), 0, 255).astype(np.uint8)
        img = B.gaussian_blur(img, 2.5)
    return img


SIZES = [(48, 40), (16, 16), (1, 1), (7, 5), (17, 33), (31, 16), (64, 64), (100, 75), (8, 24), (9, 9), (15, 47),
         (33, 17), (2, 300), (300, 2), (129, 130)]


@pytest.mark.parametrize("h,w", SIZES)
def test_default_quality_is_byte_identical_to_pillow(h, w):
    rng = np.random.default_rng(h * 1000 + w)
    for kind in ("rand", "smooth", "flat", "photo"):
        img = make_image(kind, h, w, rng)
        assert J.encode_rgb(img) == pillow_jpeg(img), (h, w, kind)


@pytest.mark.parametrize("quality", [1, 30, 50, 75, 90, 95, 100])
def test_other_qualities(quality):
    rng = np.random.default_rng(quality)
    for kind in ("rand", "photo"):
        img = make_image(kind, 40, 56, rng)
        assert J.encode_rgb(img, quality) == pillow_jpeg(img, quality=quality), (quality, kind)


def test_extreme_pixels_and_stuffing():
    """Saturated checkerboards drive the coefficients to their largest magnitudes; a long random
    image makes 0xFF bytes in the entropy stream certain (each must be followed by 0x00)."""
    yy, xx = np.mgrid[0:64, 0:64]
    chk = (((xx + yy) & 1) * 255).astype(np.uint8)
    img = np.stack([chk, 255 - chk, chk], -1)
    assert J.encode_rgb(img) == pillow_jpeg(img)
    img = np.random.default_rng(5).integers(0, 256, (96, 96, 3), dtype=np.uint8)
    data = J.encode_rgb(img, 95)
    assert data == pillow_jpeg(img, quality=95)
    scan = data[len(J.header(96, 96, 95)):-2]
    assert b"\xff\x00" in scan and all(scan[i + 1] == 0 for i in range(len(scan) - 1) if scan[i] == 0xFF)


def test_header_is_what_pillow_writes():
    img = np.zeros((24, 40, 3), np.uint8)
    want = pillow_jpeg(img)
    hdr = J.header(24, 40)
    assert len(hdr) == 623 and want[:623] == hdr and want[-2:] == b"\xff\xd9"
