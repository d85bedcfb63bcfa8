"""GPU parity tests of the JPEG half of the image path (cm_blur_batch_jpeg*): the files are
byte-identical to what the reference's calls write -- `image.filter(GaussianBlur(blur))`
(src/backend.py:322-324) then `image.save(img_io, 'JPEG')` (main.py:100-103) -- with live Pillow
as the oracle, and to oracle/jpeg_oracle.py (itself pinned against Pillow on the CPU)."""
import ctypes
import io

import numpy as np
import pytest
from PIL import Image, ImageFilter

from oracle import blur_oracle as B
from oracle import jpeg_oracle as J
from tests.test_jpeg_oracle import make_image

pytestmark = pytest.mark.gpu


def pillow_file(pil, radius, **kw) -> bytes:
    buf = io.BytesIO()
    pil.filter(ImageFilter.GaussianBlur(float(radius))).save(buf, "JPEG", **kw)
    return buf.getvalue()


def _first_diff(a: bytes, b: bytes) -> str:
    n = min(len(a), len(b))
    d = next((i for i in range(n) if a[i] != b[i]), n)
    return f"lengths {len(a)} / {len(b)}, first difference at byte {d}"


@pytest.mark.parametrize("h,w", [(16, 16), (48, 40), (1, 1), (7, 5), (17, 33), (31, 16), (64, 64), (100, 75),
                                 (8, 24), (9, 9), (15, 47), (33, 17), (2, 300), (300, 2), (129, 130), (257, 511)])
def test_files_equal_pillow_on_small_and_odd_sizes(h, w):
    from cassmantle_b200.blur import ImageBlur
    rng = np.random.default_rng(h * 1000 + w)
    blur = ImageBlur(h, w, 3, max_batch=5)                       # smaller than the batch: several calls
    for kind in ("rand", "smooth", "flat", "photo"):
        img = make_image(kind, h, w, rng)
        pil = Image.fromarray(img, "RGB")
        radii = [0.0, 0.7, 2.5, 14.9, 40.0, -1.5, float(rng.uniform(0, 15))]
        blur.set_image(pil)
        got = blur.blur_radii_jpeg(radii)
        for r, data in zip(radii, got):
            want = pillow_file(pil, r)
            assert data == want, (kind, r, _first_diff(data, want))
        assert got[0] == J.encode_rgb(img)                       # radius 0: the unblurred image, vs the numpy oracle
        assert got[2] == J.encode_rgb(B.gaussian_blur(img, 2.5))
    blur.close()


def test_full_size_reference_image():
    """1024 x 1024 RGB (the reference's SDXL image), the radius ladder of the session scores."""
    from cassmantle_b200.blur import ImageBlur
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:1024, 0:1024]
    img = np.stack([(xx + yy) % 256, (xx * 3) % 256, (yy * 7 + xx // 3) % 256], -1).astype(np.uint8)
    img = np.clip(img.astype(int) + rng.integers(-9, 9, img.shape), 0, 255).astype(np.uint8)
    pil = Image.fromarray(img, "RGB")
    blur = ImageBlur(1024, 1024, 3, max_batch=8)
    blur.set_image(pil)
    scores = [0.01, 0.4567891, 0.7071068, 0.99999994, 1.0, 0.1, 0.25, 0.9]
    files = blur.mask_images_jpeg(scores)
    for s, data in zip(scores, files):
        want = pillow_file(pil, blur.score_to_blur(s))            # src/backend.py:319-324 + main.py:100-103
        assert data == want, (s, _first_diff(data, want))
        assert Image.open(io.BytesIO(data)).size == (1024, 1024)
    again = blur.mask_images_jpeg(scores[:3])                     # scratch is reused by the next call
    assert again == files[:3]
    blur.close()


def test_incompressible_image_grows_the_output_block():
    """Noise at quality 100 is larger than the page-locked block the object starts with (a quarter
    of the raw size): the call grows it and still returns Pillow's bytes."""
    from cassmantle_b200.blur import ImageBlur
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (256, 256, 3), dtype=np.uint8)
    pil = Image.fromarray(img, "RGB")
    blur = ImageBlur(256, 256, 3, max_batch=4)
    blur.set_image(pil)
    for quality in (100, 95, 75, 30, 1):
        files = blur.blur_radii_jpeg([0.0, 0.0, 0.3, 0.0], quality=quality)
        assert files[0] == pillow_file(pil, 0.0, quality=quality), quality
        assert files[2] == pillow_file(pil, 0.3, quality=quality), quality
        assert files[1] == files[0] == files[3]
    assert len(blur.blur_radii_jpeg([0.0], quality=100)[0]) > 256 * 256 * 3 // 4
    blur.close()


def test_c_abi_copying_variant_and_errors():
    from cassmantle_b200 import _lib
    from cassmantle_b200.blur import ImageBlur
    lib = _lib.load()
    img = make_image("photo", 64, 80, np.random.default_rng(1))
    pil = Image.fromarray(img, "RGB")
    blur = ImageBlur(64, 80, 3, max_batch=3)
    radii = np.array([1.0, 0.0, 7.5], np.float32)
    offsets = np.zeros(4, np.int64)
    # before an image is set
    assert lib.cm_blur_batch_jpeg(blur._h, radii.ctypes.data, 3, 0, None, 0, offsets.ctypes.data) == -4
    blur.set_image(pil)
    # too small a destination: CM_ERR_NOMEM, offsets say how much is needed
    rc = lib.cm_blur_batch_jpeg(blur._h, radii.ctypes.data, 3, 0, None, 0, offsets.ctypes.data)
    assert rc == -3 and offsets[0] == 0 and offsets[3] > 3 * 623 and b"need" in lib.cm_last_error()
    out = np.zeros(int(offsets[3]), np.uint8)
    rc = lib.cm_blur_batch_jpeg(blur._h, radii.ctypes.data, 3, 0, out.ctypes.data, out.size, offsets.ctypes.data)
    assert rc == 0, lib.cm_last_error()
    for i, r in enumerate(radii):
        assert out[offsets[i]:offsets[i + 1]].tobytes() == pillow_file(pil, r)      # quality 0 = Pillow's default
    assert lib.cm_blur_batch_jpeg(blur._h, radii.ctypes.data, 3, 101, out.ctypes.data, out.size, offsets.ctypes.data) == -1
    assert lib.cm_blur_batch_jpeg(blur._h, radii.ctypes.data, 4, 0, out.ctypes.data, out.size, offsets.ctypes.data) == -1
    blur.close()
    gray = ImageBlur(32, 32, 1, max_batch=1)                      # Pillow would write a 1-component file: not built
    gray.set_image(np.zeros((32, 32), np.uint8))
    one = np.array([1.0], np.float32)
    assert lib.cm_blur_batch_jpeg(gray._h, one.ctypes.data, 1, 0, out.ctypes.data, out.size, offsets.ctypes.data) == -1
    assert b"RGB" in lib.cm_last_error()
    gray.close()
