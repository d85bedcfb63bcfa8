"""GPU parity tests of the image blur (cm_blur_*): bit-identical to Pillow's GaussianBlur, the
call the reference makes at src/backend.py:322-324 (Pillow is installed, so it is the oracle)."""
import os

import numpy as np
import pytest
from PIL import Image, ImageFilter

from oracle import blur_oracle as B

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_golden_outputs_of_real_pillow():
    from cassmantle_b200.blur import ImageBlur
    g = np.load(os.path.join(HERE, "golden", "blur_golden.npz"))
    blur = ImageBlur(64, 64, 3, max_batch=16)
    blur.set_image(g["image"])
    got = blur.blur_radii(g["radii"])
    assert np.array_equal(got, g["outputs"])
    view = blur.blur_radii(g["radii"], copy=False)               # cm_blur_batch_pinned: no host copy
    assert not view.flags.writeable and np.array_equal(view, g["outputs"])
    first = view[:3].copy()
    again = blur.blur_radii(g["radii"][:3], copy=False)          # the block is reused by the next call
    assert np.array_equal(again, first)
    blur.close()


@pytest.mark.parametrize("h,w,c,mode", [(37, 53, 3, "RGB"), (64, 64, 4, "RGBA"), (5, 300, 1, "L"),
                                        (300, 7, 3, "RGB"), (1, 1, 3, "RGB"), (257, 511, 3, "RGB")])
def test_matches_live_pillow(h, w, c, mode):
    from cassmantle_b200.blur import ImageBlur
    rng = np.random.default_rng(h * 1000 + w)
    img = rng.integers(0, 256, (h, w, c), dtype=np.uint8)
    pil = Image.fromarray(img[..., 0] if c == 1 else img, mode)
    # negative radii: Pillow squares the radius, and score_to_blur of a cosine that rounded to
    # 1.0000001 is slightly negative in the reference (src/backend.py:319-320)
    radii = [0.0, 0.1, 0.5, 1.0, 2.5, 7.77, 14.9, 15.0, 40.0, 400.0, -3.6e-6, -2.5] + list(rng.uniform(0, 15, 10))
    blur = ImageBlur(h, w, c, max_batch=7)                       # smaller than the batch: chunked
    blur.set_image(pil)
    got = blur.blur_radii(radii)
    for r, x in zip(radii, got):
        want = np.asarray(pil.filter(ImageFilter.GaussianBlur(float(r)))).reshape(h, w, c)
        assert np.array_equal(x, want), r
        assert np.array_equal(x, B.gaussian_blur(img, float(r))), r
    blur.close()


def test_mask_image_reference_signature_and_full_size():
    """1024 x 1024 RGB (the reference's SDXL image) for a batch of session scores."""
    from cassmantle_b200.blur import ImageBlur
    rng = np.random.default_rng(3)
    yy, xx = np.mgrid[0:1024, 0:1024]
    img = np.stack([(xx + yy) % 256, (xx * 3) % 256, (yy * 7 + xx // 3) % 256], -1).astype(np.uint8)
    img = np.clip(img.astype(int) + rng.integers(-9, 9, img.shape), 0, 255).astype(np.uint8)
    pil = Image.fromarray(img, "RGB")
    blur = ImageBlur(1024, 1024, 3, max_batch=8)
    blur.set_image(pil)
    scores = [0.01, 0.4567891, 0.7071068, 0.99999994, 1.0, 0.1, 0.25, 0.9]
    got = blur.mask_images(scores)
    assert np.array_equal(blur.mask_images(scores, copy=False), got)
    for s, x in zip(scores, got):
        want = np.asarray(pil.filter(ImageFilter.GaussianBlur(blur.score_to_blur(s))))     # src/backend.py:322-324
        assert np.array_equal(x, want), s
    one = blur.mask_image(pil, 0.5)
    assert isinstance(one, Image.Image) and one.mode == "RGB" and one.size == (1024, 1024)
    assert np.array_equal(np.asarray(one), np.asarray(pil.filter(ImageFilter.GaussianBlur(B.score_to_blur(0.5)))))
    blur.close()


def test_blur_argument_errors():
    from cassmantle_b200 import _lib
    from cassmantle_b200.blur import ImageBlur
    with pytest.raises(_lib.CassMantleError):
        ImageBlur(64, 64, 2)
    blur = ImageBlur(32, 32, 3, max_batch=2)
    with pytest.raises(RuntimeError):
        blur.blur_radii([1.0])
    with pytest.raises(_lib.CassMantleError):
        blur.set_image(np.zeros((33, 32, 3), np.uint8))
    blur.set_image(np.zeros((32, 32, 3), np.uint8))
    with pytest.raises(_lib.CassMantleError):
        blur.blur_radii([float("nan")])
    with pytest.raises(_lib.CassMantleError):
        blur.blur_radii([5000.0])
    blur.close()
