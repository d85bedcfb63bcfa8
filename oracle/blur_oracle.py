"""CPU oracle for the image blur -- TEST INFRASTRUCTURE, NOT PRODUCT (same import rule as the
other oracle modules).

PARITY PINNED.  The reference blurs with ``image.filter(ImageFilter.GaussianBlur(blur))``
(src/backend.py:322-324); the arithmetic lives in Pillow, a third-party dependency that IS
installed here (12.2.0).  This numpy restatement of Pillow's libImaging/BoxBlur.c
(``_gaussian_blur_radius`` + ``ImagingBoxBlur``: three extended-box passes per axis, uint32
arithmetic, a rounding to uint8 after every pass) was checked bit for bit against the real
``GaussianBlur`` on 2,300 radii, L / RGB / RGBA images and images smaller than the radius;
tests/test_oracle.py repeats that comparison on every run, tests/golden/blur_golden.npz holds
outputs of the REAL Pillow call (generator: tests/golden/make_blur_golden.py), and the GPU
tests compare the CUDA path with Pillow itself as well as with this file.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def gaussian_box_radius(radius: float, passes: int = 3) -> np.float32:
    """Pillow ``_gaussian_blur_radius`` (float32 variables, the sqrt/floor in double)."""
    radius = f32(radius)
    sigma2 = f32(radius * radius / f32(passes))
    L = f32(np.sqrt(12.0 * float(sigma2) + 1.0))
    l = f32(np.floor((float(L) - 1.0) / 2.0))
    a = f32(f32(f32(2) * l + f32(1)) * f32(f32(l * f32(l + f32(1))) - f32(f32(3) * sigma2)))
    a = f32(a / f32(f32(6) * f32(sigma2 - f32(f32(l + f32(1)) * f32(l + f32(1))))))
    return f32(l + a)


def box_blur_lines(img: np.ndarray, float_radius) -> np.ndarray:
    """Pillow ``ImagingHorizontalBoxBlur`` on uint8 [lines, n, C], blurring along axis 1."""
    fr = f32(float_radius)
    radius = int(fr)
    ww = int(f32(1 << 24) / f32(fr * f32(2) + f32(1)))
    fw = ((1 << 24) - (radius * 2 + 1) * ww) // 2
    n = img.shape[1]
    last = n - 1
    x = np.arange(n)
    wide = img.astype(np.int64)
    P = np.concatenate([np.zeros((img.shape[0], 1, img.shape[2]), np.int64), np.cumsum(wide, axis=1)], axis=1)
    hi, lo = np.minimum(x + radius, last), np.maximum(x - radius, 0)
    acc = P[:, hi + 1] - P[:, lo]                                    # the running accumulator, in closed form
    acc = acc + np.maximum(radius - x, 0)[None, :, None] * wide[:, :1] \
        + np.maximum(x + radius - last, 0)[None, :, None] * wide[:, last:last + 1]
    far = wide[:, np.maximum(x - radius - 1, 0)] + wide[:, np.minimum(x + radius + 1, last)]
    bulk = (acc * ww + far * fw) & 0xFFFFFFFF                        # UINT32 arithmetic
    return (((bulk + (1 << 23)) & 0xFFFFFFFF) >> 24).astype(np.uint8)


def gaussian_blur(img: np.ndarray, radius: float, passes: int = 3) -> np.ndarray:
    """``Image.filter(ImageFilter.GaussianBlur(radius))`` on uint8 [H, W, C]."""
    if radius == 0:
        return img.copy()                                            # ImageFilter.GaussianBlur.filter
    br = gaussian_box_radius(radius, passes)
    out = img
    for _ in range(passes):
        out = box_blur_lines(out, br)
    t = np.ascontiguousarray(out.transpose(1, 0, 2))                 # ImagingTranspose
    for _ in range(passes):
        t = box_blur_lines(t, br)
    return np.ascontiguousarray(t.transpose(1, 0, 2))


def score_to_blur(score: float, min_blur: float = 0.0, max_blur: float = 15) -> float:
    return min_blur + (1 - score ** 2) * (max_blur - min_blur)       # src/backend.py:319-320
