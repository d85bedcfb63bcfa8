"""ImageBlur -- the per-session image blur of the reference, batched on the GPU.

Mirrors src/backend.py:319-324:

    def score_to_blur(self, score, min_blur=0.0, max_blur=15)
    def mask_image(self, image: Image.Image, score: float) -> Image.Image

``mask_image`` in the reference runs Pillow's ``GaussianBlur`` on the round's 1024x1024 image
for every ``GET /fetch/contents`` (src/server.py:129-133, main.py:95-111) -- the heaviest
per-request CPU work it has.  Here the round's image is uploaded once and ``mask_images``
blurs it for a whole batch of sessions in one go; the bytes equal Pillow's (C ABI: cm_blur_*).
"""
from __future__ import annotations

import ctypes
from typing import Sequence

import numpy as np

from . import _lib
from ._lib import check


class ImageBlur:
    def __init__(self, max_height: int, max_width: int, channels: int = 3, max_batch: int = 64,
                 device: int = 0) -> None:
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        check(self._lib.cm_blur_create(ctypes.byref(self._h), device, max_height, max_width, channels,
                                       max_batch))
        self.channels, self.max_batch = channels, max_batch
        self.shape = None

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.cm_blur_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self) -> None:  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def set_image(self, image) -> None:
        """The round's image: a PIL image (mode L / RGB / RGBA) or a uint8 array [H, W(, C)]."""
        arr = np.ascontiguousarray(np.asarray(image), dtype=np.uint8)
        if arr.ndim == 2:
            arr = arr[..., None]
        if arr.ndim != 3 or arr.shape[2] != self.channels:
            raise ValueError(f"expected {self.channels} channel(s), got array of shape {arr.shape}")
        check(self._lib.cm_blur_set_image(self._h, arr.ctypes.data, arr.shape[0], arr.shape[1]))
        self.shape = arr.shape

    def score_to_blur(self, score: float, min_blur: float = 0.0, max_blur: float = 15):
        return min_blur + (1 - score ** 2) * (max_blur - min_blur)        # src/backend.py:319-320

    def blur_radii(self, radii: Sequence[float], copy: bool = True) -> np.ndarray:
        """uint8 [n, H, W, C]: the current image under GaussianBlur(radius) for each radius.

        ``copy=False`` (n <= max_batch) returns a read-only view of the library's page-locked
        result block instead of a fresh array: valid until the next call on this object, and
        three times cheaper for 1024x1024 images, where the host copy is most of the call."""
        if self.shape is None:
            raise RuntimeError("set_image() first")
        r = np.ascontiguousarray(radii, dtype=np.float32)                 # Pillow parses a C float
        if not copy and 0 < r.size <= self.max_batch:
            p = ctypes.c_void_p()
            check(self._lib.cm_blur_batch_pinned(self._h, r.ctypes.data, r.size, ctypes.byref(p)))
            n_bytes = r.size * int(np.prod(self.shape))
            view = np.ctypeslib.as_array(ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8)), shape=(n_bytes,))
            view = view.reshape((r.size,) + tuple(self.shape))
            view.flags.writeable = False
            return view
        out = np.empty((r.size,) + tuple(self.shape), dtype=np.uint8)
        for lo in range(0, r.size, self.max_batch):
            part = r[lo:lo + self.max_batch]
            check(self._lib.cm_blur_batch_host(self._h, part.ctypes.data, part.size,
                                               out[lo:].ctypes.data))
        return out

    def mask_images(self, scores: Sequence[float], copy: bool = True) -> np.ndarray:
        """Batched ``mask_image``: one blurred copy of the current image per session score."""
        return self.blur_radii([self.score_to_blur(float(s)) for s in scores], copy=copy)

    def blur_radii_jpeg(self, radii: Sequence[float], quality: int = 75) -> list:
        """The current image under GaussianBlur(radius), saved as JPEG, for each radius: a list of
        ``bytes`` equal to what ``image.filter(GaussianBlur(r)).save(buf, 'JPEG', quality=quality)``
        writes (main.py:100-103 uses Pillow's default quality, 75).  RGB images only."""
        if self.shape is None:
            raise RuntimeError("set_image() first")
        r = np.ascontiguousarray(radii, dtype=np.float32)
        files = []
        for lo in range(0, r.size, self.max_batch):
            part = r[lo:lo + self.max_batch]
            p, q = ctypes.c_void_p(), ctypes.c_void_p()
            check(self._lib.cm_blur_batch_jpeg_pinned(self._h, part.ctypes.data, part.size, int(quality),
                                                      ctypes.byref(p), ctypes.byref(q)))
            off = np.ctypeslib.as_array(ctypes.cast(q, ctypes.POINTER(ctypes.c_int64)), shape=(part.size + 1,))
            block = ctypes.cast(p, ctypes.POINTER(ctypes.c_uint8))
            files += [ctypes.string_at(ctypes.addressof(block.contents) + int(off[i]), int(off[i + 1] - off[i]))
                      for i in range(part.size)]
        return files

    def mask_images_jpeg(self, scores: Sequence[float], quality: int = 75) -> list:
        """Batched ``mask_image`` + ``image.save(img_io, 'JPEG')`` (src/server.py:129-133, main.py:100-103):
        the JPEG file of the blurred image for each session score."""
        return self.blur_radii_jpeg([self.score_to_blur(float(s)) for s in scores], quality=quality)

    def mask_image(self, image, score: float):
        """Same call as src/backend.py:322-324 (uploads ``image``; prefer set_image + mask_images)."""
        from PIL import Image
        self.set_image(image)
        out = self.mask_images([score])[0]
        return Image.fromarray(out[..., 0] if self.channels == 1 else out, getattr(image, "mode", None))
