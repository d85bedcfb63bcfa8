"""Measures cm_blur_batch_host (1024x1024 RGB, the reference's SDXL image) against Pillow on the host."""
import sys, time
import numpy as np
from PIL import Image, ImageFilter
sys.path.insert(0, ".")
from cassmantle_b200.blur import ImageBlur

rng = np.random.default_rng(0)
img = rng.integers(0, 256, (1024, 1024, 3), dtype=np.uint8)
pil = Image.fromarray(img, "RGB")
for nb in ((8,) if __import__('os').environ.get('NO_CPU') else (1, 8, 64)):
    blur = ImageBlur(1024, 1024, 3, max_batch=nb)
    blur.set_image(img)
    scores = rng.uniform(0.01, 1.0, nb)
    blur.mask_images(scores)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); out = blur.mask_images(scores); ts.append(time.perf_counter() - t0)
    t = np.median(ts)
    print(f"GPU batch {nb}: {t*1e3:.2f} ms per call = {t/nb*1e6:.0f} us per image incl. D2H of {out.nbytes/1e6:.0f} MB", flush=True)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); out = blur.mask_images(scores, copy=False); ts.append(time.perf_counter() - t0)
    t = np.median(ts)
    print(f"GPU batch {nb}, result left in page-locked memory: {t*1e3:.2f} ms per call = {t/nb*1e6:.0f} us per image", flush=True)
    blur.close()
import os
if os.environ.get('NO_CPU'): sys.exit(0)
t0 = time.perf_counter()
for s in scores[:8]:
    pil.filter(ImageFilter.GaussianBlur((1 - s * s) * 15))
print(f"Pillow: {(time.perf_counter() - t0) / 8 * 1e3:.2f} ms per image (1 core)")
