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
# blur + JPEG encode on the device (cm_blur_batch_jpeg_pinned): what GET /fetch/contents needs per session
import io
yy, xx = np.mgrid[0:1024, 0:1024]
photo = np.stack([(xx + yy) % 256, (xx * 3) % 256, (yy * 7 + xx // 3) % 256], -1).astype(np.uint8)
photo = np.clip(photo.astype(int) + rng.integers(-9, 9, photo.shape), 0, 255).astype(np.uint8)
for nb in (1, 8, 64):
    blur = ImageBlur(1024, 1024, 3, max_batch=nb)
    blur.set_image(photo)
    scores = rng.uniform(0.01, 1.0, nb)
    blur.mask_images_jpeg(scores)
    ts = []
    for _ in range(10):
        t0 = time.perf_counter(); files = blur.mask_images_jpeg(scores); ts.append(time.perf_counter() - t0)
    t = np.median(ts)
    kb = sum(len(f) for f in files) / nb / 1e3
    # bytes that must move per image: 3 MB of pixels read by the DCT kernel, 3 MB of coefficients written and read
    # twice, the file written: ~12.6 MB of HBM traffic + the blur's 8 passes x 6 MB
    print(f"GPU blur+JPEG batch {nb}: {t*1e3:.2f} ms per call = {t/nb*1e6:.0f} us per image, {kb:.0f} KB per file "
          f"(vs 3146 KB of raw pixels)", flush=True)
    blur.close()
import os
if os.environ.get('NO_CPU'): sys.exit(0)
pil_photo = Image.fromarray(photo, "RGB")
t0 = time.perf_counter()
for s in scores[:8]:
    b_io = io.BytesIO(); pil_photo.filter(ImageFilter.GaussianBlur((1 - s * s) * 15)).save(b_io, 'JPEG')
print(f"Pillow blur+JPEG: {(time.perf_counter() - t0) / 8 * 1e3:.2f} ms per image (1 core)")
t0 = time.perf_counter()
for s in scores[:8]:
    pil.filter(ImageFilter.GaussianBlur((1 - s * s) * 15))
print(f"Pillow: {(time.perf_counter() - t0) / 8 * 1e3:.2f} ms per image (1 core)")
