"""Developer probe: a few blur+JPEG calls on the 1024x1024 test image (for ncu launch lists / captures)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from cassmantle_b200.blur import ImageBlur

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:1024, 0:1024]
photo = np.stack([(xx + yy) % 256, (xx * 3) % 256, (yy * 7 + xx // 3) % 256], -1).astype(np.uint8)
photo = np.clip(photo.astype(int) + rng.integers(-9, 9, photo.shape), 0, 255).astype(np.uint8)
blur = ImageBlur(1024, 1024, 3, max_batch=nb)
blur.set_image(photo)
scores = rng.uniform(0.01, 1.0, nb)
for _ in range(reps):
    t0 = time.perf_counter(); files = blur.mask_images_jpeg(scores); dt = time.perf_counter() - t0
print(f"batch {nb}: {dt*1e3:.3f} ms, {sum(map(len, files))/nb/1e3:.0f} KB per file")
blur.close()
