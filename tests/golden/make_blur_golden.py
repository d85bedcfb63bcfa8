"""Regenerates tests/golden/blur_golden.npz with the REAL Pillow call the reference makes
(src/backend.py:322-324): image.filter(ImageFilter.GaussianBlur(radius)).  Run from the repo root."""
import os

import numpy as np
import PIL
from PIL import Image, ImageFilter

rng = np.random.default_rng(7)
yy, xx = np.mgrid[0:48, 0:64]
smooth = np.stack([127 + 120 * np.sin(xx / 9.0) * np.cos(yy / 7.0), (xx * 4) % 256, (yy * 5 + xx) % 256], -1)
image = np.clip(smooth + rng.integers(-20, 20, smooth.shape), 0, 255).astype(np.uint8)
scores = np.array([1.0, 0.99999994, 0.9, 0.4567891, 0.1, 0.01, 0.0], dtype=np.float64)
radii = np.array([0.0 + (1 - s ** 2) * 15 for s in scores] + [0.25, 2.0, 40.0, 100.0], dtype=np.float64)
pil = Image.fromarray(image, "RGB")
outs = np.stack([np.asarray(pil.filter(ImageFilter.GaussianBlur(float(r)))) for r in radii])
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "blur_golden.npz")
np.savez_compressed(path, image=image, radii=radii, outputs=outs, pillow=np.array(PIL.__version__))
print("wrote", path, outs.shape, "Pillow", PIL.__version__)
