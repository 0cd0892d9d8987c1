#!/usr/bin/env python3
"""Golden vectors for the colour->grey conversion from cv2 (the only OpenCV on this box): tests/golden/cv2_cvtcolor.npz.
Inputs are regenerated from the seed by the tests; only cv2's outputs are stored."""
import os

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODES = {"rgb": cv2.COLOR_RGB2GRAY, "bgr": cv2.COLOR_BGR2GRAY, "rgba": cv2.COLOR_RGBA2GRAY, "bgra": cv2.COLOR_BGRA2GRAY}


def image(seed, h, w, ch):
    return np.random.default_rng(seed).integers(0, 256, (h, w, ch), dtype=np.uint8)


out = {"cv2_version": np.array(cv2.__version__)}
for seed, (h, w) in enumerate([(48, 64), (37, 131), (5, 17), (64, 250)]):
    for name, code in CODES.items():
        img = image(seed, h, w, 4 if name.endswith("a") else 3)
        out[f"{name}_{seed}_{h}x{w}"] = cv2.cvtColor(img, code)
# every (R,G,B) on a lattice that hits all rounding boundaries cheaply: 64^3 colours
v = np.arange(0, 256, 4, dtype=np.uint8)
lat = np.stack(np.meshgrid(v, v, v, indexing="ij"), axis=-1).reshape(512, 512, 3)
out["lattice_rgb"] = cv2.cvtColor(lat, cv2.COLOR_RGB2GRAY)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cv2_cvtcolor.npz"), **out)
print("wrote tests/golden/cv2_cvtcolor.npz with", len(out), "arrays")
