"""cvtColor X2GRAY: the oracle against cv2 golden vectors (tests/golden/cv2_cvtcolor.npz,
tools/make_golden_cvtcolor.py) and, when cv2 is importable, against cv2 live."""
import os

import numpy as np
import pytest

from oracle import orbref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "cv2_cvtcolor.npz")


def image(seed, h, w, ch):
    return np.random.default_rng(seed).integers(0, 256, (h, w, ch), dtype=np.uint8)


def golden_cases():
    g = np.load(GOLDEN)
    for key in g.files:
        if key in ("cv2_version", "lattice_rgb"):
            continue
        name, seed, size = key.split("_")
        h, w = (int(x) for x in size.split("x"))
        yield key, image(int(seed), h, w, 4 if name.endswith("a") else 3), name.startswith("rgb"), g[key]
    v = np.arange(0, 256, 4, dtype=np.uint8)
    lat = np.ascontiguousarray(np.stack(np.meshgrid(v, v, v, indexing="ij"), axis=-1).reshape(512, 512, 3))
    yield "lattice_rgb", lat, True, g["lattice_rgb"]


def test_oracle_equals_cv2_golden():
    n = 0
    for key, img, rgb, want in golden_cases():
        assert np.array_equal(orbref.cvt_color_gray(img, rgb), want), key
        n += 1
    assert n == 17


def test_oracle_equals_cv2_live():
    cv2 = pytest.importorskip("cv2")
    img = image(77, 240, 321, 3)
    assert np.array_equal(orbref.cvt_color_gray(img, True), cv2.cvtColor(img, cv2.COLOR_RGB2GRAY))
    assert np.array_equal(orbref.cvt_color_gray(img, False), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))
    img4 = image(78, 100, 99, 4)
    assert np.array_equal(orbref.cvt_color_gray(img4, False), cv2.cvtColor(img4, cv2.COLOR_BGRA2GRAY))


def test_opencv24_variant_differs_only_by_rounding():
    img = image(5, 64, 64, 3)
    a, b = orbref.cvt_color_gray(img, True, 0).astype(int), orbref.cvt_color_gray(img, True, 1).astype(int)
    assert np.abs(a - b).max() <= 1
