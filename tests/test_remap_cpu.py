"""cv::remap INTER_LINEAR (stereo rectification): the oracle against cv2's golden outputs
(tests/golden/cv2_remap.npz, tools/make_golden_remap.py) and, when cv2 is importable, against cv2 live."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from make_golden_remap import CASES, maps, source  # noqa: E402
from oracle import orbref  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "cv2_remap.npz")


def test_oracle_equals_cv2_golden():
    g = np.load(GOLDEN)
    for seed, h, w in CASES:
        mx, my = maps(seed, h, w)
        assert np.array_equal(orbref.remap_linear(source(seed, h, w), mx, my), g[f"remap_{seed}"]), seed


def test_oracle_equals_cv2_live():
    cv2 = pytest.importorskip("cv2")
    for seed, h, w in [(7, 376, 1241), (8, 480, 752)]:
        src = source(seed, h, w)
        mx, my = maps(seed, h, w)
        assert np.array_equal(orbref.remap_linear(src, mx, my), cv2.remap(src, mx, my, cv2.INTER_LINEAR))
    # destination size differs from the source size
    src = source(9, 100, 140)
    mx, my = maps(9, 60, 90)
    assert np.array_equal(orbref.remap_linear(src, mx, my), cv2.remap(src, mx, my, cv2.INTER_LINEAR))
