"""GPU parity of orb_remap_linear through the C ABI: bit-exact against cv2's golden outputs and the oracle."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from ar_orbslam2_b200.color import remap  # noqa: E402
from make_golden_remap import CASES, maps, source  # noqa: E402
from oracle import orbref  # noqa: E402
from test_remap_cpu import GOLDEN  # noqa: E402

pytestmark = pytest.mark.gpu


def test_gpu_equals_cv2_golden():
    g = np.load(GOLDEN)
    for seed, h, w in CASES:
        mx, my = maps(seed, h, w)
        assert np.array_equal(remap(source(seed, h, w), mx, my), g[f"remap_{seed}"]), seed


@pytest.mark.parametrize("seed,h,w", [(7, 376, 1241), (8, 480, 752), (9, 1, 1), (10, 5, 3), (11, 33, 1023)])
def test_gpu_equals_oracle(seed, h, w):
    src = source(seed, h, w)
    mx, my = maps(seed, h, w)
    assert np.array_equal(remap(src, mx, my), orbref.remap_linear(src, mx, my))


def test_batch_shares_the_maps_and_sizes_may_differ():
    srcs = np.stack([source(20 + b, 100, 140) for b in range(3)])
    mx, my = maps(20, 60, 92)
    out = remap(srcs, mx, my)
    for b in range(3):
        assert np.array_equal(out[b], orbref.remap_linear(srcs[b], mx, my))
