"""GPU parity of orb_undistort_keypoints / orb_undistort_points through the C ABI: bit-exact against cv2's golden outputs."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from ar_orbslam2_b200._abi import KEYPOINT_DTYPE  # noqa: E402
from ar_orbslam2_b200.tracking import ComputeImageBounds, UndistortKeyPoints  # noqa: E402
from make_golden_undistort import CAMS, points  # noqa: E402
from oracle import orbref  # noqa: E402
from test_undistort_cpu import GOLDEN  # noqa: E402

pytestmark = pytest.mark.gpu


def keys_from(p):
    k = np.zeros(len(p), KEYPOINT_DTYPE)
    k["x"], k["y"] = p[:, 0], p[:, 1]
    k["size"], k["angle"], k["response"], k["octave"], k["class_id"] = 31, 12.5, 40, 2, -1
    return k


def test_gpu_equals_cv2_golden():
    g = np.load(GOLDEN)
    for i, (name, fx, fy, cx, cy, dist, w, h) in enumerate(CAMS):
        p = points(i, w, h)
        k = keys_from(p)
        un = UndistortKeyPoints(k, fx, fy, cx, cy, dist)
        got = np.stack([un["x"], un["y"]], 1)
        assert np.array_equal(got.view(np.uint32), g[name].view(np.uint32)), name
        for f in ("size", "angle", "response", "octave", "class_id"):
            assert np.array_equal(un[f], k[f])
        mnx, mxx, mny, mxy = ComputeImageBounds(w, h, fx, fy, cx, cy, dist)
        c = g[name][:4]
        assert (mnx, mxx, mny, mxy) == (min(c[0, 0], c[2, 0]), max(c[1, 0], c[3, 0]), min(c[0, 1], c[1, 1]), max(c[2, 1], c[3, 1]))


def test_zero_distortion_copies_and_large_n():
    p = points(5, 1241, 376, 200000)
    k = keys_from(p)
    assert UndistortKeyPoints(k, 718.856, 718.856, 607.1928, 185.2157, [0, 0, 0, 0]).tobytes() == k.tobytes()
    name, fx, fy, cx, cy, dist, w, h = CAMS[2]
    un = UndistortKeyPoints(k, fx, fy, cx, cy, dist)
    want = orbref.undistort_points(p, fx, fy, cx, cy, dist)
    assert np.array_equal(np.stack([un["x"], un["y"]], 1).view(np.uint32), want.view(np.uint32))
    assert ComputeImageBounds(640, 480, 500, 500, 320, 240, [0, 0, 0, 0]) == (0, 640, 0, 480)
    assert len(UndistortKeyPoints(k[:0], fx, fy, cx, cy, dist)) == 0
