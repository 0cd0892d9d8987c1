"""Frame::UndistortKeyPoints / ComputeImageBounds: the oracle's restatement of cv::undistortPoints against
cv2's golden outputs (tests/golden/cv2_undistort.npz, tools/make_golden_undistort.py), bit for bit."""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from make_golden_undistort import CAMS, points  # noqa: E402
from oracle import orbref  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "cv2_undistort.npz")


def test_oracle_equals_cv2_golden():
    g = np.load(GOLDEN)
    for i, (name, fx, fy, cx, cy, dist, w, h) in enumerate(CAMS):
        got = orbref.undistort_points(points(i, w, h), fx, fy, cx, cy, dist)
        assert np.array_equal(got.view(np.uint32), g[name].view(np.uint32)), name


def test_oracle_equals_cv2_live():
    cv2 = pytest.importorskip("cv2")
    name, fx, fy, cx, cy, dist, w, h = CAMS[0]
    p = points(99, w, h, 5000)
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    want = cv2.undistortPoints(p.reshape(-1, 1, 2), K, np.asarray(dist, np.float32), None, K).reshape(-1, 2)
    assert np.array_equal(orbref.undistort_points(p, fx, fy, cx, cy, dist).view(np.uint32), want.view(np.uint32))
