"""Frame::ComputeStereoFromRGBD (ref Frame.cc:645-667): the numpy restatement against the reference's own Frame.cc compiled
unmodified (oracle/_ref/liborbmatcher_ref.so).  The 16U -> 32F `convertTo(CV_32F, alpha)` in front of it (Tracking.cc:553-554) is
restated as OpenCV's cvtScale arithmetic `(float)src * (float)alpha`; cv2's Python binding exposes no convertTo to pin it against."""
import numpy as np
import pytest

from ar_orbslam2_b200._abi import KEYPOINT_DTYPE
from oracle import orbref


def rgbd_case(seed, w=640, h=480, n=1500):
    rng = np.random.default_rng(seed)
    keys = np.zeros(n, KEYPOINT_DTYPE)
    keys["x"] = rng.uniform(0, w - 1, n).astype(np.float32)
    keys["y"] = rng.uniform(0, h - 1, n).astype(np.float32)
    keys["x"][:4] = [0, w - 1, 17.999, 18.0]
    keys["y"][:4] = [0, h - 1, 5.5, 6.0]
    un = keys.copy()
    un["x"] += rng.uniform(-2, 2, n).astype(np.float32)
    raw = rng.integers(0, 40000, (h, w)).astype(np.uint16)
    raw[rng.random((h, w)) < 0.2] = 0                      # holes: no depth
    return keys, un, raw


@pytest.mark.parametrize("seed", [0, 1])
def test_restatement_equals_reference_frame_cc(seed):
    if not orbref.ref_matcher_available():
        pytest.skip("oracle/_ref/liborbmatcher_ref.so not built")
    keys, un, raw = rgbd_case(seed)
    depth = (raw.astype(np.float32) * np.float32(1.0 / 5000.0)).astype(np.float32)
    ur, dp = orbref.stereo_from_rgbd(keys, un, depth, 40.0)
    rur, rdp = orbref.refm_stereo_from_rgbd(keys, un, depth, 40.0)
    assert np.array_equal(ur.view(np.uint32), rur.view(np.uint32)) and np.array_equal(dp.view(np.uint32), rdp.view(np.uint32))
    assert (dp > 0).sum() > 1000 and (dp == -1).sum() > 100
