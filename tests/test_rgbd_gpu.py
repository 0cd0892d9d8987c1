"""GPU parity of orb_depth_from_rgbd (Frame::ComputeStereoFromRGBD, ref Frame.cc:645-667, + the depth convertTo of
Tracking.cc:553-554) against the restatement and the reference's compiled Frame.cc."""
import numpy as np
import pytest

from ar_orbslam2_b200.stereo import ComputeStereoFromRGBD
from oracle import orbref
from test_rgbd_cpu import rgbd_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_gpu_equals_oracle_and_reference(seed):
    keys, un, raw = rgbd_case(seed)
    f = float(np.float32(1.0 / 5000.0))
    # CV_16U map with the factor, CV_32F map with factor 1, CV_32F map with another factor
    for depth, factor in ((raw, f), ((raw.astype(np.float32) * np.float32(f)).astype(np.float32), 1.0),
                          (raw.astype(np.float32), f)):
        ur, dp = ComputeStereoFromRGBD(keys, un, depth, 40.0, factor)
        our, odp = orbref.stereo_from_rgbd(keys, un, depth, 40.0, factor)
        assert np.array_equal(ur.view(np.uint32), our.view(np.uint32)) and np.array_equal(dp.view(np.uint32), odp.view(np.uint32))
    if orbref.ref_matcher_available():
        d32 = (raw.astype(np.float32) * np.float32(f)).astype(np.float32)
        ur, dp = ComputeStereoFromRGBD(keys, un, d32, 40.0, 1.0)
        rur, rdp = orbref.refm_stereo_from_rgbd(keys, un, d32, 40.0)
        assert np.array_equal(ur.view(np.uint32), rur.view(np.uint32)) and np.array_equal(dp.view(np.uint32), rdp.view(np.uint32))


def test_pitched_and_empty():
    keys, un, raw = rgbd_case(5, 320, 240, 200)
    big = np.zeros((240, 400), np.uint16)
    big[:, :320] = raw
    view = big[:, :320]
    ur, dp = ComputeStereoFromRGBD(keys, un, view, 35.0, 0.001)
    our, odp = orbref.stereo_from_rgbd(keys, un, raw, 35.0, 0.001)
    assert np.array_equal(ur, our) and np.array_equal(dp, odp)
    ur, dp = ComputeStereoFromRGBD(keys[:0], un[:0], raw, 35.0, 0.001)
    assert len(ur) == 0 and len(dp) == 0
