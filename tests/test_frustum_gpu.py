"""GPU parity of orb_is_in_frustum through the C ABI: against the oracle and, where it travelled, the reference's compiled Frame.cc."""
import os

import numpy as np
import pytest

from ar_orbslam2_b200.tracking import isInFrustum
from frustum_cases import CX, CY, FX, FY, LOGS, MBF, cam_dict, dummy_frame, points, pose
from oracle import orbref
from test_frustum_cpu import same

pytestmark = pytest.mark.gpu


def cam_from_pose(T):
    R, t = T[:3, :3], T[:3, 3]
    Ow = (-(R.T.astype(np.float64) @ t.astype(np.float64))).astype(np.float32)
    return np.concatenate([R.reshape(9), t, Ow]).astype(np.float32)


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_gpu_equals_oracle_and_reference(seed):
    wp, nr, maxd, mind, consider = points(seed)
    if os.path.exists(orbref.REF_MATCHER_PATH):
        f = dummy_frame(orbref.ReferenceSourceBackend())
        cam15, want = orbref.refm_is_in_frustum(f, pose(seed), FX, FY, CX, CY, MBF, LOGS, wp, nr, maxd, mind, 0.5, consider)
    else:
        cam15, want = cam_from_pose(pose(seed)), None
    cam = cam_dict(cam15)
    got = isInFrustum(cam, wp, nr, maxd, mind, 0.5, consider)
    same(got, orbref.is_in_frustum(cam, wp, nr, maxd, mind, 0.5, consider))
    if want is not None:
        same(got, want)


def test_logf_on_the_device_matches_the_host_libm():
    """PredictScale's level over a dense sweep of distance ratios (every level boundary is crossed many times)."""
    n = 200000
    rng = np.random.default_rng(7)
    wp = np.zeros((n, 3), np.float32); wp[:, 2] = rng.uniform(0.5, 40, n).astype(np.float32)
    nr = np.zeros((n, 3), np.float32); nr[:, 2] = 1
    maxd = (wp[:, 2] * rng.uniform(1.0, 4.0, n)).astype(np.float32)
    mind = np.zeros(n, np.float32)
    cam = cam_dict(cam_from_pose(np.eye(4, dtype=np.float32)))
    got = isInFrustum(cam, wp, nr, maxd, mind, 0.5)
    want = orbref.is_in_frustum(cam, wp, nr, maxd, mind, 0.5)
    same(got, want)
    assert got[0].all() and got[4].max() >= 7


def test_feeds_search_by_projection():
    got = isInFrustum(cam_dict(cam_from_pose(np.eye(4, dtype=np.float32))), np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32),
                      np.zeros(0, np.float32), np.zeros(0, np.float32))
    assert all(len(a) == 0 for a in got)
