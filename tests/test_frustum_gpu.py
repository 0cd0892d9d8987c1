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


RULES = [  # (name, rule switches) = the five call sites of the reference
    ("SearchByProjection(Frame,KF) :1491-1531", dict(invz_rule=0, uv_rule=0, bounds_rule=0, reject_negative_z=0, dist_rule=0, check_view_angle=0)),
    ("SearchByProjection(KF,Scw) :311-360", dict(invz_rule=1, uv_rule=1, bounds_rule=1, reject_negative_z=1, dist_rule=0, check_view_angle=1)),
    ("Fuse(KF,points) :846-893", dict(invz_rule=1, uv_rule=1, bounds_rule=1, reject_negative_z=1, dist_rule=0, check_view_angle=1)),
    ("Fuse(KF,Scw) :1004-1048", dict(invz_rule=0, uv_rule=1, bounds_rule=1, reject_negative_z=1, dist_rule=0, check_view_angle=1)),
    ("SearchBySim3 :1150-1226", dict(invz_rule=0, uv_rule=1, bounds_rule=1, reject_negative_z=1, dist_rule=1, check_view_angle=0, two_stage=1)),
]


@pytest.mark.parametrize("seed", [0, 1, 2])
@pytest.mark.parametrize("name,rules", RULES)
def test_project_points_gpu_equals_oracle(seed, name, rules):
    """orb_project_points (the projection pre-loops of the mapping / loop matchers) against the oracle, whose rule sets for the
    first two sites are pinned to the reference's replayed lines in tests/test_reference_matcher_cpu.py; the other three are pinned
    through the whole-method tests of tests/test_m9_gpu.py.  Every output bit for bit."""
    from ar_orbslam2_b200.tracking import ProjectMapPoints
    wp, nr, maxd, mind, consider = points(seed, 6000)
    T1, T2 = pose(seed), pose(seed + 10)
    R, t = T1[:3, :3], T1[:3, 3]
    proj = dict(R1=R, t1=t, R2=(np.float32(0.9) * T2[:3, :3]).astype(np.float32), t2=T2[:3, 3], Ow=cam_from_pose(T1)[12:15], fx=FX, fy=FY, cx=CX,
                cy=CY, bf=MBF, min_x=0, max_x=752, min_y=0, max_y=480, **rules)
    mx, mn = (np.float32(1.2) * maxd).astype(np.float32), (np.float32(0.8) * mind).astype(np.float32)
    got = ProjectMapPoints(proj, wp, nr, mx, mn, consider)
    want = orbref.project_points(proj, wp, nr, mx, mn, consider)
    same(got, want)
    assert 30 < int(got[0].sum()) < len(wp) - 30, name
    if not rules.get("check_view_angle"):
        same(ProjectMapPoints(proj, wp, None, mx, mn, None), orbref.project_points(proj, wp, None, mx, mn, None))
