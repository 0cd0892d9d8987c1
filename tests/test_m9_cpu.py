"""CPU tests (no GPU) for SURVEY 8(a) row M9 -- SearchForTriangulation, Fuse x2, SearchBySim3: the oracle's restatement of
their candidate loops against the reference's OWN ORBmatcher.cc (oracle/_ref/liborbmatcher_ref.so, compiled unmodified
against stub Frame / KeyFrame / MapPoint classes).  Fuse and SearchBySim3 are compared through their complete effect
(map bookkeeping included) with the flat restatement of tests/m9_cases.py around the oracle's window search."""
import numpy as np
import pytest

import m9_cases as M
from ar_orbslam2_b200.synth import shifted_frame, synth_frame

W, H = M.W, M.H


@pytest.fixture(scope="module")
def ref(oracle):
    if not oracle.ref_matcher_available():
        pytest.skip("oracle/_ref not built and /root/reference absent")
    return oracle


@pytest.fixture(scope="module")
def scene(oracle):
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    return dict(ka=ka, da=da, kb=kb, db=db, scale=ex.tables()["scale"])


def kf_pair(scene, backend, which, u_right=None):
    from ar_orbslam2_b200.tracking import Frame
    return Frame(scene["k" + which], scene["d" + which], scene["scale"], (0.0, float(W), 0.0, float(H)), u_right=u_right, backend=backend,
                 keyframe_view=True)


@pytest.mark.parametrize("only_stereo,ori", [(False, True), (True, True), (False, False)])
def test_search_for_triangulation(scene, ref, only_stereo, ori):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.triangulation_case(scene)
    out = []
    for B in (ref.OracleBackend(), ref.ReferenceSourceBackend()):
        k1, k2 = kf_pair(scene, B, "a", c["ur1"]), kf_pair(scene, B, "b", c["ur2"])
        out.append(ORBmatcher(0.6, ori, backend=B).SearchForTriangulation(
            k1, k2, c["off1"], c["idx1"], c["off2"], c["idx2"], c["usable1"], c["usable2"], c["F12"], c["ex"], c["ey"], c["sigma2"], only_stereo))
    (no, mo), (nr, mr) = out
    assert nr > (30 if only_stereo else 200), nr
    assert no == nr and np.array_equal(mo, mr)
    assert nr == int((mr >= 0).sum())


def test_triangulation_gates_bite(scene, ref):
    """The scenario must exercise the epipolar and the epipole gates: switching them off changes the result."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.triangulation_case(scene)
    B = ref.OracleBackend()
    k1, k2 = kf_pair(scene, B, "a", c["ur1"]), kf_pair(scene, B, "b", c["ur2"])
    m = ORBmatcher(0.6, False, backend=B)
    args = (k1, k2, c["off1"], c["idx1"], c["off2"], c["idx2"], c["usable1"], c["usable2"])
    n0, m0 = m.SearchForTriangulation(*args, c["F12"], c["ex"], c["ey"], c["sigma2"])
    n_far, m_far = m.SearchForTriangulation(*args, c["F12"], np.float32(-1e6), np.float32(-1e6), c["sigma2"])
    n_wide, m_wide = m.SearchForTriangulation(*args, c["F12"], c["ex"], c["ey"], c["sigma2"] * np.float32(1e6))
    assert (m_far != m0).sum() >= 1 and n_far >= n0
    assert (m_wide != m0).sum() >= 3 and n_wide >= n0


@pytest.mark.parametrize("variant,th", [(0, 3.0), (0, 2.0), (1, 4.0)])
def test_fuse(scene, ref, variant, th):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = dict(M.fuse_case(scene))
    if variant == 1:   # the Sim3 overload dereferences every entry (:1006-1009): no NULL candidates there
        c["present"] = np.where(c["present"] == 0, 2, c["present"]).astype(np.uint8)
    Bo, Br = ref.OracleBackend(), ref.ReferenceSourceBackend()
    ko, kr = kf_pair(scene, Bo, "b", c["ur_kf"]), kf_pair(scene, Br, "b", c["ur_kf"])
    I, z, S = np.eye(3, dtype=np.float32), np.zeros(3, np.float32), np.eye(4, dtype=np.float32)
    R = ref.refm_fuse(kr, variant, I, z, S, c["world"], c["normal"], c["present"], c["maxd"], c["mind"], c["desc"], c["nobs"], c["in_kf"],
                      c["kf_mp_obs"], 1.0, 1.0, 0.0, 0.0, c["mbf"], th)
    O = M.fuse_flat(ORBmatcher(0.6, True, backend=Bo), ko, scene, c, variant, th)
    assert R["nfused"] > 300, R["nfused"]
    assert (R["pt_repl"] >= 0).sum() > 20 and (R["kf_state"] != np.where(c["kf_mp_obs"] >= 0, 1000000 + np.arange(kr.N), -1)).sum() > 100
    for key in ("nfused", "kf_state", "pt_obs", "pt_repl", "kfmp_repl"):
        assert np.array_equal(O[key], R[key]), key


def test_fuse_reprojection_gate_bites(scene, ref):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.fuse_case(scene)
    B = ref.OracleBackend()
    k = kf_pair(scene, B, "b", c["ur_kf"])
    m = ORBmatcher(0.6, True, backend=B)
    lvl = c["level"]
    radius = (np.float32(3.0) * scene["scale"][lvl]).astype(np.float32)
    inv = (np.float32(1.0) / (scene["scale"] * scene["scale"])).astype(np.float32)
    b0, _ = m.SearchWindowBest(k, c["u"], c["v"], radius, lvl - 1, lvl, c["desc"])
    b1, _ = m.SearchWindowBest(k, c["u"], c["v"], radius, lvl - 1, lvl, c["desc"], gate_ur=c["ur"], inv_level_sigma2=inv)
    assert (b0 != b1).sum() > 20


@pytest.mark.parametrize("th", [7.5, 3.0])
def test_search_by_sim3(scene, ref, th):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.sim3_case(scene)
    Bo, Br = ref.OracleBackend(), ref.ReferenceSourceBackend()
    o1, o2 = kf_pair(scene, Bo, "a"), kf_pair(scene, Bo, "b")
    r1, r2 = kf_pair(scene, Br, "a"), kf_pair(scene, Br, "b")
    I, z = np.eye(3, dtype=np.float32), np.zeros(3, np.float32)
    nr, mr = ref.refm_search_by_sim3(r1, r2, 1.0, I, c["t12"], I, z, I, z, c["has1"], c["world1"], c["maxd1"], c["mind1"], c["desc1"],
                                     c["has2"], c["world2"], c["maxd2"], c["mind2"], c["desc2"], c["pre12"], 1.0, 1.0, 0.0, 0.0, th)
    no, mo = M.sim3_flat(ORBmatcher(0.6, True, backend=Bo), o1, o2, scene, c, th)
    assert nr > 200, nr
    assert no == nr and np.array_equal(mo, mr)
