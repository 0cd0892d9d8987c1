"""GPU parity for SURVEY 8(a) row M9 (SearchForTriangulation, Fuse x2, SearchBySim3) through the C ABI:
orb_search_for_triangulation and orb_search_window_best against the oracle on the scenarios of tests/m9_cases.py
(the same ones tests/test_m9_cpu.py pins to the reference's compiled ORBmatcher.cc), and the C++ shim bodies
(ar_orbslam2_b200/shim/ORBmatcher_gpu.cc, compiled against the reference's headers) against the reference's own methods,
map bookkeeping included."""
import os

import numpy as np
import pytest

import m9_cases as M
from ar_orbslam2_b200.synth import shifted_frame, synth_frame

pytestmark = pytest.mark.gpu
W, H = M.W, M.H


@pytest.fixture(scope="module")
def scene(oracle):
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    return dict(ka=ka, da=da, kb=kb, db=db, scale=ex.tables()["scale"])


def kf(scene, backend, which, u_right=None, bounds=None):
    from ar_orbslam2_b200.tracking import Frame
    return Frame(scene["k" + which], scene["d" + which], scene["scale"], bounds or (0.0, float(W), 0.0, float(H)), u_right=u_right,
                 backend=backend, keyframe_view=True)


@pytest.mark.parametrize("only_stereo,ori", [(False, True), (True, True), (False, False)])
def test_search_for_triangulation(scene, oracle, gpu_lib, only_stereo, ori):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.triangulation_case(scene)
    out = []
    for B in (oracle.OracleBackend(), None):
        k1, k2 = kf(scene, B, "a", c["ur1"]), kf(scene, B, "b", c["ur2"])
        out.append(ORBmatcher(0.6, ori, backend=B).SearchForTriangulation(
            k1, k2, c["off1"], c["idx1"], c["off2"], c["idx2"], c["usable1"], c["usable2"], c["F12"], c["ex"], c["ey"], c["sigma2"], only_stereo))
    (no, mo), (ng, mg) = out
    assert no > (30 if only_stereo else 200)
    assert ng == no and np.array_equal(mg, mo)


def test_window_best_random_queries(scene, oracle, gpu_lib):
    """orb_search_window_best on arbitrary queries (windows partly outside the image, empty windows, every level range, the
    reprojection gate on and off, fractional key-frame bounds): index and distance of every query equal the oracle's."""
    from ar_orbslam2_b200.synth import SplitMix64
    from ar_orbslam2_b200.tracking import ORBmatcher
    rng = SplitMix64(77)
    n = 3000
    k = scene["kb"]
    pick = np.array([rng.next() % len(k) for _ in range(n)])
    x = (k["x"][pick] + np.array([(rng.next() % 41 - 20) / 4.0 for _ in range(n)])).astype(np.float32)
    y = (k["y"][pick] + np.array([(rng.next() % 41 - 20) / 4.0 for _ in range(n)])).astype(np.float32)
    x[::50] = -30.0; y[1::50] = H + 25.0
    radius = np.array([1.0 + (rng.next() % 200) / 8.0 for _ in range(n)], np.float32)
    lo = np.array([rng.next() % 9 - 1 for _ in range(n)], np.int32)
    hi = (lo + np.array([rng.next() % 3 for _ in range(n)])).astype(np.int32)
    valid = np.array([rng.next() % 11 != 0 for _ in range(n)], np.uint8)
    desc = scene["da"][np.array([rng.next() % len(scene["da"]) for _ in range(n)])]
    desc[::3] = scene["db"][pick[::3]]
    ur_kf = np.where(np.arange(len(k)) % 3 == 0, k["x"] - np.float32(5.0), -1.0).astype(np.float32)
    inv = (np.float32(1.0) / (scene["scale"] * scene["scale"])).astype(np.float32)
    gur = (x - np.float32(5.0) + np.array([(rng.next() % 9 - 4) / 2.0 for _ in range(n)])).astype(np.float32)
    for bounds in ((0.0, float(W), 0.0, float(H)), (-3.7, W + 2.4, -1.2, H + 5.9)):
        for gate in ({}, dict(gate_ur=gur, inv_level_sigma2=inv), dict(gate_ur=gur, inv_level_sigma2=inv, chi2_mono=0.8, chi2_stereo=1.1)):
            res = []
            for B in (oracle.OracleBackend(), None):
                f = kf(scene, B, "b", ur_kf, bounds)
                res.append(ORBmatcher(0.6, True, backend=B).SearchWindowBest(f, x, y, radius, lo, hi, desc, valid=valid, **gate))
            (bo, do), (bg, dg) = res
            assert (bo >= 0).sum() > 100
            assert np.array_equal(bg, bo) and np.array_equal(dg, do), (bounds, list(gate))


@pytest.mark.parametrize("variant,th", [(0, 3.0), (0, 2.0), (1, 4.0)])
def test_fuse_flat(scene, oracle, gpu_lib, variant, th):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = dict(M.fuse_case(scene))
    if variant == 1:
        c["present"] = np.where(c["present"] == 0, 2, c["present"]).astype(np.uint8)
    Bo = oracle.OracleBackend()
    O = M.fuse_flat(ORBmatcher(0.6, True, backend=Bo), kf(scene, Bo, "b", c["ur_kf"]), scene, c, variant, th)
    G = M.fuse_flat(ORBmatcher(0.6, True), kf(scene, None, "b", c["ur_kf"]), scene, c, variant, th)
    assert O["nfused"] > 300
    for key in ("nfused", "kf_state", "pt_obs", "pt_repl", "kfmp_repl"):
        assert np.array_equal(G[key], O[key]), key


@pytest.mark.parametrize("th", [7.5, 3.0])
def test_search_by_sim3_flat(scene, oracle, gpu_lib, th):
    from ar_orbslam2_b200.tracking import ORBmatcher
    c = M.sim3_case(scene)
    Bo = oracle.OracleBackend()
    no, mo = M.sim3_flat(ORBmatcher(0.6, True, backend=Bo), kf(scene, Bo, "a"), kf(scene, Bo, "b"), scene, c, th)
    ng, mg = M.sim3_flat(ORBmatcher(0.6, True), kf(scene, None, "a"), kf(scene, None, "b"), scene, c, th)
    assert no > 200 and ng == no and np.array_equal(mg, mo)


def test_cpp_m9_shim_equals_reference(scene, oracle, gpu_lib):
    """The shim's SearchForTriangulation / Fuse / Fuse(Scw) / SearchBySim3 bodies, compiled against the reference's headers
    (oracle/_ref/liborbmatcher_shim.so), against the reference's compiled methods (liborbmatcher_ref.so) behind the same driver:
    identical match lists, identical key-frame and map-point state after the fusion bookkeeping."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    rb, sb = oracle.ReferenceSourceBackend(), oracle.ShimBackend()
    I, z, S = np.eye(3, dtype=np.float32), np.zeros(3, np.float32), np.eye(4, dtype=np.float32)
    # SearchForTriangulation
    c = M.triangulation_case(scene)
    for only_stereo, ori in ((False, True), (True, False)):
        out = []
        for B in (rb, sb):
            out.append(ORBmatcher(0.6, ori, backend=B).SearchForTriangulation(
                kf(scene, B, "a", c["ur1"]), kf(scene, B, "b", c["ur2"]), c["off1"], c["idx1"], c["off2"], c["idx2"], c["usable1"], c["usable2"],
                c["F12"], c["ex"], c["ey"], c["sigma2"], only_stereo))
        (nr, mr), (ns, ms) = out
        assert nr > 30 and ns == nr and np.array_equal(ms, mr)
    # Fuse, both overloads
    for variant, th in ((0, 3.0), (1, 4.0)):
        c = dict(M.fuse_case(scene))
        if variant == 1:
            c["present"] = np.where(c["present"] == 0, 2, c["present"]).astype(np.uint8)
        res = [oracle.refm_fuse(kf(scene, B, "b", c["ur_kf"]), variant, I, z, S, c["world"], c["normal"], c["present"], c["maxd"], c["mind"], c["desc"],
                                c["nobs"], c["in_kf"], c["kf_mp_obs"], 1.0, 1.0, 0.0, 0.0, c["mbf"], th) for B in (rb, sb)]
        assert res[0]["nfused"] > 300
        for key in ("nfused", "kf_state", "pt_obs", "pt_repl", "kfmp_repl"):
            assert np.array_equal(res[1][key], res[0][key]), (variant, key)
    # SearchBySim3
    c = M.sim3_case(scene)
    res = [oracle.refm_search_by_sim3(kf(scene, B, "a"), kf(scene, B, "b"), 1.0, I, c["t12"], I, z, I, z, c["has1"], c["world1"], c["maxd1"],
                                      c["mind1"], c["desc1"], c["has2"], c["world2"], c["maxd2"], c["mind2"], c["desc2"], c["pre12"], 1.0, 1.0,
                                      0.0, 0.0, 7.5) for B in (rb, sb)]
    assert res[0][0] > 200 and res[1][0] == res[0][0] and np.array_equal(res[1][1], res[0][1])
