"""GPU parity for the tracking matchers (ref: ORB_SLAM2/src/ORBmatcher.cc) on the C3 scenario of SURVEY 8(d):
752x480 frames, frame B = frame A translated by (+7,+4) px.  Match lists must be identical to the oracle."""
import numpy as np
import pytest

from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_frame

pytestmark = pytest.mark.gpu
W, H = 752, 480


@pytest.fixture(scope="module")
def scene(oracle):
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    return dict(ka=ka, da=da, kb=kb, db=db, scale=ex.tables()["scale"])


def both(scene, oracle, which):
    from ar_orbslam2_b200.tracking import Frame
    k, d = scene["k" + which], scene["d" + which]
    bounds = (0.0, float(W), 0.0, float(H))
    return (Frame(k, d, scene["scale"], bounds), Frame(k, d, scene["scale"], bounds, backend=oracle.OracleBackend()))


def test_features_in_area(scene, oracle, gpu_lib):
    g, r = both(scene, oracle, "a")
    rng = SplitMix64(5)
    for _ in range(60):
        x, y = rng.next() % (W + 40) - 20.0, rng.next() % (H + 40) - 20.0
        rad = 3.0 + rng.next() % 120
        lo = int(rng.next() % 9) - 1
        hi = int(rng.next() % 9) - 1
        a, b = g.GetFeaturesInArea(x, y, rad, lo, hi), r.GetFeaturesInArea(x, y, rad, lo, hi)
        assert np.array_equal(a, b), (x, y, rad, lo, hi)


@pytest.mark.parametrize("ori", [True, False])
def test_search_for_initialization(scene, oracle, gpu_lib, ori):
    from ar_orbslam2_b200.tracking import ORBmatcher
    g1, r1 = both(scene, oracle, "a")
    g2, r2 = both(scene, oracle, "b")
    prev = np.stack([scene["ka"]["x"], scene["ka"]["y"]], 1).astype(np.float32).copy()
    pg, pr = prev.copy(), prev.copy()
    ng, mg = ORBmatcher(0.9, ori).SearchForInitialization(g1, g2, pg, 100)
    nr, mr = ORBmatcher(0.9, ori, backend=oracle.OracleBackend()).SearchForInitialization(r1, r2, pr, 100)
    assert nr > 100, "scenario should produce matches"
    assert ng == nr and np.array_equal(mg, mr) and np.array_equal(pg, pr)
    # second call, as Tracking does on the next frame, starts from the updated vbPrevMatched
    ng, mg = ORBmatcher(0.9, ori).SearchForInitialization(g1, g2, pg, 100)
    nr, mr = ORBmatcher(0.9, ori, backend=oracle.OracleBackend()).SearchForInitialization(r1, r2, pr, 100)
    assert ng == nr and np.array_equal(mg, mr) and np.array_equal(pg, pr)


def _last_points(scene):
    """Every Last keypoint gets a synthetic map point at depth z = 2 + (hash(i) mod 1000)/100 m; Last pose = I,
    Cur pose = translation (0.05, 0.02, 0) m; EuRoC intrinsics."""
    fx = fy = np.float32(435.2); cx, cy = np.float32(367.2), np.float32(252.2)
    k = scene["ka"]
    n = len(k)
    z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
    X = (k["x"] - cx) * z / fx
    Y = (k["y"] - cy) * z / fy
    xc, yc = X + np.float32(0.05), Y + np.float32(0.02)
    invz = (np.float32(1.0) / z).astype(np.float32)
    u = (fx * xc * invz + cx).astype(np.float32)
    v = (fy * yc * invz + cy).astype(np.float32)
    valid = ((u >= 0) & (u <= W) & (v >= 0) & (v <= H)).astype(np.uint8)
    valid[::17] = 0   # some features without a map point
    return valid, u, v, invz, k["octave"].astype(np.int32), k["angle"].astype(np.float32), scene["da"]


@pytest.mark.parametrize("mode", ["mono", "forward", "backward"])
def test_search_by_projection_last(scene, oracle, gpu_lib, mode):
    from ar_orbslam2_b200.tracking import ORBmatcher
    valid, u, v, invz, octv, ang, desc = _last_points(scene)
    # the "current" frame is frame A itself seen from the translated pose: reuse A's keypoints, with stereo info
    ur = np.where(np.arange(len(u)) % 3 == 0, scene["ka"]["x"] - 5.0, -1.0).astype(np.float32)
    from ar_orbslam2_b200.tracking import Frame
    bounds = (0.0, float(W), 0.0, float(H))
    g = Frame(scene["ka"], scene["da"], scene["scale"], bounds, u_right=ur)
    r = Frame(scene["ka"], scene["da"], scene["scale"], bounds, u_right=ur, backend=oracle.OracleBackend())
    claimed = (np.arange(g.N) % 11 == 0).astype(np.uint8)
    obs = (np.arange(len(u)) % 5 != 0).astype(np.uint8)
    kw = dict(th=15.0, forward=mode == "forward", backward=mode == "backward", mbf=40.0, claimed=claimed, obs_positive=obs)
    ng, ag = ORBmatcher(0.9, True).SearchByProjectionLast(g, valid, u, v, invz, octv, ang, desc, **kw)
    nr, ar = ORBmatcher(0.9, True, backend=oracle.OracleBackend()).SearchByProjectionLast(r, valid, u, v, invz, octv, ang, desc, **kw)
    assert nr > 50
    assert ng == nr and np.array_equal(ag, ar)


def test_search_by_projection_points(scene, oracle, gpu_lib):
    from ar_orbslam2_b200.tracking import ORBmatcher
    g, r = both(scene, oracle, "b")
    k = scene["ka"]
    n = len(k)
    rng = SplitMix64(9)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (k["x"] + 7).astype(np.float32), (k["y"] + 4).astype(np.float32)
    pxr = (px - 3).astype(np.float32)
    level = np.clip(k["octave"] + np.array([rng.next() % 3 - 1 for _ in range(n)]), 0, 7).astype(np.int32)
    vc = np.array([0.999 if rng.next() % 2 else 0.9 for _ in range(n)], np.float32)
    claimed = (np.arange(g.N) % 13 == 0).astype(np.uint8)
    obs = (np.arange(n) % 4 != 0).astype(np.uint8)
    for th in (1.0, 3.0):
        ng, ag = ORBmatcher(0.8).SearchByProjectionPoints(g, in_view, px, py, pxr, level, vc, scene["da"], th, claimed, obs)
        nr, ar = ORBmatcher(0.8, backend=oracle.OracleBackend()).SearchByProjectionPoints(r, in_view, px, py, pxr, level, vc, scene["da"], th, claimed, obs)
        assert nr > 100
        assert ng == nr and np.array_equal(ag, ar)


def test_search_windowed_keyframe_style(scene, oracle, gpu_lib):
    """strict level filter + TH_LOW + every assignment blocks: SearchByProjection(KeyFrame,Scw,...) / Fuse loops."""
    from ar_orbslam2_b200.tracking import ORB_SELECT_BEST, ORBmatcher
    g, r = both(scene, oracle, "b")
    k = scene["ka"]
    x, y = (k["x"] + 7).astype(np.float32), (k["y"] + 4).astype(np.float32)
    pred = k["octave"].astype(np.int32)
    rad = (4.0 * scene["scale"][pred]).astype(np.float32)
    kw = dict(select=ORB_SELECT_BEST, th_dist=50, strict_levels=True, check_orientation=False)
    ng, ag = ORBmatcher(0.6).SearchWindowed(g, x, y, rad, pred - 1, pred, scene["da"], **kw)
    nr, ar = ORBmatcher(0.6, backend=oracle.OracleBackend()).SearchWindowed(r, x, y, rad, pred - 1, pred, scene["da"], **kw)
    assert nr > 100 and ng == nr and np.array_equal(ag, ar)


@pytest.mark.parametrize("kfkf", [False, True])
def test_search_by_bow(scene, oracle, gpu_lib, kfkf):
    """Synthetic FeatureVectors: features are grouped into ~220 'vocabulary nodes' by image position, so that
    matching features of A and B mostly share a node (the vocabulary file is not in the reference repo)."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    gk, rk = both(scene, oracle, "a")
    gf, rf = both(scene, oracle, "b")

    def groups(k, dx, dy):
        node = ((k["y"] - dy) // 40).astype(np.int64) * 19 + ((k["x"] - dx) // 40).astype(np.int64)
        order = np.argsort(node, kind="stable")
        return node, order
    na, oa = groups(scene["ka"], 0, 0)
    nb, ob = groups(scene["kb"], 7, 4)
    common = np.intersect1d(na, nb)
    kf_off, f_off, kf_idx, f_idx = [0], [0], [], []
    for c in common:
        kf_idx += oa[na[oa] == c].tolist()
        f_idx += ob[nb[ob] == c].tolist()
        kf_off.append(len(kf_idx)); f_off.append(len(f_idx))
    kv = (np.arange(gk.N) % 7 != 0).astype(np.uint8)
    fv = (np.arange(gf.N) % 9 != 0).astype(np.uint8) if kfkf else None
    args = (kf_off, kf_idx, f_off, f_idx, kv, fv, kfkf)
    ng, mg = ORBmatcher(0.7, True).SearchByBoW(gk, gf, *args)
    nr, mr = ORBmatcher(0.7, True, backend=oracle.OracleBackend()).SearchByBoW(rk, rf, *args)
    assert nr > 100 and ng == nr and np.array_equal(mg, mr)


def test_rot_hist_filter(oracle, gpu_lib):
    from ar_orbslam2_b200.tracking import rot_hist_filter
    rng = np.random.default_rng(3)
    a = rng.uniform(0, 360, 3000).astype(np.float32)
    b = (a + rng.choice([0, 0, 0, 31, 95, 200], 3000) + rng.normal(0, 4, 3000)).astype(np.float32) % np.float32(360)
    assert np.array_equal(rot_hist_filter(a, b), oracle.rot_hist_filter(a, b))
    # the reference's 1/30 factor: only bins 0..12 can ever be hit
    assert np.array_equal(rot_hist_filter(a[:5], a[:5]), np.ones(5, bool))


def test_gpu_equals_reference_source_matchers(scene, oracle, gpu_lib):
    """liborbgpu directly against the reference's compiled ORBmatcher.cc (oracle/_ref), see tests/test_reference_matcher_cpu.py."""
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    if not oracle.ref_matcher_available():
        pytest.skip("oracle/_ref not built")
    bounds = (0.0, float(W), 0.0, float(H))
    mk = lambda which, be: Frame(scene["k" + which], scene["d" + which], scene["scale"], bounds, backend=be)
    g1, g2 = mk("a", None), mk("b", None)
    rb = oracle.ReferenceSourceBackend()
    r1, r2 = mk("a", rb), mk("b", rb)
    prev = np.stack([scene["ka"]["x"], scene["ka"]["y"]], 1).astype(np.float32).copy()
    pg, pr = prev.copy(), prev.copy()
    ng, mg = ORBmatcher(0.9, True).SearchForInitialization(g1, g2, pg, 100)
    nr, mr = ORBmatcher(0.9, True, backend=rb).SearchForInitialization(r1, r2, pr, 100)
    assert nr > 100 and ng == nr and np.array_equal(mg, mr) and np.array_equal(pg, pr)
    k = scene["ka"]
    n = len(k)
    in_view = (np.arange(n) % 10 != 0).astype(np.uint8)
    px, py = (k["x"] + 7).astype(np.float32), (k["y"] + 4).astype(np.float32)
    args = (in_view, px, py, (px - 6).astype(np.float32), k["octave"].astype(np.int32), np.full(n, 0.9, np.float32), scene["da"], 3.0,
            (np.arange(g2.N) % 13 == 0).astype(np.uint8), (np.arange(n) % 4 != 0).astype(np.uint8))
    ng, ag = ORBmatcher(0.8).SearchByProjectionPoints(g2, *args)
    nr, ar = ORBmatcher(0.8, backend=rb).SearchByProjectionPoints(r2, *args)
    assert nr > 100 and ng == nr and np.array_equal(ag, ar)


def test_gpu_windowed_equals_reference_kf_and_sim3(scene, oracle, gpu_lib):
    """orb_search_windowed on the GPU against the reference's SearchByProjection(Frame&, KeyFrame*, ...) (ORBmatcher.cc:1475) and
    SearchByProjection(KeyFrame*, Scw, ...) (:290), both run from the compiled reference source (oracle/_ref)."""
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.tracking import ORB_SELECT_BEST, Frame, ORBmatcher
    if not oracle.ref_matcher_available():
        pytest.skip("oracle/_ref not built")
    bounds = (0.0, float(W), 0.0, float(H))
    rb = oracle.ReferenceSourceBackend()
    g = Frame(scene["ka"], scene["da"], scene["scale"], bounds)
    rc = Frame(scene["ka"], scene["da"], scene["scale"], bounds, backend=rb)
    rk = Frame(scene["ka"], scene["da"], scene["scale"], bounds, backend=rb)
    k = scene["ka"]
    c = T.kf_projection_case(scene, oracle, None)
    for th, orbdist, ori in ((10.0, 100, True), (3.0, 64, True)):
        R = oracle.refm_search_by_projection_kf(rc, rk, c["xyz"], c["usable"], c["maxd"], c["mind"], c["claimed"], c["Tc"], T.FX, T.FY, T.CX, T.CY,
                                                th, orbdist, 0.9, ori)
        lvl = R["level"]
        radius = (np.float32(th) * scene["scale"][np.clip(lvl, 0, 7)]).astype(np.float32)
        ng, ag = ORBmatcher(0.9, ori).SearchWindowed(g, R["u"], R["v"], radius, lvl - 1, lvl + 1, scene["da"], ORB_SELECT_BEST, orbdist,
                                                     valid=R["valid"], angle=k["angle"], claimed=c["claimed"])
        assert R["nmatches"] > 100 and ng == R["nmatches"] and np.array_equal(ag, R["assigned"])
    c = T.sim3_projection_case(scene)
    for th in (10, 4):
        R = oracle.refm_search_by_projection_sim3(rk, c["S"], c["xyz"], c["nrm"], c["usable"], c["maxd"], c["mind"], c["claimed"], T.FX, T.FY, T.CX,
                                                  T.CY, th, 0.9)
        lvl = R["level"]
        radius = (np.float32(th) * scene["scale"][np.clip(lvl, 0, 7)]).astype(np.float32)
        ng, ag = ORBmatcher(0.9, False).SearchWindowed(g, R["u"], R["v"], radius, lvl - 1, lvl, scene["da"], ORB_SELECT_BEST, 50, valid=R["valid"],
                                                       strict_levels=True, claimed=c["claimed"], check_orientation=False)
        assert R["nmatches"] > 100 and ng == R["nmatches"] and np.array_equal(ag, R["assigned"])


# ---- round 2: device-built frame index, contention stress for the sweep-based claim resolution ----------------------

def test_frame_from_extractor_equals_host_frame(scene, oracle, gpu_lib):
    """orb_frame_from_extractor (grid built on the device from the extractor's resident results) must behave exactly like
    the frame built from host arrays: same GetFeaturesInArea lists, same SearchForInitialization result."""
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    bounds = (0.0, float(W), 0.0, float(H))
    ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H)
    ka, da = ex(a)
    assert ka.tobytes() == scene["ka"].tobytes() and np.array_equal(da, scene["da"])
    fa = Frame.from_extractor(ex, 0, bounds)
    fa.sync()
    kb, db = ex(b)
    fb = Frame.from_extractor(ex, 0, bounds)
    assert fa.N == len(ka) and fb.N == len(kb)
    ha, hb = Frame(ka, da, scene["scale"], bounds), Frame(kb, db, scene["scale"], bounds)
    rng = SplitMix64(77)
    for _ in range(40):
        x, y = rng.next() % (W + 40) - 20.0, rng.next() % (H + 40) - 20.0
        rad = 3.0 + rng.next() % 120
        lo, hi = int(rng.next() % 9) - 1, int(rng.next() % 9) - 1
        assert np.array_equal(fa.GetFeaturesInArea(x, y, rad, lo, hi), ha.GetFeaturesInArea(x, y, rad, lo, hi)), (x, y, rad, lo, hi)
    prev = np.stack([ka["x"], ka["y"]], 1).astype(np.float32).copy()
    p1, p2 = prev.copy(), prev.copy()
    n1, m1 = ORBmatcher(0.9, True).SearchForInitialization(fa, fb, p1, 100)
    n2, m2 = ORBmatcher(0.9, True).SearchForInitialization(ha, hb, p2, 100)
    assert n2 > 100 and n1 == n2 and np.array_equal(m1, m2) and np.array_equal(p1, p2)
    cand, sweeps = fb.last_stats()
    assert cand > 0 and sweeps >= 1
    for f in (fa, fb, ha, hb):
        f.close()
    ex.close()


def test_frame_from_extractor_stereo_and_distortion(oracle, gpu_lib):
    """Stereo frame: mvuRight handed over on the device (orb_extractor_stereo_result); distorted camera: key points
    undistorted on the device.  Both against the host-array frame fed with the same numbers."""
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.stereo import StereoFrame
    from ar_orbslam2_b200.synth import synth_stereo_pair
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher, UndistortKeyPoints, ComputeImageBounds
    import ctypes as C
    w, h = 1241, 376
    l, r = synth_stereo_pair(w, h, 9)
    mbf, mb = 386.1448, float(np.float32(386.1448) / np.float32(718.856))
    ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=2)
    kl, dl, kr, dr, ur, dp = StereoFrame(ex, l, r, mbf, mb)
    d_ur = C.c_void_p()
    assert gpu_lib.orb_extractor_stereo_result(ex._h, C.byref(d_ur), None) == 0
    cam = (718.856, 718.856, 607.1928, 185.2157)
    dist = np.array([-0.28, 0.07, 0.0002, 0.00002], np.float32)
    bounds = ComputeImageBounds(w, h, *cam, dist)
    fdev = Frame.from_extractor(ex, 0, bounds, camera=cam, dist_coef=dist, u_right_device=d_ur.value)
    kun = UndistortKeyPoints(kl, *cam, dist)
    fhost = Frame(kun, dl, ex.GetScaleFactors(), bounds, u_right=ur)
    n = len(kl)
    rng = SplitMix64(3)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (kun["x"] + 1.5).astype(np.float32), (kun["y"] - 1.0).astype(np.float32)
    pxr = np.where(ur > 0, ur + 1.5, px - 30).astype(np.float32)
    lvl = kl["octave"].astype(np.int32)
    vc = np.full(n, 0.95, np.float32)
    a = ORBmatcher(0.8).SearchByProjectionPoints(fdev, in_view, px, py, pxr, lvl, vc, dl, 3.0)
    b = ORBmatcher(0.8).SearchByProjectionPoints(fhost, in_view, px, py, pxr, lvl, vc, dl, 3.0)
    assert b[0] > 200 and a[0] == b[0] and np.array_equal(a[1], b[1])
    fdev.close(); fhost.close(); ex.close()


@pytest.mark.parametrize("select", [0, 1])
def test_windowed_contention_stress(scene, oracle, gpu_lib, select):
    """Three copies of every query (slightly moved, alternating `claims`) fight for the same features: long chains of
    'an earlier query took it' dependencies.  The sweep-based resolution must leave the reference's sequential result."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    g, r = both(scene, oracle, "b")
    k = scene["ka"]
    n = len(k)
    reps = 3
    x = np.concatenate([(k["x"] + 7 + 0.5 * j).astype(np.float32) for j in range(reps)])
    y = np.concatenate([(k["y"] + 4 - 0.5 * j).astype(np.float32) for j in range(reps)])
    # interleave so that copies of one point are far apart in query order as well as adjacent ones
    perm = np.argsort((np.arange(reps * n) * 7919) % (reps * n), kind="stable")
    x, y = x[perm], y[perm]
    oct_ = np.tile(k["octave"].astype(np.int32), reps)[perm]
    desc = np.tile(scene["da"], (reps, 1))[perm]
    ang = np.tile(k["angle"].astype(np.float32), reps)[perm]
    rad = (6.0 * scene["scale"][oct_]).astype(np.float32)
    claims = (np.arange(reps * n) % 3 != 1).astype(np.uint8)
    claimed = (np.arange(g.N) % 17 == 0).astype(np.uint8)
    kw = dict(select=select, th_dist=100, angle=ang, claims=claims, claimed=claimed, check_orientation=True, raw=True)
    ng, ag = ORBmatcher(0.9).SearchWindowed(g, x, y, rad, oct_ - 1, oct_ + 1, desc, **kw)
    nr, ar = ORBmatcher(0.9, backend=oracle.OracleBackend()).SearchWindowed(r, x, y, rad, oct_ - 1, oct_ + 1, desc, **kw)
    assert nr > 500 and ng == nr and np.array_equal(ag, ar)
    assert g.last_stats()[1] >= 3, "the scenario is meant to need several sweeps"


def test_initialization_contention_stress(scene, oracle, gpu_lib):
    """F1 holds every key point of A twice (identical descriptors): the second copy meets vMatchedDistance <= dist (ref
    ORBmatcher.cc:444-445) on its best candidate and must fall back exactly as the sequential loop does."""
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    bounds = (0.0, float(W), 0.0, float(H))
    k2 = np.concatenate([scene["ka"], scene["ka"]])
    k2["x"][len(scene["ka"]):] += 1.0
    d2 = np.concatenate([scene["da"], scene["da"]])
    d2[len(scene["da"]):, 0] ^= 1      # one bit off: the copy is slightly worse or better on its rivals
    mk = lambda be: (Frame(k2, d2, scene["scale"], bounds, backend=be), Frame(scene["kb"], scene["db"], scene["scale"], bounds, backend=be))
    g1, g2 = mk(None)
    r1, r2 = mk(oracle.OracleBackend())
    prev = np.stack([k2["x"], k2["y"]], 1).astype(np.float32).copy()
    for ori in (True, False):
        pg, pr = prev.copy(), prev.copy()
        ng, mg = ORBmatcher(0.9, ori).SearchForInitialization(g1, g2, pg, 100)
        nr, mr = ORBmatcher(0.9, ori, backend=oracle.OracleBackend()).SearchForInitialization(r1, r2, pr, 100)
        assert nr > 100 and ng == nr and np.array_equal(mg, mr) and np.array_equal(pg, pr)


def test_bow_contention_stress(scene, oracle, gpu_lib):
    """Few, large vocabulary nodes: many KF features of a node want the same F feature (ref ORBmatcher.cc:209-210)."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    gk, rk = both(scene, oracle, "a")
    gf, rf = both(scene, oracle, "a")      # the same frame on both sides: every feature has a perfect rival-free match ...
    ka = scene["ka"]
    node = (ka["y"] // 160).astype(np.int64) * 5 + (ka["x"] // 160).astype(np.int64)
    order = np.argsort(node, kind="stable")
    kf_off, f_off, kf_idx, f_idx = [0], [0], [], []
    for c in np.unique(node):
        members = order[node[order] == c].tolist()
        kf_idx += members + members[: len(members) // 2]    # ... and half of them are listed twice: the repeat must lose it
        f_idx += members
        kf_off.append(len(kf_idx)); f_off.append(len(f_idx))
    kv = np.ones(gk.N, np.uint8)
    args = (kf_off, kf_idx, f_off, f_idx, kv, None, False)
    ng, mg = ORBmatcher(0.9, True).SearchByBoW(gk, gf, *args)
    nr, mr = ORBmatcher(0.9, True, backend=oracle.OracleBackend()).SearchByBoW(rk, rf, *args)
    assert nr > 1000 and ng == nr and np.array_equal(mg, mr)
