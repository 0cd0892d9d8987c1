"""CPU test (no GPU): the flat-array matcher oracle against oracle/_ref/liborbmatcher_ref.so -- the reference's OWN
ORBmatcher.cc compiled unmodified against stub Frame / KeyFrame / MapPoint classes (oracle/refstub, oracle/Makefile.ref).
Match lists, counts and the updated vbPrevMatched must be identical.  This pins M1, M2, M3, M6, M7, M8 and the
GetFeaturesInArea traversal (S1) of SURVEY 8(a) to the reference's source."""
import numpy as np
import pytest

from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_descriptors, synth_frame

W, H = 752, 480
FX = FY = np.float32(435.2)
CX, CY = np.float32(367.2), np.float32(252.2)


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


def frames(scene, ref, which, u_right=None):
    from ar_orbslam2_b200.tracking import Frame
    k, d = scene["k" + which], scene["d" + which]
    bounds = (0.0, float(W), 0.0, float(H))
    return (Frame(k, d, scene["scale"], bounds, u_right=u_right, backend=ref.OracleBackend()),
            Frame(k, d, scene["scale"], bounds, u_right=u_right, backend=ref.ReferenceSourceBackend()))


def test_descriptor_distance(ref):
    a, b = synth_descriptors(4000, 1), synth_descriptors(4000, 2)
    assert np.array_equal(ref.descriptor_distance(a, b), ref.refm_descriptor_distance(a, b))


def test_features_in_area(scene, ref):
    o, r = frames(scene, ref, "a")
    rng = SplitMix64(5)
    for _ in range(200):
        x, y = rng.next() % (W + 40) - 20.0, rng.next() % (H + 40) - 20.0
        rad = 3.0 + rng.next() % 120
        lo, hi = int(rng.next() % 9) - 1, int(rng.next() % 9) - 1
        assert np.array_equal(o.GetFeaturesInArea(x, y, rad, lo, hi), r.GetFeaturesInArea(x, y, rad, lo, hi))


@pytest.mark.parametrize("ori,ratio", [(True, 0.9), (False, 0.9), (True, 0.6)])
def test_search_for_initialization(scene, ref, ori, ratio):
    from ar_orbslam2_b200.tracking import ORBmatcher
    o1, r1 = frames(scene, ref, "a")
    o2, r2 = frames(scene, ref, "b")
    prev = np.stack([scene["ka"]["x"], scene["ka"]["y"]], 1).astype(np.float32).copy()
    po, pr = prev.copy(), prev.copy()
    for _ in range(2):   # second call starts from the updated vbPrevMatched, as Tracking does
        no, mo = ORBmatcher(ratio, ori, backend=ref.OracleBackend()).SearchForInitialization(o1, o2, po, 100)
        nr, mr = ORBmatcher(ratio, ori, backend=ref.ReferenceSourceBackend()).SearchForInitialization(r1, r2, pr, 100)
        assert nr > 50 and no == nr and np.array_equal(mo, mr) and np.array_equal(po, pr)


def test_search_by_projection_points(scene, ref):
    from ar_orbslam2_b200.tracking import ORBmatcher
    ur = np.where(np.arange(len(scene["kb"])) % 3 == 0, scene["kb"]["x"] - 5.0, -1.0).astype(np.float32)
    o, r = frames(scene, ref, "b", ur)
    k = scene["ka"]
    n = len(k)
    rng = SplitMix64(9)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (k["x"] + 7).astype(np.float32), (k["y"] + 4).astype(np.float32)
    pxr = (px - 6).astype(np.float32)
    level = np.clip(k["octave"] + np.array([rng.next() % 3 - 1 for _ in range(n)]), 0, 7).astype(np.int32)
    vc = np.array([0.999 if rng.next() % 2 else 0.9 for _ in range(n)], np.float32)
    claimed = (np.arange(o.N) % 13 == 0).astype(np.uint8)
    obs = (np.arange(n) % 4 != 0).astype(np.uint8)
    for th, ratio in ((1.0, 0.8), (3.0, 0.8), (3.0, 0.6)):
        no, ao = ORBmatcher(ratio, backend=ref.OracleBackend()).SearchByProjectionPoints(o, in_view, px, py, pxr, level, vc, scene["da"], th, claimed, obs)
        nr, ar = ORBmatcher(ratio, backend=ref.ReferenceSourceBackend()).SearchByProjectionPoints(r, in_view, px, py, pxr, level, vc, scene["da"], th, claimed, obs)
        assert nr > 100 and no == nr and np.array_equal(ao, ar), (th, ratio)


@pytest.mark.parametrize("mono,tz", [(True, 0.0), (False, 0.3), (False, -0.3), (False, 0.0)])
def test_search_by_projection_last(scene, ref, mono, tz):
    """Last = frame A with a synthetic map point per feature, Cur = frame A's features seen from a translated
    pose; forward / backward / neither by the z translation vs the baseline mb."""
    from ar_orbslam2_b200.tracking import ORBmatcher
    k = scene["ka"]
    n = len(k)
    ur = np.where(np.arange(n) % 3 == 0, k["x"] - 5.0, -1.0).astype(np.float32)
    oc, rc = frames(scene, ref, "a", ur)
    ol, rl = frames(scene, ref, "a")
    z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
    xyz = np.stack([(k["x"] - CX) * z / FX, (k["y"] - CY) * z / FY, z], 1).astype(np.float32)
    has = (np.arange(n) % 17 != 0).astype(np.uint8)
    obs = (np.arange(n) % 5 != 0).astype(np.uint8)
    claimed = (np.arange(n) % 11 == 0).astype(np.uint8)
    Tc = np.eye(4, dtype=np.float32); Tc[:3, 3] = (0.05, 0.02, tz)
    Tl = np.eye(4, dtype=np.float32)
    mbf, mb, th = 40.0, 0.11, 15.0
    R = ref.refm_search_by_projection_last(rc, rl, xyz, has, obs, claimed, Tc, Tl, FX, FY, CX, CY, mbf, mb, th, mono, 0.9, True)
    assert R["nmatches"] > 50
    assert R["forward"] == ((not mono) and -tz > mb) and R["backward"] == ((not mono) and tz > mb)   # ref :1351-1352
    no, ao = ORBmatcher(0.9, True, backend=ref.OracleBackend()).SearchByProjectionLast(
        oc, R["valid"], R["u"], R["v"], R["invz"], k["octave"].astype(np.int32), k["angle"], scene["da"], th,
        forward=R["forward"], backward=R["backward"], mbf=mbf, claimed=claimed, obs_positive=obs)
    assert no == R["nmatches"] and np.array_equal(ao, R["assigned"])


@pytest.mark.parametrize("kfkf", [False, True])
def test_search_by_bow(scene, ref, kfkf):
    from ar_orbslam2_b200.tracking import ORBmatcher
    ok, rk = frames(scene, ref, "a")
    of, rf = frames(scene, ref, "b")

    def groups(k, dx, dy):
        node = ((k["y"] - dy) // 40).astype(np.int64) * 19 + ((k["x"] - dx) // 40).astype(np.int64)
        return node, np.argsort(node, kind="stable")
    na, oa = groups(scene["ka"], 0, 0)
    nb, ob = groups(scene["kb"], 7, 4)
    kf_off, f_off, kf_idx, f_idx = [0], [0], [], []
    for c in np.intersect1d(na, nb):
        kf_idx += oa[na[oa] == c].tolist(); f_idx += ob[nb[ob] == c].tolist()
        kf_off.append(len(kf_idx)); f_off.append(len(f_idx))
    kv = (np.arange(ok.N) % 7 != 0).astype(np.uint8)
    fv = (np.arange(of.N) % 9 != 0).astype(np.uint8) if kfkf else None
    for ori in (True, False):
        no, mo = ORBmatcher(0.7, ori, backend=ref.OracleBackend()).SearchByBoW(ok, of, kf_off, kf_idx, f_off, f_idx, kv, fv, kfkf)
        nr, mr = ref.refm_search_by_bow(rk, rf, kf_off, kf_idx, f_off, f_idx, kv, fv, kfkf, 0.7, ori)
        assert nr > 100 and no == nr and np.array_equal(mo, mr)


@pytest.mark.parametrize("w,h,nf,seed", [(1241, 376, 2000, 0), (1241, 376, 2000, 7), (640, 480, 1000, 3)])
def test_stereo_oracle_equals_reference_frame_cc(ref, w, h, nf, seed):
    """Frame::ComputeStereoMatches: the oracle's restatement against the reference's own Frame.cc member function
    (compiled into oracle/_ref with the reference's Frame.h and ORBextractor): mvuRight and mvDepth bit-identical."""
    from ar_orbslam2_b200.synth import synth_stereo_pair
    left, right = synth_stereo_pair(w, h, seed)
    a, b = ref.RefExtractor(nf), ref.RefExtractor(nf)
    kl, dl = a.extract(left)
    kr, dr = b.extract(right)
    mbf = np.float32(386.1448)
    mb = mbf / np.float32(718.856)
    rc, ur, dp, _, nm = ref.stereo_match(a, b, kl, dl, kr, dr, mbf, mb)
    assert rc == 0 and nm > 100
    rur, rdp = ref.refm_stereo_match(left, right, nf, kl, dl, kr, dr, mbf, mb)
    assert np.array_equal(ur.view(np.uint32), rur.view(np.uint32)), "mvuRight differs from the reference's Frame.cc"
    assert np.array_equal(dp.view(np.uint32), rdp.view(np.uint32)), "mvDepth differs from the reference's Frame.cc"


def _world_points(k, depth_seed=0):
    n = len(k)
    z = np.array([2 + (SplitMix64(i + depth_seed).next() % 1000) / 100 for i in range(n)], np.float32)
    xyz = np.stack([(k["x"] - CX) * z / FX, (k["y"] - CY) * z / FY, z], 1).astype(np.float32)
    return xyz, z


def kf_projection_case(scene, ref, backend_frames):
    """SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) scenario shared by the CPU and GPU tests: KF = frame A
    with a map point per feature, current frame = frame A's features seen from a translated pose."""
    k = scene["ka"]
    n = len(k)
    xyz, z = _world_points(k)
    usable = (np.arange(n) % 9 != 0).astype(np.uint8)
    # distance-invariance range chosen so that PredictScale lands on the feature's own octave +-1
    maxd = (z * np.float32(1.2) ** k["octave"]).astype(np.float32)
    mind = (maxd / np.float32(1.2) ** 8).astype(np.float32)
    claimed = (np.arange(n) % 14 == 0).astype(np.uint8)
    Tc = np.eye(4, dtype=np.float32); Tc[:3, 3] = (0.03, -0.02, 0.1)
    return dict(xyz=xyz, usable=usable, maxd=maxd, mind=mind, claimed=claimed, Tc=Tc)


@pytest.mark.parametrize("th,orbdist,ori", [(10.0, 100, True), (3.0, 64, True), (10.0, 100, False)])
def test_search_by_projection_keyframe(scene, ref, th, orbdist, ori):
    from ar_orbslam2_b200.tracking import ORB_SELECT_BEST, ORBmatcher
    oc, rc = frames(scene, ref, "a")
    _, rk = frames(scene, ref, "a")
    c = kf_projection_case(scene, ref, None)
    R = ref.refm_search_by_projection_kf(rc, rk, c["xyz"], c["usable"], c["maxd"], c["mind"], c["claimed"], c["Tc"], FX, FY, CX, CY, th, orbdist, 0.9, ori)
    assert R["nmatches"] > 100 and R["valid"].sum() > 500
    lvl = R["level"]
    assert lvl[R["valid"] == 1].min() >= 0 and lvl[R["valid"] == 1].max() <= 7
    radius = (np.float32(th) * scene["scale"][np.clip(lvl, 0, 7)]).astype(np.float32)
    k = scene["ka"]
    no, ao = ORBmatcher(0.9, ori, backend=ref.OracleBackend()).SearchWindowed(
        oc, R["u"], R["v"], radius, lvl - 1, lvl + 1, scene["da"], ORB_SELECT_BEST, orbdist, valid=R["valid"], angle=k["angle"], claimed=c["claimed"])
    assert no == R["nmatches"] and np.array_equal(ao, R["assigned"])


def sim3_projection_case(scene):
    k = scene["ka"]
    n = len(k)
    xyz, z = _world_points(k, 7)
    nrm = -xyz / np.linalg.norm(xyz, axis=1, keepdims=True)          # facing the camera at the origin ...
    nrm[::6] *= -1                                                    # ... except some, which fail the 60-degree test
    usable = (np.arange(n) % 8 != 0).astype(np.uint8)
    maxd = (z * np.float32(1.2) ** k["octave"] * np.float32(0.999)).astype(np.float32)
    mind = (maxd / np.float32(1.2) ** 8).astype(np.float32)
    claimed = (np.arange(n) % 12 == 0).astype(np.uint8)
    S = np.eye(4, dtype=np.float32) * np.float32(1.0); S[:3, 3] = (0.02, 0.01, 0.05)
    S[:3, :] *= np.float32(1.1)                                       # a similarity with scale 1.1
    return dict(xyz=xyz, nrm=nrm.astype(np.float32), usable=usable, maxd=maxd, mind=mind, claimed=claimed, S=S)


@pytest.mark.parametrize("th", [10, 4])
def test_search_by_projection_sim3(scene, ref, th):
    from ar_orbslam2_b200.tracking import ORB_SELECT_BEST, ORBmatcher
    ok, rk = frames(scene, ref, "a")
    c = sim3_projection_case(scene)
    R = ref.refm_search_by_projection_sim3(rk, c["S"], c["xyz"], c["nrm"], c["usable"], c["maxd"], c["mind"], c["claimed"], FX, FY, CX, CY, th, 0.9)
    assert R["nmatches"] > 100
    lvl = R["level"]
    assert lvl[R["valid"] == 1].min() >= 0 and lvl[R["valid"] == 1].max() <= 7
    radius = (np.float32(th) * scene["scale"][np.clip(lvl, 0, 7)]).astype(np.float32)
    no, ao = ORBmatcher(0.9, False, backend=ref.OracleBackend()).SearchWindowed(
        ok, R["u"], R["v"], radius, lvl - 1, lvl, scene["da"], ORB_SELECT_BEST, 50, valid=R["valid"], strict_levels=True, claimed=c["claimed"],
        check_orientation=False)
    assert no == R["nmatches"] and np.array_equal(ao, R["assigned"])


# ---- the projection pre-loops as one flat routine (orb_project_points / orbref_project_points) ----
def _neg_Rt_t(R, t):
    """-R^T t with cv::Mat's rounding: the product accumulates in double, the result is float (negation is exact)."""
    return (-(R.T.astype(np.float64) @ t.astype(np.float64))).astype(np.float32)


def _predict_scale(maxd_raw, dist, logs):
    """MapPoint::PredictScale (MapPoint.cc:357-366) with the host libm's logf."""
    import ctypes as C
    libm = C.CDLL("libm.so.6")
    libm.logf.restype = C.c_float
    libm.logf.argtypes = [C.c_float]
    out = np.zeros(len(dist), np.int32)
    for i in range(len(dist)):
        if dist[i] > 0:
            ratio = np.float32(maxd_raw[i]) / np.float32(dist[i])
            out[i] = int(np.ceil(np.float32(libm.logf(float(ratio))) / np.float32(logs)))
    return out


def _rot(seed, amount):
    rng = np.random.default_rng(seed)
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    return (np.eye(3) + np.sin(amount) * K + (1 - np.cos(amount)) * K @ K).astype(np.float32)


@pytest.mark.parametrize("rot", [0.0, 0.05])
def test_project_points_equals_reference_lines_kf(scene, ref, rot):
    """orbref_project_points with the rules of SearchByProjection(Frame&, KeyFrame*) against the reference's lines :1491-1531 as
    the driver replays them: valid set, u and v bit for bit."""
    _, rc = frames(scene, ref, "a")
    _, rk = frames(scene, ref, "a")
    c = kf_projection_case(scene, ref, None)
    Tc = c["Tc"].copy(); Tc[:3, :3] = _rot(3, rot)
    R = ref.refm_search_by_projection_kf(rc, rk, c["xyz"], c["usable"], c["maxd"], c["mind"], c["claimed"], Tc, FX, FY, CX, CY, 10.0, 100, 0.9, True)
    Rcw, tcw = Tc[:3, :3], Tc[:3, 3]
    proj = dict(R1=Rcw, t1=tcw, Ow=_neg_Rt_t(Rcw, tcw), fx=FX, fy=FY, cx=CX, cy=CY, min_x=0, max_x=W, min_y=0, max_y=H,
                invz_rule=0, uv_rule=0, bounds_rule=0, reject_negative_z=0, dist_rule=0, check_view_angle=0)
    ok, u, v, ur, dist = ref.project_points(proj, c["xyz"], None, np.float32(1.2) * c["maxd"], np.float32(0.8) * c["mind"], c["usable"])
    lvl = _predict_scale(c["maxd"], dist, np.log(np.float32(1.2)))
    valid = (ok == 1) & (lvl >= 0) & (lvl < 8)
    assert np.array_equal(valid.astype(np.uint8), R["valid"]) and valid.sum() > 500
    assert np.array_equal(u[valid].view(np.uint32), R["u"][valid].view(np.uint32))
    assert np.array_equal(v[valid].view(np.uint32), R["v"][valid].view(np.uint32))
    assert np.array_equal(lvl[valid], R["level"][valid])


@pytest.mark.parametrize("rot", [0.0, 0.04])
def test_project_points_equals_reference_lines_sim3(scene, ref, rot):
    """... with the rules of SearchByProjection(KeyFrame*, Scw, ...) against :298-360 (Sim3 decomposition included)."""
    _, rk = frames(scene, ref, "a")
    c = sim3_projection_case(scene)
    S = c["S"].copy(); S[:3, :3] = (np.float32(1.1) * _rot(5, rot)).astype(np.float32)
    R = ref.refm_search_by_projection_sim3(rk, S, c["xyz"], c["nrm"], c["usable"], c["maxd"], c["mind"], c["claimed"], FX, FY, CX, CY, 10, 0.9)
    sR = S[:3, :3]
    scw = np.float32(np.sqrt(np.sum(sR[0].astype(np.float64) * sR[0].astype(np.float64))))     # const float scw = sqrt(row.dot(row))
    Rcw = (sR.astype(np.float64) / np.float64(scw)).astype(np.float32)                          # Mat / scalar: double, then float
    tcw = (S[:3, 3].astype(np.float64) / np.float64(scw)).astype(np.float32)
    proj = dict(R1=Rcw, t1=tcw, Ow=_neg_Rt_t(Rcw, tcw), fx=FX, fy=FY, cx=CX, cy=CY, min_x=0, max_x=W, min_y=0, max_y=H,
                invz_rule=1, uv_rule=1, bounds_rule=1, reject_negative_z=1, dist_rule=0, check_view_angle=1)
    ok, u, v, ur, dist = ref.project_points(proj, c["xyz"], c["nrm"], np.float32(1.2) * c["maxd"], np.float32(0.8) * c["mind"], c["usable"])
    lvl = _predict_scale(c["maxd"], dist, np.log(np.float32(1.2)))
    valid = (ok == 1) & (lvl >= 0) & (lvl < 8)
    assert np.array_equal(valid.astype(np.uint8), R["valid"]) and valid.sum() > 150
    assert np.array_equal(u[valid].view(np.uint32), R["u"][valid].view(np.uint32))
    assert np.array_equal(v[valid].view(np.uint32), R["v"][valid].view(np.uint32))
