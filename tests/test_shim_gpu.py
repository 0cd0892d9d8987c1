"""GPU test of the C++ drop-in shim (ar_orbslam2_b200/shim/ORBextractor.{h,cc}): the reference's class interface --
ORBextractor(int,float,int,int,int), operator()(InputArray, InputArray, vector<KeyPoint>&, OutputArray), the getters,
mvImagePyramid -- driven as Frame's stereo constructor drives it (two extractors on two threads, ref Frame.cc:83-95),
compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

from ar_orbslam2_b200._abi import KEYPOINT_DTYPE
from ar_orbslam2_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_shim_stereo_frame(oracle, gpu_lib, tmp_path):
    exe = os.path.join(ROOT, "ar_orbslam2_b200", "shim", "shim_selftest")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(exe)])
    w, h, nf = 1241, 376, 2000
    left, right = synth_stereo_pair(w, h, 77)
    left.tofile(tmp_path / "l.raw"); right.tofile(tmp_path / "r.raw")
    out = subprocess.run([exe, str(w), str(h), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(nf), str(tmp_path / "o")],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr + out.stdout
    assert "shim ok=1" in out.stdout and "levels=8" in out.stdout
    rl, rr = oracle.RefExtractor(nf), oracle.RefExtractor(nf)
    kl, dl = rl.extract(left)
    kr, dr = rr.extract(right)
    mbf = np.float32(386.1448)
    rc, ur, dp, _, _ = oracle.stereo_match(rl, rr, kl, dl, kr, dr, mbf, mbf / np.float32(718.856))
    g = lambda name, dt: np.fromfile(tmp_path / f"o_{name}.bin", dt)
    assert g("kl", KEYPOINT_DTYPE).tobytes() == kl.tobytes() and g("kr", KEYPOINT_DTYPE).tobytes() == kr.tobytes()
    assert np.array_equal(g("dl", np.uint8).reshape(-1, 32), dl) and np.array_equal(g("dr", np.uint8).reshape(-1, 32), dr)
    assert np.array_equal(g("ur", np.uint32), ur.view(np.uint32)) and np.array_equal(g("dp", np.uint32), dp.view(np.uint32))
    # ORBGPU_ExtractStereo: the single-call front end leaves the same six results
    assert "fused ok=1" in out.stdout
    assert g("fkl", KEYPOINT_DTYPE).tobytes() == kl.tobytes() and g("fkr", KEYPOINT_DTYPE).tobytes() == kr.tobytes()
    assert np.array_equal(g("fdl", np.uint8).reshape(-1, 32), dl) and np.array_equal(g("fdr", np.uint8).reshape(-1, 32), dr)
    assert np.array_equal(g("fur", np.uint32), ur.view(np.uint32)) and np.array_equal(g("fdp", np.uint32), dp.view(np.uint32))
    # mvImagePyramid mirror keeps the reference's padded-plane layout (19-px reflect border around the ROI)
    assert np.array_equal(g("l3", np.uint8), rl.pyramid_level(3, True).ravel())
    # the marker pipeline's calls (Marker.cc:83-84,107-121,314-349) through orbgpu_shim::ORB / BruteForceMatchHamming / SearchMatches
    assert "marker ok" in out.stdout
    mk1, md1 = oracle.cv_orb(left)
    mk2, md2 = oracle.cv_orb(right)
    assert g("mk1", KEYPOINT_DTYPE).tobytes() == mk1.tobytes() and g("mk2", KEYPOINT_DTYPE).tobytes() == mk2.tobytes()
    assert np.array_equal(g("md1", np.uint8).reshape(-1, 32), md1) and np.array_equal(g("md2", np.uint8).reshape(-1, 32), md2)
    dmatch = np.dtype([("queryIdx", "<i4"), ("trainIdx", "<i4"), ("imgIdx", "<i4"), ("distance", "<f4")])
    bf = g("mbf", dmatch)
    bd, bi, _, _ = oracle.hamming_top2(md1, md2)
    assert np.array_equal(bf["queryIdx"], np.arange(len(md1))) and np.array_equal(bf["trainIdx"], bi) and np.array_equal(bf["distance"], bd)
    rm = g("mrm", dmatch)
    tidx, dist = oracle.search_matches_ratio(md1, md2)
    keep = np.nonzero(tidx >= 0)[0]
    assert np.array_equal(rm["queryIdx"], keep) and np.array_equal(rm["trainIdx"], tidx[keep]) and np.array_equal(rm["distance"], dist[keep])
    # AR-1.3's naive_nn_search (nearest marker row per frame row, kept when <= 50) and naive_nn_search2 (= the ratio matcher) + minD / maxD
    nd, ni, _, _ = oracle.hamming_top2(md2, md1)
    k1 = np.nonzero(nd <= 50)[0]
    n1 = g("mn1", dmatch)
    assert np.array_equal(n1["queryIdx"], k1) and np.array_equal(n1["trainIdx"], ni[k1]) and np.array_equal(n1["distance"], nd[k1])
    assert g("mn2", dmatch).tobytes() == rm.tobytes()
    alld = np.unpackbits(md2[:, None, :] ^ md1[None, :, :], axis=2).sum(axis=2)
    import re
    m = re.search(r"minD=(\d+) maxD=(\d+)", out.stdout)
    assert m and int(m.group(1)) == min(100, int(alld.min())) and int(m.group(2)) == int(alld.max())


def test_cpp_matcher_shim_equals_reference(oracle, gpu_lib):
    """ar_orbslam2_b200/shim/ORBmatcher_gpu.cc -- the reference-side binding of INTEGRATION.md section 4 -- compiled against the
    reference's own Frame.h (oracle/_ref/liborbmatcher_shim.so) and driven exactly like the reference's compiled ORBmatcher.cc
    (liborbmatcher_ref.so): SearchForInitialization, SearchByProjection(Frame,MapPoints), SearchByProjection(Cur,Last),
    SearchByBoW must leave the same results in the reference's data structures."""
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_frame
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    W, H = T.W, T.H
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    scale = ex.tables()["scale"]
    bounds = (0.0, float(W), 0.0, float(H))
    rb, sb = oracle.ReferenceSourceBackend(), oracle.ShimBackend()
    mk = lambda k, d, be, ur=None: Frame(k, d, scale, bounds, u_right=ur, backend=be)
    # SearchForInitialization
    r1, r2, s1, s2 = mk(ka, da, rb), mk(kb, db, rb), mk(ka, da, sb), mk(kb, db, sb)
    prev = np.stack([ka["x"], ka["y"]], 1).astype(np.float32).copy()
    pr, ps = prev.copy(), prev.copy()
    nr, mr = ORBmatcher(0.9, True, backend=rb).SearchForInitialization(r1, r2, pr, 100)
    ns, ms = ORBmatcher(0.9, True, backend=sb).SearchForInitialization(s1, s2, ps, 100)
    assert nr > 100 and ns == nr and np.array_equal(ms, mr) and np.array_equal(ps, pr)
    # SearchByProjection(Frame, MapPoints)
    n = len(ka)
    ur = np.where(np.arange(len(kb)) % 3 == 0, kb["x"] - 5.0, -1.0).astype(np.float32)
    rf, sf = mk(kb, db, rb, ur), mk(kb, db, sb, ur)
    rng = SplitMix64(9)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (ka["x"] + 7).astype(np.float32), (ka["y"] + 4).astype(np.float32)
    args = (in_view, px, py, (px - 6).astype(np.float32), ka["octave"].astype(np.int32), np.full(n, 0.9, np.float32), da, 3.0,
            (np.arange(len(kb)) % 13 == 0).astype(np.uint8), (np.arange(n) % 4 != 0).astype(np.uint8))
    nr, ar = ORBmatcher(0.8, backend=rb).SearchByProjectionPoints(rf, *args)
    ns, as_ = ORBmatcher(0.8, backend=sb).SearchByProjectionPoints(sf, *args)
    assert nr > 100 and ns == nr and np.array_equal(as_, ar)
    # SearchByProjection(Cur, Last): the shim runs the reference's pose algebra itself
    ura = np.where(np.arange(n) % 3 == 0, ka["x"] - 5.0, -1.0).astype(np.float32)
    z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
    xyz = np.stack([(ka["x"] - T.CX) * z / T.FX, (ka["y"] - T.CY) * z / T.FY, z], 1).astype(np.float32)
    has, obs, claimed = (np.arange(n) % 17 != 0).astype(np.uint8), (np.arange(n) % 5 != 0).astype(np.uint8), (np.arange(n) % 11 == 0).astype(np.uint8)
    Tl = np.eye(4, dtype=np.float32)
    for mono, tz in ((True, 0.0), (False, 0.3), (False, -0.3)):
        Tc = np.eye(4, dtype=np.float32); Tc[:3, 3] = (0.05, 0.02, tz)
        R = oracle.refm_search_by_projection_last(mk(ka, da, rb, ura), mk(ka, da, rb), xyz, has, obs, claimed, Tc, Tl, T.FX, T.FY, T.CX, T.CY, 40.0, 0.11,
                                                  15.0, mono, 0.9, True)
        S = oracle.refm_search_by_projection_last(mk(ka, da, sb, ura), mk(ka, da, sb), xyz, has, obs, claimed, Tc, Tl, T.FX, T.FY, T.CX, T.CY, 40.0, 0.11,
                                                  15.0, mono, 0.9, True)
        assert R["nmatches"] > 50 and S["nmatches"] == R["nmatches"] and np.array_equal(S["assigned"], R["assigned"])
    # SearchByBoW(KeyFrame, Frame)
    def groups(k, dx, dy):
        node = ((k["y"] - dy) // 40).astype(np.int64) * 19 + ((k["x"] - dx) // 40).astype(np.int64)
        return node, np.argsort(node, kind="stable")
    na, oa = groups(ka, 0, 0)
    nb, ob = groups(kb, 7, 4)
    kf_off, f_off, kf_idx, f_idx = [0], [0], [], []
    for c in np.intersect1d(na, nb):
        kf_idx += oa[na[oa] == c].tolist(); f_idx += ob[nb[ob] == c].tolist()
        kf_off.append(len(kf_idx)); f_off.append(len(f_idx))
    kv = (np.arange(n) % 7 != 0).astype(np.uint8)
    nr, mr = oracle.refm_search_by_bow(mk(ka, da, rb), mk(kb, db, rb), kf_off, kf_idx, f_off, f_idx, kv, None, False, 0.7, True)
    ns, ms = oracle.refm_search_by_bow(mk(ka, da, sb), mk(kb, db, sb), kf_off, kf_idx, f_off, f_idx, kv, None, False, 0.7, True)
    assert nr > 100 and ns == nr and np.array_equal(ms, mr)
    # SearchByBoW(KeyFrame, KeyFrame): strict `< TH_LOW`, second-side validity
    fv = (np.arange(len(kb)) % 9 != 0).astype(np.uint8)
    for ori in (True, False):
        nr, mr = oracle.refm_search_by_bow(mk(ka, da, rb), mk(kb, db, rb), kf_off, kf_idx, f_off, f_idx, kv, fv, True, 0.7, ori)
        ns, ms = oracle.refm_search_by_bow(mk(ka, da, sb), mk(kb, db, sb), kf_off, kf_idx, f_off, f_idx, kv, fv, True, 0.7, ori)
        assert nr > 100 and ns == nr and np.array_equal(ms, mr)


def test_cpp_projection_shims_m4_m5(oracle, gpu_lib):
    """The shim bodies of SearchByProjection(Frame&, KeyFrame*, set, th, ORBdist) (relocalisation) and
    SearchByProjection(KeyFrame*, Scw, points, matched, th) (loop closing), compiled against the reference's headers, against
    the reference's own compiled methods behind the same driver: same assignments and counts."""
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.synth import shifted_frame, synth_frame
    from ar_orbslam2_b200.tracking import Frame
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    W, H = T.W, T.H
    a = synth_frame(W, H, 31)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    scene = dict(ka=ka, da=da, scale=ex.tables()["scale"])
    bounds = (0.0, float(W), 0.0, float(H))
    rb, sb = oracle.ReferenceSourceBackend(), oracle.ShimBackend()
    mk = lambda be: Frame(ka, da, scene["scale"], bounds, backend=be)
    c = T.kf_projection_case(scene, oracle, None)
    for th, orbdist, ori in ((10.0, 100, True), (3.0, 64, False)):
        R, S = [oracle.refm_search_by_projection_kf(mk(be), mk(be), c["xyz"], c["usable"], c["maxd"], c["mind"], c["claimed"], c["Tc"],
                                                    T.FX, T.FY, T.CX, T.CY, th, orbdist, 0.9, ori) for be in (rb, sb)]
        assert R["nmatches"] > 100 and S["nmatches"] == R["nmatches"] and np.array_equal(S["assigned"], R["assigned"]), (th, orbdist, ori)
    c = T.sim3_projection_case(scene)
    for th in (10, 4):
        R, S = [oracle.refm_search_by_projection_sim3(mk(be), c["S"], c["xyz"], c["nrm"], c["usable"], c["maxd"], c["mind"], c["claimed"],
                                                      T.FX, T.FY, T.CX, T.CY, th, 0.9) for be in (rb, sb)]
        assert R["nmatches"] > 100 and S["nmatches"] == R["nmatches"] and np.array_equal(S["assigned"], R["assigned"]), th


def test_cpp_frame_stereo_dropin(oracle, gpu_lib):
    """The reference's Frame (its own Frame.h / Frame.cc, unmodified) linked with the product's ORBextractor class and
    shim/Frame_stereo.cc: `F.ComputeStereoMatches()` -- the call the stereo constructor makes (ref Frame.cc:95) -- leaves the
    mvuRight / mvDepth of the reference's own member function."""
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    w, h, nf = 1241, 376, 2000
    left, right = synth_stereo_pair(w, h, 41)
    rl, rr = oracle.RefExtractor(nf), oracle.RefExtractor(nf)
    kl, dl = rl.extract(left)
    kr, dr = rr.extract(right)
    mbf = np.float32(386.1448)
    mb = mbf / np.float32(718.856)
    ur_ref, dp_ref = oracle.refm_stereo_match(left, right, nf, kl, dl, kr, dr, mbf, mb)
    ur, dp = oracle.refm_stereo_match(left, right, nf, kl, dl, kr, dr, mbf, mb, shim=True)
    assert (ur_ref >= 0).sum() > 500
    assert np.array_equal(ur.view(np.uint32), ur_ref.view(np.uint32)) and np.array_equal(dp.view(np.uint32), dp_ref.view(np.uint32))


def test_cpp_matcher_class_frame_cache(oracle, gpu_lib):
    """The per-thread cache of device frame indices inside the shim: repeated matcher calls on the same Frame objects (what
    Tracking does with mCurrentFrame) reuse one index and keep returning the reference's result."""
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.synth import shifted_frame, synth_frame
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    W, H = T.W, T.H
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    scale = ex.tables()["scale"]
    bounds = (0.0, float(W), 0.0, float(H))
    rb, sb = oracle.ReferenceSourceBackend(), oracle.ShimBackend()
    r1, r2 = Frame(ka, da, scale, bounds, backend=rb), Frame(kb, db, scale, bounds, backend=rb)
    s1, s2 = Frame(ka, da, scale, bounds, backend=sb), Frame(kb, db, scale, bounds, backend=sb)
    prev = np.stack([ka["x"], ka["y"]], 1).astype(np.float32).copy()
    pr, ps = prev.copy(), prev.copy()
    us = []
    for it in range(6):   # vbPrevMatched evolves call by call, as in Tracking::MonocularInitialization
        nr, mr = ORBmatcher(0.9, True, backend=rb).SearchForInitialization(r1, r2, pr, 100)
        ns, ms = ORBmatcher(0.9, True, backend=sb).SearchForInitialization(s1, s2, ps, 100)
        assert ns == nr and np.array_equal(ms, mr) and np.array_equal(ps, pr)
        sb.L.refm_last_call_us.restype = __import__("ctypes").c_double
        us.append(sb.L.refm_last_call_us())
    assert min(us[1:]) < us[0], f"cached calls should be faster than the first: {us}"
