"""The thread contract of the drop-in classes on the GPU: the reference runs its matchers from three threads at once -- Tracking
(main), LocalMapping and LoopClosing (ref ORB_SLAM2/src/System.cc:91,95) -- and extracts on two (ref Frame.cc:83-86).  Here >= 3
host threads go through the C++ shim (ar_orbslam2_b200/shim/ORBmatcher_gpu.cc compiled against the reference's headers,
oracle/_ref/liborbmatcher_shim.so; ctypes releases the GIL around every call, so the native code does run concurrently):

  thread A (Tracking)      ORBextractor::operator() + SearchByProjection(Cur, Last) + SearchByProjection(Frame, MapPoints)
  thread B (LocalMapping)  SearchForTriangulation + Fuse
  thread C (LoopClosing)   SearchByBoW(KF, KF) + SearchBySim3 + SearchByProjection(KF, Scw)
  threads D, E             the stereo constructor's two extractions, one extractor object each

Every iteration's result must be byte-identical to the same thread body run alone beforehand.  Every thread works on its own
host-side Frame / KeyFrame objects built from the SAME key points and descriptors (the test driver writes match state into those
objects, as the reference does into Frame::mvpMapPoints); what the threads share is liborbgpu itself: the device, the frame-handle
pool, the per-thread call streams and caches."""
import os
import threading

import numpy as np
import pytest

import m9_cases as M
from ar_orbslam2_b200.synth import SplitMix64, shifted_frame, synth_frame, synth_stereo_pair

pytestmark = pytest.mark.gpu
W, H = M.W, M.H
ITERATIONS = 200


@pytest.fixture(scope="module")
def scene(oracle):
    a = synth_frame(W, H, 31)
    b = shifted_frame(a, 7, 4, 32)
    ex = oracle.RefExtractor(2000)
    ka, da = ex.extract(a)
    kb, db = ex.extract(b)
    return dict(ka=ka, da=da, kb=kb, db=db, scale=ex.tables()["scale"], img_a=a)


def _flat(x):
    """anything a matcher wrapper returns -> bytes"""
    if isinstance(x, dict):
        return b"".join(_flat(x[k]) for k in sorted(x))
    if isinstance(x, (tuple, list)):
        return b"".join(_flat(v) for v in x)
    return np.ascontiguousarray(x).tobytes()


def _tracking_body(scene, oracle):
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.extractor import ORBextractor
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    sb = oracle.ShimBackend()
    ka, da, kb, db = scene["ka"], scene["da"], scene["kb"], scene["db"]
    bounds = (0.0, float(W), 0.0, float(H))
    n = len(ka)
    ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=W, max_height=H)
    ur_b = np.where(np.arange(len(kb)) % 3 == 0, kb["x"] - 5.0, -1.0).astype(np.float32)
    ur_a = np.where(np.arange(n) % 3 == 0, ka["x"] - 5.0, -1.0).astype(np.float32)
    f_pts = Frame(kb, db, scene["scale"], bounds, u_right=ur_b, backend=sb)
    f_cur, f_last = Frame(ka, da, scene["scale"], bounds, u_right=ur_a, backend=sb), Frame(ka, da, scene["scale"], bounds, backend=sb)
    rng = SplitMix64(9)
    in_view = np.array([rng.next() % 10 != 0 for _ in range(n)], np.uint8)
    px, py = (ka["x"] + 7).astype(np.float32), (ka["y"] + 4).astype(np.float32)
    args = (in_view, px, py, (px - 6).astype(np.float32), ka["octave"].astype(np.int32), np.full(n, 0.9, np.float32), da, 3.0,
            (np.arange(len(kb)) % 13 == 0).astype(np.uint8), (np.arange(n) % 4 != 0).astype(np.uint8))
    z = np.array([2 + (SplitMix64(i).next() % 1000) / 100 for i in range(n)], np.float32)
    xyz = np.stack([(ka["x"] - T.CX) * z / T.FX, (ka["y"] - T.CY) * z / T.FY, z], 1).astype(np.float32)
    has, obs, claimed = (np.arange(n) % 17 != 0).astype(np.uint8), (np.arange(n) % 5 != 0).astype(np.uint8), (np.arange(n) % 11 == 0).astype(np.uint8)
    Tl, Tc = np.eye(4, dtype=np.float32), np.eye(4, dtype=np.float32)
    Tc[:3, 3] = (0.05, 0.02, 0.3)

    def once():
        k, d = ex(scene["img_a"])
        S = oracle.refm_search_by_projection_last(f_cur, f_last, xyz, has, obs, claimed, Tc, Tl, T.FX, T.FY, T.CX, T.CY, 40.0, 0.11, 15.0, False, 0.9, True)
        P = ORBmatcher(0.8, backend=sb).SearchByProjectionPoints(f_pts, *args)
        return _flat((k, d, S["nmatches"], S["assigned"], P))
    return once


def _mapping_body(scene, oracle):
    from ar_orbslam2_b200.tracking import Frame, ORBmatcher
    sb = oracle.ShimBackend()
    bounds = (0.0, float(W), 0.0, float(H))
    mk = lambda which, ur=None: Frame(scene["k" + which], scene["d" + which], scene["scale"], bounds, u_right=ur, backend=sb, keyframe_view=True)
    ct, cf = M.triangulation_case(scene), dict(M.fuse_case(scene))
    k1, k2, kfuse = mk("a", ct["ur1"]), mk("b", ct["ur2"]), mk("b", cf["ur_kf"])
    I, z, S = np.eye(3, dtype=np.float32), np.zeros(3, np.float32), np.eye(4, dtype=np.float32)

    def once():
        t = ORBmatcher(0.6, True, backend=sb).SearchForTriangulation(k1, k2, ct["off1"], ct["idx1"], ct["off2"], ct["idx2"], ct["usable1"],
                                                                     ct["usable2"], ct["F12"], ct["ex"], ct["ey"], ct["sigma2"], False)
        f = oracle.refm_fuse(kfuse, 0, I, z, S, cf["world"], cf["normal"], cf["present"], cf["maxd"], cf["mind"], cf["desc"], cf["nobs"],
                             cf["in_kf"], cf["kf_mp_obs"], 1.0, 1.0, 0.0, 0.0, cf["mbf"], 3.0)
        return _flat((t, [f[k] for k in ("nfused", "kf_state", "pt_obs", "pt_repl", "kfmp_repl")]))
    return once


def _loop_body(scene, oracle):
    import test_reference_matcher_cpu as T
    from ar_orbslam2_b200.tracking import Frame
    sb = oracle.ShimBackend()
    bounds = (0.0, float(W), 0.0, float(H))
    mk = lambda which, kfv: Frame(scene["k" + which], scene["d" + which], scene["scale"], bounds, backend=sb, keyframe_view=kfv)
    ka, kb = scene["ka"], scene["kb"]
    kf_off, kf_idx, f_off, f_idx = M.bow_groups(ka, kb, 7, 4)
    kv, fv = (np.arange(len(ka)) % 7 != 0).astype(np.uint8), (np.arange(len(kb)) % 9 != 0).astype(np.uint8)
    b1, b2, s1, s2, kp = mk("a", False), mk("b", False), mk("a", True), mk("b", True), mk("a", False)
    cs, cp = M.sim3_case(scene), T.sim3_projection_case(scene)
    I, z = np.eye(3, dtype=np.float32), np.zeros(3, np.float32)

    def once():
        bow = oracle.refm_search_by_bow(b1, b2, kf_off, kf_idx, f_off, f_idx, kv, fv, True, 0.7, True)
        sim = oracle.refm_search_by_sim3(s1, s2, 1.0, I, cs["t12"], I, z, I, z, cs["has1"], cs["world1"], cs["maxd1"], cs["mind1"], cs["desc1"],
                                         cs["has2"], cs["world2"], cs["maxd2"], cs["mind2"], cs["desc2"], cs["pre12"], 1.0, 1.0, 0.0, 0.0, 7.5)
        prj = oracle.refm_search_by_projection_sim3(kp, cp["S"], cp["xyz"], cp["nrm"], cp["usable"], cp["maxd"], cp["mind"], cp["claimed"], T.FX, T.FY,
                                                    T.CX, T.CY, 10, 0.9)
        return _flat((bow, sim, prj["nmatches"], prj["assigned"]))
    return once


def _extract_body(image):
    def make(scene, oracle):
        from ar_orbslam2_b200.extractor import ORBextractor
        ex = ORBextractor(2000, 1.2, 8, 20, 7, max_width=image.shape[1], max_height=image.shape[0])
        return lambda: _flat(ex(image))
    return make


def test_matchers_and_extractors_from_concurrent_host_threads(scene, oracle, gpu_lib):
    if not (oracle.ref_matcher_available() and os.path.exists(oracle.REF_SHIM_PATH)):
        pytest.skip("oracle/_ref shim library not built")
    left, right = synth_stereo_pair(1241, 376, 77)
    bodies = {"A tracking": _tracking_body, "B local mapping": _mapping_body, "C loop closing": _loop_body,
              "D extract left": _extract_body(left), "E extract right": _extract_body(right)}
    # each body alone, on the main thread: the results every concurrent iteration has to reproduce
    want = {name: make(scene, oracle)() for name, make in bodies.items()}
    assert all(len(v) > 1000 for v in want.values())
    failures, start = [], threading.Barrier(len(bodies))

    def worker(name, make):
        try:
            once = make(scene, oracle)      # handles, frames and caches of this thread are created on this thread
            start.wait()
            for it in range(ITERATIONS):
                if once() != want[name]:
                    failures.append(f"{name}: iteration {it} differs from the serial result")
                    return
        except Exception as e:              # noqa: BLE001 -- reported through the list, the thread must not die silently
            failures.append(f"{name}: {type(e).__name__}: {e}")
            start.abort()

    threads = [threading.Thread(target=worker, args=(n, m), name=n) for n, m in bodies.items()]
    for t in threads:
        t.start()
    for t in threads:
        t.join(600)
    assert not any(t.is_alive() for t in threads), "a worker thread did not finish"
    assert not failures, failures
