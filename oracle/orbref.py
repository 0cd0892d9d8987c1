"""ORACLE -- TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liborbref.so.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from ar_orbslam2_b200._abi import CANDIDATE_DTYPE, KEYPOINT_DTYPE, OrbParams, ptr

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liborbref.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
    return _LIB


def _check(rc, what):
    if rc != 0:
        raise RuntimeError(f"orbref {what} failed: {rc}")


class RefExtractor:
    """CPU restatement of ORB_SLAM2::ORBextractor (ORB_SLAM2/src/ORBextractor.cc)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, blur_taps=0):
        self.params = OrbParams(nfeatures, scale_factor, nlevels, ini_th, min_th, 0, 0, 1, -1, blur_taps)
        self.h = C.c_void_p()
        _check(lib().orbref_extractor_create(C.byref(self.params), C.byref(self.h)), "create")
        self.nlevels = nlevels

    def __del__(self):
        if getattr(self, "h", None):
            lib().orbref_extractor_destroy(self.h)
            self.h = None

    def set_trig_mode(self, mode):
        lib().orbref_extractor_set_trig_mode(self.h, mode)

    def tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.zeros(n, np.float32) for _ in range(4))
        per = np.zeros(n, np.int32)
        umax = np.zeros(16, np.int32)
        _check(lib().orbref_extractor_tables(self.h, ptr(sc), ptr(isc), ptr(s2), ptr(is2), ptr(per), ptr(umax)), "tables")
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, per_level=per, umax=umax)

    def extract(self, img: np.ndarray, capacity=None, check=True):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape if img.ndim == 2 else (0, 0)
        cap = capacity or (self.params.nfeatures + 64 * self.nlevels + 4096)
        kps = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        rc = lib().orbref_extract(self.h, ptr(img), w, h, img.strides[0] if img.ndim == 2 and h else 0,
                                  ptr(kps), ptr(desc), cap, C.byref(n))
        if check:
            _check(rc, "extract")
            return kps[: n.value].copy(), desc[: n.value].copy()
        return rc, kps[: n.value].copy(), desc[: n.value].copy()

    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _check(lib().orbref_level_size(self.h, level, C.byref(w), C.byref(h)), "level_size")
        return w.value, h.value

    def pyramid_level(self, level, with_border=False):
        w, h = self.level_size(level)
        if with_border:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        _check(lib().orbref_pyramid_download(self.h, level, int(with_border), ptr(out), w), "pyramid")
        return out

    def candidates(self, level):
        cap = 1 << 20
        out = np.zeros(cap, CANDIDATE_DTYPE)
        n = C.c_int()
        _check(lib().orbref_debug_candidates(self.h, level, ptr(out), cap, C.byref(n)), "candidates")
        return out[: n.value].copy()

    def blurred_level(self, level):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        _check(lib().orbref_debug_blur_level(self.h, level, ptr(out), w), "blur")
        return out


def resize_linear(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orbref_resize_linear(ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), dw, dh, dw)
    return dst


def gaussian_blur7(src: np.ndarray, taps=0) -> np.ndarray:
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    lib().orbref_gaussian_blur7(ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), dst.strides[0], taps)
    return dst


def fast(img: np.ndarray, threshold: int) -> np.ndarray:
    img = np.ascontiguousarray(img, np.uint8)
    cap = img.size // 4 + 16
    out = np.zeros(cap, CANDIDATE_DTYPE)
    n = C.c_int()
    _check(lib().orbref_fast(ptr(img), img.shape[1], img.shape[0], img.strides[0], threshold, ptr(out), cap, C.byref(n)), "fast")
    return out[: n.value].copy()


def fast_atan2(y: np.ndarray, x: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(y, np.float32)
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(y)
    lib().orbref_fast_atan2(ptr(y), ptr(x), y.size, ptr(out))
    return out


def sincos(angle_rad: np.ndarray, mode=0):
    a = np.ascontiguousarray(angle_rad, np.float32)
    s = np.zeros_like(a)
    c = np.zeros_like(a)
    lib().orbref_sincos(ptr(a), a.size, ptr(s), ptr(c), mode)
    return s, c


def make_border(interior: np.ndarray, border=19) -> np.ndarray:
    h, w = interior.shape
    plane = np.zeros((h + 2 * border, w + 2 * border), np.uint8)
    plane[border:border + h, border:border + w] = interior
    lib().orbref_make_border(ptr(plane), w, h, plane.strides[0], border)
    return plane


def stereo_match(left: RefExtractor, right: RefExtractor, keys_l, desc_l, keys_r, desc_r, mbf, mb):
    """Frame::ComputeStereoMatches on the pyramids held by the two oracle extractors.
    Returns (rc, uRight, depth, sad, n_matched)."""
    n_l, n_r = len(keys_l), len(keys_r)
    kl, kr = np.ascontiguousarray(keys_l), np.ascontiguousarray(keys_r)
    dl, dr = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    ur = np.zeros(n_l, np.float32)
    dp = np.zeros(n_l, np.float32)
    sad = np.zeros(n_l, np.int32)
    nm = C.c_int()
    rc = lib().orbref_stereo_match(left.h, right.h, ptr(kl), ptr(dl), n_l, ptr(kr), ptr(dr), n_r, C.c_float(mbf), C.c_float(mb),
                                   ptr(ur), ptr(dp), ptr(sad), C.byref(nm))
    return rc, ur, dp, sad, nm.value


def descriptor_distance(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.zeros(a.shape[0], np.int32)
    lib().orbref_descriptor_distance(ptr(a), ptr(b), a.shape[0], ptr(out))
    return out


def hamming_top2(queries, db, idx_base=0, fast=False, threads=1):
    q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
    d = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
    nq = q.shape[0]
    d1, d2 = np.zeros(nq, np.int32), np.zeros(nq, np.int32)
    i1, i2 = np.zeros(nq, np.int64), np.zeros(nq, np.int64)
    lib().orbref_hamming_top2(ptr(q), nq, ptr(d), C.c_int64(d.shape[0]), C.c_int64(idx_base), ptr(d1), ptr(i1), ptr(d2), ptr(i2),
                              int(fast), threads)
    return d1, i1, d2, i2


def db_keyframe_counts(queries, db, kf_rows, th_low=50, ratio=0.75, threads=1):
    """Per keyframe (kf_rows consecutive rows of db): queries accepted by ORBmatcher.cc:216-231's rule inside that keyframe."""
    q = np.ascontiguousarray(queries, np.uint8).reshape(-1, 32)
    d = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
    counts = np.zeros(len(d) // kf_rows, np.int32)
    _check(lib().orbref_db_keyframe_counts(ptr(q), len(q), ptr(d), C.c_int64(len(d)), kf_rows, th_low, C.c_float(ratio), ptr(counts),
                                           threads), "db_keyframe_counts")
    return counts


def rot_hist_filter(angle_a, angle_b):
    a = np.ascontiguousarray(angle_a, np.float32)
    b = np.ascontiguousarray(angle_b, np.float32)
    keep = np.zeros(a.size, np.uint8)
    lib().orbref_rot_hist_filter(ptr(a), ptr(b), a.size, ptr(keep))
    return keep.astype(bool)


class OracleBackend:
    """Lets ar_orbslam2_b200.tracking.Frame / ORBmatcher drive the oracle's flat-array matchers (tests only)."""
    prefix = "orbref_"

    def __init__(self):
        self.L = lib()

    def check(self, rc, what):
        _check(rc, what)


# ---- oracle/_ref: the reference's own ORBextractor.cc compiled against the stub OpenCV (oracle/Makefile.ref) ----
_REF = None
REF_LIB_PATH = os.path.join(_HERE, "_ref", "liborbslam_ref.so")


def ref_available() -> bool:
    if os.path.exists(REF_LIB_PATH):
        return True
    if os.path.isdir("/root/reference/ORB_SLAM2/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-f", "Makefile.ref"])
        return True
    return False


def ref_lib():
    global _REF
    if _REF is None:
        _REF = C.CDLL(REF_LIB_PATH)
    return _REF


class ReferenceExtractor:
    """ORB_SLAM2::ORBextractor from the reference's own source (control flow verbatim), OpenCV primitives = the
    cv2-pinned models, monotonic allocator so its pointer tie-break equals creation order."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.h = C.c_void_p()
        ref_lib().ref_extractor_create(nfeatures, C.c_float(scale_factor), nlevels, ini_th, min_th, C.byref(self.h))
        self.cap = nfeatures + 64 * nlevels + 4096
        self.nlevels = nlevels

    def __del__(self):
        if getattr(self, "h", None):
            ref_lib().ref_extractor_destroy(self.h)
            self.h = None

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        kps = np.zeros(self.cap, KEYPOINT_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int()
        _check(ref_lib().ref_extract(self.h, ptr(img), img.shape[1], img.shape[0], img.strides[0], ptr(kps), ptr(desc), self.cap,
                                     C.byref(n)), "ref_extract")
        return kps[: n.value].copy(), desc[: n.value].copy()

    def pyramid_level(self, level):
        w, h = C.c_int(), C.c_int()
        ref_lib().ref_pyramid_level(self.h, level, None, 0, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        ref_lib().ref_pyramid_level(self.h, level, ptr(out), w.value, C.byref(w), C.byref(h))
        return out


# ---- oracle/_ref matcher: the reference's own ORBmatcher.cc compiled against stub Frame/KeyFrame/MapPoint ----
REF_MATCHER_PATH = os.path.join(_HERE, "_ref", "liborbmatcher_ref.so")
_REFM = None


def ref_matcher_available() -> bool:
    if os.path.exists(REF_MATCHER_PATH):
        return True
    if os.path.isdir("/root/reference/ORB_SLAM2/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-f", "Makefile.ref"])
        return True
    return False


class ReferenceSourceBackend:
    """Drives ar_orbslam2_b200.tracking.Frame / ORBmatcher through the reference's compiled ORBmatcher.cc for the entry
    points whose flat signature is identical (frame_create, features_in_area, search_for_initialization,
    search_by_projection_points)."""
    prefix = "refm_"

    def __init__(self):
        global _REFM
        if _REFM is None:
            _REFM = C.CDLL(REF_MATCHER_PATH)
        self.L = _REFM

    def check(self, rc, what):
        _check(rc, what)


def refm_search_by_bow(kf, f, kf_off, kf_idx, f_off, f_idx, kf_valid, f_valid, kfkf, nnratio, check_ori):
    """kf, f: tracking.Frame objects created with ReferenceSourceBackend."""
    L = kf.B.L
    ko, ki, fo, fi = (np.ascontiguousarray(a, np.int32) for a in (kf_off, kf_idx, f_off, f_idx))
    kv = np.ascontiguousarray(kf_valid, np.uint8)
    matches = np.full(f.N, -1, np.int32)
    nm = C.c_int()
    if kfkf:
        fv = None if f_valid is None else np.ascontiguousarray(f_valid, np.uint8)
        L.refm_search_by_bow_kfkf(kf._h, f._h, len(ko) - 1, ptr(ko), ptr(ki), ptr(fo), ptr(fi), ptr(kv),
                                  None if fv is None else ptr(fv), C.c_float(nnratio), int(check_ori), ptr(matches), C.byref(nm))
    else:
        L.refm_search_by_bow(kf._h, f._h, len(ko) - 1, ptr(ko), ptr(ki), ptr(fo), ptr(fi), ptr(kv), C.c_float(nnratio), int(check_ori),
                             ptr(matches), C.byref(nm))
    return nm.value, matches


def refm_search_by_projection_last(cur, last, world_xyz, has_point, obs_positive, claimed, Tcw_cur, Tcw_last, fx, fy, cx, cy, mbf, mb, th,
                                   mono, nnratio, check_ori):
    """Reference SearchByProjection(Cur, Last, th, bMono).  Returns dict with assigned, nmatches and the projection the
    reference computed (u, v, invz, valid, forward, backward)."""
    L = cur.B.L
    n = last.N
    xyz = np.ascontiguousarray(world_xyz, np.float32)
    hp = np.ascontiguousarray(has_point, np.uint8)
    ob = None if obs_positive is None else np.ascontiguousarray(obs_positive, np.uint8)
    cl = None if claimed is None else np.ascontiguousarray(claimed, np.uint8)
    tc, tl = np.ascontiguousarray(Tcw_cur, np.float32), np.ascontiguousarray(Tcw_last, np.float32)
    assigned = np.full(cur.N, -1, np.int32)
    u, v, iz = (np.zeros(n, np.float32) for _ in range(3))
    valid = np.zeros(n, np.uint8)
    nm, fw, bw = C.c_int(), C.c_int(), C.c_int()
    f = C.c_float
    L.refm_search_by_projection_last(cur._h, last._h, ptr(xyz), ptr(hp), None if ob is None else ptr(ob), None if cl is None else ptr(cl),
                                     ptr(tc), ptr(tl), f(fx), f(fy), f(cx), f(cy), f(mbf), f(mb), f(th), int(mono), f(nnratio), int(check_ori),
                                     ptr(assigned), C.byref(nm), ptr(u), ptr(v), ptr(iz), ptr(valid), C.byref(fw), C.byref(bw))
    return dict(assigned=assigned, nmatches=nm.value, u=u, v=v, invz=iz, valid=valid, forward=bool(fw.value), backward=bool(bw.value))


def refm_descriptor_distance(a, b):
    global _REFM
    if _REFM is None:
        _REFM = C.CDLL(REF_MATCHER_PATH)
    a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
    out = np.zeros(a.shape[0], np.int32)
    _REFM.refm_descriptor_distance(ptr(a), ptr(b), a.shape[0], ptr(out))
    return out


def refm_stereo_match(img_left, img_right, nfeatures, keys_l, desc_l, keys_r, desc_r, mbf, mb, scale=1.2, nlevels=8, ini_th=20, min_th=7,
                      shim=False):
    """The reference's own Frame::ComputeStereoMatches (Frame.cc compiled into oracle/_ref) on the reference's own
    pyramids.  Returns (mvuRight, mvDepth).  shim=True: the same driver lines through liborbmatcher_shim.so, where
    ORBextractor is the product's drop-in class and Frame::ComputeStereoMatches is shim/Frame_stereo.cc (needs a GPU)."""
    global _REFM
    if shim:
        _REFM_saved, _REFM = _REFM, ShimBackend().L
        try:
            return refm_stereo_match(img_left, img_right, nfeatures, keys_l, desc_l, keys_r, desc_r, mbf, mb, scale, nlevels, ini_th, min_th)
        finally:
            _REFM = _REFM_saved
    if _REFM is None:
        _REFM = C.CDLL(REF_MATCHER_PATH)
    il, ir = np.ascontiguousarray(img_left, np.uint8), np.ascontiguousarray(img_right, np.uint8)
    kl, kr = np.ascontiguousarray(keys_l), np.ascontiguousarray(keys_r)
    dl, dr = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    ur, dp = np.zeros(len(kl), np.float32), np.zeros(len(kl), np.float32)
    f = C.c_float
    _check(_REFM.refm_stereo_match(ptr(il), ptr(ir), il.shape[1], il.shape[0], nfeatures, f(scale), nlevels, ini_th, min_th, ptr(kl), ptr(dl),
                                   len(kl), ptr(kr), ptr(dr), len(kr), f(mbf), f(mb), ptr(ur), ptr(dp)), "refm_stereo_match")
    return ur, dp


def refm_search_by_projection_kf(cur, kf, world_xyz, usable, max_dist, min_dist, claimed, Tcw_cur, fx, fy, cx, cy, th, orb_dist, nnratio, check_ori):
    """Reference SearchByProjection(Frame&, KeyFrame*, set<MapPoint*>, th, ORBdist), ORBmatcher.cc:1475-1602."""
    L = cur.B.L
    n = kf.N
    a = [np.ascontiguousarray(world_xyz, np.float32), np.ascontiguousarray(usable, np.uint8), np.ascontiguousarray(max_dist, np.float32),
         np.ascontiguousarray(min_dist, np.float32), np.ascontiguousarray(claimed, np.uint8), np.ascontiguousarray(Tcw_cur, np.float32)]
    assigned = np.full(cur.N, -1, np.int32)
    u, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    lvl, valid = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    nm = C.c_int()
    f = C.c_float
    _check(L.refm_search_by_projection_kf(cur._h, kf._h, *[ptr(x) for x in a], f(fx), f(fy), f(cx), f(cy), f(th), int(orb_dist), f(nnratio),
                                          int(check_ori), ptr(assigned), C.byref(nm), ptr(u), ptr(v), ptr(lvl), ptr(valid)), "refm kf")
    return dict(assigned=assigned, nmatches=nm.value, u=u, v=v, level=lvl, valid=valid)


def refm_search_by_projection_sim3(kf, Scw, world_xyz, normal_xyz, usable, max_dist, min_dist, claimed, fx, fy, cx, cy, th, nnratio):
    """Reference SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th), ORBmatcher.cc:290-403."""
    L = kf.B.L
    xyz = np.ascontiguousarray(world_xyz, np.float32)
    n = xyz.shape[0]
    a = [np.ascontiguousarray(normal_xyz, np.float32), np.ascontiguousarray(usable, np.uint8), np.ascontiguousarray(max_dist, np.float32),
         np.ascontiguousarray(min_dist, np.float32), np.ascontiguousarray(claimed, np.uint8)]
    S = np.ascontiguousarray(Scw, np.float32)
    assigned = np.full(kf.N, -1, np.int32)
    u, v = np.zeros(n, np.float32), np.zeros(n, np.float32)
    lvl, valid = np.zeros(n, np.int32), np.zeros(n, np.uint8)
    nm = C.c_int()
    f = C.c_float
    _check(L.refm_search_by_projection_sim3(kf._h, ptr(S), n, ptr(xyz), *[ptr(x) for x in a], f(fx), f(fy), f(cx), f(cy), int(th), f(nnratio),
                                            ptr(assigned), C.byref(nm), ptr(u), ptr(v), ptr(lvl), ptr(valid)), "refm sim3")
    return dict(assigned=assigned, nmatches=nm.value, u=u, v=v, level=lvl, valid=valid)


REF_SHIM_PATH = os.path.join(_HERE, "_ref", "liborbmatcher_shim.so")
_SHIM = None


class ShimBackend:
    """Same driver as ReferenceSourceBackend, but SearchForInitialization / SearchByProjection x2 / SearchByBoW go through the
    PRODUCT's C++ shim (ar_orbslam2_b200/shim/ORBmatcher_gpu.cc over liborbgpu) compiled against the reference's Frame.h.
    Needs a GPU."""
    prefix = "refm_"

    def __init__(self):
        global _SHIM
        if _SHIM is None:
            _SHIM = C.CDLL(REF_SHIM_PATH)
        self.L = _SHIM

    def check(self, rc, what):
        _check(rc, what)


def _c(a, dt):
    return np.ascontiguousarray(a, dt)


def refm_fuse(kf, variant, Rcw, tcw, Scw, world_xyz, normal_xyz, present, max_dist, min_dist, desc, nobs, in_kf, kf_mp_obs, fx, fy, cx, cy,
              mbf, th):
    """ORBmatcher::Fuse (variant 0: :828-978, variant 1: the Sim3 overload :980-1103) with its map bookkeeping, through the
    library of kf's backend (reference source or the product shim).  Returns the final state as a dict of arrays."""
    L = kf.B.L
    xyz = _c(world_xyz, np.float32)
    n = xyz.shape[0]
    a = [_c(Rcw, np.float32), _c(tcw, np.float32), _c(Scw, np.float32)]
    b = [xyz, _c(normal_xyz, np.float32), _c(present, np.uint8), _c(max_dist, np.float32), _c(min_dist, np.float32), _c(desc, np.uint8),
         _c(nobs, np.int32), _c(in_kf, np.int32), _c(kf_mp_obs, np.int32)]
    kf_state, kfmp_repl = np.zeros(kf.N, np.int32), np.zeros(kf.N, np.int32)
    pt_obs, pt_repl = np.zeros(n, np.int32), np.zeros(n, np.int32)
    nf = C.c_int()
    f = C.c_float
    _check(L.refm_fuse(kf._h, int(variant), *[ptr(x) for x in a], n, *[ptr(x) for x in b], f(fx), f(fy), f(cx), f(cy), f(mbf), f(th),
                       ptr(kf_state), ptr(pt_obs), ptr(pt_repl), ptr(kfmp_repl), C.byref(nf)), "refm_fuse")
    return dict(kf_state=kf_state, pt_obs=pt_obs, pt_repl=pt_repl, kfmp_repl=kfmp_repl, nfused=nf.value)


def refm_search_by_sim3(kf1, kf2, s12, R12, t12, R1w, t1w, R2w, t2w, has1, world1, maxd1, mind1, desc1, has2, world2, maxd2, mind2, desc2,
                        pre12, fx, fy, cx, cy, th):
    """ORBmatcher::SearchBySim3, ORBmatcher.cc:1105-1329, through the library of kf1's backend.  Returns (nFound, matches12)."""
    L = kf1.B.L
    poses = [_c(x, np.float32) for x in (R12, t12, R1w, t1w, R2w, t2w)]
    s1 = [_c(has1, np.uint8), _c(world1, np.float32), _c(maxd1, np.float32), _c(mind1, np.float32), _c(desc1, np.uint8)]
    s2 = [_c(has2, np.uint8), _c(world2, np.float32), _c(maxd2, np.float32), _c(mind2, np.float32), _c(desc2, np.uint8)]
    pre = _c(pre12, np.int32)
    m12 = np.full(kf1.N, -1, np.int32)
    nf = C.c_int()
    f = C.c_float
    _check(L.refm_search_by_sim3(kf1._h, kf2._h, f(s12), *[ptr(x) for x in poses], *[ptr(x) for x in s1], *[ptr(x) for x in s2], ptr(pre),
                                 f(fx), f(fy), f(cx), f(cy), f(th), ptr(m12), C.byref(nf)), "refm_search_by_sim3")
    return nf.value, m12


# ---- DBoW2 vocabulary: the restated oracle (orbref_bow.cpp) and the reference's own DBoW2 (oracle/_ref/libdbow2_ref.so) ----
class _VocabDesc(C.Structure):
    _fields_ = [("k", C.c_int32), ("L", C.c_int32), ("scoring", C.c_int32), ("weighting", C.c_int32), ("n_nodes", C.c_int32),
                ("parent", C.c_void_p), ("descriptors", C.c_void_p), ("weight", C.c_void_p), ("word_id", C.c_void_p)]


def _bow_outputs(cap):
    return (np.zeros(cap, np.int32), np.zeros(cap, np.float64), C.c_int32(0), np.zeros(cap, np.int32), np.zeros(cap + 1, np.int32),
            np.zeros(cap, np.int32), C.c_int32(0))


def _bow_pack(o):
    bw, bv, nb, fn, fo, ff, nf = o
    g = nf.value
    return bw[: nb.value].copy(), bv[: nb.value].copy(), fn[:g].copy(), fo[: g + 1].copy(), ff[: fo[g]].copy()


class RefVocabulary:
    """CPU restatement of DBoW2::TemplatedVocabulary<FORB> (text loader, descent, BowVector/FeatureVector assembly)."""

    def __init__(self, path=None, arrays=None, scoring=0, weighting=0):
        self.h = C.c_void_p()
        if path is not None:
            _check(lib().orbref_vocab_load_text(str(path).encode(), C.byref(self.h)), "vocab_load_text")
        else:
            a = arrays
            self._keep = [np.ascontiguousarray(a["parent"], np.int32), np.ascontiguousarray(a["descriptors"], np.uint8),
                          np.ascontiguousarray(a["weight"], np.float64), np.ascontiguousarray(a["word_id"], np.int32)]
            d = _VocabDesc(a["k"], a["L"], scoring, weighting, self._keep[0].shape[0], *[x.ctypes.data for x in self._keep])
            _check(lib().orbref_vocab_create(C.byref(d), C.byref(self.h)), "vocab_create")

    def __del__(self):
        if getattr(self, "h", None):
            lib().orbref_vocab_destroy(self.h)
            self.h = None

    def info(self):
        a = np.zeros(6, np.int32)
        lib().orbref_vocab_info(self.h, ptr(a))
        return dict(k=int(a[0]), L=int(a[1]), scoring=int(a[2]), weighting=int(a[3]), nodes=int(a[4]), words=int(a[5]))

    def nodes(self):
        n = self.info()["nodes"]
        parent, word = np.zeros(n, np.int32), np.zeros(n, np.int32)
        desc, weight = np.zeros((n, 32), np.uint8), np.zeros(n, np.float64)
        lib().orbref_vocab_nodes(self.h, ptr(parent), ptr(desc), ptr(weight), ptr(word))
        return parent, desc, weight, word

    def transform_features(self, desc, levelsup=0):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = d.shape[0]
        word, node, w = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
        _check(lib().orbref_vocab_transform_features(self.h, ptr(d), n, levelsup, ptr(word), ptr(node), ptr(w)), "vocab_transform_features")
        return word, node, w

    def transform(self, desc, levelsup=4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = d.shape[0]
        o = _bow_outputs(max(n, 1))
        _check(lib().orbref_vocab_transform(self.h, ptr(d), n, levelsup, max(n, 1), ptr(o[0]), ptr(o[1]), C.byref(o[2]), ptr(o[3]),
                                            ptr(o[4]), ptr(o[5]), C.byref(o[6])), "vocab_transform")
        return _bow_pack(o)


REF_DBOW_PATH = os.path.join(_HERE, "_ref", "libdbow2_ref.so")
_REFB = None


def ref_dbow_available() -> bool:
    if os.path.exists(REF_DBOW_PATH):
        return True
    if os.path.isdir("/root/reference/ORB_SLAM2/Thirdparty/DBoW2"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-f", "Makefile.ref"])
        return True
    return False


class ReferenceVocabulary:
    """The reference's own DBoW2 (TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp) compiled unmodified."""

    def __init__(self, path):
        global _REFB
        if _REFB is None:
            _REFB = C.CDLL(REF_DBOW_PATH)
            _REFB.ref_bow_load_text.restype = C.c_void_p
            _REFB.ref_bow_free.argtypes = [C.c_void_p]
        self.L = _REFB
        self.h = C.c_void_p(self.L.ref_bow_load_text(str(path).encode()))
        if not self.h:
            raise RuntimeError("reference loadFromTextFile returned false")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_bow_free(self.h)
            self.h = None

    def info(self):
        a = np.zeros(6, np.int32)
        self.L.ref_bow_info(self.h, ptr(a))
        return dict(k=int(a[0]), L=int(a[1]), scoring=int(a[2]), weighting=int(a[3]), nodes=int(a[4]), words=int(a[5]))

    def nodes(self):
        n = self.info()["nodes"]
        parent, word, nch = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.int32)
        desc, weight = np.zeros((n, 32), np.uint8), np.zeros(n, np.float64)
        self.L.ref_bow_nodes(self.h, ptr(parent), ptr(desc), ptr(weight), ptr(word), ptr(nch))
        return parent, desc, weight, word

    def transform_features(self, desc, levelsup=0):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = d.shape[0]
        word, node, w = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
        self.L.ref_bow_transform_features(self.h, ptr(d), n, levelsup, ptr(word), ptr(node), ptr(w))
        return word, node, w

    def transform(self, desc, levelsup=4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = d.shape[0]
        o = _bow_outputs(max(n, 1))
        _check(self.L.ref_bow_transform(self.h, ptr(d), n, levelsup, max(n, 1), ptr(o[0]), ptr(o[1]), C.byref(o[2]), ptr(o[3]), ptr(o[4]),
                                        ptr(o[5]), C.byref(o[6])), "ref_bow_transform")
        return _bow_pack(o)


# ---- MapPoint::ComputeDistinctiveDescriptors: the restated oracle and the reference's own MapPoint.cc ----
def distinctive_descriptors(desc, offsets):
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    off = np.ascontiguousarray(offsets, np.int32)
    n = off.shape[0] - 1
    row, med = np.zeros(n, np.int32), np.zeros(n, np.int32)
    _check(lib().orbref_distinctive_descriptors(ptr(d), ptr(off), n, ptr(row), ptr(med)), "distinctive_descriptors")
    return row, med


REF_MAPPOINT_PATH = os.path.join(_HERE, "_ref", "libmappoint_ref.so")
_REFMP = None


def ref_mappoint_available() -> bool:
    if os.path.exists(REF_MAPPOINT_PATH):
        return True
    if os.path.isdir("/root/reference/ORB_SLAM2/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-f", "Makefile.ref"])
        return True
    return False


def ref_distinctive_descriptors(desc, offsets, bad=None):
    """The reference's compiled MapPoint::ComputeDistinctiveDescriptors: returns the chosen descriptor (32 bytes) per map point."""
    global _REFMP
    if _REFMP is None:
        _REFMP = C.CDLL(REF_MAPPOINT_PATH)
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    off = np.ascontiguousarray(offsets, np.int32)
    n = off.shape[0] - 1
    out = np.zeros((n, 32), np.uint8)
    b = None if bad is None else np.ascontiguousarray(bad, np.uint8)
    _check(_REFMP.ref_distinctive_descriptors(ptr(d), ptr(off), n, None if b is None else ptr(b), ptr(out)), "ref_distinctive_descriptors")
    return out


def cvt_color_gray(img, rgb_order=True, variant=0):
    """cv::cvtColor(img, CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY), 8-bit."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w, ch = img.shape
    out = np.zeros((h, w), np.uint8)
    _check(lib().orbref_cvt_color_gray(ptr(img), w, h, img.strides[0], ch, int(rgb_order), variant, ptr(out), w), "cvt_color_gray")
    return out


def undistort_points(xy, fx, fy, cx, cy, dist):
    """cv::undistortPoints(xy, K, dist, None, K) as Frame::UndistortKeyPoints calls it."""
    p = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    d = np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(p)
    _check(lib().orbref_undistort_points(ptr(p), p.shape[0], C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), ptr(d),
                                         d.shape[0], ptr(out)), "undistort_points")
    return out


def search_matches_ratio(train, query, ratio=0.8, max_dist=50):
    """Marker::searchMatches / naive_nn_search2 (restatement; pinned to the reference's compiled Marker::searchMatches by
    tests/test_reference_cpu.py::test_search_matches_ratio_equals_marker_cc through ref_marker_search_matches below)."""
    t = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
    q = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
    idx, dist = np.zeros(len(q), np.int32), np.zeros(len(q), np.int32)
    _check(lib().orbref_search_matches_ratio(ptr(t), len(t), ptr(q), len(q), C.c_double(ratio), max_dist, ptr(idx), ptr(dist)),
           "search_matches_ratio")
    return idx, dist


REF_MARKER_PATH = os.path.join(_HERE, "_ref", "libmarker_ref.so")
_REFMK = None


def ref_marker_available() -> bool:
    if os.path.exists(REF_MARKER_PATH):
        return True
    if os.path.isdir("/root/reference/ORB_SLAM2/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-f", "Makefile.ref"])
        return os.path.exists(REF_MARKER_PATH)
    return False


def ref_marker_search_matches(train, query):
    """The reference's own Marker::searchMatches (src/Marker.cc:314-349, compiled by Makefile.ref into _ref/libmarker_ref.so):
    DMatch list as (queryIdx, trainIdx, distance)."""
    global _REFMK
    if _REFMK is None:
        _REFMK = C.CDLL(REF_MARKER_PATH)
    t = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
    q = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
    cap = max(len(q), 1)
    qi, ti, d = np.zeros(cap, np.int32), np.zeros(cap, np.int32), np.zeros(cap, np.float32)
    n = C.c_int()
    _check(_REFMK.refmk_search_matches(ptr(t), len(t), ptr(q), len(q), ptr(qi), ptr(ti), ptr(d), cap, C.byref(n)), "refmk_search_matches")
    return qi[: n.value].copy(), ti[: n.value].copy(), d[: n.value].copy()


# ---- Frame::isInFrustum: the restated oracle and the reference's compiled Frame.cc ----
def is_in_frustum(cam, world_pos, normal, max_distance, min_distance, limit=0.5, consider=None):
    from ar_orbslam2_b200._abi import frustum_args, frustum_outputs
    c, p, keep = frustum_args(cam, world_pos, normal, max_distance, min_distance, consider)
    o = frustum_outputs(p.n)
    _check(lib().orbref_is_in_frustum(C.byref(c), C.byref(p), C.c_float(limit), *[ptr(a) for a in o]), "is_in_frustum")
    return o


def project_points(proj, world_pos, normal, max_distance, min_distance, consider=None):
    """The projection pre-loops of the mapping / loop matchers (restatement; rules as include/orbgpu.h: orb_projection)."""
    from ar_orbslam2_b200._abi import projection_args, projection_outputs
    c, p, keep = projection_args(proj, world_pos, normal, max_distance, min_distance, consider)
    o = projection_outputs(p.n)
    _check(lib().orbref_project_points(C.byref(c), C.byref(p), *[ptr(a) for a in o]), "project_points")
    return o


def refm_is_in_frustum(frame, Tcw, fx, fy, cx, cy, mbf, log_scale_factor, world_pos, normal, max_distance, min_distance, limit=0.5,
                       consider=None):
    """The reference's own Frame::isInFrustum (compiled Frame.cc).  `frame` = a Frame built with ReferenceSourceBackend.
    Returns (cam15 = mRcw|mtcw|mOw as the reference derived them, outputs)."""
    from ar_orbslam2_b200._abi import frustum_outputs
    L = ReferenceSourceBackend().L
    wp = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
    nr = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    mx, mn = np.ascontiguousarray(max_distance, np.float32), np.ascontiguousarray(min_distance, np.float32)
    co = None if consider is None else np.ascontiguousarray(consider, np.uint8)
    T = np.ascontiguousarray(Tcw, np.float32).reshape(4, 4)
    cam = np.zeros(15, np.float32)
    o = frustum_outputs(wp.shape[0])
    f = C.c_float
    _check(L.refm_is_in_frustum(frame._h, ptr(T), f(fx), f(fy), f(cx), f(cy), f(mbf), f(log_scale_factor), wp.shape[0], ptr(wp), ptr(nr),
                                ptr(mx), ptr(mn), None if co is None else ptr(co), f(limit), ptr(cam), *[ptr(a) for a in o]),
           "refm_is_in_frustum")
    return cam, o


def remap_linear(src, map_x, map_y):
    """cv::remap(src, map_x, map_y, INTER_LINEAR), 8-bit single channel, BORDER_CONSTANT 0."""
    s = np.ascontiguousarray(src, np.uint8)
    mx, my = np.ascontiguousarray(map_x, np.float32), np.ascontiguousarray(map_y, np.float32)
    dh, dw = mx.shape
    out = np.zeros((dh, dw), np.uint8)
    _check(lib().orbref_remap_linear(ptr(s), s.shape[1], s.shape[0], s.strides[0], ptr(mx), ptr(my), dw, ptr(out), dw, dh, dw), "remap_linear")
    return out


def resize_linear_exact(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    """cv::resize(..., INTER_LINEAR_EXACT), 8UC1."""
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros((dh, dw), np.uint8)
    _check(lib().orbref_resize_linear_exact(ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), dw, dh, dw), "resize_exact")
    return dst


def cv_orb(img, nfeatures=500, scale_factor=1.2, nlevels=8, edge_threshold=31, score_type=0, fast_threshold=20):
    """OpenCV's cv::ORB (Harris score) detectAndCompute as Marker.cc:83-84,107-108 calls it."""
    from ar_orbslam2_b200._abi import OrbCvOrbParams
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    p = OrbCvOrbParams(nfeatures, scale_factor, nlevels, edge_threshold, score_type, fast_threshold, w, h, -1)
    cap = nfeatures + 4096
    while True:
        kps = np.zeros(cap, KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        rc = lib().orbref_cvorb_detect_and_compute(C.byref(p), ptr(img), w, h, img.strides[0], ptr(kps), ptr(desc), cap, C.byref(n))
        if rc == -4:
            cap = n.value
            continue
        _check(rc, "cvorb")
        return kps[: n.value].copy(), desc[: n.value].copy()


def cv_orb_level(img, level, blurred=False, nfeatures=500, scale_factor=1.2, nlevels=8, edge_threshold=31, score_type=0,
                 fast_threshold=20):
    from ar_orbslam2_b200._abi import OrbCvOrbParams
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    p = OrbCvOrbParams(nfeatures, scale_factor, nlevels, edge_threshold, score_type, fast_threshold, w, h, -1)
    lw, lh = C.c_int(), C.c_int()
    _check(lib().orbref_cvorb_level(C.byref(p), ptr(img), w, h, img.strides[0], level, int(blurred), None, 0, C.byref(lw), C.byref(lh)), "lvl")
    out = np.zeros((lh.value, lw.value), np.uint8)
    _check(lib().orbref_cvorb_level(C.byref(p), ptr(img), w, h, img.strides[0], level, int(blurred), ptr(out), lw.value, None, None), "lvl")
    return out


def gaussian_blur7_sep32f(src: np.ndarray) -> np.ndarray:
    """GaussianBlur(7x7, sigma 2, REFLECT_101) through OpenCV's float sepFilter2D path (what cv::ORB's layers get)."""
    src = np.ascontiguousarray(src, np.uint8)
    dst = np.zeros_like(src)
    _check(lib().orbref_gaussian_blur7_sep32f(ptr(src), src.shape[1], src.shape[0], src.strides[0], ptr(dst), dst.strides[0]), "blur32f")
    return dst


def retain_best(responses, n_points, depth_limit=-1):
    """KeyPointsFilter::retainBest: surviving original indices in the order std::nth_element + std::partition leave."""
    r = np.ascontiguousarray(responses, np.float32)
    order = np.zeros(max(len(r), 1), np.int32)
    n = C.c_int(0)
    _check(lib().orbref_retain_best(ptr(r), len(r), int(n_points), int(depth_limit), ptr(order), C.byref(n)), "retain_best")
    return order[: n.value].copy()


def refm_stereo_from_rgbd(keys, keys_un, depth_f32, mbf):
    """Frame::ComputeStereoFromRGBD run from the reference's own Frame.cc (oracle/_ref/liborbmatcher_ref.so)."""
    lib_ = ReferenceSourceBackend().L
    k = np.ascontiguousarray(keys, KEYPOINT_DTYPE); ku = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
    d = np.ascontiguousarray(depth_f32, np.float32)
    ur, dp = np.zeros(len(k), np.float32), np.zeros(len(k), np.float32)
    _check(lib_.refm_stereo_from_rgbd(ptr(k), ptr(ku), len(k), ptr(d), d.shape[1], d.shape[0], C.c_float(mbf), ptr(ur), ptr(dp)), "refm rgbd")
    return ur, dp


def stereo_from_rgbd(keys, keys_un, im_depth, mbf, depth_factor=1.0):
    """Restatement of Frame::ComputeStereoFromRGBD (Frame.cc:645-667) + the convertTo of Tracking.cc:553-554 in numpy fp32."""
    k = np.ascontiguousarray(keys, KEYPOINT_DTYPE); ku = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
    d = im_depth
    if d.dtype != np.float32 or np.float32(depth_factor) != np.float32(1):
        d = (d.astype(np.float32) * np.float32(depth_factor)).astype(np.float32)   # convertTo(CV_32F, alpha): (float)src * alpha
    v, u = k["y"].astype(np.int32), k["x"].astype(np.int32)                        # at<float>(float, float): truncation
    dd = d[v, u]
    ok = dd > 0
    ur, dp = np.full(len(k), -1, np.float32), np.full(len(k), -1, np.float32)
    dp[ok] = dd[ok]
    ur[ok] = ku["x"][ok] - (np.float32(mbf) / dd[ok]).astype(np.float32)
    return ur, dp
