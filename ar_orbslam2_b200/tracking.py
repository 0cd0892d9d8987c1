"""Host-side mirror of the tracking-time entry points of ORB_SLAM2::ORBmatcher (ref:
ORB_SLAM2/include/ORBmatcher.h:42-94) and of the Frame members they read, over liborbgpu.

`Frame` holds what the matchers use of ORB_SLAM2::Frame (mvKeysUn, mDescriptors, mvuRight, the 64x48 grid);
`ORBmatcher` keeps the reference's method names.  MapPoint / KeyFrame object graphs stay with the caller:
the methods take the flat arrays the C++ shim gathers under the reference's getters.

The same classes drive the CPU oracle when constructed with `backend=oracle_module` (tests only)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import (KEYPOINT_DTYPE, OrbFrameDesc, OrbFrameSource, OrbLastPoints, OrbReprojGate, OrbTrackPoints, OrbWindowQueries, optr,
                   ptr)

ORB_SELECT_BEST, ORB_SELECT_RATIO_LEVEL = 0, 1


class _GpuBackend:
    prefix = "orb_"

    def __init__(self):
        self.L = _lib.lib()

    def check(self, rc, what):
        _lib.check(rc, what)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def _i32(a):
    return None if a is None else np.ascontiguousarray(a, np.int32)


def _u8(a):
    return None if a is None else np.ascontiguousarray(a, np.uint8)


class Frame:
    """The matcher-facing part of ORB_SLAM2::Frame (ref Frame.h:39-40,161-163; Frame.cc:235-250,332-397)."""

    def __init__(self, keys_un, descriptors, scale_factors, bounds, u_right=None, backend=None, keyframe_view=False):
        self.B = backend or _GpuBackend()
        self.keys = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
        self.desc = _u8(descriptors).reshape(-1, 32)
        self.scale = _f32(scale_factors)
        self.u_right = _f32(u_right)
        self.N = len(self.keys)
        self.bounds = tuple(float(b) for b in bounds)   # mnMinX, mnMaxX, mnMinY, mnMaxY
        d = OrbFrameDesc(self.N, ptr(self.keys), ptr(self.desc), optr(self.u_right), *self.bounds, len(self.scale), ptr(self.scale),
                         1 if keyframe_view else 0)
        self._h = C.c_void_p()
        self.B.check(getattr(self.B.L, self.B.prefix + "frame_create")(C.byref(d), C.byref(self._h)), "frame_create")
        _lib.register(self)

    @classmethod
    def from_extractor(cls, ex, slot, bounds, camera=None, dist_coef=None, u_right=None, u_right_device=None, keyframe_view=False):
        """The Frame constructor's tail (ref Frame.cc:196-232: UndistortKeyPoints + AssignFeaturesToGrid) on the device, from the
        results the extractor's last call left in HBM (orb_frame_from_extractor): no key point or descriptor crosses the host.
        camera = (fx, fy, cx, cy) and dist_coef only matter for a distorted camera; u_right: host array, or u_right_device: device
        pointer (int) of mvuRight."""
        self = cls.__new__(cls)
        self.B = _GpuBackend()
        s = OrbFrameSource()
        s.min_x, s.max_x, s.min_y, s.max_y = [float(b) for b in bounds]
        if camera is not None:
            s.fx, s.fy, s.cx, s.cy = [float(np.float32(c)) for c in camera]
        d = np.zeros(5, np.float32) if dist_coef is None else np.ascontiguousarray(dist_coef, np.float32)
        s.ndist = 0 if dist_coef is None else len(d)
        for i in range(min(len(d), 5)):
            s.dist_coef[i] = float(d[i])
        self.u_right = _f32(u_right)
        if self.u_right is not None:
            s.u_right, s.u_right_loc = self.u_right.ctypes.data, 0
        elif u_right_device:
            s.u_right, s.u_right_loc = int(u_right_device), 1
        s.keyframe_view = 1 if keyframe_view else 0
        self._h = C.c_void_p()
        self.B.check(self.B.L.orb_frame_from_extractor(ex._h, int(slot), C.byref(s), C.byref(self._h)), "orb_frame_from_extractor")
        self.N = int(self.B.L.orb_frame_size(self._h))
        self.scale = ex.GetScaleFactors()
        self.bounds = tuple(float(b) for b in bounds)
        self.keys = self.desc = None
        _lib.register(self)
        return self

    def sync(self):
        self.B.check(self.B.L.orb_frame_sync(self._h), "orb_frame_sync")

    def last_stats(self):
        """(candidate pairs whose distance was computed, claim-resolution sweeps) of the last matcher call on this frame."""
        c, s = C.c_int64(), C.c_int32()
        self.B.check(self.B.L.orb_frame_last_stats(self._h, C.byref(c), C.byref(s)), "orb_frame_last_stats")
        return c.value, s.value

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            fn = getattr(self.B.L, self.B.prefix + "frame_destroy")
            fn.argtypes = [C.c_void_p]
            fn.restype = None
            fn(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def GetFeaturesInArea(self, x, y, r, minLevel=-1, maxLevel=-1):
        out = np.zeros(max(self.N, 1), np.int32)
        n = C.c_int()
        self.B.check(getattr(self.B.L, self.B.prefix + "frame_features_in_area")(
            self._h, C.c_float(x), C.c_float(y), C.c_float(r), int(minLevel), int(maxLevel), ptr(out), len(out), C.byref(n)),
            "frame_features_in_area")
        return out[: n.value].copy()


class ORBmatcher:
    """ref ORBmatcher(float nnratio=0.6, bool checkOri=true), ORBmatcher.h:37"""
    TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30

    def __init__(self, nnratio=0.6, checkOri=True, backend=None):
        self.mfNNratio = float(np.float32(nnratio))
        self.mbCheckOrientation = bool(checkOri)
        self.B = backend or _GpuBackend()

    def _fn(self, name):
        return getattr(self.B.L, self.B.prefix + name)

    def SearchForInitialization(self, F1: Frame, F2: Frame, vbPrevMatched: np.ndarray, windowSize=10):
        """ref ORBmatcher.cc:405-523.  Returns (nmatches, vnMatches12); vbPrevMatched [n1,2] is updated in place."""
        assert vbPrevMatched.dtype == np.float32 and vbPrevMatched.flags.c_contiguous and vbPrevMatched.shape == (F1.N, 2)
        m12 = np.full(F1.N, -1, np.int32)
        nm = C.c_int()
        self.B.check(self._fn("search_for_initialization")(F1._h, F2._h, ptr(vbPrevMatched), ptr(m12), int(windowSize),
                                                           C.c_float(self.mfNNratio), int(self.mbCheckOrientation), C.byref(nm)),
                     "search_for_initialization")
        return nm.value, m12

    def SearchByProjectionPoints(self, F: Frame, in_view, proj_x, proj_y, proj_xr, level, view_cos, descriptors, th=3.0,
                                 claimed=None, obs_positive=None):
        """ref SearchByProjection(Frame&, const vector<MapPoint*>&, th), ORBmatcher.cc:45-129.
        Returns (nmatches, assigned[F.N]) with assigned[i] = index of the map point put in F.mvpMapPoints[i] or -1."""
        a = [_u8(in_view), _f32(proj_x), _f32(proj_y), _f32(proj_xr), _i32(level), _f32(view_cos), _u8(descriptors), _u8(obs_positive)]
        P = OrbTrackPoints(len(a[0]), *[optr(x) for x in a])
        assigned = np.full(F.N, -1, np.int32)
        nm = C.c_int()
        cl = _u8(claimed)
        self.B.check(self._fn("search_by_projection_points")(F._h, C.byref(P), optr(cl), C.c_float(th), C.c_float(self.mfNNratio),
                                                             ptr(assigned), C.byref(nm)), "search_by_projection_points")
        return nm.value, assigned

    def SearchByProjectionLast(self, Cur: Frame, valid, u, v, invz, octave, angle, descriptors, th, forward=False, backward=False,
                               mbf=0.0, claimed=None, obs_positive=None, raw=False):
        """ref SearchByProjection(Frame &Cur, const Frame &Last, th, bMono), ORBmatcher.cc:1331-1473 (after the projection)."""
        a = [_u8(valid), _f32(u), _f32(v), _f32(invz), _i32(octave), _f32(angle), _u8(descriptors), _u8(obs_positive)]
        P = OrbLastPoints(len(a[0]), *[optr(x) for x in a])
        assigned = np.full(Cur.N, -1, np.int32)
        nm = C.c_int()
        cl = _u8(claimed)
        self.B.check(self._fn("search_by_projection_last")(Cur._h, C.byref(P), optr(cl), C.c_float(th), int(forward), int(backward),
                                                           C.c_float(mbf), int(self.mbCheckOrientation), ptr(assigned), C.byref(nm)),
                     "search_by_projection_last")
        return nm.value, self._fold(assigned, raw)

    @staticmethod
    def _fold(assigned, raw):
        """The C ABI reports -2 where the rotation filter reset an assignment to NULL and -1 where the feature was never
        written; callers that only want the final holder get both as -1."""
        return assigned if raw else np.where(assigned < 0, -1, assigned).astype(np.int32)

    def SearchWindowed(self, F: Frame, x, y, radius, min_level, max_level, descriptors, select, th_dist, valid=None, strict_levels=False,
                       stereo_ref=None, stereo_tol=None, angle=None, claims=None, claimed=None, check_orientation=None, raw=False):
        """The generic projection-matcher loop (orb_search_windowed): SearchByProjection(Frame,KeyFrame,...) :1475-1602,
        SearchByProjection(KeyFrame,Scw,...) :290-403, Fuse :828-1103 and SearchBySim3 :1105-1329 map onto it."""
        a = dict(valid=_u8(valid), x=_f32(x), y=_f32(y), radius=_f32(radius), min_level=_i32(min_level), max_level=_i32(max_level),
                 stereo_ref=_f32(stereo_ref), stereo_tol=_f32(stereo_tol), descriptors=_u8(descriptors), angle=_f32(angle), claims=_u8(claims))
        Q = OrbWindowQueries(len(a["x"]), optr(a["valid"]), optr(a["x"]), optr(a["y"]), optr(a["radius"]), optr(a["min_level"]),
                             optr(a["max_level"]), int(strict_levels), optr(a["stereo_ref"]), optr(a["stereo_tol"]),
                             optr(a["descriptors"]), optr(a["angle"]), optr(a["claims"]))
        assigned = np.full(F.N, -1, np.int32)
        nm = C.c_int()
        cl = _u8(claimed)
        ori = self.mbCheckOrientation if check_orientation is None else check_orientation
        self.B.check(self._fn("search_windowed")(F._h, C.byref(Q), optr(cl), int(select), int(th_dist), C.c_float(self.mfNNratio),
                                                 int(ori), ptr(assigned), C.byref(nm)), "search_windowed")
        return nm.value, self._fold(assigned, raw)

    def SearchByBoW(self, KF: Frame, F: Frame, kf_off, kf_idx, f_off, f_idx, kf_valid, f_valid=None, strict_low=False):
        """ref SearchByBoW(KeyFrame*, Frame&, ...) ORBmatcher.cc:159-288 (strict_low/f_valid: the KF-KF form :525-658).
        The FeatureVector merge-join is given as CSR groups over the shared vocabulary nodes."""
        ko, ki, fo, fi = _i32(kf_off), _i32(kf_idx), _i32(f_off), _i32(f_idx)
        kv, fv = _u8(kf_valid), _u8(f_valid)
        matches = np.full(F.N, -1, np.int32)
        nm = C.c_int()
        self.B.check(self._fn("search_by_bow")(KF._h, F._h, len(ko) - 1, ptr(ko), ptr(ki), ptr(fo), ptr(fi), ptr(kv), optr(fv),
                                               int(strict_low), C.c_float(self.mfNNratio), int(self.mbCheckOrientation), ptr(matches),
                                               C.byref(nm)), "search_by_bow")
        return nm.value, matches

    def SearchWindowBest(self, F: Frame, x, y, radius, min_level, max_level, descriptors, valid=None, strict_levels=True,
                         gate_ur=None, inv_level_sigma2=None, chi2_mono=5.99, chi2_stereo=7.8):
        """The candidate loop of Fuse (ORBmatcher.cc:895-951, 1052-1072) and SearchBySim3 (:1191-1213, 1271-1293): independent
        queries, closest descriptor of the window.  gate_ur + inv_level_sigma2 switch on Fuse's reprojection-error gate
        (:917-945).  Returns (best_idx[n], best_dist[n]); the caller applies TH_LOW / TH_HIGH and the map bookkeeping."""
        a = dict(valid=_u8(valid), x=_f32(x), y=_f32(y), radius=_f32(radius), min_level=_i32(min_level), max_level=_i32(max_level),
                 descriptors=_u8(descriptors))
        n = len(a["x"])
        Q = OrbWindowQueries(n, optr(a["valid"]), optr(a["x"]), optr(a["y"]), optr(a["radius"]), optr(a["min_level"]),
                             optr(a["max_level"]), int(strict_levels), None, None, optr(a["descriptors"]), None, None)
        best_idx, best_dist = np.full(n, -1, np.int32), np.zeros(n, np.int32)
        gur, inv = _f32(gate_ur), _f32(inv_level_sigma2)
        G = OrbReprojGate(optr(gur), optr(inv), float(chi2_mono), float(chi2_stereo)) if inv is not None else None
        self.B.check(self._fn("search_window_best")(F._h, C.byref(Q), C.byref(G) if G is not None else None, ptr(best_idx),
                                                    ptr(best_dist)), "search_window_best")
        return best_idx, best_dist

    def SearchForTriangulation(self, KF1: Frame, KF2: Frame, off1, idx1, off2, idx2, usable1, usable2, F12, ex, ey, level_sigma2,
                               bOnlyStereo=False):
        """ref SearchForTriangulation, ORBmatcher.cc:660-826 (+ CheckDistEpipolarLine :140-157).  Returns (nmatches, vMatches12)."""
        o1, i1, o2, i2 = _i32(off1), _i32(idx1), _i32(off2), _i32(idx2)
        u1, u2, F, s2 = _u8(usable1), _u8(usable2), _f32(F12).reshape(9), _f32(level_sigma2)
        m12 = np.full(KF1.N, -1, np.int32)
        nm = C.c_int()
        self.B.check(self._fn("search_for_triangulation")(KF1._h, KF2._h, len(o1) - 1, ptr(o1), ptr(i1), ptr(o2), ptr(i2), ptr(u1), ptr(u2),
                                                          ptr(F), C.c_float(ex), C.c_float(ey), ptr(s2), int(bOnlyStereo),
                                                          int(self.mbCheckOrientation), ptr(m12), C.byref(nm)), "search_for_triangulation")
        return nm.value, m12


def rot_hist_filter(angle_a, angle_b) -> np.ndarray:
    """Rotation-consistency filter (30-bin histogram with the reference's 1/HISTO_LENGTH factor + ComputeThreeMaxima)."""
    a, b = _f32(angle_a), _f32(angle_b)
    keep = np.zeros(a.size, np.uint8)
    _lib.check(_lib.lib().orb_rot_hist_filter(ptr(a), ptr(b), a.size, ptr(keep)), "orb_rot_hist_filter")
    return keep.astype(bool)


def UndistortKeyPoints(keys: np.ndarray, fx: float, fy: float, cx: float, cy: float, dist_coef) -> np.ndarray:
    """Frame::UndistortKeyPoints (ref Frame.cc:409-439) on the GPU: keys is a KEYPOINT_DTYPE array (mvKeys); returns mvKeysUn."""
    import ctypes as C
    k = np.ascontiguousarray(keys)
    d = np.ascontiguousarray(dist_coef, np.float32)
    out = np.zeros_like(k)
    _lib.check(_lib.lib().orb_undistort_keypoints(ptr(k), k.shape[0], C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), ptr(d),
                                                  d.shape[0], ptr(out), 0), "orb_undistort_keypoints")
    return out


def ComputeImageBounds(width: int, height: int, fx: float, fy: float, cx: float, cy: float, dist_coef):
    """Frame::ComputeImageBounds (ref Frame.cc:441-468): (mnMinX, mnMaxX, mnMinY, mnMaxY); the corner undistortion runs on the GPU."""
    import ctypes as C
    d = np.ascontiguousarray(dist_coef, np.float32)
    if d[0] == 0.0:
        return np.float32(0), np.float32(width), np.float32(0), np.float32(height)
    m = np.array([[0, 0], [width, 0], [0, height], [width, height]], np.float32)
    o = np.zeros_like(m)
    _lib.check(_lib.lib().orb_undistort_points(ptr(m), 4, C.c_float(fx), C.c_float(fy), C.c_float(cx), C.c_float(cy), ptr(d), d.shape[0],
                                               ptr(o), 0), "orb_undistort_points")
    return min(o[0, 0], o[2, 0]), max(o[1, 0], o[3, 0]), min(o[0, 1], o[1, 1]), max(o[2, 1], o[3, 1])


def isInFrustum(cam: dict, world_pos, normal, max_distance, min_distance, viewing_cos_limit: float = 0.5, consider=None):
    """Frame::isInFrustum (ref Frame.cc:274-330) for all local map points of a frame on the GPU (the loop of
    Tracking::SearchLocalPoints, ref Tracking.cc:1527-1539).  cam: dict with Rcw (3x3), tcw, Ow, fx, fy, cx, cy, min_x, max_x,
    min_y, max_y, mbf, log_scale_factor.  Returns (in_view, proj_x, proj_y, proj_xr, level, view_cos) = the MapPoint tracking
    fields, ready for ORBmatcher.SearchByProjectionPoints."""
    from ._abi import frustum_args, frustum_outputs
    c, p, keep = frustum_args(cam, world_pos, normal, max_distance, min_distance, consider)
    o = frustum_outputs(p.n)
    _lib.check(_lib.lib().orb_is_in_frustum(C.byref(c), C.byref(p), C.c_float(viewing_cos_limit), 0, *[ptr(a) for a in o]),
               "orb_is_in_frustum")
    return o


def ProjectMapPoints(proj: dict, world_pos, normal, max_distance, min_distance, consider=None):
    """The projection pre-loop of SearchByProjection(Frame&, KeyFrame*) / SearchByProjection(KeyFrame*, Scw) / Fuse / SearchBySim3
    (ref ORBmatcher.cc:1491-1531, :311-360, :846-893, :1004-1048, :1150-1226) for all map points at once on the GPU.  proj: dict
    with R1, t1 (R2, t2, Ow where used), fx, fy, cx, cy, bf, the bounds and the rule switches of include/orbgpu.h: orb_projection;
    max_distance / min_distance = MapPoint::GetMaxDistanceInvariance / GetMinDistanceInvariance.  Returns (ok, u, v, u_right, dist);
    the caller applies MapPoint::PredictScale to dist."""
    from ._abi import projection_args, projection_outputs
    c, p, keep = projection_args(proj, world_pos, normal, max_distance, min_distance, consider)
    o = projection_outputs(p.n)
    _lib.check(_lib.lib().orb_project_points(C.byref(c), C.byref(p), 0, *[ptr(a) for a in o]), "orb_project_points")
    return o
