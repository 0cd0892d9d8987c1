"""ctypes mirrors of the POD types in include/orbgpu.h (kept field-for-field identical)."""
import ctypes as C

import numpy as np

ORB_OK = 0
ORB_ERR_INVALID = -1
ORB_ERR_CUDA = -2
ORB_ERR_TOO_SMALL = -3
ORB_ERR_CAPACITY = -4
ORB_ERR_EMPTY = -5
ORB_ERR_NCCL = -6
ORB_HOST = 0
ORB_DEVICE = 1
ORB_MAX_LEVELS = 16
EDGE_THRESHOLD = 19


class OrbParams(C.Structure):
    _fields_ = [
        ("nfeatures", C.c_int32),
        ("scale_factor", C.c_float),
        ("nlevels", C.c_int32),
        ("ini_th_fast", C.c_int32),
        ("min_th_fast", C.c_int32),
        ("max_width", C.c_int32),
        ("max_height", C.c_int32),
        ("max_batch", C.c_int32),
        ("device", C.c_int32),
        ("blur_taps", C.c_int32),
    ]


# cv::KeyPoint layout, 28 bytes
KEYPOINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
     ("octave", "<i4"), ("class_id", "<i4")]
)
assert KEYPOINT_DTYPE.itemsize == 28

CANDIDATE_DTYPE = np.dtype([("x", "<i2"), ("y", "<i2"), ("response", "<i4")])
assert CANDIDATE_DTYPE.itemsize == 8


def ptr(a: np.ndarray, ctype=C.c_void_p):
    return a.ctypes.data_as(ctype)


class OrbFrameDesc(C.Structure):
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("descriptors", C.c_void_p), ("u_right", C.c_void_p),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("nlevels", C.c_int32), ("scale_factors", C.c_void_p), ("keyframe_view", C.c_int32)]


class OrbFrameSource(C.Structure):
    _fields_ = [("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("dist_coef", C.c_float * 5),
                ("ndist", C.c_int32), ("u_right", C.c_void_p), ("u_right_loc", C.c_int32), ("keyframe_view", C.c_int32)]


class OrbReprojGate(C.Structure):
    _fields_ = [("ur", C.c_void_p), ("inv_level_sigma2", C.c_void_p), ("chi2_mono", C.c_double), ("chi2_stereo", C.c_double)]


class OrbWindowQueries(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("x", C.c_void_p), ("y", C.c_void_p), ("radius", C.c_void_p),
                ("min_level", C.c_void_p), ("max_level", C.c_void_p), ("strict_levels", C.c_int32),
                ("stereo_ref", C.c_void_p), ("stereo_tol", C.c_void_p), ("descriptors", C.c_void_p), ("angle", C.c_void_p),
                ("claims", C.c_void_p)]


class OrbTrackPoints(C.Structure):
    _fields_ = [("n", C.c_int32), ("in_view", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("level", C.c_void_p), ("view_cos", C.c_void_p), ("descriptors", C.c_void_p), ("obs_positive", C.c_void_p)]


class OrbLastPoints(C.Structure):
    _fields_ = [("n", C.c_int32), ("valid", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("invz", C.c_void_p),
                ("octave", C.c_void_p), ("angle", C.c_void_p), ("descriptors", C.c_void_p), ("obs_positive", C.c_void_p)]


class OrbFrustumCamera(C.Structure):
    _fields_ = [("Rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("Ow", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float),
                ("cx", C.c_float), ("cy", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float),
                ("max_y", C.c_float), ("mbf", C.c_float), ("log_scale_factor", C.c_float)]


class OrbFrustumPoints(C.Structure):
    _fields_ = [("n", C.c_int32), ("world_pos", C.c_void_p), ("normal", C.c_void_p), ("max_distance", C.c_void_p),
                ("min_distance", C.c_void_p), ("consider", C.c_void_p)]


def frustum_args(cam: dict, world_pos, normal, max_distance, min_distance, consider=None):
    """(OrbFrustumCamera, OrbFrustumPoints, keep-alive list) from numpy inputs; cam has Rcw[9], tcw[3], Ow[3] and the scalars."""
    c = OrbFrustumCamera()
    c.Rcw[:] = [float(x) for x in np.asarray(cam["Rcw"], np.float32).reshape(9)]
    c.tcw[:] = [float(x) for x in np.asarray(cam["tcw"], np.float32).reshape(3)]
    c.Ow[:] = [float(x) for x in np.asarray(cam["Ow"], np.float32).reshape(3)]
    for k in ("fx", "fy", "cx", "cy", "min_x", "max_x", "min_y", "max_y", "mbf", "log_scale_factor"):
        setattr(c, k, float(np.float32(cam[k])))
    wp = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
    nr = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    mx = np.ascontiguousarray(max_distance, np.float32)
    mn = np.ascontiguousarray(min_distance, np.float32)
    co = None if consider is None else np.ascontiguousarray(consider, np.uint8)
    p = OrbFrustumPoints(wp.shape[0], wp.ctypes.data, nr.ctypes.data, mx.ctypes.data, mn.ctypes.data, None if co is None else co.ctypes.data)
    return c, p, [wp, nr, mx, mn, co]


class OrbProjection(C.Structure):
    _fields_ = [("R1", C.c_float * 9), ("t1", C.c_float * 3), ("R2", C.c_float * 9), ("t2", C.c_float * 3), ("Ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("min_x", C.c_float), ("max_x", C.c_float), ("min_y", C.c_float), ("max_y", C.c_float),
                ("two_stage", C.c_int32), ("reject_negative_z", C.c_int32), ("invz_rule", C.c_int32), ("uv_rule", C.c_int32),
                ("bounds_rule", C.c_int32), ("dist_rule", C.c_int32), ("check_view_angle", C.c_int32)]


def projection_args(proj: dict, world_pos, normal, max_distance, min_distance, consider=None):
    """(OrbProjection, OrbFrustumPoints, keep-alive list): proj holds R1[9], t1[3] (and R2, t2, Ow where used), the camera scalars
    and the rule switches of include/orbgpu.h: orb_projection."""
    c = OrbProjection()
    for k, m in (("R1", 9), ("t1", 3), ("R2", 9), ("t2", 3), ("Ow", 3)):
        getattr(c, k)[:] = [float(x) for x in np.asarray(proj.get(k, np.zeros(m)), np.float32).reshape(m)]
    for k in ("fx", "fy", "cx", "cy", "bf", "min_x", "max_x", "min_y", "max_y"):
        setattr(c, k, float(np.float32(proj.get(k, 0.0))))
    for k in ("two_stage", "reject_negative_z", "invz_rule", "uv_rule", "bounds_rule", "dist_rule", "check_view_angle"):
        setattr(c, k, int(proj.get(k, 0)))
    wp = np.ascontiguousarray(world_pos, np.float32).reshape(-1, 3)
    nr = None if normal is None else np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    mx = np.ascontiguousarray(max_distance, np.float32)
    mn = np.ascontiguousarray(min_distance, np.float32)
    co = None if consider is None else np.ascontiguousarray(consider, np.uint8)
    p = OrbFrustumPoints(wp.shape[0], wp.ctypes.data, None if nr is None else nr.ctypes.data, mx.ctypes.data, mn.ctypes.data,
                         None if co is None else co.ctypes.data)
    return c, p, [wp, nr, mx, mn, co]


def projection_outputs(n):
    return (np.zeros(n, np.uint8), np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32))


def frustum_outputs(n):
    return (np.zeros(n, np.uint8), np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32),
            np.zeros(n, np.float32))


def optr(a):
    """pointer or NULL"""
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OrbCvOrbParams(C.Structure):
    """orb_cvorb_params (include/orbgpu.h): the arguments of OpenCV's cv::ORB constructor the marker path uses."""
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32), ("edge_threshold", C.c_int32),
                ("score_type", C.c_int32), ("fast_threshold", C.c_int32), ("max_width", C.c_int32), ("max_height", C.c_int32),
                ("device", C.c_int32)]
