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
