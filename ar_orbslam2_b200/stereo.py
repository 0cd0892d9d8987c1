"""Host-side mirror of Frame::ComputeStereoMatches (ref: ORB_SLAM2/src/Frame.cc:471-643) over liborbgpu.

The reference reads `mpORBextractorLeft/Right->mvImagePyramid` directly; here the two extractor
handles keep their pyramids on the device and `orb_stereo_match` consumes them in place."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import ORB_ERR_EMPTY, ptr


def ComputeStereoMatches(left, right, keys_left, desc_left, keys_right, desc_right, mbf: float, mb: float, slot: int = 0,
                         allow_empty: bool = False):
    """Returns (mvuRight, mvDepth) float32 arrays of len(keys_left); -1 where unmatched.

    `mb` is Frame::mb (= mbf/fx).  The reference reads it before the constructor assigns it
    (Frame.cc:95 vs :119); the intended value is passed explicitly here."""
    L = _lib.lib()
    n_l, n_r = len(keys_left), len(keys_right)
    u_right = np.full(n_l, -1, np.float32)
    depth = np.full(n_l, -1, np.float32)
    nm = C.c_int(0)
    kl = np.ascontiguousarray(keys_left)
    kr = np.ascontiguousarray(keys_right)
    dl = np.ascontiguousarray(desc_left, np.uint8)
    dr = np.ascontiguousarray(desc_right, np.uint8)
    rc = L.orb_stereo_match(left._h, right._h, slot, ptr(kl), ptr(dl), n_l, ptr(kr), ptr(dr), n_r, C.c_float(mbf),
                            C.c_float(mb), ptr(u_right), ptr(depth), C.byref(nm))
    if rc == ORB_ERR_EMPTY and allow_empty:
        return u_right, depth
    _lib.check(rc, "orb_stereo_match")
    return u_right, depth


def stereo_match_batch_device(left, left_slot, right, right_slot, batch, d_keys_l, d_desc_l, d_n_l, d_keys_r, d_desc_r,
                              d_n_r, capacity, mbf, mb, d_u_right, d_depth, d_n_matched):
    """Asynchronous device-resident form (raw device pointers as ints); see orb_stereo_match_batch."""
    L = _lib.lib()
    vp = C.c_void_p
    _lib.check(L.orb_stereo_match_batch(left._h, left_slot, right._h, right_slot, batch, vp(d_keys_l), vp(d_desc_l), vp(d_n_l),
                                        vp(d_keys_r), vp(d_desc_r), vp(d_n_r), capacity, C.c_float(mbf), C.c_float(mb),
                                        vp(d_u_right), vp(d_depth), vp(d_n_matched), 1), "orb_stereo_match_batch")
