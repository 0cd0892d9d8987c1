"""Host-side mirror of Frame::ComputeStereoMatches (ref: ORB_SLAM2/src/Frame.cc:471-643) over liborbgpu.

The reference reads `mpORBextractorLeft/Right->mvImagePyramid` directly; here the two extractor
handles keep their pyramids on the device and `orb_stereo_match` consumes them in place."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import KEYPOINT_DTYPE, ORB_ERR_EMPTY, ptr


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


def StereoFrame(extractor, image_left, image_right, mbf: float, mb: float, capacity: int | None = None, allow_empty: bool = False):
    """The stereo Frame constructor's front end in one call (ref: ORB_SLAM2/src/Frame.cc:83-95: ExtractORB left and right,
    then ComputeStereoMatches) on ONE handle created with max_batch >= 2: orb_stereo_frame.
    Returns (keys_left, desc_left, keys_right, desc_right, mvuRight, mvDepth)."""
    L = _lib.lib()
    il, ir = image_left, image_right
    assert il.dtype == np.uint8 and ir.dtype == np.uint8 and il.ndim == 2 and il.shape == ir.shape, "two CV_8UC1 images of one size required"
    if il.strides[1] != 1:
        il = np.ascontiguousarray(il)
    if ir.strides[1] != 1:
        ir = np.ascontiguousarray(ir)
    h, w = il.shape
    cap = capacity or extractor.default_capacity
    land = extractor._landing.get(("stereo", cap))
    if land is None:   # landing arrays and their ctypes pointers are made once per capacity: the call is latency-critical
        arrs = (np.zeros(cap, KEYPOINT_DTYPE), np.zeros((cap, 32), np.uint8), np.zeros(cap, KEYPOINT_DTYPE), np.zeros((cap, 32), np.uint8),
                np.full(cap, -1, np.float32), np.full(cap, -1, np.float32))
        cnt = (C.c_int(0), C.c_int(0), C.c_int(0))
        land = extractor._landing[("stereo", cap)] = (arrs, cnt, tuple(ptr(a) for a in arrs), tuple(C.byref(c) for c in cnt))
    (kl, dl, kr, dr, ur, dp), (nl, nr, nm), (p_kl, p_dl, p_kr, p_dr, p_ur, p_dp), (p_nl, p_nr, p_nm) = land
    rc = L.orb_stereo_frame(extractor._h, C.c_void_p(il.ctypes.data), C.c_void_p(ir.ctypes.data), w, h, il.strides[0], ir.strides[0],
                            p_kl, p_dl, p_nl, p_kr, p_dr, p_nr, cap, C.c_float(mbf), C.c_float(mb), p_ur, p_dp, p_nm)
    if not (rc == ORB_ERR_EMPTY and allow_empty):
        _lib.check(rc, "orb_stereo_frame")
    a, b = nl.value, nr.value
    return kl[:a].copy(), dl[:a].copy(), kr[:b].copy(), dr[:b].copy(), ur[:a].copy(), dp[:a].copy()


def ComputeStereoFromRGBD(keys, keys_un, im_depth: np.ndarray, mbf: float, depth_factor: float = 1.0):
    """Frame::ComputeStereoFromRGBD (ref Frame.cc:645-667) with the convertTo of Tracking.cc:553-554 folded in: im_depth is the
    CV_32F or CV_16U depth image, depth_factor = mDepthMapFactor.  Returns (mvuRight, mvDepth)."""
    import ctypes as C
    from . import _lib
    from ._abi import KEYPOINT_DTYPE, ORB_HOST, ptr
    k = np.ascontiguousarray(keys, KEYPOINT_DTYPE)
    ku = np.ascontiguousarray(keys_un, KEYPOINT_DTYPE)
    assert im_depth.ndim == 2 and im_depth.dtype in (np.float32, np.uint16) and im_depth.strides[1] == im_depth.itemsize
    ur, dp = np.zeros(len(k), np.float32), np.zeros(len(k), np.float32)
    _lib.check(_lib.lib().orb_depth_from_rgbd(ptr(k), ptr(ku), len(k), C.c_void_p(im_depth.ctypes.data), int(im_depth.dtype == np.uint16),
                                              im_depth.shape[1], im_depth.shape[0], im_depth.strides[0], C.c_float(depth_factor),
                                              C.c_float(mbf), ORB_HOST, ptr(ur), ptr(dp)), "orb_depth_from_rgbd")
    return ur, dp
