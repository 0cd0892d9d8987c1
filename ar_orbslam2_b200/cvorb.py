"""Host-side mirror of OpenCV's `cv::ORB` as the reference's marker pipeline uses it -- `ORB orb; orb(image, Mat(), keys,
descriptors)` in Marker::setTargetImage / Marker::Match (ref: ORB_SLAM2/src/Marker.cc:83-84, 107-108) -- on top of the
liborbgpu C ABI (orb_cvorb_*, include/orbgpu.h).  Constructor arguments are cv::ORB's, with its defaults; firstLevel, WTA_K
and patchSize are fixed at 0 / 2 / 31 (what `ORB orb;` uses)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import KEYPOINT_DTYPE, ORB_DEVICE, ORB_HOST, OrbCvOrbParams, ptr

HARRIS_SCORE = 0
FAST_SCORE = 1


class ORB:
    def __init__(self, nfeatures: int = 500, scaleFactor: float = 1.2, nlevels: int = 8, edgeThreshold: int = 31, firstLevel: int = 0,
                 WTA_K: int = 2, scoreType: int = HARRIS_SCORE, patchSize: int = 31, fastThreshold: int = 20,
                 max_width: int = 1920, max_height: int = 1080, device: int = -1):
        if firstLevel != 0 or WTA_K != 2 or patchSize != 31:
            raise ValueError("cv::ORB mirror: firstLevel / WTA_K / patchSize are fixed at 0 / 2 / 31")
        self._L = _lib.lib()
        self._L.orb_cvorb_launch_count.restype = C.c_int64
        self._L.orb_cvorb_launch_count.argtypes = [C.c_void_p]
        self._L.orb_cvorb_destroy.argtypes = [C.c_void_p]
        self._L.orb_cvorb_destroy.restype = None
        self._L.orb_cvorb_capacity.argtypes = [C.c_void_p]
        self.params = OrbCvOrbParams(nfeatures, scaleFactor, nlevels, edgeThreshold, scoreType, fastThreshold, max_width, max_height, device)
        self._h = C.c_void_p()
        _lib.check(self._L.orb_cvorb_create(C.byref(self.params), C.byref(self._h)), "orb_cvorb_create")
        self.capacity = self._L.orb_cvorb_capacity(self._h)
        self._kps = np.zeros(self.capacity, KEYPOINT_DTYPE)
        self._desc = np.zeros((self.capacity, 32), np.uint8)
        _lib.register(self)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orb_cvorb_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self):
        return self._L.orb_cvorb_launch_count(self._h)

    def __call__(self, image: np.ndarray, mask=None):
        """(keypoints[KEYPOINT_DTYPE] in OpenCV's order, descriptors[n, 32]); the marker path passes an empty mask."""
        if mask is not None:
            raise ValueError("cv::ORB mirror: masks are not supported (the reference passes Mat())")
        if image is None or image.size == 0:
            return np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "8-bit single-channel image required"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        n = C.c_int(0)
        _lib.check(self._L.orb_cvorb_detect_and_compute(self._h, C.c_void_p(image.ctypes.data), w, h, image.strides[0], ORB_HOST,
                                                        ptr(self._kps), ptr(self._desc), self.capacity, C.byref(n)),
                   "orb_cvorb_detect_and_compute")
        return self._kps[: n.value].copy(), self._desc[: n.value].copy()

    detectAndCompute = __call__

    def detect_and_compute_device(self, dev_ptr: int, width: int, height: int, pitch: int):
        """The same call on an 8-bit image that already lives on the device (e.g. the output of orb_cvt_color_gray /
        orb_remap_linear); results come back to the host."""
        n = C.c_int(0)
        _lib.check(self._L.orb_cvorb_detect_and_compute(self._h, C.c_void_p(dev_ptr), width, height, pitch, ORB_DEVICE,
                                                        ptr(self._kps), ptr(self._desc), self.capacity, C.byref(n)),
                   "orb_cvorb_detect_and_compute")
        return self._kps[: n.value].copy(), self._desc[: n.value].copy()

    def blurred_level(self, level: int, which: int = 0) -> np.ndarray:
        """Layer `level` of the last call as the descriptors sampled it (tests); which = 1: unblurred, 2: FAST score map."""
        level |= which << 8
        w, h = C.c_int(), C.c_int()
        _lib.check(self._L.orb_cvorb_level(self._h, level, None, 0, C.byref(w), C.byref(h)), "orb_cvorb_level")
        out = np.zeros((h.value, w.value), np.uint8)
        _lib.check(self._L.orb_cvorb_level(self._h, level, ptr(out), w.value, None, None), "orb_cvorb_level")
        return out


def retain_best(responses: np.ndarray, n_points: int, depth_limit: int = -1) -> np.ndarray:
    """KeyPointsFilter::retainBest on the device (tests): the surviving original indices in libstdc++'s order."""
    r = np.ascontiguousarray(responses, np.float32)
    order = np.zeros(max(len(r), 1), np.int32)
    n = C.c_int(0)
    _lib.check(_lib.lib().orb_debug_retain_best(ptr(r), len(r), int(n_points), int(depth_limit), ptr(order), C.byref(n)), "orb_debug_retain_best")
    return order[: n.value].copy()
