"""Host-side mirror of ORB_SLAM2::ORBextractor (ref: ORB_SLAM2/include/ORBextractor.h:45-118) on top of
the liborbgpu C ABI.  Same constructor arguments, same call semantics (`mask` is ignored, as in the
reference), same getters; `mvImagePyramid` is materialised lazily from the device pyramid."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import CANDIDATE_DTYPE, KEYPOINT_DTYPE, ORB_DEVICE, ORB_HOST, OrbParams, ptr


class ORBextractor:
    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int, minThFAST: int,
                 max_width: int = 1920, max_height: int = 1080, max_batch: int = 1, device: int = -1, blur_taps: int = 0):
        self._L = _lib.lib()
        self.params = OrbParams(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width, max_height,
                                max_batch, device, blur_taps)
        self._h = C.c_void_p()
        _lib.check(self._L.orb_extractor_create(C.byref(self.params), C.byref(self._h)), "orb_extractor_create")
        self.nlevels = nlevels
        self.max_batch = max_batch
        n = nlevels
        self._scale, self._inv_scale, self._sigma2, self._inv_sigma2 = (np.zeros(n, np.float32) for _ in range(4))
        self._per_level = np.zeros(n, np.int32)
        _lib.check(self._L.orb_extractor_tables(self._h, ptr(self._scale), ptr(self._inv_scale), ptr(self._sigma2),
                                                ptr(self._inv_sigma2), ptr(self._per_level)), "orb_extractor_tables")
        self.default_capacity = nfeatures + 4 * nlevels + 256
        self._landing = {}
        _lib.register(self)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orb_extractor_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- reference getters (ORBextractor.h:64-86)
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return float(np.float32(np.float64(np.float32(self.params.scale_factor))))
    def GetScaleFactors(self): return self._scale.copy()
    def GetInverseScaleFactors(self): return self._inv_scale.copy()
    def GetScaleSigmaSquares(self): return self._sigma2.copy()
    def GetInverseScaleSigmaSquares(self): return self._inv_sigma2.copy()
    @property
    def mnFeaturesPerLevel(self): return self._per_level.copy()

    # --- operator() (ORBextractor.cc:985-1045)
    def __call__(self, image: np.ndarray, mask=None, capacity: int | None = None):
        """Returns (keypoints[KEYPOINT_DTYPE], descriptors[n,32] uint8).  `mask` is ignored (ref :60)."""
        if image is None or image.size == 0:
            return np.zeros(0, KEYPOINT_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 image required (ref ORBextractor.cc:991)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        cap = capacity or self.default_capacity
        land = self._landing.get(cap)
        if land is None:   # landing arrays (and their ctypes pointers) are made once per capacity: the call is latency-critical
            kps, desc, n = np.zeros(cap, KEYPOINT_DTYPE), np.zeros((cap, 32), np.uint8), C.c_int(0)
            land = self._landing[cap] = (kps, desc, n, ptr(kps), ptr(desc), C.byref(n))
        kps, desc, n, p_kps, p_desc, p_n = land
        h, w = image.shape
        _lib.check(self._L.orb_extract(self._h, C.c_void_p(image.ctypes.data), w, h, image.strides[0], ORB_HOST, p_kps, p_desc, cap, p_n),
                   "orb_extract")
        return kps[: n.value].copy(), desc[: n.value].copy()

    def extract_batch(self, images: np.ndarray, capacity: int | None = None):
        """images: [B,H,W] uint8 (host).  Returns per-frame lists of (keypoints, descriptors)."""
        assert images.dtype == np.uint8 and images.ndim == 3
        images = np.ascontiguousarray(images)
        B, h, w = images.shape
        cap = capacity or self.default_capacity
        kps = np.zeros((B, cap), KEYPOINT_DTYPE)
        desc = np.zeros((B, cap, 32), np.uint8)
        n = np.zeros(B, np.int32)
        _lib.check(self._L.orb_extract_batch(self._h, ptr(images), B, w, h, w, C.c_size_t(w * h), ORB_HOST, ptr(kps),
                                             ptr(desc), cap, ptr(n), ORB_HOST), "orb_extract_batch")
        return [(kps[b, : n[b]].copy(), desc[b, : n[b]].copy()) for b in range(B)]

    def extract_batch_device(self, d_images: int, batch: int, width: int, height: int, pitch: int, frame_stride: int,
                             d_kps: int, d_desc: int, capacity: int, d_n: int):
        """Asynchronous device-to-device form (raw device pointers as ints); see orb_extract_batch."""
        _lib.check(self._L.orb_extract_batch(self._h, C.c_void_p(d_images), batch, width, height, pitch,
                                             C.c_size_t(frame_stride), ORB_DEVICE, C.c_void_p(d_kps), C.c_void_p(d_desc),
                                             capacity, C.c_void_p(d_n), ORB_DEVICE), "orb_extract_batch")

    def sync(self):
        _lib.check(self._L.orb_extractor_sync(self._h), "orb_extractor_sync")

    def check(self):
        _lib.check(self._L.orb_extractor_check(self._h), "orb_extractor_check")

    @property
    def stream(self) -> int:
        return self._L.orb_extractor_stream(self._h)

    @property
    def launch_count(self) -> int:
        return self._L.orb_extractor_launch_count(self._h)

    @property
    def graph_replays(self) -> int:
        """Extractions that ran as one CUDA-graph launch (repeated calls with the same device buffers and geometry)."""
        return self._L.orb_extractor_graph_replays(self._h)

    # --- pyramid (ORBextractor.h:88 mvImagePyramid)
    def level_size(self, level: int, slot: int = 0):
        w, h, p = C.c_int(), C.c_int(), C.c_int()
        dp = C.c_void_p()
        _lib.check(self._L.orb_pyramid_level(self._h, slot, level, C.byref(dp), C.byref(w), C.byref(h), C.byref(p)),
                   "orb_pyramid_level")
        return w.value, h.value

    def pyramid_level(self, level: int, with_border: bool = False, slot: int = 0) -> np.ndarray:
        w, h = self.level_size(level, slot)
        if with_border:
            w, h = w + 38, h + 38
        out = np.zeros((h, w), np.uint8)
        _lib.check(self._L.orb_pyramid_download(self._h, slot, level, int(with_border), ptr(out), w), "orb_pyramid_download")
        return out

    @property
    def mvImagePyramid(self):
        return [self.pyramid_level(l) for l in range(self.nlevels)]

    # --- stage outputs for parity tests
    def candidates(self, level: int, slot: int = 0) -> np.ndarray:
        cap = 1 << 20
        out = np.zeros(cap, CANDIDATE_DTYPE)
        n = C.c_int()
        _lib.check(self._L.orb_debug_candidates(self._h, slot, level, ptr(out), cap, C.byref(n)), "orb_debug_candidates")
        return out[: n.value].copy()

    def blurred_level(self, level: int, slot: int = 0) -> np.ndarray:
        w, h = self.level_size(level, slot)
        out = np.zeros((h, w), np.uint8)
        _lib.check(self._L.orb_debug_blur_level(self._h, slot, level, ptr(out), w), "orb_debug_blur_level")
        return out
