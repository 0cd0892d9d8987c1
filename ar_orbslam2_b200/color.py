"""Host-side mirror of the cvtColor calls in front of the extractor (ref: Tracking::GrabImageStereo / GrabImageRGBD /
GrabImageMonocular, ORB_SLAM2/src/Tracking.cc:510-526,545-556,567-577) over liborbgpu.  No CPU fallback."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._abi import ORB_DEVICE, ORB_HOST, ptr


def cvtColorGray(img: np.ndarray, rgb: bool = True, variant: int = 0) -> np.ndarray:
    """img: H x W x {3,4} uint8 (or B x H x W x C).  rgb=True is CV_RGB(A)2GRAY (mbRGB), False CV_BGR(A)2GRAY."""
    a = np.ascontiguousarray(img, np.uint8)
    batched = a.ndim == 4
    if not batched:
        a = a[None]
    B, h, w, ch = a.shape
    out = np.zeros((B, h, w), np.uint8)
    _lib.check(_lib.lib().orb_cvt_color_gray(ptr(a), w, h, w * ch, C.c_size_t(h * w * ch), ch, int(rgb), variant, B, ptr(out), w,
                                             C.c_size_t(h * w), ORB_HOST, None), "orb_cvt_color_gray")
    return out if batched else out[0]


def cvtColorGray_device(d_src: int, width: int, height: int, src_pitch: int, src_stride: int, channels: int, rgb: bool, batch: int,
                        d_dst: int, dst_pitch: int, dst_stride: int, stream: int = 0, variant: int = 0):
    """Device-resident form (raw addresses), asynchronous on `stream`."""
    _lib.check(_lib.lib().orb_cvt_color_gray(C.c_void_p(d_src), width, height, src_pitch, C.c_size_t(src_stride), channels, int(rgb),
                                             variant, batch, C.c_void_p(d_dst), dst_pitch, C.c_size_t(dst_stride), ORB_DEVICE,
                                             C.c_void_p(stream)), "orb_cvt_color_gray")


def remap(src: np.ndarray, map_x: np.ndarray, map_y: np.ndarray) -> np.ndarray:
    """cv::remap(src, dst, M1, M2, cv::INTER_LINEAR) of the stereo rectification (ref Examples/ROS/ORB_SLAM2/src/ros_stereo.cc:163-164)
    on the GPU.  src: H x W (or B x H x W) uint8; map_x / map_y: float32 of the destination size."""
    a = np.ascontiguousarray(src, np.uint8)
    batched = a.ndim == 3
    if not batched:
        a = a[None]
    B, sh, sw = a.shape
    mx, my = np.ascontiguousarray(map_x, np.float32), np.ascontiguousarray(map_y, np.float32)
    dh, dw = mx.shape
    out = np.zeros((B, dh, dw), np.uint8)
    _lib.check(_lib.lib().orb_remap_linear(ptr(a), sw, sh, sw, C.c_size_t(sh * sw), ptr(mx), ptr(my), dw, ptr(out), dw, dh, dw,
                                           C.c_size_t(dh * dw), B, ORB_HOST, None), "orb_remap_linear")
    return out if batched else out[0]
