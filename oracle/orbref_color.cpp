// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of cv::cvtColor(..., CV_RGB2GRAY / BGR2GRAY / RGBA2GRAY / BGRA2GRAY) on 8-bit images, the call the
// reference makes in Tracking::GrabImage* (ORB_SLAM2/src/Tracking.cc:510-526,545-556,567-577).  OpenCV is an un-vendored
// dependency; the arithmetic is pinned against cv2 4.13.0 by tests/golden/cv2_cvtcolor.npz (tools/make_golden_cvtcolor.py).
#include "orbref.h"

extern "C" int orbref_cvt_color_gray(const uint8_t *src, int width, int height, int src_pitch, int channels, int rgb_order,
                                     int variant, uint8_t *dst, int dst_pitch) {
    if (channels != 3 && channels != 4) return ORB_ERR_INVALID;
    const int wr = variant ? 4899 : 9798, wg = variant ? 9617 : 19235, wb = variant ? 1868 : 3735;   // R2Y, G2Y, B2Y
    const int shift = variant ? 14 : 15;
    for (int y = 0; y < height; y++)
        for (int x = 0; x < width; x++) {
            const uint8_t *p = src + (size_t)y * src_pitch + (size_t)x * channels;
            const int r = rgb_order ? p[0] : p[2], g = p[1], b = rgb_order ? p[2] : p[0];
            dst[(size_t)y * dst_pitch + x] = (uint8_t)((r * wr + g * wg + b * wb + (1 << (shift - 1))) >> shift);
        }
    return ORB_OK;
}
