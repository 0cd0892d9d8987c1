// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of cv::remap(src, dst, map1, map2, INTER_LINEAR) for 8-bit single-channel images and CV_32F maps, as
// Examples/ROS/ORB_SLAM2/src/ros_stereo.cc:163-164 calls it.  OpenCV is an un-vendored dependency; the arithmetic
// (coordinates rounded to 1/32 pixel, 15-bit BilinearTab_i weights incl. its one saturated entry, (s + 2^14) >> 15,
// BORDER_CONSTANT 0) is pinned against cv2 4.13.0 by tests/golden/cv2_remap.npz (tools/make_golden_remap.py).
#include "orbref.h"

#include <cmath>

extern "C" int orbref_remap_linear(const uint8_t *src, int sw, int sh, int spitch, const float *mapx, const float *mapy, int mpitch,
                                   uint8_t *dst, int dw, int dh, int dpitch) {
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            const int sx = (int)std::nearbyint((double)(mapx[(size_t)y * mpitch + x] * 32.0f));   // cvRound (round half to even)
            const int sy = (int)std::nearbyint((double)(mapy[(size_t)y * mpitch + x] * 32.0f));
            const int fx = sx & 31, fy = sy & 31, ix = sx >> 5, iy = sy >> 5;
            int w[4] = {(32 - fx) * (32 - fy) * 32, fx * (32 - fy) * 32, (32 - fx) * fy * 32, fx * fy * 32};
            if (fx == 0 && fy == 0) { w[0] = 32767; w[3] = 1; }   // saturate_cast<short>(32768), then the sum correction of initInterTab2D
            int s = 0;
            for (int k = 0; k < 4; k++) {
                const int xx = ix + (k & 1), yy = iy + (k >> 1);
                const int p = (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * spitch + xx] : 0;
                s += p * w[k];
            }
            int v = (s + (1 << 14)) >> 15;
            dst[(size_t)y * dpitch + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    return ORB_OK;
}
