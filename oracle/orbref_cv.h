// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path
// (ar_orbslam2_b200/, include/); only tests/, __graft_entry__.smoke() and bench.py's CPU
// baseline legs may use it.
//
// CPU models of the OpenCV primitives the reference's hot path calls.  OpenCV's source is
// NOT part of /root/reference (un-vendored dependency: OpenCV >= 2.4.3, "tested with
// 2.4.11", ORB_SLAM2/CMakeLists.txt:31, ORB_SLAM2/README.md:63), so each routine restates
// the published algorithm and is pinned against cv2 4.13.0 by the golden vectors in
// tests/golden/ (made by tools/make_golden_cv2.py).  Reference call sites are cited per
// function as ORB_SLAM2/src/<file>:<line>.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <emmintrin.h>
#include <cstring>
#include <vector>

namespace orbref {

// cvRound: round-half-to-even (SSE cvtsd2si).  ref call sites: ORBextractor.cc:76,106,110,111,433,451,1050.
static inline int cv_round(double v) { return _mm_cvtsd_si32(_mm_set_sd(v)); }   // what OpenCV's cvRound compiles to on x86-64
static inline int cv_floor(double v) { return (int)std::floor(v); }
static inline int cv_ceil(double v) { return (int)std::ceil(v); }

static inline int reflect101(int p, int n) {
    // BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for 8UC1.  ref: ORBextractor.cc:1060-1061.
// Fixed point: 11-bit coefficients, horizontal pass in int32, vertical pass with the
// >>4 / >>16 / +2 >>2 staging of OpenCV's VResizeLinear<uchar,int,short>.
static inline void resize_linear_u8(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst,
                                    int dw, int dh, int dpitch) {
    const double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
    const double scale_x = 1.0 / inv_scale_x, scale_y = 1.0 / inv_scale_y;
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> ialpha(2 * dw), ibeta(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        ialpha[2 * dx] = (short)cv_round((1.f - fx) * 2048.f);
        ialpha[2 * dx + 1] = (short)cv_round(fx * 2048.f);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor(fy);
        fy -= sy;
        if (sy < 0) { fy = 0; sy = 0; }
        if (sy >= sh - 1) { fy = 0; sy = sh - 1; }
        yofs[dy] = sy;
        ibeta[2 * dy] = (short)cv_round((1.f - fy) * 2048.f);
        ibeta[2 * dy + 1] = (short)cv_round(fy * 2048.f);
    }
    std::vector<int> r0(dw), r1(dw);
    for (int dy = 0; dy < dh; dy++) {
        const int sy0 = yofs[dy], sy1 = sy0 + 1 < sh ? sy0 + 1 : sh - 1;
        const uint8_t *s0 = src + (size_t)sy0 * spitch, *s1 = src + (size_t)sy1 * spitch;
        for (int dx = 0; dx < dw; dx++) {
            const int sx0 = xofs[dx], sx1 = sx0 + 1 < sw ? sx0 + 1 : sw - 1;
            const int a0 = ialpha[2 * dx], a1 = ialpha[2 * dx + 1];
            r0[dx] = s0[sx0] * a0 + s0[sx1] * a1;
            r1[dx] = s1[sx0] * a0 + s1[sx1] * a1;
        }
        const int b0 = ibeta[2 * dy], b1 = ibeta[2 * dy + 1];
        uint8_t *d = dst + (size_t)dy * dpitch;
        for (int dx = 0; dx < dw; dx++) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            d[dx] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
}

// cv::copyMakeBorder(..., BORDER_REFLECT_101) around an interior that already sits inside
// the padded plane.  ref: ORBextractor.cc:1063-1068.  `interior` points at pixel (0,0).
static inline void make_border_reflect101(uint8_t *interior, int w, int h, int pitch, int border) {
    for (int y = 0; y < h; y++) {
        uint8_t *row = interior + (std::ptrdiff_t)y * pitch;
        for (int x = 1; x <= border; x++) {
            row[-x] = row[reflect101(-x, w)];
            row[w - 1 + x] = row[reflect101(w - 1 + x, w)];
        }
    }
    for (int y = 1; y <= border; y++) {
        std::memcpy(interior + (std::ptrdiff_t)(-y) * pitch - border,
                    interior + (std::ptrdiff_t)reflect101(-y, h) * pitch - border, w + 2 * border);
        std::memcpy(interior + (std::ptrdiff_t)(h - 1 + y) * pitch - border,
                    interior + (std::ptrdiff_t)reflect101(h - 1 + y, h) * pitch - border, w + 2 * border);
    }
}

// cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for 8UC1.
// ref: ORBextractor.cc:1024-1026 (applied to an un-padded clone of the level).
// taps_variant 0: OpenCV 4.x fixed-point 8.8 taps; 1: OpenCV 2.4 rounded-float taps
// (unverifiable here).
static inline const int *gauss_taps(int variant) {
    static const int k4[7] = {18, 34, 48, 56, 48, 34, 18};
    static const int k2[7] = {18, 34, 49, 55, 49, 34, 18};
    return variant == 1 ? k2 : k4;
}
static inline void gaussian_blur7_u8(const uint8_t *src, int w, int h, int spitch, uint8_t *dst,
                                     int dpitch, int taps_variant) {
    const int *k = gauss_taps(taps_variant);
    std::vector<uint16_t> hbuf((size_t)w * h);   // row sums fit 16 bits (256*255)
    for (int y = 0; y < h; y++) {
        const uint8_t *s = src + (size_t)y * spitch;
        uint16_t *hb = &hbuf[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            int acc = 0;
            if (x >= 3 && x < w - 3) {
                for (int i = 0; i < 7; i++) acc += k[i] * s[x + i - 3];
            } else {
                for (int i = 0; i < 7; i++) acc += k[i] * s[reflect101(x + i - 3, w)];
            }
            hb[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t *d = dst + (size_t)y * dpitch;
        const uint16_t *r[7];
        for (int j = 0; j < 7; j++) r[j] = &hbuf[(size_t)reflect101(y + j - 3, h) * w];
        for (int x = 0; x < w; x++) {
            int acc = 0;
            for (int j = 0; j < 7; j++) acc += k[j] * r[j][x];
            int v = (acc + 32768) >> 16;
            d[x] = (uint8_t)(v > 255 ? 255 : v);
        }
    }
}

// cv::fastAtan2(y, x) in degrees.  ref: ORBextractor.cc:97.  fp32, no FMA (build with
// -ffp-contract=off).
static inline float fast_atan2(float y, float x) {
    const float p1 = 57.283627f, p3 = -18.667446f, p5 = 8.9140005f, p7 = -2.5397246f;
    const float eps = 2.2204460492503131e-16f;
    float ax = std::fabs(x), ay = std::fabs(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + eps);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + eps);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

struct FastPoint {
    int x, y;        // ROI coordinates
    int response;    // V - 1
};

// FAST-9/16 arc strength V of pixel p: max over the 16 arcs of 9 consecutive circle pixels
// of min(I(p)-I(q)) and of min(I(q)-I(p)).  Corner at threshold t  <=>  V > t.
static inline int fast_arc_strength(const uint8_t *p, int pitch) {
    static const int ox[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
    static const int oy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; k++) d[k] = v - p[oy[k] * pitch + ox[k]];
    for (int k = 16; k < 25; k++) d[k] = d[k - 16];
    int best = -256;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int j = 1; j < 9; j++) {
            if (d[k + j] < mn) mn = d[k + j];
            if (d[k + j] > mx) mx = d[k + j];
        }
        if (mn > best) best = mn;
        if (-mx > best) best = -mx;
    }
    return best;
}

// cv::FAST(roi, keypoints, threshold, nonmaxSuppression=true), TYPE_9_16.
// ref: ORBextractor.cc:776-783.  Pixels within 3 of the ROI edge are not tested; the score
// of a non-corner (or untested) neighbour is 0 in the 3x3 non-maximum suppression; output is
// row-major; response = V - 1.
static inline void fast9_16_nms(const uint8_t *roi, int w, int h, int pitch, int threshold,
                                std::vector<FastPoint> &out) {
    out.clear();
    if (w < 7 || h < 7) return;
    std::vector<int> score((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const uint8_t *p = roi + (size_t)y * pitch + x;
            // exact early-out (every 9-arc of the 16-circle holds two adjacent compass points), the
            // same high-speed test OpenCV's FAST applies before scoring
            const int c = p[0], d0 = c - p[3 * pitch], d4 = c - p[3], d8 = c - p[-3 * pitch], d12 = c - p[-3];
            const int hi = (d0 > threshold) + (d4 > threshold) + (d8 > threshold) + (d12 > threshold);
            const int lo = (d0 < -threshold) + (d4 < -threshold) + (d8 < -threshold) + (d12 < -threshold);
            if (hi < 2 && lo < 2) continue;
            int V = fast_arc_strength(p, pitch);
            score[(size_t)y * w + x] = V > threshold ? V - 1 : 0;
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            const int s = score[(size_t)y * w + x];
            if (s == 0) continue;
            const int *r = &score[(size_t)y * w + x];
            if (s > r[-1] && s > r[1] && s > r[-w - 1] && s > r[-w] && s > r[-w + 1] && s > r[w - 1] &&
                s > r[w] && s > r[w + 1])
                out.push_back(FastPoint{x, y, s});
        }
}

// cosf / sinf as the reference's host libm computes them (ref: ORBextractor.cc:103-104,
// `(float)cos(angle)` on a float argument resolves to cosf).  Restatement of glibc's
// sincosf algorithm (Szabolcs Nagy's optimized-routines, glibc >= 2.28): double-precision
// quadrant reduction + degree-7/8 polynomials, no FMA.  Verified bit-identical to glibc
// 2.39 cosf/sinf on ALL 1 135 869 952 fp32 angles a*pi/180, a in [0,360)
// (tools/check_sincos_exhaustive.cpp; 0 mismatches).
struct SinCosTab { double sign[4]; double hpi_inv, hpi, c0, c1, c2, c3, c4, s1, s2, s3; };
static inline const SinCosTab *sincos_tab() {
    static const SinCosTab T[2] = {
        {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, 0x1.921FB54442D18p0, 0x1p0,
         -0x1.ffffffd0c621cp-2, 0x1.55553e1068f19p-5, -0x1.6c087e89a359dp-10,
         0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7,
         -0x1.994eb3774cf24p-13},
        {{1.0, -1.0, -1.0, 1.0}, 0x1.45F306DC9C883p+23, -0x1.921FB54442D18p0, -0x1p0,
         0x1.ffffffd0c621cp-2, -0x1.55553e1068f19p-5, 0x1.6c087e89a359dp-10,
         -0x1.99343027bf8c3p-16, -0x1.555545995a603p-3, 0x1.1107605230bc4p-7,
         -0x1.994eb3774cf24p-13}};
    return T;
}
static inline float sincos_poly(double x, double x2, const SinCosTab *p, int n) {
    if ((n & 1) == 0) {
        double x3 = x * x2, s1 = p->s2 + x2 * p->s3, x7 = x3 * x2, s = x + x3 * p->s1;
        return (float)(s + x7 * s1);
    }
    double x4 = x2 * x2, c2 = p->c3 + x2 * p->c4, c1 = p->c0 + x2 * p->c1, x6 = x4 * x2,
           c = c1 + x4 * p->c2;
    return (float)(c + x6 * c2);
}
static inline uint32_t abstop12(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return (u >> 20) & 0x7ff;
}
// valid for |y| < 120 (the descriptor only needs [0, 2*pi))
static inline void sincosf_libm(float y, float *sn, float *cs) {
    double x = y;
    const SinCosTab *p = sincos_tab();
    if (abstop12(y) < abstop12(0x1.921FB54442D18p-1f)) {
        double x2 = x * x;
        if (abstop12(y) < abstop12(0x1p-12f)) { *sn = y; *cs = 1.0f; return; }
        *sn = sincos_poly(x, x2, p, 0);
        *cs = sincos_poly(x, x2, p, 1);
        return;
    }
    double r = x * p->hpi_inv;
    int n = ((int32_t)r + 0x800000) >> 24;
    x = x - n * p->hpi;
    double s = p->sign[n & 3];
    if (n & 2) p = p + 1;
    *sn = sincos_poly(x * s, x * x, p, n);
    *cs = sincos_poly(x * s, x * x, p, n ^ 1);
}

}  // namespace orbref
