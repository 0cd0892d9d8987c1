// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of OpenCV's `cv::ORB` feature detector/descriptor as the reference's marker
// pipeline uses it: `ORB orb; orb(image, Mat(), keys, descriptors);` in Marker::setTargetImage
// (ORB_SLAM2/src/Marker.cc:83-84) and Marker::Match (:107-108) -- default parameters: 500
// features, scale 1.2, 8 levels, edgeThreshold 31, firstLevel 0, WTA_K 2, HARRIS_SCORE, patch 31
// (the commented-out constructor call at Marker.cc:80 spells them), FAST threshold 20.
//
// OpenCV is an un-vendored dependency of the reference (OpenCV >= 2.4.3, CMakeLists.txt:31); its
// source is not under /root/reference.  This file restates the published algorithm of
// modules/features2d/src/orb.cpp (ORB_Impl::detectAndCompute, computeKeyPoints, HarrisResponses,
// ICAngles, computeOrbDescriptors), KeyPointsFilter::{runByImageBorder, retainBest}
// (keypoint.cpp) and resize(INTER_LINEAR_EXACT) (resize.cpp: fixed-point 8.8 bit-exact path), and
// is PINNED against cv2 4.13.0 -- the only OpenCV on this box -- by tests/golden/cv2_orb.npz
// (tools/make_golden_cvorb.py): key points (all seven fields, in order) and descriptors equal.
// retainBest's output ORDER is whatever std::nth_element + std::partition leave behind; the
// oracle calls the same libstdc++ algorithms (the device restates them).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "orbref.h"
#include "orbref_cv.h"
#include "../ar_orbslam2_b200/csrc/orb_pattern.h"

namespace {
using namespace orbref;

struct KP {
    float x, y, size, angle, response;
    int octave, class_id;
};

// resize.cpp, interpolation_linear<ufixedpoint16>::getCoeffs: 8.8 fixed-point coefficient pairs
struct LinExact {
    std::vector<int> ofs;
    std::vector<uint16_t> c0, c1;
    int mn, mx;   // [mn, mx): outputs interpolated between two source samples
    LinExact(int srcsize, int dstsize) : ofs(dstsize), c0(dstsize), c1(dstsize), mn(0), mx(dstsize) {
        const double inv_scale = (double)dstsize / srcsize;
        const double scale = 1.0 / inv_scale;
        for (int v = 0; v < dstsize; v++) {
            volatile double t = scale * ((double)v + 0.5);   // softdouble: every operation rounded to double, never fused
            const double fval = t - 0.5;
            const int ival = cv_floor(fval);
            ofs[v] = 0;
            c0[v] = c1[v] = 0;
            if (ival >= 0 && srcsize > 1) {
                if (ival < srcsize - 1) {
                    ofs[v] = ival;
                    c1[v] = (uint16_t)cv_round((fval - (double)ival) * 256.0);
                    c0[v] = (uint16_t)(256 - c1[v]);
                } else {
                    ofs[v] = srcsize - 1;
                    mx = std::min(mx, v);
                }
            } else {
                mn = std::max(mn, v + 1);
            }
        }
        if (mn > mx) mn = mx;
    }
};

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR_EXACT), 8UC1 (resize_bitExact<uint8_t, ufixedpoint16, 2>)
void resize_linear_exact_u8(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst, int dw, int dh, int dpitch) {
    LinExact X(sw, dw), Y(sh, dh);
    std::vector<uint16_t> row0(dw), row1(dw);
    auto hline = [&](const uint8_t *s, uint16_t *d) {
        int i = 0;
        for (; i < X.mn; i++) d[i] = (uint16_t)(s[0] << 8);
        for (; i < X.mx; i++) d[i] = (uint16_t)(X.c0[i] * s[X.ofs[i]] + X.c1[i] * s[X.ofs[i] + 1]);
        for (; i < dw; i++) d[i] = (uint16_t)(s[sw - 1] << 8);
    };
    for (int dy = 0; dy < dh; dy++) {
        uint8_t *d = dst + (size_t)dy * dpitch;
        if (dy < Y.mn || dy >= Y.mx) {
            hline(src + (size_t)(dy < Y.mn ? 0 : sh - 1) * spitch, row0.data());
            for (int i = 0; i < dw; i++) d[i] = (uint8_t)std::min(255, (row0[i] + 128) >> 8);
        } else {
            hline(src + (size_t)Y.ofs[dy] * spitch, row0.data());
            hline(src + (size_t)(Y.ofs[dy] + 1) * spitch, row1.data());
            const uint32_t m0 = Y.c0[dy], m1 = Y.c1[dy];
            for (int i = 0; i < dw; i++) {
                const uint32_t v = row0[i] * m0 + row1[i] * m1;
                d[i] = (uint8_t)std::min<uint32_t>(255u, (v + 32768u) >> 16);
            }
        }
    }
}

// GaussianBlur(layer, layer, Size(7,7), 2, 2, BORDER_REFLECT_101) as orb.cpp applies it: the layer is a SUBMATRIX of the pyramid
// buffer and the border type is not BORDER_ISOLATED, so OpenCV 4.x leaves its bit-exact fixed-point blur (the one
// ORBextractor.cc:1025 gets on its cloned level, orbref_cv.h gaussian_blur7_u8) and runs sepFilter2D with the CV_32F kernel
// getGaussianKernel(7, 2) in fp32.  The arithmetic is what cv2 4.13.0 executes on this box (x86-64, AVX2/FMA3 dispatch of
// filter.simd.hpp, found by matching all 16 order/fusion variants against cv2.sepFilter2D over 3.7 M pixels: this one has 0
// mismatches, every other one 7..21): row filter RowVec_32f  s = k[0]*p[x-3]; s = fma(p[x-3+j], k[j], s), j = 1..6;
// column filter SymmColumnVec_32f  s = k[3]*r[y]; s = fma(r[y+j] + r[y-j], k[3+j], s), j = 1..3;  saturate_cast<uchar> =
// round-half-even.  (A build of OpenCV without FMA rounds ~1 pixel in 10^5 differently.)  Pinned through the descriptors of
// cv2's ORB (0 differing; the fixed-point blur leaves 25 % of them different) and directly against cv2.sepFilter2D.
static const float kGauss7[7] = {0x1.1f5f62p-4f, 0x1.0c70fcp-3f, 0x1.869472p-3f, 0x1.ba95c0p-3f, 0x1.869472p-3f, 0x1.0c70fcp-3f, 0x1.1f5f62p-4f};
void gaussian_blur7_sep32f_u8(const uint8_t *src, int w, int h, int spitch, uint8_t *dst, int dpitch) {
    std::vector<float> rows((size_t)w * h);
    for (int y = 0; y < h; y++) {
        const uint8_t *s = src + (size_t)y * spitch;
        float *r = &rows[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            float acc = kGauss7[0] * (float)s[reflect101(x - 3, w)];
            for (int j = 1; j < 7; j++) acc = std::fmaf((float)s[reflect101(x - 3 + j, w)], kGauss7[j], acc);
            r[x] = acc;
        }
    }
    for (int y = 0; y < h; y++) {
        uint8_t *d = dst + (size_t)y * dpitch;
        const float *c = &rows[(size_t)y * w];
        const float *rp[4], *rm[4];
        for (int j = 1; j <= 3; j++) {
            rp[j] = &rows[(size_t)reflect101(y + j, h) * w];
            rm[j] = &rows[(size_t)reflect101(y - j, h) * w];
        }
        for (int x = 0; x < w; x++) {
            float acc = kGauss7[3] * c[x];
            for (int j = 1; j <= 3; j++) acc = std::fmaf(rp[j][x] + rm[j][x], kGauss7[3 + j], acc);
            const int v = cv_round(acc);
            d[x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
}

struct Level {
    int w, h;
    float scale;
    std::vector<uint8_t> img;   // w x h, pitch w (orb.cpp keeps all layers in one buffer with a 32-px reflect-101 frame;
                                // no read of the path leaves a layer's interior -- FAST and the Harris / moment / descriptor
                                // windows stay >= 31 px inside, the blur reflects at the layer edge -- so the frame is not stored)
};

struct ResponseGreater {
    bool operator()(const KP &a, const KP &b) const { return a.response > b.response; }
};

// KeyPointsFilter::retainBest (keypoint.cpp)
void retain_best(std::vector<KP> &k, int n_points) {
    if (n_points >= 0 && k.size() > (size_t)n_points) {
        if (n_points == 0) { k.clear(); return; }
        std::nth_element(k.begin(), k.begin() + n_points - 1, k.end(), ResponseGreater());
        const float amb = k[n_points - 1].response;
        auto new_end = std::partition(k.begin() + n_points, k.end(), [amb](const KP &p) { return p.response >= amb; });
        k.resize(new_end - k.begin());
    }
}

int run(const orb_cvorb_params *p, const uint8_t *image, int width, int height, int pitch, std::vector<KP> &all,
        std::vector<uint8_t> &desc, std::vector<Level> *levels_out) {
    const int nlevels = p->nlevels, patch = 31, half = 15, edge = p->edge_threshold;
    if (nlevels < 1 || nlevels > 16 || p->nfeatures < 0 || edge < 31 || !(p->scale_factor > 1.f)) return ORB_ERR_INVALID;
    all.clear();
    desc.clear();
    if (!image || width <= 0 || height <= 0) return ORB_OK;
    const double scale_factor = (double)p->scale_factor;   // ORB_Impl keeps the float argument in a double member
    std::vector<Level> L(nlevels);
    for (int l = 0; l < nlevels; l++) {
        const float scale = (float)std::pow(scale_factor, (double)l);   // getScale(level, firstLevel = 0, scaleFactor)
        const float inv = 1.0f / scale;
        L[l].scale = scale;
        L[l].w = cv_round((float)width * inv);
        L[l].h = cv_round((float)height * inv);
        if (L[l].w < 1 || L[l].h < 1) return ORB_ERR_INVALID;
        L[l].img.resize((size_t)L[l].w * L[l].h);
        if (l == 0) {
            for (int y = 0; y < height; y++) std::memcpy(&L[0].img[(size_t)y * width], image + (size_t)y * pitch, width);
        } else {
            // prevImg stays the input for level 1 and is the previous layer afterwards: a cascade
            resize_linear_exact_u8(L[l - 1].img.data(), L[l - 1].w, L[l - 1].h, L[l - 1].w, L[l].img.data(), L[l].w, L[l].h, L[l].w);
        }
    }
    // computeKeyPoints: per-level quotas
    std::vector<int> quota(nlevels);
    {
        const float factor = (float)(1.0 / scale_factor);
        float nd = p->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
        int sum = 0;
        for (int l = 0; l < nlevels - 1; l++) {
            quota[l] = cv_round(nd);
            sum += quota[l];
            nd *= factor;
        }
        quota[nlevels - 1] = std::max(p->nfeatures - sum, 0);
    }
    int umax[17];
    {
        const int vmax = cv_floor(half * std::sqrt(2.f) / 2 + 1), vmin = cv_ceil(half * std::sqrt(2.f) / 2);
        for (int v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt((double)half * half - v * v));
        for (int v = half, v0 = 0; v >= vmin; --v) {
            while (umax[v0] == umax[v0 + 1]) ++v0;
            umax[v] = v0;
            ++v0;
        }
    }
    std::vector<int> counters(nlevels);
    std::vector<FastPoint> fp;
    const bool harris = p->score_type == 0;
    for (int l = 0; l < nlevels; l++) {
        const int w = L[l].w, h = L[l].h;
        fast9_16_nms(L[l].img.data(), w, h, w, p->fast_threshold, fp);
        std::vector<KP> k;
        // runByImageBorder(keypoints, img.size(), edgeThreshold)
        if (!(h <= edge * 2 || w <= edge * 2))
            for (const FastPoint &q : fp)
                if (q.x >= edge && q.x < w - edge && q.y >= edge && q.y < h - edge)
                    k.push_back(KP{(float)q.x, (float)q.y, 7.f, -1.f, (float)q.response, 0, -1});
        retain_best(k, harris ? 2 * quota[l] : quota[l]);
        counters[l] = (int)k.size();
        for (KP &q : k) {
            q.octave = l;
            q.size = patch * L[l].scale;
        }
        all.insert(all.end(), k.begin(), k.end());
    }
    if (all.empty()) {
        if (levels_out) *levels_out = std::move(L);
        return ORB_OK;
    }
    if (harris) {
        // HarrisResponses(imagePyramid, layerInfo, allKeypoints, 7, 0.04f)
        const int bs = 7, r = bs / 2;
        const float harris_k = 0.04f;
        const float scale = 1.f / ((1 << 2) * bs * 255.f);
        const float scale_sq_sq = scale * scale * scale * scale;
        for (KP &q : all) {
            const Level &lv = L[q.octave];
            const int step = lv.w, x0 = cv_round(q.x), y0 = cv_round(q.y);
            const uint8_t *ptr0 = lv.img.data() + (size_t)(y0 - r) * step + (x0 - r);
            int a = 0, b = 0, c = 0;
            for (int i = 0; i < bs; i++)
                for (int j = 0; j < bs; j++) {
                    const uint8_t *ptr = ptr0 + i * step + j;
                    const int Ix = (ptr[1] - ptr[-1]) * 2 + (ptr[-step + 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[step - 1]);
                    const int Iy = (ptr[step] - ptr[-step]) * 2 + (ptr[step - 1] - ptr[-step - 1]) + (ptr[step + 1] - ptr[-step + 1]);
                    a += Ix * Ix;
                    b += Iy * Iy;
                    c += Ix * Iy;
                }
            q.response = ((float)a * b - (float)c * c - harris_k * ((float)a + b) * ((float)a + b)) * scale_sq_sq;
        }
        std::vector<KP> nall;
        int offset = 0;
        for (int l = 0; l < nlevels; l++) {
            std::vector<KP> k(all.begin() + offset, all.begin() + offset + counters[l]);
            offset += counters[l];
            retain_best(k, quota[l]);
            nall.insert(nall.end(), k.begin(), k.end());
        }
        all.swap(nall);
    }
    // ICAngles
    for (KP &q : all) {
        const Level &lv = L[q.octave];
        const int step = lv.w;
        const uint8_t *center = lv.img.data() + (size_t)cv_round(q.y) * step + cv_round(q.x);
        int m_01 = 0, m_10 = 0;
        for (int u = -half; u <= half; ++u) m_10 += u * center[u];
        for (int v = 1; v <= half; ++v) {
            int v_sum = 0;
            const int d = umax[v];
            for (int u = -d; u <= d; ++u) {
                const int val_plus = center[u + v * step], val_minus = center[u - v * step];
                v_sum += (val_plus - val_minus);
                m_10 += u * (val_plus + val_minus);
            }
            m_01 += v * v_sum;
        }
        q.angle = fast_atan2((float)m_01, (float)m_10);
    }
    for (KP &q : all) {
        const float scale = L[q.octave].scale;
        q.x *= scale;
        q.y *= scale;
    }
    // descriptors: GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) of every layer in place, then computeOrbDescriptors (WTA_K = 2)
    for (int l = 0; l < nlevels; l++) {
        std::vector<uint8_t> b((size_t)L[l].w * L[l].h);
        gaussian_blur7_sep32f_u8(L[l].img.data(), L[l].w, L[l].h, L[l].w, b.data(), L[l].w);
        L[l].img.swap(b);
    }
    desc.assign(all.size() * 32, 0);
    static const int pat[ORB_PATTERN_INTS] = {ORB_PATTERN_VALUES};   // OpenCV's bit_pattern_31_ is the table ORB-SLAM2 copied
    for (size_t j = 0; j < all.size(); j++) {
        const KP &q = all[j];
        const Level &lv = L[q.octave];
        const int step = lv.w;
        const float scale = 1.f / lv.scale;
        float angle = q.angle;
        angle *= (float)(3.1415926535897932384626433832795 / 180.f);
        float a, b;   // `(float)cos(angle)` on a float argument is cosf (libstdc++'s global overloads), as in ORBextractor.cc:103-104
        sincosf_libm(angle, &b, &a);
        const uint8_t *center = lv.img.data() + (size_t)cv_round(q.y * scale) * step + cv_round(q.x * scale);
        uint8_t *d = &desc[j * 32];
        for (int i = 0; i < 32; i++) {
            int val = 0;
            for (int k = 0; k < 8; k++) {
                const int *pp = pat + (i * 16 + k * 2) * 2;
                int t[2];
                for (int s = 0; s < 2; s++) {
                    const float px = (float)pp[s * 2], py = (float)pp[s * 2 + 1];
                    const float x = px * a - py * b, y = px * b + py * a;
                    t[s] = center[cv_round(y) * step + cv_round(x)];
                }
                val |= (t[0] < t[1]) << k;
            }
            d[i] = (uint8_t)val;
        }
    }
    if (levels_out) *levels_out = std::move(L);
    return ORB_OK;
}
}  // namespace

extern "C" {

int orbref_gaussian_blur7_sep32f(const uint8_t *src, int w, int h, int spitch, uint8_t *dst, int dpitch) {
    if (!src || !dst || w < 1 || h < 1) return ORB_ERR_INVALID;
    gaussian_blur7_sep32f_u8(src, w, h, spitch, dst, dpitch);
    return ORB_OK;
}

int orbref_resize_linear_exact(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst, int dw, int dh, int dpitch) {
    if (!src || !dst || sw < 1 || sh < 1 || dw < 1 || dh < 1) return ORB_ERR_INVALID;
    resize_linear_exact_u8(src, sw, sh, spitch, dst, dw, dh, dpitch);
    return ORB_OK;
}

int orbref_cvorb_detect_and_compute(const orb_cvorb_params *p, const uint8_t *image, int width, int height, int pitch,
                                    orb_keypoint *kps, uint8_t *desc, int capacity, int *n_out) {
    if (!p || !n_out) return ORB_ERR_INVALID;
    std::vector<KP> all;
    std::vector<uint8_t> d;
    const int rc = run(p, image, width, height, pitch, all, d, nullptr);
    if (rc) return rc;
    *n_out = (int)all.size();
    if ((int)all.size() > capacity) return ORB_ERR_CAPACITY;
    for (size_t i = 0; i < all.size(); i++) {
        kps[i].x = all[i].x; kps[i].y = all[i].y; kps[i].size = all[i].size; kps[i].angle = all[i].angle;
        kps[i].response = all[i].response; kps[i].octave = all[i].octave; kps[i].class_id = all[i].class_id;
    }
    if (!d.empty()) std::memcpy(desc, d.data(), d.size());
    return ORB_OK;
}

/* KeyPointsFilter::retainBest on bare responses: std::nth_element + std::partition of this libstdc++ (depth_limit < 0), or
 * libstdc++'s own __introselect with a forced depth limit (the heap-select fallback the device restates). */
int orbref_retain_best(const float *responses, int n, int n_points, int depth_limit, int32_t *order_out, int *n_out) {
    std::vector<KP> k(n);
    for (int i = 0; i < n; i++) k[i] = KP{0, 0, 0, 0, responses[i], i, 0};
    if (depth_limit < 0) retain_best(k, n_points);
    else if (n_points >= 0 && k.size() > (size_t)n_points) {
        if (n_points == 0) k.clear();
        else {
            std::__introselect(k.begin(), k.begin() + n_points - 1, k.end(), depth_limit, __gnu_cxx::__ops::__iter_comp_iter(ResponseGreater()));
            const float amb = k[n_points - 1].response;
            auto new_end = std::partition(k.begin() + n_points, k.end(), [amb](const KP &p) { return p.response >= amb; });
            k.resize(new_end - k.begin());
        }
    }
    for (size_t i = 0; i < k.size(); i++) order_out[i] = k[i].octave;
    *n_out = (int)k.size();
    return ORB_OK;
}

/* layer `level` after detectAndCompute (blurred = 1) or of the pyramid alone (blurred = 0) */
int orbref_cvorb_level(const orb_cvorb_params *p, const uint8_t *image, int width, int height, int pitch, int level, int blurred,
                       uint8_t *dst, int dst_pitch, int *w, int *h) {
    if (!p || level < 0 || level >= p->nlevels) return ORB_ERR_INVALID;
    std::vector<KP> all;
    std::vector<uint8_t> d;
    std::vector<Level> L;
    orb_cvorb_params q = *p;
    if (!blurred) q.nfeatures = 0;   // no key points -> the layers are returned before the blur
    const int rc = run(&q, image, width, height, pitch, all, d, &L);
    if (rc) return rc;
    if (w) *w = L[level].w;
    if (h) *h = L[level].h;
    if (dst)
        for (int y = 0; y < L[level].h; y++) std::memcpy(dst + (size_t)y * dst_pitch, &L[level].img[(size_t)y * L[level].w], L[level].w);
    return ORB_OK;
}
}
