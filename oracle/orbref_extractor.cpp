// ORACLE -- TEST INFRASTRUCTURE ONLY (see orbref_cv.h).  CPU restatement of the reference's
// ORBextractor (ORB_SLAM2/src/ORBextractor.cc, ORB_SLAM2/include/ORBextractor.h), written to
// follow the reference's control flow so its output order and tie-breaks are the reference's.
// Build: g++ -O2 -ffp-contract=off (no FMA contraction).
//
// Two choices the reference leaves to the platform are FIXED here:
//   * DistributeOctTree sorts (size, node pointer); ties are broken by heap address in the
//     reference.  Here the tie key is the node's creation sequence number, which is what the
//     reference produces under a monotonic allocator (oracle/_ref proves that).
//   * cosf/sinf are the glibc algorithm restated in orbref_cv.h (bit-identical to glibc 2.39).
#include "orbref.h"
#include "orbref_cv.h"
#include "orbref_internal.h"
#include "../ar_orbslam2_b200/csrc/orb_pattern.h"

#include <algorithm>
#include <list>
#include <utility>

using namespace orbref;

namespace {

const int PATCH_SIZE = 31;        // ref ORBextractor.cc:69
const int HALF_PATCH_SIZE = 15;   // ref :70

const int kPattern[ORB_PATTERN_INTS] = {ORB_PATTERN_VALUES};

}  // namespace


// ref ExtractorNode::DivideNode, ORBextractor.cc:470-523
void orbref_extractor::divide(const Node &n, const std::vector<Cand> &c, Node out[4]) {
    const int halfX = (int)std::ceil(static_cast<float>(n.URx - n.ULx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(n.BRy - n.ULy) / 2);
    Node &n1 = out[0], &n2 = out[1], &n3 = out[2], &n4 = out[3];
    n1.ULx = n.ULx; n1.ULy = n.ULy;
    n1.URx = n.ULx + halfX; n1.URy = n.ULy;
    n1.BLx = n.ULx; n1.BLy = n.ULy + halfY;
    n1.BRx = n.ULx + halfX; n1.BRy = n.ULy + halfY;

    n2.ULx = n1.URx; n2.ULy = n1.URy;
    n2.URx = n.URx; n2.URy = n.URy;
    n2.BLx = n1.BRx; n2.BLy = n1.BRy;
    n2.BRx = n.URx; n2.BRy = n.ULy + halfY;

    n3.ULx = n1.BLx; n3.ULy = n1.BLy;
    n3.URx = n1.BRx; n3.URy = n1.BRy;
    n3.BLx = n.BLx; n3.BLy = n.BLy;
    n3.BRx = n1.BRx; n3.BRy = n.BLy;

    n4.ULx = n3.URx; n4.ULy = n3.URy;
    n4.URx = n2.BRx; n4.URy = n2.BRy;
    n4.BLx = n3.BRx; n4.BLy = n3.BRy;
    n4.BRx = n.BRx; n4.BRy = n.BRy;

    for (size_t i = 0; i < n.keys.size(); i++) {
        const Cand &kp = c[n.keys[i]];
        if (kp.x < n1.URx) {                      // float vs int compare, ref :503
            if (kp.y < n1.BRy) n1.keys.push_back(n.keys[i]);
            else n3.keys.push_back(n.keys[i]);
        } else if (kp.y < n1.BRy)
            n2.keys.push_back(n.keys[i]);
        else
            n4.keys.push_back(n.keys[i]);
    }
    for (int k = 0; k < 4; k++)
        if (out[k].keys.size() == 1) out[k].no_more = true;
}

// ref ORBextractor::DistributeOctTree, ORBextractor.cc:525-733
std::vector<Cand> orbref_extractor::distribute(const std::vector<Cand> &c, int minX, int maxX, int minY,
                                               int maxY, int N) {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;

    std::list<Node> nodes;
    std::vector<Node *> ini(nIni);
    node_seq = 0;
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * static_cast<float>(i)); ni.ULy = 0;
        ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
        ni.BLx = ni.ULx; ni.BLy = maxY - minY;
        ni.BRx = ni.URx; ni.BRy = maxY - minY;
        ni.seq = node_seq++;
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (size_t i = 0; i < c.size(); i++) ini[(int)(c[i].x / hX)]->keys.push_back((int)i);

    for (auto lit = nodes.begin(); lit != nodes.end();) {
        if (lit->keys.size() == 1) { lit->no_more = true; ++lit; }
        else if (lit->keys.empty()) lit = nodes.erase(lit);
        else ++lit;
    }

    bool finish = false;
    // (size, creation seq, node): the reference keys on (size, pointer)
    typedef std::pair<std::pair<int, long>, Node *> SizeNode;
    std::vector<SizeNode> size_and_node;

    auto add_children = [&](Node ch[4], int *n_to_expand) {
        for (int k = 0; k < 4; k++) {
            if (ch[k].keys.size() > 0) {
                ch[k].seq = node_seq++;
                nodes.push_front(ch[k]);
                if (ch[k].keys.size() > 1) {
                    if (n_to_expand) (*n_to_expand)++;
                    size_and_node.push_back(SizeNode(std::make_pair((int)ch[k].keys.size(), nodes.front().seq), &nodes.front()));
                    nodes.front().lit = nodes.begin();
                }
            }
        }
    };

    while (!finish) {
        int prev_size = (int)nodes.size();
        auto lit = nodes.begin();
        int n_to_expand = 0;
        size_and_node.clear();

        while (lit != nodes.end()) {
            if (lit->no_more) { ++lit; continue; }
            Node ch[4];
            divide(*lit, c, ch);
            add_children(ch, &n_to_expand);
            lit = nodes.erase(lit);
        }

        if ((int)nodes.size() >= N || (int)nodes.size() == prev_size) {
            finish = true;
        } else if (((int)nodes.size() + n_to_expand * 3) > N) {
            while (!finish) {
                prev_size = (int)nodes.size();
                std::vector<SizeNode> prev = size_and_node;
                size_and_node.clear();
                std::sort(prev.begin(), prev.end());
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node ch[4];
                    divide(*prev[j].second, c, ch);
                    add_children(ch, nullptr);
                    nodes.erase(prev[j].second->lit);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prev_size) finish = true;
            }
        }
    }

    std::vector<Cand> result;
    for (auto lit = nodes.begin(); lit != nodes.end(); ++lit) {
        const std::vector<int> &k = lit->keys;
        int best = k[0];
        float max_response = c[best].response;
        for (size_t i = 1; i < k.size(); i++)
            if (c[k[i]].response > max_response) { best = k[i]; max_response = c[k[i]].response; }
        result.push_back(c[best]);
    }
    return result;
}

// ref IC_Angle, ORBextractor.cc:73-98
static float ic_angle(const uint8_t *center, int step, const std::vector<int> &u_max) {
    int m_01 = 0, m_10 = 0;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0, d = u_max[v];
        for (int u = -d; u <= d; ++u) {
            int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
}

// ref computeOrbDescriptor, ORBextractor.cc:100-144
static void orb_descriptor(float kp_angle, const uint8_t *center, int step, int trig_mode, uint8_t *desc) {
    const float factorPI = (float)(M_PI / 180.f);
    float angle = kp_angle * factorPI, a, b;
    if (trig_mode == 1) { a = cosf(angle); b = sinf(angle); }
    else sincosf_libm(angle, &b, &a);
    const int *pat = kPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int k = 0; k < 8; k++) {
            const float x0 = (float)pat[4 * k], y0 = (float)pat[4 * k + 1];
            const float x1 = (float)pat[4 * k + 2], y1 = (float)pat[4 * k + 3];
            int t0 = center[cv_round(x0 * b + y0 * a) * step + cv_round(x0 * a - y0 * b)];
            int t1 = center[cv_round(x1 * b + y1 * a) * step + cv_round(x1 * a - y1 * b)];
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

extern "C" {

// ref ORBextractor::ORBextractor, ORBextractor.cc:404-460
int orbref_extractor_create(const orb_params *p, orbref_extractor **out) {
    if (!p || !out || p->nlevels < 1 || p->nlevels > ORB_MAX_LEVELS || p->nfeatures < 1) return ORB_ERR_INVALID;
    orbref_extractor *ex = new orbref_extractor();
    ex->prm = *p;
    const int nlevels = p->nlevels;
    const double scaleFactor = p->scale_factor;   // member is double, ORBextractor.h:105
    ex->scale.resize(nlevels); ex->sigma2.resize(nlevels);
    ex->inv_scale.resize(nlevels); ex->inv_sigma2.resize(nlevels);
    ex->scale[0] = 1.0f; ex->sigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        ex->scale[i] = (float)(ex->scale[i - 1] * scaleFactor);
        ex->sigma2[i] = ex->scale[i] * ex->scale[i];
    }
    for (int i = 0; i < nlevels; i++) {
        ex->inv_scale[i] = 1.0f / ex->scale[i];
        ex->inv_sigma2[i] = 1.0f / ex->sigma2[i];
    }
    ex->per_level.resize(nlevels);
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = p->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        ex->per_level[level] = cv_round(nDesired);
        sum += ex->per_level[level];
        nDesired *= factor;
    }
    ex->per_level[nlevels - 1] = std::max(p->nfeatures - sum, 0);

    ex->umax.assign(HALF_PATCH_SIZE + 1, 0);
    int v, v0, vmax = cv_floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = cv_ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) ex->umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (ex->umax[v0] == ex->umax[v0 + 1]) ++v0;
        ex->umax[v] = v0;
        ++v0;
    }
    ex->lv.resize(nlevels);
    *out = ex;
    return ORB_OK;
}

void orbref_extractor_destroy(orbref_extractor *ex) { delete ex; }

int orbref_extractor_set_trig_mode(orbref_extractor *ex, int mode) { ex->trig_mode = mode; return ORB_OK; }

int orbref_extractor_tables(const orbref_extractor *ex, float *scale, float *inv_scale, float *sigma2,
                            float *inv_sigma2, int32_t *per_level, int32_t *umax16) {
    for (int i = 0; i < ex->prm.nlevels; i++) {
        if (scale) scale[i] = ex->scale[i];
        if (inv_scale) inv_scale[i] = ex->inv_scale[i];
        if (sigma2) sigma2[i] = ex->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = ex->inv_sigma2[i];
        if (per_level) per_level[i] = ex->per_level[i];
    }
    if (umax16) for (int i = 0; i < 16; i++) umax16[i] = ex->umax[i];
    return ORB_OK;
}

// Level geometry check shared with the GPU library's ORB_ERR_TOO_SMALL rule: the reference
// divides by zero when a level's detection window is under 30 px (nCols/nRows = 0,
// ORBextractor.cc:753-756) or nIni rounds to 0 (:530-532).
static int level_geometry_ok(int w, int h) {
    const int minB = EDGE_THRESHOLD - 3, maxBX = w - EDGE_THRESHOLD + 3, maxBY = h - EDGE_THRESHOLD + 3;
    const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
    if (width < 30.f || height < 30.f) return 0;
    const int nIni = (int)std::round(static_cast<float>(maxBX - minB) / (maxBY - minB));
    return nIni >= 1;
}

// ref ORBextractor::operator(), ORBextractor.cc:985-1045
int orbref_extract(orbref_extractor *ex, const uint8_t *image, int width, int height, int pitch,
                   orb_keypoint *kps, uint8_t *desc, int capacity, int *n_out) {
    if (!ex || !n_out) return ORB_ERR_INVALID;
    *n_out = 0;
    if (width == 0 || height == 0 || !image) return ORB_OK;   // ref :987-988
    const int nlevels = ex->prm.nlevels;

    // ---- ComputePyramid, ref :1047-1072
    for (int level = 0; level < nlevels; ++level) {
        Level &L = ex->lv[level];
        float scale = ex->inv_scale[level];
        L.w = cv_round((float)width * scale);
        L.h = cv_round((float)height * scale);
        if (!level_geometry_ok(L.w, L.h)) return ORB_ERR_TOO_SMALL;
        L.pitch = L.w + 2 * EDGE_THRESHOLD;
        L.plane.assign((size_t)L.pitch * (L.h + 2 * EDGE_THRESHOLD), 0);
        L.cands.clear(); L.kept.clear(); L.blurred.clear();
        if (level != 0) {
            Level &P = ex->lv[level - 1];
            resize_linear_u8(P.interior(), P.w, P.h, P.pitch, L.interior(), L.w, L.h, L.pitch);
        } else {
            for (int y = 0; y < height; y++) std::memcpy(L.interior() + (size_t)y * L.pitch, image + (size_t)y * pitch, width);
        }
        make_border_reflect101(L.interior(), L.w, L.h, L.pitch, EDGE_THRESHOLD);
    }

    // ---- ComputeKeyPointsOctTree, ref :735-819
    const float W = 30;
    std::vector<FastPoint> cell;
    for (int level = 0; level < nlevels; ++level) {
        Level &L = ex->lv[level];
        const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
        const int maxBorderX = L.w - EDGE_THRESHOLD + 3, maxBorderY = L.h - EDGE_THRESHOLD + 3;
        const float width_f = (float)(maxBorderX - minBorderX), height_f = (float)(maxBorderY - minBorderY);
        const int nCols = (int)(width_f / W), nRows = (int)(height_f / W);
        const int wCell = (int)std::ceil(width_f / nCols), hCell = (int)std::ceil(height_f / nRows);
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBorderY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3) continue;
            if (maxY > maxBorderY) maxY = (float)maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBorderX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6) continue;
                if (maxX > maxBorderX) maxX = (float)maxBorderX;
                const int x0 = (int)iniX, y0 = (int)iniY, x1 = (int)maxX, y1 = (int)maxY;
                const uint8_t *roi = L.interior() + (ptrdiff_t)y0 * L.pitch + x0;
                fast9_16_nms(roi, x1 - x0, y1 - y0, L.pitch, ex->prm.ini_th_fast, cell);
                if (cell.empty()) fast9_16_nms(roi, x1 - x0, y1 - y0, L.pitch, ex->prm.min_th_fast, cell);
                for (const FastPoint &fp : cell) {
                    Cand c;
                    c.x = (float)fp.x + j * wCell;
                    c.y = (float)fp.y + i * hCell;
                    c.response = (float)fp.response;
                    c.angle = -1.f;
                    L.cands.push_back(c);
                }
            }
        }
        L.kept = ex->distribute(L.cands, minBorderX, maxBorderX, minBorderY, maxBorderY, ex->per_level[level]);
        for (Cand &k : L.kept) { k.x += minBorderX; k.y += minBorderY; }
    }
    for (int level = 0; level < nlevels; ++level) {   // ref :816-818, computeOrientation :462-468
        Level &L = ex->lv[level];
        for (Cand &k : L.kept)
            k.angle = ic_angle(L.interior() + (ptrdiff_t)cv_round(k.y) * L.pitch + cv_round(k.x), L.pitch, ex->umax);
    }

    int total = 0;
    for (int level = 0; level < nlevels; ++level) total += (int)ex->lv[level].kept.size();
    *n_out = total;
    if (total > capacity) return ORB_ERR_CAPACITY;

    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
        Level &L = ex->lv[level];
        if (L.kept.empty()) continue;                                    // ref :1020-1021
        // clone + GaussianBlur on the un-padded level, ref :1024-1026
        L.blurred.assign((size_t)L.w * L.h, 0);
        gaussian_blur7_u8(L.interior(), L.w, L.h, L.pitch, L.blurred.data(), L.w, ex->prm.blur_taps);
        const int scaledPatchSize = (int)(PATCH_SIZE * ex->scale[level]);   // ref :804
        for (size_t i = 0; i < L.kept.size(); i++) {
            const Cand &k = L.kept[i];
            if (desc)
                orb_descriptor(k.angle, L.blurred.data() + (size_t)cv_round(k.y) * L.w + cv_round(k.x), L.w,
                               ex->trig_mode, desc + (size_t)(offset + i) * 32);
            if (kps) {
                orb_keypoint &o = kps[offset + i];
                o.x = k.x; o.y = k.y;
                if (level != 0) { float s = ex->scale[level]; o.x *= s; o.y *= s; }   // ref :1035-1041
                o.size = (float)scaledPatchSize;
                o.angle = k.angle;
                o.response = k.response;
                o.octave = level;
                o.class_id = -1;
            }
        }
        offset += (int)L.kept.size();
    }
    return ORB_OK;
}

int orbref_level_size(const orbref_extractor *ex, int level, int *w, int *h) {
    if (!ex || level < 0 || level >= ex->prm.nlevels) return ORB_ERR_INVALID;
    *w = ex->lv[level].w; *h = ex->lv[level].h;
    return ORB_OK;
}

int orbref_pyramid_download(orbref_extractor *ex, int level, int with_border, uint8_t *dst, int dst_pitch) {
    if (!ex || level < 0 || level >= ex->prm.nlevels) return ORB_ERR_INVALID;
    Level &L = ex->lv[level];
    if (with_border) {
        for (int y = 0; y < L.h + 2 * EDGE_THRESHOLD; y++)
            std::memcpy(dst + (size_t)y * dst_pitch, L.plane.data() + (size_t)y * L.pitch, L.pitch);
    } else {
        for (int y = 0; y < L.h; y++) std::memcpy(dst + (size_t)y * dst_pitch, L.interior() + (size_t)y * L.pitch, L.w);
    }
    return ORB_OK;
}

int orbref_debug_candidates(orbref_extractor *ex, int level, orb_candidate *out, int capacity, int *n_out) {
    if (!ex || level < 0 || level >= ex->prm.nlevels) return ORB_ERR_INVALID;
    Level &L = ex->lv[level];
    *n_out = (int)L.cands.size();
    if (*n_out > capacity) return ORB_ERR_CAPACITY;
    for (size_t i = 0; i < L.cands.size(); i++) {
        out[i].x = (int16_t)L.cands[i].x; out[i].y = (int16_t)L.cands[i].y;
        out[i].response = (int32_t)L.cands[i].response;
    }
    return ORB_OK;
}

int orbref_debug_blur_level(orbref_extractor *ex, int level, uint8_t *dst, int dst_pitch) {
    if (!ex || level < 0 || level >= ex->prm.nlevels) return ORB_ERR_INVALID;
    Level &L = ex->lv[level];
    std::vector<uint8_t> tmp((size_t)L.w * L.h);
    gaussian_blur7_u8(L.interior(), L.w, L.h, L.pitch, tmp.data(), L.w, ex->prm.blur_taps);
    for (int y = 0; y < L.h; y++) std::memcpy(dst + (size_t)y * dst_pitch, tmp.data() + (size_t)y * L.w, L.w);
    return ORB_OK;
}

// Primitive entry points for the cv2 golden-vector tests.
int orbref_resize_linear(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst, int dw, int dh, int dpitch) {
    resize_linear_u8(src, sw, sh, spitch, dst, dw, dh, dpitch);
    return ORB_OK;
}
int orbref_gaussian_blur7(const uint8_t *src, int w, int h, int spitch, uint8_t *dst, int dpitch, int taps) {
    gaussian_blur7_u8(src, w, h, spitch, dst, dpitch, taps);
    return ORB_OK;
}
int orbref_fast(const uint8_t *img, int w, int h, int pitch, int threshold, orb_candidate *out, int capacity, int *n_out) {
    std::vector<FastPoint> pts;
    fast9_16_nms(img, w, h, pitch, threshold, pts);
    *n_out = (int)pts.size();
    if (*n_out > capacity) return ORB_ERR_CAPACITY;
    for (size_t i = 0; i < pts.size(); i++) { out[i].x = (int16_t)pts[i].x; out[i].y = (int16_t)pts[i].y; out[i].response = pts[i].response; }
    return ORB_OK;
}
int orbref_fast_atan2(const float *y, const float *x, int n, float *out) {
    for (int i = 0; i < n; i++) out[i] = fast_atan2(y[i], x[i]);
    return ORB_OK;
}
int orbref_sincos(const float *angle_rad, int n, float *s, float *c, int mode) {
    for (int i = 0; i < n; i++) {
        if (mode == 1) { s[i] = sinf(angle_rad[i]); c[i] = cosf(angle_rad[i]); }
        else sincosf_libm(angle_rad[i], &s[i], &c[i]);
    }
    return ORB_OK;
}
int orbref_make_border(uint8_t *plane, int w, int h, int pitch, int border) {
    make_border_reflect101(plane + (size_t)border * pitch + border, w, h, pitch, border);
    return ORB_OK;
}

}  // extern "C"
