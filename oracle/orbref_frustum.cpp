// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of Frame::isInFrustum (ORB_SLAM2/src/Frame.cc:274-330) + MapPoint::PredictScale (src/MapPoint.cc:357-366)
// on flat arrays.  cv::Mat algebra = double accumulation rounded to float (cv::gemm / cv::norm / Mat::dot); log = the host
// libm's logf (the reference calls std::log(float)).  Pinned against the reference's own Frame.cc compiled unmodified
// (oracle/_ref/liborbmatcher_ref.so, refm_is_in_frustum) by tests/test_frustum_cpu.py.
#include "orbref.h"

#include <cmath>

extern "C" int orbref_is_in_frustum(const orb_frustum_camera *cam, const orb_frustum_points *pts, float limit, uint8_t *in_view,
                                    float *proj_x, float *proj_y, float *proj_xr, int32_t *level, float *view_cos) {
    for (int i = 0; i < pts->n; i++) {
        in_view[i] = 0; proj_x[i] = proj_y[i] = proj_xr[i] = view_cos[i] = 0.0f; level[i] = 0;
        if (pts->consider && !pts->consider[i]) continue;
        const float *P = pts->world_pos + 3 * i, *Pn = pts->normal + 3 * i;
        float Pc[3];
        for (int r = 0; r < 3; r++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += (double)cam->Rcw[3 * r + k] * (double)P[k];
            Pc[r] = (float)s + cam->tcw[r];
        }
        if (Pc[2] < 0.0f) continue;
        const float invz = 1.0f / Pc[2];
        const float u = cam->fx * Pc[0] * invz + cam->cx, v = cam->fy * Pc[1] * invz + cam->cy;
        if (u < cam->min_x || u > cam->max_x) continue;
        if (v < cam->min_y || v > cam->max_y) continue;
        const float maxDistance = 1.2f * pts->max_distance[i], minDistance = 0.8f * pts->min_distance[i];
        const float PO[3] = {P[0] - cam->Ow[0], P[1] - cam->Ow[1], P[2] - cam->Ow[2]};
        double s2 = 0, sd = 0;
        for (int k = 0; k < 3; k++) { s2 += (double)PO[k] * (double)PO[k]; sd += (double)PO[k] * (double)Pn[k]; }
        const float dist = (float)std::sqrt(s2);
        if (dist < minDistance || dist > maxDistance) continue;
        const float viewCos = (float)(sd / dist);
        if (viewCos < limit) continue;
        const float ratio = pts->max_distance[i] / dist;
        in_view[i] = 1; proj_x[i] = u; proj_y[i] = v; proj_xr[i] = u - cam->mbf * invz; view_cos[i] = viewCos;
        level[i] = (int)std::ceil(std::log(ratio) / cam->log_scale_factor);
    }
    return ORB_OK;
}

// The projection pre-loops of SearchByProjection(Frame&, KeyFrame*) (src/ORBmatcher.cc:1491-1531), SearchByProjection(KeyFrame*,
// Scw) (:311-360), Fuse (:846-893, :1004-1048) and SearchBySim3 (:1150-1226) on flat arrays, the sites' differences as the rules
// of orb_projection (include/orbgpu.h).  Pinned against the reference's lines replayed by oracle/ref_matcher_driver.cpp
// (refm_search_by_projection_kf / _sim3 return u, v, valid of every point) in tests/test_projection_cpu.py.
extern "C" int orbref_project_points(const orb_projection *P, const orb_projection_points *pts, uint8_t *ok, float *pu, float *pv,
                                     float *pur, float *pdist) {
    for (int i = 0; i < pts->n; i++) {
        ok[i] = 0; pu[i] = pv[i] = pur[i] = pdist[i] = 0.0f;
        if (pts->consider && !pts->consider[i]) continue;
        const float *X = pts->world_pos + 3 * i;
        float a[3], c[3];
        for (int r = 0; r < 3; r++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += (double)P->R1[3 * r + k] * (double)X[k];
            a[r] = (float)s + P->t1[r];
        }
        for (int r = 0; r < 3; r++) {
            if (!P->two_stage) { c[r] = a[r]; continue; }
            double s = 0;
            for (int k = 0; k < 3; k++) s += (double)P->R2[3 * r + k] * (double)a[k];
            c[r] = (float)s + P->t2[r];
        }
        if (P->reject_negative_z && c[2] < 0.0f) continue;
        const float invz = P->invz_rule == 0 ? (float)(1.0 / c[2]) : 1.0f / c[2];
        float u, v;
        if (P->uv_rule == 0) { u = P->fx * c[0] * invz + P->cx; v = P->fy * c[1] * invz + P->cy; }
        else { const float x = c[0] * invz, y = c[1] * invz; u = P->fx * x + P->cx; v = P->fy * y + P->cy; }
        if (P->bounds_rule == 0) { if (u < P->min_x || u > P->max_x || v < P->min_y || v > P->max_y) continue; }
        else if (!(u >= P->min_x && u < P->max_x && v >= P->min_y && v < P->max_y)) continue;
        double s2 = 0, sd = 0;
        if (P->dist_rule == 0) {
            const float PO[3] = {X[0] - P->Ow[0], X[1] - P->Ow[1], X[2] - P->Ow[2]};
            for (int k = 0; k < 3; k++) { s2 += (double)PO[k] * (double)PO[k]; if (P->check_view_angle) sd += (double)PO[k] * (double)pts->normal[3 * i + k]; }
        } else {
            for (int k = 0; k < 3; k++) s2 += (double)c[k] * (double)c[k];
        }
        const float dist = (float)std::sqrt(s2);
        if (dist < pts->min_distance[i] || dist > pts->max_distance[i]) continue;
        if (P->check_view_angle && sd < 0.5 * dist) continue;
        ok[i] = 1; pu[i] = u; pv[i] = v; pur[i] = u - P->bf * invz; pdist[i] = dist;
    }
    return ORB_OK;
}
