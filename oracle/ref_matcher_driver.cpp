// ORACLE -- TEST INFRASTRUCTURE ONLY.  C entry points around the reference's own ORB_SLAM2::ORBmatcher, compiled from
// /root/reference/ORB_SLAM2/src/ORBmatcher.cc (unmodified) against oracle/refstub/slam_stub.h + oracle/cvstub
// (see oracle/Makefile.ref).  Each function builds the stub Frame / KeyFrame / MapPoint objects from the same flat
// arrays the C ABI of include/orbgpu.h takes, calls the reference method, and flattens the result the same way, so
// tests can compare reference source == restated oracle == GPU.
// With -DUSE_GPU_SHIM the SAME driver is compiled with ar_orbslam2_b200/shim in front of the reference's include directory:
// "ORBmatcher.h" / "ORBextractor.h" then resolve to the product's drop-in classes and every call below -- unchanged --
// runs on liborbgpu (oracle/_ref/liborbmatcher_shim.so).  That is the drop-in claim, tested.
#include "ORBmatcher.h"
#include "../include/orbgpu.h"

#include <chrono>
#include <cstring>
#include <memory>

using namespace ORB_SLAM2;

#ifndef ORBREF_REAL_FRAME
float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;
#endif

// bench.py's C3 leg: wall time of the last matcher call alone (the stub object set-up around it is not the reference's cost)
static double g_call_us = 0;
struct CallTimer {
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    ~CallTimer() { g_call_us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count(); }
};

struct refm_frame {
    Frame F;
    orb_frame_desc d;
};

static cv::Mat desc_row(const uint8_t *p) {
    cv::Mat m(1, 32, CV_8UC1);
    memcpy(m.data, p, 32);
    return m;
}

static void set_statics(const refm_frame *f) {   // the reference keeps the image bounds / grid scale in statics (Frame.h:161-162,191-194)
    Frame::mnMinX = f->d.min_x; Frame::mnMaxX = f->d.max_x; Frame::mnMinY = f->d.min_y; Frame::mnMaxY = f->d.max_y;
    Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(Frame::mnMaxX - Frame::mnMinX);   // ref Frame.cc:217-218
    Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(Frame::mnMaxY - Frame::mnMinY);
}

// KeyFrame view of the same data (ref KeyFrame.cc:31-63 copies these members from the Frame)
static std::unique_ptr<KeyFrame> as_keyframe(const refm_frame *f) {
    std::unique_ptr<KeyFrame> k(new KeyFrame());
    k->N = f->F.N; k->mvKeysUn = f->F.mvKeysUn; k->mvuRight = f->F.mvuRight; k->mDescriptors = f->F.mDescriptors;
    k->mvScaleFactors = f->F.mvScaleFactors; k->mfLogScaleFactor = f->F.mfLogScaleFactor;
    for (size_t i = 0; i < k->mvScaleFactors.size(); i++) {   // ref ORBextractor.cc:417-428 via KeyFrame.cc:38-40
        k->mvLevelSigma2.push_back(k->mvScaleFactors[i] * k->mvScaleFactors[i]);
        k->mvInvLevelSigma2.push_back(1.0f / k->mvLevelSigma2[i]);
    }
    k->mvpMapPoints.assign(k->N, nullptr);
    k->mnMinX = (int)f->d.min_x; k->mnMaxX = (int)f->d.max_x; k->mnMinY = (int)f->d.min_y; k->mnMaxY = (int)f->d.max_y;
    k->mfGridElementWidthInv = Frame::mfGridElementWidthInv; k->mfGridElementHeightInv = Frame::mfGridElementHeightInv;
    k->mGrid.resize(FRAME_GRID_COLS);
    for (int i = 0; i < FRAME_GRID_COLS; i++) {
        k->mGrid[i].resize(FRAME_GRID_ROWS);
        for (int j = 0; j < FRAME_GRID_ROWS; j++) k->mGrid[i][j] = f->F.mGrid[i][j];
    }
    return k;
}

extern "C" {

int refm_frame_create(const orb_frame_desc *d, refm_frame **out) {
    refm_frame *f = new refm_frame();
    f->d = *d;
    Frame &F = f->F;
    F.N = d->n;
    F.mvKeysUn.resize(d->n);
    for (int i = 0; i < d->n; i++) {
        const orb_keypoint &k = d->keys_un[i];
        cv::KeyPoint kp(k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id);
        F.mvKeysUn[i] = kp;
    }
    F.mvKeys = F.mvKeysUn;
    F.mvuRight.assign(d->n, -1.0f);
    if (d->u_right) F.mvuRight.assign(d->u_right, d->u_right + d->n);
    F.mDescriptors = cv::Mat(d->n > 0 ? d->n : 1, 32, CV_8UC1);
    if (d->n) memcpy(F.mDescriptors.data, d->descriptors, (size_t)d->n * 32);
    F.mvpMapPoints.assign(d->n, nullptr);
    F.mvbOutlier.assign(d->n, false);
    F.mvScaleFactors.assign(d->scale_factors, d->scale_factors + d->nlevels);
    F.mfLogScaleFactor = log(d->nlevels > 1 ? d->scale_factors[1] : 1.2f);
    set_statics(f);
    F.AssignFeaturesToGrid();
    *out = f;
    return 0;
}
void refm_frame_destroy(refm_frame *f) { delete f; }
double refm_last_call_us(void) { return g_call_us; }

int refm_frame_features_in_area(refm_frame *f, float x, float y, float r, int min_level, int max_level, int32_t *indices, int capacity,
                                int *n_out) {
    set_statics(f);
    std::vector<size_t> v = f->F.GetFeaturesInArea(x, y, r, min_level, max_level);
    *n_out = (int)v.size();
    if ((int)v.size() > capacity) return ORB_ERR_CAPACITY;
    for (size_t i = 0; i < v.size(); i++) indices[i] = (int32_t)v[i];
    return 0;
}

// ref ORBmatcher::SearchForInitialization, ORBmatcher.cc:405-523
int refm_search_for_initialization(refm_frame *f1, refm_frame *f2, float *prev_matched, int32_t *matches12, int window_size, float nnratio,
                                   int check_orientation, int *nmatches) {
    set_statics(f2);
    std::vector<cv::Point2f> prev(f1->F.N);
    for (int i = 0; i < f1->F.N; i++) prev[i] = cv::Point2f(prev_matched[2 * i], prev_matched[2 * i + 1]);
    std::vector<int> m12;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    {
        CallTimer t;
        *nmatches = matcher.SearchForInitialization(f1->F, f2->F, prev, m12, window_size);
    }
    for (int i = 0; i < f1->F.N; i++) {
        matches12[i] = m12[i];
        prev_matched[2 * i] = prev[i].x; prev_matched[2 * i + 1] = prev[i].y;
    }
    return 0;
}

// ref ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th), ORBmatcher.cc:45-129
int refm_search_by_projection_points(refm_frame *f, const orb_track_points *P, const uint8_t *claimed, float th, float nnratio,
                                     int32_t *assigned, int *nmatches) {
    set_statics(f);
    Frame &F = f->F;
    std::vector<std::unique_ptr<MapPoint>> store;
    std::vector<MapPoint *> pts(P->n);
    for (int i = 0; i < P->n; i++) {
        std::unique_ptr<MapPoint> m(new MapPoint());
        m->id = i;
        m->mbTrackInView = P->in_view[i] != 0;
        m->mTrackProjX = P->proj_x[i]; m->mTrackProjY = P->proj_y[i]; m->mTrackProjXR = P->proj_xr[i];
        m->mnTrackScaleLevel = P->level[i]; m->mTrackViewCos = P->view_cos[i];
        m->descriptor = desc_row(P->descriptors + (size_t)i * 32);
        m->nObs = P->obs_positive ? (P->obs_positive[i] ? 1 : 0) : 1;
        pts[i] = m.get();
        store.push_back(std::move(m));
    }
    MapPoint pre;      // a map point that already sits in the frame with observations (blocks the feature, ref :87-89)
    pre.id = -2; pre.nObs = 1;
    for (int i = 0; i < F.N; i++) F.mvpMapPoints[i] = (claimed && claimed[i]) ? &pre : nullptr;
    ORBmatcher matcher(nnratio, true);
    {
        CallTimer t;
        *nmatches = matcher.SearchByProjection(F, pts, th);
    }
    for (int i = 0; i < F.N; i++) {
        MapPoint *m = F.mvpMapPoints[i];
        assigned[i] = (m && m->id >= 0) ? m->id : -1;
        F.mvpMapPoints[i] = nullptr;
    }
    return 0;
}

// ref ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono), ORBmatcher.cc:1331-1473.
// world_xyz: n x 3 map point positions of the Last features (has_point[i]=0: no map point / outlier); Tcw_cur, Tcw_last:
// row-major 4x4.  Also returns what the reference computes on the way -- u, v, invz, valid (ref :1362-1379) and the
// forward/backward flags (:1351-1352) -- so the same numbers can be handed to the flat C ABI.
int refm_search_by_projection_last(refm_frame *cur, refm_frame *last, const float *world_xyz, const uint8_t *has_point,
                                   const uint8_t *obs_positive, const uint8_t *claimed, const float *Tcw_cur, const float *Tcw_last,
                                   float fx, float fy, float cx, float cy, float mbf, float mb, float th, int mono, float nnratio,
                                   int check_orientation, int32_t *assigned, int *nmatches, float *u_out, float *v_out,
                                   float *invz_out, uint8_t *valid_out, int *forward_out, int *backward_out) {
    set_statics(cur);
    Frame &C = cur->F;
    Frame &L = last->F;
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    C.mTcw = cv::Mat(4, 4, CV_32F); L.mTcw = cv::Mat(4, 4, CV_32F);
    memcpy(C.mTcw.data, Tcw_cur, 64); memcpy(L.mTcw.data, Tcw_last, 64);
    C.mbf = mbf; C.mb = mb;
    const int n = L.N;
    std::vector<std::unique_ptr<MapPoint>> store;
    for (int i = 0; i < n; i++) {
        L.mvpMapPoints[i] = nullptr;
        L.mvbOutlier[i] = false;
        if (!has_point[i]) continue;
        std::unique_ptr<MapPoint> m(new MapPoint());
        m->id = i;
        m->worldPos = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) m->worldPos.at<float>(k) = world_xyz[3 * i + k];
        m->descriptor = desc_row(L.mDescriptors.ptr(i));
        m->nObs = obs_positive ? (obs_positive[i] ? 1 : 0) : 1;
        L.mvpMapPoints[i] = m.get();
        store.push_back(std::move(m));
    }
    MapPoint pre;
    pre.id = -2; pre.nObs = 1;
    for (int i = 0; i < C.N; i++) C.mvpMapPoints[i] = (claimed && claimed[i]) ? &pre : nullptr;

    // the projection exactly as the reference writes it (ORBmatcher.cc:1341-1379), for the flat interface
    {
        const cv::Mat Rcw = C.mTcw.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tcw = C.mTcw.rowRange(0, 3).col(3);
        const cv::Mat twc = -Rcw.t() * tcw;
        const cv::Mat Rlw = L.mTcw.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tlw = L.mTcw.rowRange(0, 3).col(3);
        const cv::Mat tlc = Rlw * twc + tlw;
        const bool bMono = mono != 0;
        *forward_out = tlc.at<float>(2) > C.mb && !bMono;
        *backward_out = -tlc.at<float>(2) > C.mb && !bMono;
        for (int i = 0; i < n; i++) {
            valid_out[i] = 0; u_out[i] = v_out[i] = invz_out[i] = 0;
            MapPoint *pMP = L.mvpMapPoints[i];
            if (!pMP) continue;
            cv::Mat x3Dw = pMP->GetWorldPos();
            cv::Mat x3Dc = Rcw * x3Dw + tcw;
            const float xc = x3Dc.at<float>(0);
            const float yc = x3Dc.at<float>(1);
            const float invzc = 1.0 / x3Dc.at<float>(2);
            if (invzc < 0) continue;
            float u = Frame::fx * xc * invzc + Frame::cx;
            float v = Frame::fy * yc * invzc + Frame::cy;
            if (u < Frame::mnMinX || u > Frame::mnMaxX) continue;
            if (v < Frame::mnMinY || v > Frame::mnMaxY) continue;
            valid_out[i] = 1; u_out[i] = u; v_out[i] = v; invz_out[i] = invzc;
        }
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    {
        CallTimer t;
        *nmatches = matcher.SearchByProjection(C, L, th, mono != 0);
    }
    for (int i = 0; i < C.N; i++) {
        MapPoint *m = C.mvpMapPoints[i];
        assigned[i] = (m && m->id >= 0) ? m->id : -1;
        C.mvpMapPoints[i] = nullptr;
    }
    for (int i = 0; i < n; i++) L.mvpMapPoints[i] = nullptr;
    return 0;
}

// ref ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist),
// ORBmatcher.cc:1475-1602 (relocalisation refinement).  KF map point i sits at world_xyz[i] with distance-invariance range
// [min_dist, max_dist]; usable[i]=0: no point / bad / already found.  Also returns the per-point quantities the reference
// computes before its candidate loop (:1499-1531) for the flat interface: valid, u, v, predicted level.
int refm_search_by_projection_kf(refm_frame *cur, refm_frame *kf, const float *world_xyz, const uint8_t *usable, const float *max_dist,
                                 const float *min_dist, const uint8_t *claimed, const float *Tcw_cur, float fx, float fy, float cx, float cy,
                                 float th, int orb_dist, float nnratio, int check_orientation, int32_t *assigned, int *nmatches,
                                 float *u_out, float *v_out, int32_t *level_out, uint8_t *valid_out) {
    set_statics(cur);
    Frame &C = cur->F;
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    C.mTcw = cv::Mat(4, 4, CV_32F);
    memcpy(C.mTcw.data, Tcw_cur, 64);
    std::unique_ptr<KeyFrame> K = as_keyframe(kf);
    std::vector<std::unique_ptr<MapPoint>> store;
    std::set<MapPoint *> already;
    for (int i = 0; i < K->N; i++) {
        if (!usable[i]) continue;
        std::unique_ptr<MapPoint> m(new MapPoint());
        m->id = i;
        m->worldPos = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) m->worldPos.at<float>(k) = world_xyz[3 * i + k];
        m->mfMaxDistance = max_dist[i]; m->mfMinDistance = min_dist[i];
        m->descriptor = desc_row(K->mDescriptors.ptr(i));
        K->mvpMapPoints[i] = m.get();
        store.push_back(std::move(m));
    }
    MapPoint pre;
    pre.id = -2; pre.nObs = 1;
    for (int i = 0; i < C.N; i++) C.mvpMapPoints[i] = (claimed && claimed[i]) ? &pre : nullptr;
    {   // ref :1479-1531
        const cv::Mat Rcw = C.mTcw.rowRange(0, 3).colRange(0, 3);
        const cv::Mat tcw = C.mTcw.rowRange(0, 3).col(3);
        const cv::Mat Ow = -Rcw.t() * tcw;
        for (int i = 0; i < K->N; i++) {
            valid_out[i] = 0; u_out[i] = v_out[i] = 0; level_out[i] = 0;
            MapPoint *pMP = K->mvpMapPoints[i];
            if (!pMP) continue;
            cv::Mat x3Dw = pMP->GetWorldPos();
            cv::Mat x3Dc = Rcw * x3Dw + tcw;
            const float xc = x3Dc.at<float>(0);
            const float yc = x3Dc.at<float>(1);
            const float invzc = 1.0 / x3Dc.at<float>(2);
            const float u = Frame::fx * xc * invzc + Frame::cx;
            const float v = Frame::fy * yc * invzc + Frame::cy;
            if (u < Frame::mnMinX || u > Frame::mnMaxX) continue;
            if (v < Frame::mnMinY || v > Frame::mnMaxY) continue;
            cv::Mat PO = x3Dw - Ow;
            float dist3D = cv::norm(PO);
            const float maxDistance = pMP->GetMaxDistanceInvariance();
            const float minDistance = pMP->GetMinDistanceInvariance();
            if (dist3D < minDistance || dist3D > maxDistance) continue;
            valid_out[i] = 1; u_out[i] = u; v_out[i] = v;
            level_out[i] = pMP->PredictScale(dist3D, C.mfLogScaleFactor);
        }
    }
    ORBmatcher matcher(nnratio, check_orientation != 0);
    *nmatches = matcher.SearchByProjection(C, K.get(), already, th, orb_dist);
    for (int i = 0; i < C.N; i++) {
        MapPoint *m = C.mvpMapPoints[i];
        assigned[i] = (m && m->id >= 0) ? m->id : -1;
        C.mvpMapPoints[i] = nullptr;
    }
    return 0;
}

// ref ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*> &vpMatched, th),
// ORBmatcher.cc:290-403 (loop closing).  n candidate points (position, normal, distance range); claimed[idx] = vpMatched[idx] != NULL
// on entry.  Returns the reference's pre-loop quantities (:311-360) for the flat interface.
int refm_search_by_projection_sim3(refm_frame *kf, const float *Scw, int n, const float *world_xyz, const float *normal_xyz,
                                   const uint8_t *usable, const float *max_dist, const float *min_dist, const uint8_t *claimed,
                                   float fx, float fy, float cx, float cy, int th, float nnratio, int32_t *assigned, int *nmatches,
                                   float *u_out, float *v_out, int32_t *level_out, uint8_t *valid_out) {
    set_statics(kf);
    std::unique_ptr<KeyFrame> K = as_keyframe(kf);
    K->fx = fx; K->fy = fy; K->cx = cx; K->cy = cy;
    std::vector<std::unique_ptr<MapPoint>> store;
    std::vector<MapPoint *> pts(n);
    MapPoint badpt;
    badpt.bad = true;
    for (int i = 0; i < n; i++) {
        if (!usable[i]) { pts[i] = &badpt; continue; }
        std::unique_ptr<MapPoint> m(new MapPoint());
        m->id = i;
        m->worldPos = cv::Mat(3, 1, CV_32F); m->normal = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) { m->worldPos.at<float>(k) = world_xyz[3 * i + k]; m->normal.at<float>(k) = normal_xyz[3 * i + k]; }
        m->mfMaxDistance = max_dist[i]; m->mfMinDistance = min_dist[i];
        m->descriptor = desc_row(kf->F.mDescriptors.ptr(i % kf->F.N));   // descriptor of the matching feature, as the scenario plants it
        pts[i] = m.get();
        store.push_back(std::move(m));
    }
    MapPoint pre;
    pre.id = -2;
    std::vector<MapPoint *> matched(K->N, nullptr);
    for (int i = 0; i < K->N; i++) if (claimed && claimed[i]) matched[i] = &pre;
    cv::Mat S(4, 4, CV_32F);
    memcpy(S.data, Scw, 64);
    {   // ref :298-360
        cv::Mat sRcw = S.rowRange(0, 3).colRange(0, 3);
        const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
        cv::Mat Rcw = sRcw / scw;
        cv::Mat tcw = S.rowRange(0, 3).col(3) / scw;
        cv::Mat Ow = -Rcw.t() * tcw;
        for (int i = 0; i < n; i++) {
            valid_out[i] = 0; u_out[i] = v_out[i] = 0; level_out[i] = 0;
            MapPoint *pMP = pts[i];
            if (pMP->isBad()) continue;
            cv::Mat p3Dw = pMP->GetWorldPos();
            cv::Mat p3Dc = Rcw * p3Dw + tcw;
            if (p3Dc.at<float>(2) < 0.0) continue;
            const float invz = 1 / p3Dc.at<float>(2);
            const float x = p3Dc.at<float>(0) * invz;
            const float y = p3Dc.at<float>(1) * invz;
            const float u = fx * x + cx;
            const float v = fy * y + cy;
            if (!K->IsInImage(u, v)) continue;
            const float maxDistance = pMP->GetMaxDistanceInvariance();
            const float minDistance = pMP->GetMinDistanceInvariance();
            cv::Mat PO = p3Dw - Ow;
            const float dist = cv::norm(PO);
            if (dist < minDistance || dist > maxDistance) continue;
            cv::Mat Pn = pMP->GetNormal();
            if (PO.dot(Pn) < 0.5 * dist) continue;
            valid_out[i] = 1; u_out[i] = u; v_out[i] = v;
            level_out[i] = pMP->PredictScale(dist, K->mfLogScaleFactor);
        }
    }
    ORBmatcher matcher(nnratio, true);
    *nmatches = matcher.SearchByProjection(K.get(), S, pts, matched, th);
    for (int i = 0; i < K->N; i++) assigned[i] = (matched[i] && matched[i]->id >= 0) ? matched[i]->id : -1;
    return 0;
}

static void fill_featvec(DBoW2::FeatureVector &fv, int ngroups, const int32_t *off, const int32_t *idx) {
    fv.clear();
    for (int g = 0; g < ngroups; g++) {
        std::vector<unsigned int> v(idx + off[g], idx + off[g + 1]);
        if (!v.empty()) fv[(DBoW2::NodeId)g] = v;
    }
}

// ref ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&), ORBmatcher.cc:159-288
int refm_search_by_bow(refm_frame *kf, refm_frame *f, int ngroups, const int32_t *kf_off, const int32_t *kf_idx, const int32_t *f_off,
                       const int32_t *f_idx, const uint8_t *kf_valid, float nnratio, int check_orientation, int32_t *matches,
                       int *nmatches) {
    set_statics(f);
    std::unique_ptr<KeyFrame> K = as_keyframe(kf);
    std::vector<std::unique_ptr<MapPoint>> store;
    for (int i = 0; i < K->N; i++)
        if (kf_valid[i]) {
            std::unique_ptr<MapPoint> m(new MapPoint());
            m->id = i;
            K->mvpMapPoints[i] = m.get();
            store.push_back(std::move(m));
        }
    fill_featvec(K->mFeatVec, ngroups, kf_off, kf_idx);
    fill_featvec(f->F.mFeatVec, ngroups, f_off, f_idx);
    std::vector<MapPoint *> out;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    {
        CallTimer t;
        *nmatches = matcher.SearchByBoW(K.get(), f->F, out);
    }
    for (int i = 0; i < f->F.N; i++) matches[i] = out[i] ? out[i]->id : -1;
    return 0;
}

// ref ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&), ORBmatcher.cc:525-658.  Output flattened the way
// orb_search_by_bow reports it: matches[idx2] = idx1.
int refm_search_by_bow_kfkf(refm_frame *kf1, refm_frame *kf2, int ngroups, const int32_t *off1, const int32_t *idx1, const int32_t *off2,
                            const int32_t *idx2, const uint8_t *valid1, const uint8_t *valid2, float nnratio, int check_orientation,
                            int32_t *matches, int *nmatches) {
    set_statics(kf2);
    std::unique_ptr<KeyFrame> K1 = as_keyframe(kf1), K2 = as_keyframe(kf2);
    std::vector<std::unique_ptr<MapPoint>> store;
    for (int i = 0; i < K1->N; i++)
        if (valid1[i]) { std::unique_ptr<MapPoint> m(new MapPoint()); m->id = i; K1->mvpMapPoints[i] = m.get(); store.push_back(std::move(m)); }
    for (int i = 0; i < K2->N; i++)
        if (!valid2 || valid2[i]) { std::unique_ptr<MapPoint> m(new MapPoint()); m->id = i; K2->mvpMapPoints[i] = m.get(); store.push_back(std::move(m)); }
    fill_featvec(K1->mFeatVec, ngroups, off1, idx1);
    fill_featvec(K2->mFeatVec, ngroups, off2, idx2);
    std::vector<MapPoint *> out;
    ORBmatcher matcher(nnratio, check_orientation != 0);
    *nmatches = matcher.SearchByBoW(K1.get(), K2.get(), out);
    for (int i = 0; i < K2->N; i++) matches[i] = -1;
    for (int i1 = 0; i1 < K1->N; i1++)
        if (out[i1]) matches[out[i1]->id] = i1;
    return 0;
}

#ifdef ORBREF_REAL_FRAME
// ref Frame::ComputeStereoMatches, Frame.cc:471-643 -- the reference's own member function on the reference's own Frame.
// The two reference ORBextractor objects are run on the images only to fill mvImagePyramid (ComputePyramid is
// deterministic); keypoints and descriptors are the caller's.  Returns ORB_ERR_EMPTY where the reference would index an
// empty vector (Frame.cc:630), detected beforehand from the outputs being all -1 is impossible, so the call is made on
// a guarded copy: the reference's UB case is reported by the oracle instead and never reaches this function in tests.
int refm_stereo_match(const uint8_t *img_left, const uint8_t *img_right, int width, int height, int nfeatures, float scale, int nlevels,
                      int ini_th, int min_th, const orb_keypoint *keysL, const uint8_t *descL, int N, const orb_keypoint *keysR,
                      const uint8_t *descR, int Nr, float mbf, float mb, float *uRight, float *depth) {
    ORBextractor exL(nfeatures, scale, nlevels, ini_th, min_th), exR(nfeatures, scale, nlevels, ini_th, min_th);
    cv::Mat imL(height, width, CV_8UC1), imR(height, width, CV_8UC1);
    memcpy(imL.data, img_left, (size_t)width * height);
    memcpy(imR.data, img_right, (size_t)width * height);
    {
        std::vector<cv::KeyPoint> k;
        cv::Mat d;
        exL(imL, cv::Mat(), k, d);
        exR(imR, cv::Mat(), k, d);
    }
    Frame F;
    F.mpORBextractorLeft = &exL; F.mpORBextractorRight = &exR;
    F.N = N;
    auto conv = [](const orb_keypoint *k, int n) {
        std::vector<cv::KeyPoint> v(n);
        for (int i = 0; i < n; i++) v[i] = cv::KeyPoint(k[i].x, k[i].y, k[i].size, k[i].angle, k[i].response, k[i].octave, k[i].class_id);
        return v;
    };
    F.mvKeys = conv(keysL, N); F.mvKeysRight = conv(keysR, Nr);
    F.mDescriptors = cv::Mat(N > 0 ? N : 1, 32, CV_8UC1); F.mDescriptorsRight = cv::Mat(Nr > 0 ? Nr : 1, 32, CV_8UC1);
    if (N) memcpy(F.mDescriptors.data, descL, (size_t)N * 32);
    if (Nr) memcpy(F.mDescriptorsRight.data, descR, (size_t)Nr * 32);
    F.mvScaleFactors = exL.GetScaleFactors(); F.mvInvScaleFactors = exL.GetInverseScaleFactors();
    F.mbf = mbf; F.mb = mb;   // the reference reads mb before its constructor assigns it (Frame.cc:95 vs :119); intended value
    F.ComputeStereoMatches();
    for (int i = 0; i < N; i++) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    return 0;
}
// ref Frame::ComputeStereoFromRGBD, Frame.cc:645-667, run from the reference's own Frame.cc on a CV_32F depth image
int refm_stereo_from_rgbd(const orb_keypoint *keys, const orb_keypoint *keys_un, int N, const float *depth_map, int width, int height,
                          float mbf, float *uRight, float *depth) {
    Frame F;
    F.N = N;
    auto conv = [](const orb_keypoint *k, int n) {
        std::vector<cv::KeyPoint> v(n);
        for (int i = 0; i < n; i++) v[i] = cv::KeyPoint(k[i].x, k[i].y, k[i].size, k[i].angle, k[i].response, k[i].octave, k[i].class_id);
        return v;
    };
    F.mvKeys = conv(keys, N); F.mvKeysUn = conv(keys_un, N);
    F.mbf = mbf;
    cv::Mat imDepth(height, width, CV_32F);
    memcpy(imDepth.data, depth_map, (size_t)width * height * 4);
    F.ComputeStereoFromRGBD(imDepth);
    for (int i = 0; i < N; i++) { uRight[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
    return 0;
}
#endif

static void set_pose(KeyFrame *K, const float *R9, const float *t3) {
    K->Rcw = cv::Mat(3, 3, CV_32F); K->tcw = cv::Mat(3, 1, CV_32F);
    memcpy(K->Rcw.data, R9, 36); memcpy(K->tcw.data, t3, 12);
    K->Ow = -K->Rcw.t() * K->tcw;   // ref KeyFrame.cc:79
}
static cv::Mat vec3(const float *p) { cv::Mat m(3, 1, CV_32F); memcpy(m.data, p, 12); return m; }

// ref ORBmatcher::SearchForTriangulation, ORBmatcher.cc:660-826.  Same signature as orb_search_for_triangulation; KF2 is given an
// identity pose and unit intrinsics and KF1's centre is put at (ex, ey, 1), so the reference's own epipole arithmetic (:667-673)
// returns exactly (ex, ey).
int refm_search_for_triangulation(refm_frame *kf1, refm_frame *kf2, int ngroups, const int32_t *off1, const int32_t *idx1, const int32_t *off2,
                                  const int32_t *idx2, const uint8_t *usable1, const uint8_t *usable2, const float *F12, float ex, float ey,
                                  const float *level_sigma2, int only_stereo, int check_orientation, int32_t *matches12, int *nmatches) {
    set_statics(kf2);
    std::unique_ptr<KeyFrame> K1 = as_keyframe(kf1), K2 = as_keyframe(kf2);
    std::vector<std::unique_ptr<MapPoint>> store;
    for (int i = 0; i < K1->N; i++)
        if (!usable1[i]) { store.emplace_back(new MapPoint()); K1->mvpMapPoints[i] = store.back().get(); }
    for (int i = 0; i < K2->N; i++)
        if (!usable2[i]) { store.emplace_back(new MapPoint()); K2->mvpMapPoints[i] = store.back().get(); }
    fill_featvec(K1->mFeatVec, ngroups, off1, idx1);
    fill_featvec(K2->mFeatVec, ngroups, off2, idx2);
    const float I9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z3[3] = {0, 0, 0};
    set_pose(K1.get(), I9, z3); set_pose(K2.get(), I9, z3);
    K1->Ow = cv::Mat(3, 1, CV_32F);
    K1->Ow.at<float>(0) = ex; K1->Ow.at<float>(1) = ey; K1->Ow.at<float>(2) = 1.0f;
    K2->fx = K2->fy = 1.0f; K2->cx = K2->cy = 0.0f;
    K2->mvLevelSigma2.assign(level_sigma2, level_sigma2 + K2->mvScaleFactors.size());
    cv::Mat F(3, 3, CV_32F);
    memcpy(F.data, F12, 36);
    std::vector<std::pair<size_t, size_t>> pairs;
    ORBmatcher matcher(0.6f, check_orientation != 0);
    *nmatches = matcher.SearchForTriangulation(K1.get(), K2.get(), F, pairs, only_stereo != 0);
    for (int i = 0; i < K1->N; i++) matches12[i] = -1;
    for (size_t k = 0; k < pairs.size(); k++) matches12[pairs[k].first] = (int32_t)pairs[k].second;
    return 0;
}

// Scenario for the two ORBmatcher::Fuse overloads (ORBmatcher.cc:828-978 and :980-1103) with the map bookkeeping they do.
// Points: present[i] 0 = NULL entry, 1 = good, 2 = bad; in_kf[i] >= 0 = the point is already observed by the key frame at that feature.
// kf_mp_obs[idx] >= 0 = the key frame holds some other map point there with that many observations (id 1000000+idx).
// Outputs: kf_state[idx] = id at that feature afterwards (-1 none); pt_obs[i] = feature the point observes in the key frame (-1 none);
// pt_repl[i] = id that replaced point i (variant 0) / vpReplacePoint[i] (variant 1); kfmp_repl[idx] = id that replaced the pre-existing point.
int refm_fuse(refm_frame *kf, int variant, const float *Rcw9, const float *tcw3, const float *Scw16, int n, const float *world_xyz,
              const float *normal_xyz, const uint8_t *present, const float *max_dist, const float *min_dist, const uint8_t *desc,
              const int32_t *nobs, const int32_t *in_kf, const int32_t *kf_mp_obs, float fx, float fy, float cx, float cy, float mbf, float th,
              int32_t *kf_state, int32_t *pt_obs, int32_t *pt_repl, int32_t *kfmp_repl, int *nfused) {
    set_statics(kf);
    std::unique_ptr<KeyFrame> K = as_keyframe(kf);
    K->fx = fx; K->fy = fy; K->cx = cx; K->cy = cy; K->mbf = mbf;
    set_pose(K.get(), Rcw9, tcw3);
    std::vector<std::unique_ptr<MapPoint>> store;
    std::vector<MapPoint *> pts(n, nullptr), pre(K->N, nullptr);
    for (int idx = 0; idx < K->N; idx++)
        if (kf_mp_obs[idx] >= 0) {
            store.emplace_back(new MapPoint());
            MapPoint *m = store.back().get();
            m->id = 1000000 + idx; m->nObs = kf_mp_obs[idx]; m->observations[K.get()] = idx;
            K->mvpMapPoints[idx] = m; pre[idx] = m;
        }
    for (int i = 0; i < n; i++) {
        if (!present[i]) continue;
        store.emplace_back(new MapPoint());
        MapPoint *m = store.back().get();
        m->id = i; m->bad = present[i] == 2; m->nObs = nobs[i];
        m->worldPos = vec3(world_xyz + 3 * i); m->normal = vec3(normal_xyz + 3 * i);
        m->mfMaxDistance = max_dist[i]; m->mfMinDistance = min_dist[i];
        m->descriptor = desc_row(desc + (size_t)i * 32);
        if (in_kf[i] >= 0) { m->observations[K.get()] = in_kf[i]; K->mvpMapPoints[in_kf[i]] = m; }
        pts[i] = m;
    }
    std::vector<MapPoint *> repl(n, nullptr);
    ORBmatcher matcher(0.6f, true);
    if (variant == 0) {
        *nfused = matcher.Fuse(K.get(), pts, th);
    } else {
        cv::Mat S(4, 4, CV_32F);
        memcpy(S.data, Scw16, 64);
        *nfused = matcher.Fuse(K.get(), S, pts, th, repl);
    }
    for (int idx = 0; idx < K->N; idx++) {
        kf_state[idx] = K->mvpMapPoints[idx] ? K->mvpMapPoints[idx]->id : -1;
        kfmp_repl[idx] = (pre[idx] && pre[idx]->replaced) ? pre[idx]->replaced->id : -1;
    }
    for (int i = 0; i < n; i++) {
        pt_obs[i] = -1; pt_repl[i] = -1;
        if (!pts[i]) continue;
        if (pts[i]->observations.count(K.get())) pt_obs[i] = (int32_t)pts[i]->observations[K.get()];
        MapPoint *r = variant == 0 ? pts[i]->replaced : repl[i];
        if (r) pt_repl[i] = r->id;
    }
    return 0;
}

// ref ORBmatcher::SearchBySim3, ORBmatcher.cc:1105-1329.  has1/has2: 0 = no map point at the feature, 1 = good, 2 = bad.
// pre12[i1] >= 0: vpMatches12[i1] holds KF2's map point of that feature on entry.  matches12[i1] = feature of KF2 whose map point
// ends up in vpMatches12[i1], or -1.
int refm_search_by_sim3(refm_frame *kf1, refm_frame *kf2, float s12, const float *R12, const float *t12, const float *R1w, const float *t1w,
                        const float *R2w, const float *t2w, const uint8_t *has1, const float *world1, const float *maxd1, const float *mind1,
                        const uint8_t *desc1, const uint8_t *has2, const float *world2, const float *maxd2, const float *mind2,
                        const uint8_t *desc2, const int32_t *pre12, float fx, float fy, float cx, float cy, float th, int32_t *matches12,
                        int *nfound) {
    set_statics(kf1);
    std::unique_ptr<KeyFrame> K1 = as_keyframe(kf1), K2 = as_keyframe(kf2);
    K1->fx = fx; K1->fy = fy; K1->cx = cx; K1->cy = cy;
    K2->fx = fx; K2->fy = fy; K2->cx = cx; K2->cy = cy;
    set_pose(K1.get(), R1w, t1w); set_pose(K2.get(), R2w, t2w);
    std::vector<std::unique_ptr<MapPoint>> store;
    auto fill = [&](KeyFrame *K, const uint8_t *has, const float *world, const float *maxd, const float *mind, const uint8_t *desc) {
        for (int i = 0; i < K->N; i++) {
            if (!has[i]) continue;
            store.emplace_back(new MapPoint());
            MapPoint *m = store.back().get();
            m->id = i; m->bad = has[i] == 2;
            m->worldPos = vec3(world + 3 * i);
            m->mfMaxDistance = maxd[i]; m->mfMinDistance = mind[i];
            m->descriptor = desc_row(desc + (size_t)i * 32);
            m->observations[K] = i;
            K->mvpMapPoints[i] = m;
        }
    };
    fill(K1.get(), has1, world1, maxd1, mind1, desc1);
    fill(K2.get(), has2, world2, maxd2, mind2, desc2);
    std::vector<MapPoint *> m12(K1->N, nullptr);
    for (int i = 0; i < K1->N; i++)
        if (pre12[i] >= 0) m12[i] = K2->mvpMapPoints[pre12[i]];
    cv::Mat R(3, 3, CV_32F), t = vec3(t12);
    memcpy(R.data, R12, 36);
    ORBmatcher matcher(0.6f, true);
    *nfound = matcher.SearchBySim3(K1.get(), K2.get(), m12, s12, R, t, th);
    for (int i = 0; i < K1->N; i++) matches12[i] = m12[i] ? m12[i]->id : -1;
    return 0;
}

// ref ORBmatcher::DescriptorDistance, ORBmatcher.cc:1650-1666
// Frame::isInFrustum (src/Frame.cc:274-330) from the reference's compiled Frame.cc, one call per map point like
// Tracking::SearchLocalPoints (src/Tracking.cc:1527-1539).  The pose goes through the reference's Frame::SetPose /
// UpdatePoseMatrices (:260-272); cam_out receives the mRcw / mtcw / mOw it derived so that the flat interfaces see the same
// camera.  MapPoint::PredictScale is the stub's restatement of src/MapPoint.cc:357-366.
int refm_is_in_frustum(refm_frame *f, const float *Tcw16, float fx, float fy, float cx, float cy, float mbf, float log_scale_factor,
                       int n, const float *world_xyz, const float *normal_xyz, const float *max_dist, const float *min_dist,
                       const uint8_t *consider, float limit, float *cam_out15, uint8_t *in_view, float *proj_x, float *proj_y,
                       float *proj_xr, int32_t *level, float *view_cos) {
    set_statics(f);
    Frame &F = f->F;
    Frame::fx = fx; Frame::fy = fy; Frame::cx = cx; Frame::cy = cy;
    F.mbf = mbf; F.mfLogScaleFactor = log_scale_factor;
    cv::Mat T(4, 4, CV_32F);
    memcpy(T.data, Tcw16, 64);
    F.SetPose(T);
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) cam_out15[3 * r + c] = F.mRcw.at<float>(r, c);
        cam_out15[9 + r] = F.mtcw.at<float>(r);
        cam_out15[12 + r] = F.mOw.at<float>(r);
    }
    for (int i = 0; i < n; i++) {
        in_view[i] = 0; proj_x[i] = proj_y[i] = proj_xr[i] = view_cos[i] = 0.0f; level[i] = 0;
        if (consider && !consider[i]) continue;
        MapPoint m;
        m.worldPos = cv::Mat(3, 1, CV_32F); m.normal = cv::Mat(3, 1, CV_32F);
        for (int k = 0; k < 3; k++) { m.worldPos.at<float>(k) = world_xyz[3 * i + k]; m.normal.at<float>(k) = normal_xyz[3 * i + k]; }
        m.mfMaxDistance = max_dist[i]; m.mfMinDistance = min_dist[i];
        if (F.isInFrustum(&m, limit)) {
            in_view[i] = m.mbTrackInView; proj_x[i] = m.mTrackProjX; proj_y[i] = m.mTrackProjY; proj_xr[i] = m.mTrackProjXR;
            level[i] = m.mnTrackScaleLevel; view_cos[i] = m.mTrackViewCos;
        }
    }
    return 0;
}

int refm_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist) {
    for (int i = 0; i < n; i++) dist[i] = ORBmatcher::DescriptorDistance(desc_row(a + (size_t)i * 32), desc_row(b + (size_t)i * 32));
    return 0;
}

}  // extern "C"
