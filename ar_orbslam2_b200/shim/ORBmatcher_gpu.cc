// See ORBmatcher_gpu.h.  Line references are to ORB_SLAM2/src/ORBmatcher.cc of the reference.
//
// PROVENANCE.  This file is meant to be compiled INSIDE the ORB-SLAM2 tree as the body of ORB_SLAM2::ORBmatcher (see
// ORBmatcher.cc next to it).  The candidate loops -- the hot path -- are liborbgpu calls.  What remains on the host is the
// reference's per-point set-up in front of each loop (projection of a map point with the frame pose, frustum / distance /
// PredictScale gates) and the map bookkeeping behind it (AddObservation / Replace / AddMapPoint).  Those statements operate on
// the reference's own cv::Mat members and must produce the same floats as ORBmatcher.cc to keep the match lists bit-identical,
// so they follow ORB_SLAM2/src/ORBmatcher.cc (Raul Mur-Artal et al., GPLv3) statement for statement; each block cites the
// lines it restates.  They are derived work of that file and carry its licence (GPLv3) when distributed with it.
#include "ORBmatcher_gpu.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <set>
#include <stdexcept>
#include <string>

#include "orbgpu.h"

namespace ORB_SLAM2 {
namespace orbgpu_shim {

static void Check(int rc, const char *what) {
    if (rc != ORB_OK) throw std::runtime_error(std::string("liborbgpu: ") + what + ": " + orb_last_error());
}

// The part of a Frame / KeyFrame the matchers read (mvKeysUn, mDescriptors, mvuRight, the image bounds that define the 64x48
// grid, mvScaleFactors) as an orb_frame on the device.  Built the first time a matcher sees the object and kept in a small
// per-thread LRU cache afterwards: Tracking calls two or three matchers on mCurrentFrame per camera frame, LocalMapping and
// LoopClosing revisit the same key frames many times.  The cache key is the identity of the object's buffers (the vectors'
// and the descriptor matrix's data pointers -- a Frame copy gets new vectors, a KeyFrame keeps its own for life) plus the
// view flags; a hit is confirmed by a 64-bit fingerprint of the WHOLE contents, so an address reused by a later frame (Tracking::Reset
// restarts the ids, frames come and go) cannot alias.  Key points, descriptors and mvuRight never change after the Frame /
// KeyFrame constructor in the reference (Frame.cc:62-233, KeyFrame.cc:31-63).  ORBGPU_SHIM_NO_CACHE=1 builds per call.
namespace {

struct CacheEntry {
    const void *keys, *desc, *ur;
    int flags, n;
    uint64_t fp, stamp;
    orb_frame *h;
};
struct FrameCache {
    std::vector<CacheEntry> e;
    uint64_t clock = 0;
};
// One cache per thread, held through a pointer that is never deleted: a thread_local destructor would call cudaFree /
// cudaStreamDestroy at thread or process exit, possibly after the CUDA runtime has been torn down (liborbgpu's own per-thread
// state is left alone at exit for the same reason, csrc/orb_runtime.cu).  A thread that ends while the process goes on calls
// ReleaseThreadFrameCache() first; otherwise its <= CACHE_MAX device frames (about 0.4 MB each at 2000 key points: SoA key
// points, descriptors, the 64x48 grid) stay allocated until the process exits.
FrameCache &thread_cache() {
    thread_local FrameCache *c = new FrameCache;
    return *c;
}
const size_t CACHE_MAX = 48;

inline uint64_t mix(uint64_t h, const void *p, size_t n) {   // FNV-1a
    const unsigned char *b = static_cast<const unsigned char *>(p);
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

// the same for whole arrays: four independent multiply-xorshift lanes over 8-byte words (about 0.1 ns per byte on one core,
// ~10 us for a 2000-point frame), the tail through the byte loop above
inline uint64_t mix_block(uint64_t h, const void *p, size_t n) {
    const unsigned char *b = static_cast<const unsigned char *>(p);
    uint64_t l[4] = {h, h ^ 0x9E3779B97F4A7C15ull, h ^ 0xBF58476D1CE4E5B9ull, h ^ 0x94D049BB133111EBull};
    size_t i = 0;
    for (; i + 32 <= n; i += 32) {
        uint64_t w[4];
        std::memcpy(w, b + i, 32);
        for (int k = 0; k < 4; k++) { l[k] = (l[k] ^ w[k]) * 0xD6E8FEB86659FD93ull; l[k] ^= l[k] >> 32; }
    }
    h = l[0];
    for (int k = 1; k < 4; k++) h = (h ^ l[k]) * 1099511628211ull;
    return mix(h, b + i, n - i);
}

}  // namespace

struct GpuFrame {
    orb_frame *h = nullptr;
    bool owned = false;
    template <class KeyVec>
    GpuFrame(const KeyVec &keysUn, const cv::Mat &descriptors, const std::vector<float> &uRight, const std::vector<float> &scaleFactors,
             bool keyframeView = false) {
        const int n = (int)keysUn.size();
        const bool withUr = (int)uRight.size() == n && n > 0;
        static const bool noCache = std::getenv("ORBGPU_SHIM_NO_CACHE") != nullptr;
        const void *kp = n ? (const void *)&keysUn[0] : nullptr, *dp = n ? (const void *)descriptors.ptr(0) : nullptr;
        const void *up = withUr ? (const void *)uRight.data() : nullptr;
        const int flags = (keyframeView ? 1 : 0) | (withUr ? 2 : 0);
        uint64_t fp = 1469598103934665603ull;
        if (n) {
            // EVERY key point, descriptor and uRight goes into the fingerprint: buffer addresses are reused as a matter of course
            // (Frame copy-assignment keeps the vectors' storage, equal-sized cv::Mat clones get the same block back), so the
            // contents are the only guard against matching a later frame with a stale device index
            static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint is hashed as raw bytes: it must have no padding");
            fp = mix_block(fp, &keysUn[0], (size_t)n * sizeof(cv::KeyPoint));
            if (descriptors.isContinuous()) fp = mix_block(fp, descriptors.ptr(0), (size_t)n * 32);
            else for (int i = 0; i < n; i++) fp = mix_block(fp, descriptors.ptr(i), 32);
            if (withUr) fp = mix_block(fp, uRight.data(), (size_t)n * 4);
        }
        const float bounds[4] = {Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY};
        fp = mix(fp, bounds, sizeof(bounds));
        FrameCache &C = thread_cache();
        if (!noCache)
            for (size_t i = 0; i < C.e.size(); i++) {
                CacheEntry &c = C.e[i];
                if (c.keys == kp && c.desc == dp && c.ur == up && c.flags == flags && c.n == n && c.fp == fp) {
                    c.stamp = ++C.clock;
                    h = c.h;
                    return;
                }
            }
        std::vector<orb_keypoint> keys(n);
        std::vector<uint8_t> desc((size_t)n * 32 + 32);
        for (int i = 0; i < n; i++) {
            const cv::KeyPoint &k = keysUn[i];
            keys[i].x = k.pt.x; keys[i].y = k.pt.y; keys[i].size = k.size; keys[i].angle = k.angle; keys[i].response = k.response;
            keys[i].octave = k.octave; keys[i].class_id = k.class_id;
            std::memcpy(&desc[(size_t)i * 32], descriptors.ptr(i), 32);
        }
        orb_frame_desc d;
        d.n = n; d.keys_un = keys.data(); d.descriptors = desc.data();
        d.u_right = withUr ? uRight.data() : nullptr;
        d.min_x = Frame::mnMinX; d.max_x = Frame::mnMaxX; d.min_y = Frame::mnMinY; d.max_y = Frame::mnMaxY;
        d.nlevels = (int)scaleFactors.size(); d.scale_factors = scaleFactors.data();
        d.keyframe_view = keyframeView ? 1 : 0;
        Check(orb_frame_create(&d, &h), "orb_frame_create");
        if (noCache) { owned = true; return; }
        if (C.e.size() >= CACHE_MAX) {   // drop the least recently used entry
            size_t lru = 0;
            for (size_t i = 1; i < C.e.size(); i++) if (C.e[i].stamp < C.e[lru].stamp) lru = i;
            orb_frame_destroy(C.e[lru].h);
            C.e.erase(C.e.begin() + lru);
        }
        CacheEntry c = {kp, dp, up, flags, n, fp, ++C.clock, h};
        C.e.push_back(c);
    }
    ~GpuFrame() { if (owned) orb_frame_destroy(h); }
    GpuFrame(const GpuFrame &) = delete;
    GpuFrame &operator=(const GpuFrame &) = delete;
};

void ReleaseThreadFrameCache() {
    FrameCache &C = thread_cache();
    for (size_t i = 0; i < C.e.size(); i++) orb_frame_destroy(C.e[i].h);
    C.e.clear();
}

int SearchByProjection(Frame &F, const std::vector<MapPoint *> &vpMapPoints, float th, float mfNNratio) {
    const int n = (int)vpMapPoints.size();
    if (n == 0 || F.N == 0) return 0;
    std::vector<uint8_t> inView(n), obs(n), desc((size_t)n * 32), claimed(F.N, 0);
    std::vector<float> px(n), py(n), pxr(n), vc(n);
    std::vector<int32_t> lvl(n), assigned(F.N);
    for (int i = 0; i < n; i++) {                                   // the reads of :53-74
        MapPoint *p = vpMapPoints[i];
        inView[i] = p->mbTrackInView && !p->isBad();
        px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; pxr[i] = p->mTrackProjXR;
        lvl[i] = p->mnTrackScaleLevel; vc[i] = p->mTrackViewCos;
        obs[i] = p->Observations() > 0;
        if (inView[i]) { cv::Mat d = p->GetDescriptor(); std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32); }
    }
    for (int i = 0; i < F.N; i++) claimed[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;   // :87-89
    GpuFrame g(F.mvKeysUn, F.mDescriptors, F.mvuRight, F.mvScaleFactors);
    orb_track_points P = {n, inView.data(), px.data(), py.data(), pxr.data(), lvl.data(), vc.data(), desc.data(), obs.data()};
    int nmatches = 0;
    Check(orb_search_by_projection_points(g.h, &P, claimed.data(), th, mfNNratio, assigned.data(), &nmatches), "orb_search_by_projection_points");
    for (int i = 0; i < F.N; i++)
        if (assigned[i] >= 0) F.mvpMapPoints[i] = vpMapPoints[assigned[i]];   // :123
    return nmatches;
}

int SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, float th, bool bMono, bool mbCheckOrientation) {
    const int n = LastFrame.N;
    if (n == 0 || CurrentFrame.N == 0) return 0;
    // pose algebra of :1341-1352, verbatim
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat twc = -Rcw.t() * tcw;
    const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat tlc = Rlw * twc + tlw;
    const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
    const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;

    std::vector<uint8_t> valid(n, 0), obs(n, 0), desc((size_t)n * 32), claimed(CurrentFrame.N, 0);
    std::vector<float> u(n, 0.f), v(n, 0.f), invz(n, 0.f), angle(n, 0.f);
    std::vector<int32_t> octave(n, 0), assigned(CurrentFrame.N);
    for (int i = 0; i < n; i++) {                                   // :1354-1381
        MapPoint *pMP = LastFrame.mvpMapPoints[i];
        octave[i] = LastFrame.mvKeys[i].octave;
        angle[i] = LastFrame.mvKeysUn[i].angle;
        if (!pMP || LastFrame.mvbOutlier[i]) continue;
        cv::Mat x3Dw = pMP->GetWorldPos();
        cv::Mat x3Dc = Rcw * x3Dw + tcw;
        const float xc = x3Dc.at<float>(0);
        const float yc = x3Dc.at<float>(1);
        const float invzc = 1.0 / x3Dc.at<float>(2);
        if (invzc < 0) continue;
        float uu = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
        float vv = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
        if (uu < CurrentFrame.mnMinX || uu > CurrentFrame.mnMaxX) continue;
        if (vv < CurrentFrame.mnMinY || vv > CurrentFrame.mnMaxY) continue;
        valid[i] = 1; u[i] = uu; v[i] = vv; invz[i] = invzc;
        obs[i] = pMP->Observations() > 0;
        cv::Mat d = pMP->GetDescriptor();
        std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
    for (int i = 0; i < CurrentFrame.N; i++)
        claimed[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;   // :1406-1408
    GpuFrame g(CurrentFrame.mvKeysUn, CurrentFrame.mDescriptors, CurrentFrame.mvuRight, CurrentFrame.mvScaleFactors);
    orb_last_points P = {n, valid.data(), u.data(), v.data(), invz.data(), octave.data(), angle.data(), desc.data(), obs.data()};
    int nmatches = 0;
    Check(orb_search_by_projection_last(g.h, &P, claimed.data(), th, bForward, bBackward, CurrentFrame.mbf, mbCheckOrientation,
                                        assigned.data(), &nmatches), "orb_search_by_projection_last");
    // what the reference leaves in CurrentFrame.mvpMapPoints: Last's map point where a query won (:1431), NULL where the
    // rotation filter reset a written entry (:1465, reported as -2), everything else untouched (-1)
    for (int i = 0; i < CurrentFrame.N; i++) {
        if (assigned[i] >= 0) CurrentFrame.mvpMapPoints[i] = LastFrame.mvpMapPoints[assigned[i]];
        else if (assigned[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint *>(NULL);
    }
    return nmatches;
}

int SearchForInitialization(Frame &F1, Frame &F2, std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12, int windowSize,
                            float mfNNratio, bool mbCheckOrientation) {
    const int n1 = (int)F1.mvKeysUn.size();
    vnMatches12 = std::vector<int>(n1, -1);                          // :408
    if (n1 == 0 || F2.mvKeysUn.empty()) return 0;
    GpuFrame g1(F1.mvKeysUn, F1.mDescriptors, F1.mvuRight, F1.mvScaleFactors);
    GpuFrame g2(F2.mvKeysUn, F2.mDescriptors, F2.mvuRight, F2.mvScaleFactors);
    std::vector<float> prev((size_t)n1 * 2);
    for (int i = 0; i < n1; i++) { prev[2 * i] = vbPrevMatched[i].x; prev[2 * i + 1] = vbPrevMatched[i].y; }
    std::vector<int32_t> m12(n1);
    int nmatches = 0;
    Check(orb_search_for_initialization(g1.h, g2.h, prev.data(), m12.data(), windowSize, mfNNratio, mbCheckOrientation, &nmatches),
          "orb_search_for_initialization");
    for (int i = 0; i < n1; i++) {
        vnMatches12[i] = m12[i];
        vbPrevMatched[i] = cv::Point2f(prev[2 * i], prev[2 * i + 1]);   // :518-520
    }
    return nmatches;
}

int SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint *> &vpMapPointMatches, float mfNNratio, bool mbCheckOrientation) {
    const std::vector<MapPoint *> vpMapPointsKF = pKF->GetMapPointMatches();
    vpMapPointMatches = std::vector<MapPoint *>(F.N, static_cast<MapPoint *>(NULL));   // :163
    if (F.N == 0 || vpMapPointsKF.empty()) return 0;
    // the FeatureVector merge-join of :175-264, flattened to CSR over the shared vocabulary nodes
    std::vector<int32_t> kfOff(1, 0), fOff(1, 0), kfIdx, fIdx;
    DBoW2::FeatureVector::const_iterator KFit = pKF->mFeatVec.begin(), Fit = F.mFeatVec.begin();
    const DBoW2::FeatureVector::const_iterator KFend = pKF->mFeatVec.end(), Fend = F.mFeatVec.end();
    while (KFit != KFend && Fit != Fend) {
        if (KFit->first == Fit->first) {
            kfIdx.insert(kfIdx.end(), KFit->second.begin(), KFit->second.end());
            fIdx.insert(fIdx.end(), Fit->second.begin(), Fit->second.end());
            kfOff.push_back((int32_t)kfIdx.size()); fOff.push_back((int32_t)fIdx.size());
            KFit++; Fit++;
        } else if (KFit->first < Fit->first) KFit = pKF->mFeatVec.lower_bound(Fit->first);
        else Fit = F.mFeatVec.lower_bound(KFit->first);
    }
    const int nKF = (int)vpMapPointsKF.size();
    std::vector<uint8_t> kfValid(nKF);
    for (int i = 0; i < nKF; i++) kfValid[i] = vpMapPointsKF[i] && !vpMapPointsKF[i]->isBad();   // :191-197
    const std::vector<float> none;
    GpuFrame gk(pKF->mvKeysUn, pKF->mDescriptors, none, pKF->mvScaleFactors);
    GpuFrame gf(F.mvKeys, F.mDescriptors, none, F.mvScaleFactors);   // the reference reads F.mvKeys[..].angle here (:238)
    std::vector<int32_t> matches(F.N);
    int nmatches = 0;
    Check(orb_search_by_bow(gk.h, gf.h, (int)kfOff.size() - 1, kfOff.data(), kfIdx.data(), fOff.data(), fIdx.data(), kfValid.data(), nullptr, 0,
                            mfNNratio, mbCheckOrientation, matches.data(), &nmatches), "orb_search_by_bow");
    for (int i = 0; i < F.N; i++)
        if (matches[i] >= 0) vpMapPointMatches[i] = vpMapPointsKF[matches[i]];   // :232
    return nmatches;
}

int SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, float mfNNratio, bool mbCheckOrientation) {
    const std::vector<MapPoint *> vpMapPoints1 = pKF1->GetMapPointMatches();
    const std::vector<MapPoint *> vpMapPoints2 = pKF2->GetMapPointMatches();
    vpMatches12 = std::vector<MapPoint *>(vpMapPoints1.size(), static_cast<MapPoint *>(NULL));   // :538
    const int n1 = (int)vpMapPoints1.size(), n2 = (int)vpMapPoints2.size();
    if (n1 == 0 || n2 == 0) return 0;
    std::vector<int32_t> off1(1, 0), off2(1, 0), idx1, idx2;           // the merge-join of :551-627 as CSR
    DBoW2::FeatureVector::const_iterator f1it = pKF1->mFeatVec.begin(), f2it = pKF2->mFeatVec.begin();
    const DBoW2::FeatureVector::const_iterator f1end = pKF1->mFeatVec.end(), f2end = pKF2->mFeatVec.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            idx1.insert(idx1.end(), f1it->second.begin(), f1it->second.end());
            idx2.insert(idx2.end(), f2it->second.begin(), f2it->second.end());
            off1.push_back((int32_t)idx1.size()); off2.push_back((int32_t)idx2.size());
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) f1it = pKF1->mFeatVec.lower_bound(f2it->first);
        else f2it = pKF2->mFeatVec.lower_bound(f1it->first);
    }
    std::vector<uint8_t> valid1(n1), valid2(n2);
    for (int i = 0; i < n1; i++) valid1[i] = vpMapPoints1[i] && !vpMapPoints1[i]->isBad();   // :561-566
    for (int i = 0; i < n2; i++) valid2[i] = vpMapPoints2[i] && !vpMapPoints2[i]->isBad();   // :577-583
    const std::vector<float> none;
    GpuFrame g1(pKF1->mvKeysUn, pKF1->mDescriptors, none, pKF1->mvScaleFactors, true);
    GpuFrame g2(pKF2->mvKeysUn, pKF2->mDescriptors, none, pKF2->mvScaleFactors, true);
    std::vector<int32_t> matches(n2);
    int nmatches = 0;
    Check(orb_search_by_bow(g1.h, g2.h, (int)off1.size() - 1, off1.data(), idx1.data(), off2.data(), idx2.data(), valid1.data(), valid2.data(),
                            1 /* best < TH_LOW, :601 */, mfNNratio, mbCheckOrientation, matches.data(), &nmatches), "orb_search_by_bow");
    for (int i2 = 0; i2 < n2; i2++)
        if (matches[i2] >= 0) vpMatches12[matches[i2]] = vpMapPoints2[i2];   // :605
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------
// Geometry of a batch of map points on the device (orb_project_points): transform, projection, image test, distance range and
// viewing angle with the rounding of the site that is being replaced (the rule switches).  What stays here is what only the
// reference's objects can answer: which points take part, their positions / normals / ranges through the public getters, and
// MapPoint::PredictScale.
// ---------------------------------------------------------------------------------------------------------------
struct Projected {
    std::vector<uint8_t> ok;
    std::vector<float> u, v, ur, dist;
};

static void SetPose(float *R9, float *t3, const cv::Mat &R, const cv::Mat &t) {
    for (int r = 0; r < 3; r++) {
        for (int c = 0; c < 3; c++) R9[3 * r + c] = R.at<float>(r, c);
        t3[r] = t.at<float>(r);
    }
}

static orb_projection MakeProjection(float fx, float fy, float cx, float cy, float bf, float minX, float maxX, float minY, float maxY) {
    orb_projection P;
    std::memset(&P, 0, sizeof(P));
    P.fx = fx; P.fy = fy; P.cx = cx; P.cy = cy; P.bf = bf;
    P.min_x = minX; P.max_x = maxX; P.min_y = minY; P.max_y = maxY;
    return P;
}

static void Project(const orb_projection &P, const std::vector<MapPoint *> &pts, const std::vector<uint8_t> &consider, Projected &out) {
    const int n = (int)pts.size();
    out.ok.assign(n, 0); out.u.assign(n, 0.f); out.v.assign(n, 0.f); out.ur.assign(n, 0.f); out.dist.assign(n, 0.f);
    if (n == 0) return;
    std::vector<float> X((size_t)n * 3, 0.f), Nrm((size_t)n * 3, 0.f), maxD(n, 0.f), minD(n, 0.f);
    for (int i = 0; i < n; i++) {
        if (!consider[i]) continue;
        MapPoint *pMP = pts[i];
        const cv::Mat x = pMP->GetWorldPos();
        for (int k = 0; k < 3; k++) X[3 * i + k] = x.at<float>(k);
        if (P.check_view_angle) {
            const cv::Mat nn = pMP->GetNormal();
            for (int k = 0; k < 3; k++) Nrm[3 * i + k] = nn.at<float>(k);
        }
        maxD[i] = pMP->GetMaxDistanceInvariance();
        minD[i] = pMP->GetMinDistanceInvariance();
    }
    orb_projection_points Q = {n, X.data(), P.check_view_angle ? Nrm.data() : NULL, maxD.data(), minD.data(), consider.data()};
    Check(orb_project_points(&P, &Q, ORB_HOST, out.ok.data(), out.u.data(), out.v.data(), out.ur.data(), out.dist.data()), "orb_project_points");
}

// Relocalisation refinement (Tracking::Relocalization): project the key frame's map points with the current pose and search
// the current frame; any map point already in the frame blocks its feature (:1544-1545).
int SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const std::set<MapPoint *> &sAlreadyFound, float th, int ORBdist,
                       bool mbCheckOrientation) {
    const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
    const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
    const cv::Mat Ow = -Rcw.t() * tcw;
    const std::vector<MapPoint *> vpMPs = pKF->GetMapPointMatches();
    const int n = (int)vpMPs.size();
    if (n == 0 || CurrentFrame.N == 0) return 0;
    const int nlev = (int)CurrentFrame.mvScaleFactors.size();
    std::vector<uint8_t> valid(n, 0), consider(n, 0), desc((size_t)n * 32 + 32, 0), claimed(CurrentFrame.N, 0);
    std::vector<float> radius(n, 0.f), angle(n, 0.f);
    std::vector<int32_t> minL(n, 0), maxL(n, 0), assigned(CurrentFrame.N);
    for (int i = 0; i < n; i++) {                                      // who takes part, :1491-1497
        MapPoint *pMP = vpMPs[i];
        angle[i] = pKF->mvKeysUn[i].angle;
        consider[i] = pMP && !pMP->isBad() && !sAlreadyFound.count(pMP);
    }
    // :1499-1524 on the device: 1.0/z in double, fx*xc*invz, the frame's bounds, no z test, no viewing angle
    orb_projection P = MakeProjection(CurrentFrame.fx, CurrentFrame.fy, CurrentFrame.cx, CurrentFrame.cy, 0.f, CurrentFrame.mnMinX,
                                      CurrentFrame.mnMaxX, CurrentFrame.mnMinY, CurrentFrame.mnMaxY);
    SetPose(P.R1, P.t1, Rcw, tcw);
    for (int k = 0; k < 3; k++) P.Ow[k] = Ow.at<float>(k);
    Projected G;
    Project(P, vpMPs, consider, G);
    for (int i = 0; i < n; i++) {
        if (!G.ok[i]) continue;
        MapPoint *pMP = vpMPs[i];
        const int nPredictedLevel = pMP->PredictScale(G.dist[i], CurrentFrame.mfLogScaleFactor);
        if (nPredictedLevel < 0 || nPredictedLevel >= nlev) continue;   // the reference indexes mvScaleFactors out of range here
        valid[i] = 1;
        radius[i] = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
        minL[i] = nPredictedLevel - 1; maxL[i] = nPredictedLevel + 1;   // GetFeaturesInArea's own level arguments (:1533)
        cv::Mat d = pMP->GetDescriptor();
        std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
    const std::vector<float> &u = G.u, &v = G.v;
    for (int i = 0; i < CurrentFrame.N; i++) claimed[i] = CurrentFrame.mvpMapPoints[i] != NULL;   // :1544-1545
    GpuFrame g(CurrentFrame.mvKeysUn, CurrentFrame.mDescriptors, CurrentFrame.mvuRight, CurrentFrame.mvScaleFactors);
    orb_window_queries Q;
    std::memset(&Q, 0, sizeof(Q));
    Q.n = n; Q.valid = valid.data(); Q.x = u.data(); Q.y = v.data(); Q.radius = radius.data();
    Q.min_level = minL.data(); Q.max_level = maxL.data(); Q.strict_levels = 0; Q.descriptors = desc.data(); Q.angle = angle.data();
    int nmatches = 0;
    Check(orb_search_windowed(g.h, &Q, claimed.data(), ORB_SELECT_BEST, ORBdist, 0.f, mbCheckOrientation, assigned.data(), &nmatches),
          "orb_search_windowed");
    for (int i = 0; i < CurrentFrame.N; i++) {
        if (assigned[i] >= 0) CurrentFrame.mvpMapPoints[i] = vpMPs[assigned[i]];             // :1559
        else if (assigned[i] == -2) CurrentFrame.mvpMapPoints[i] = static_cast<MapPoint *>(NULL);   // :1594
    }
    return nmatches;
}

// Loop closing (LoopClosing::ComputeSim3 / SearchAndFuse): project the loop candidates' points with the Sim3 and search the
// key frame; a feature that already holds a match is skipped (:374-375).
int SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, std::vector<MapPoint *> &vpMatched, int th) {
    const float &fx = pKF->fx;
    const float &fy = pKF->fy;
    const float &cx = pKF->cx;
    const float &cy = pKF->cy;
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);                  // :298-303
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    std::set<MapPoint *> spAlreadyFound(vpMatched.begin(), vpMatched.end());
    spAlreadyFound.erase(static_cast<MapPoint *>(NULL));
    const int n = (int)vpPoints.size();
    if (n == 0 || pKF->N == 0) return 0;
    const int nlev = (int)pKF->mvScaleFactors.size();
    std::vector<uint8_t> valid(n, 0), consider(n, 0), desc((size_t)n * 32 + 32, 0), claimed(pKF->N, 0);
    std::vector<float> radius(n, 0.f);
    std::vector<int32_t> minL(n, 0), maxL(n, 0), assigned(pKF->N);
    for (int iMP = 0; iMP < n; iMP++) consider[iMP] = !vpPoints[iMP]->isBad() && !spAlreadyFound.count(vpPoints[iMP]);   // :316
    // :318-353 on the device: z < 0 rejected, 1/z in float, x = xc*invz, IsInImage, viewing angle
    orb_projection P = MakeProjection(fx, fy, cx, cy, 0.f, pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY, pKF->mnMaxY);
    SetPose(P.R1, P.t1, Rcw, tcw);
    for (int k = 0; k < 3; k++) P.Ow[k] = Ow.at<float>(k);
    P.reject_negative_z = 1; P.invz_rule = 1; P.uv_rule = 1; P.bounds_rule = 1; P.check_view_angle = 1;
    Projected G;
    Project(P, vpPoints, consider, G);
    for (int iMP = 0; iMP < n; iMP++) {
        if (!G.ok[iMP]) continue;
        MapPoint *pMP = vpPoints[iMP];
        const int nPredictedLevel = pMP->PredictScale(G.dist[iMP], pKF->mfLogScaleFactor);
        if (nPredictedLevel < 0 || nPredictedLevel >= nlev) continue;
        valid[iMP] = 1;
        radius[iMP] = th * pKF->mvScaleFactors[nPredictedLevel];
        minL[iMP] = nPredictedLevel - 1; maxL[iMP] = nPredictedLevel;   // the in-loop filter of :378-381
        cv::Mat d = pMP->GetDescriptor();
        std::memcpy(&desc[(size_t)iMP * 32], d.ptr(0), 32);
    }
    const std::vector<float> &u = G.u, &v = G.v;
    for (int i = 0; i < pKF->N; i++) claimed[i] = vpMatched[i] != NULL;
    GpuFrame g(pKF->mvKeysUn, pKF->mDescriptors, pKF->mvuRight, pKF->mvScaleFactors, true);
    orb_window_queries Q;
    std::memset(&Q, 0, sizeof(Q));
    Q.n = n; Q.valid = valid.data(); Q.x = u.data(); Q.y = v.data(); Q.radius = radius.data(); Q.min_level = minL.data();
    Q.max_level = maxL.data(); Q.strict_levels = 1; Q.descriptors = desc.data();
    int nmatches = 0;
    Check(orb_search_windowed(g.h, &Q, claimed.data(), ORB_SELECT_BEST, 50 /* TH_LOW */, 0.f, 0, assigned.data(), &nmatches),
          "orb_search_windowed");
    for (int i = 0; i < pKF->N; i++)
        if (assigned[i] >= 0) vpMatched[i] = vpPoints[assigned[i]];   // :396
    return nmatches;
}

// ---------------------------------------------------------------------------------------------------------------
// LocalMapping / LoopClosing matchers.  The candidate loops run on the GPU (orb_search_window_best,
// orb_search_for_triangulation); the projection algebra before them and the map bookkeeping after them are the
// reference's statements, kept in the reference's order.
// ---------------------------------------------------------------------------------------------------------------
static const int TH_HIGH = 100, TH_LOW = 50;   // :37-38

// Window queries gathered for the points that pass the geometric tests.
struct WindowSet {
    std::vector<uint8_t> valid, desc;
    std::vector<float> u, v, ur, radius;
    std::vector<int32_t> minL, maxL, bestIdx, bestDist;
    explicit WindowSet(int n) : valid(n, 0), desc((size_t)n * 32 + 32, 0), u(n, 0.f), v(n, 0.f), ur(n, 0.f), radius(n, 0.f), minL(n, 0),
                                maxL(n, 0), bestIdx(n, -1), bestDist(n, 0x7fffffff) {}
    void Set(int i, float uu, float vv, float uur, int level, float r, MapPoint *pMP) {
        valid[i] = 1; u[i] = uu; v[i] = vv; ur[i] = uur; radius[i] = r; minL[i] = level - 1; maxL[i] = level;
        cv::Mat d = pMP->GetDescriptor();
        std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
    void Search(KeyFrame *pKF, bool reprojGate) {
        const int n = (int)valid.size();
        if (n == 0 || pKF->N == 0) return;
        GpuFrame g(pKF->mvKeysUn, pKF->mDescriptors, pKF->mvuRight, pKF->mvScaleFactors, true);
        orb_window_queries Q;
        std::memset(&Q, 0, sizeof(Q));
        Q.n = n; Q.valid = valid.data(); Q.x = u.data(); Q.y = v.data(); Q.radius = radius.data(); Q.min_level = minL.data();
        Q.max_level = maxL.data(); Q.strict_levels = 1; Q.descriptors = desc.data();
        orb_reproj_gate G = {ur.data(), pKF->mvInvLevelSigma2.data(), 5.99, 7.8};   // :930, :943
        Check(orb_search_window_best(g.h, &Q, reprojGate ? &G : NULL, bestIdx.data(), bestDist.data()), "orb_search_window_best");
    }
};

// PredictScale is not clamped in this fork (MapPoint.cc:357-366); the reference then indexes mvScaleFactors out of range.
static bool LevelOk(int level, const KeyFrame *pKF) { return level >= 0 && level < (int)pKF->mvScaleFactors.size(); }

int SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> > &vMatchedPairs,
                           bool bOnlyStereo, bool mbCheckOrientation) {
    const DBoW2::FeatureVector &vFeatVec1 = pKF1->mFeatVec;
    const DBoW2::FeatureVector &vFeatVec2 = pKF2->mFeatVec;
    // epipole in the second image, :667-673
    cv::Mat Cw = pKF1->GetCameraCenter();
    cv::Mat R2w = pKF2->GetRotation();
    cv::Mat t2w = pKF2->GetTranslation();
    cv::Mat C2 = R2w * Cw + t2w;
    const float invz = 1.0f / C2.at<float>(2);
    const float ex = pKF2->fx * C2.at<float>(0) * invz + pKF2->cx;
    const float ey = pKF2->fy * C2.at<float>(1) * invz + pKF2->cy;

    vMatchedPairs.clear();
    if (pKF1->N == 0 || pKF2->N == 0) return 0;
    std::vector<int32_t> off1(1, 0), off2(1, 0), idx1, idx2;   // the merge-join of :695-786 as CSR
    DBoW2::FeatureVector::const_iterator f1it = vFeatVec1.begin(), f2it = vFeatVec2.begin();
    const DBoW2::FeatureVector::const_iterator f1end = vFeatVec1.end(), f2end = vFeatVec2.end();
    while (f1it != f1end && f2it != f2end) {
        if (f1it->first == f2it->first) {
            idx1.insert(idx1.end(), f1it->second.begin(), f1it->second.end());
            idx2.insert(idx2.end(), f2it->second.begin(), f2it->second.end());
            off1.push_back((int32_t)idx1.size()); off2.push_back((int32_t)idx2.size());
            f1it++; f2it++;
        } else if (f1it->first < f2it->first) f1it = vFeatVec1.lower_bound(f2it->first);
        else f2it = vFeatVec2.lower_bound(f1it->first);
    }
    std::vector<uint8_t> usable1(pKF1->N), usable2(pKF2->N);
    for (int i = 0; i < pKF1->N; i++) usable1[i] = pKF1->GetMapPoint(i) == NULL;   // :704-708
    for (int i = 0; i < pKF2->N; i++) usable2[i] = pKF2->GetMapPoint(i) == NULL;   // :726-730
    float F[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) F[r * 3 + c] = F12.at<float>(r, c);
    GpuFrame g1(pKF1->mvKeysUn, pKF1->mDescriptors, pKF1->mvuRight, pKF1->mvScaleFactors, true);
    GpuFrame g2(pKF2->mvKeysUn, pKF2->mDescriptors, pKF2->mvuRight, pKF2->mvScaleFactors, true);
    std::vector<int32_t> m12(pKF1->N, -1);
    int nmatches = 0;
    Check(orb_search_for_triangulation(g1.h, g2.h, (int)off1.size() - 1, off1.data(), idx1.data(), off2.data(), idx2.data(), usable1.data(),
                                       usable2.data(), F, ex, ey, pKF2->mvLevelSigma2.data(), bOnlyStereo, mbCheckOrientation, m12.data(),
                                       &nmatches), "orb_search_for_triangulation");
    vMatchedPairs.reserve(nmatches);                                   // :817-825
    for (size_t i = 0, iend = m12.size(); i < iend; i++) {
        if (m12[i] < 0) continue;
        vMatchedPairs.push_back(std::make_pair(i, (size_t)m12[i]));
    }
    return nmatches;
}

int Fuse(KeyFrame *pKF, const std::vector<MapPoint *> &vpMapPoints, float th) {
    cv::Mat Rcw = pKF->GetRotation();
    cv::Mat tcw = pKF->GetTranslation();
    const float &fx = pKF->fx;
    const float &fy = pKF->fy;
    const float &cx = pKF->cx;
    const float &cy = pKF->cy;
    const float &bf = pKF->mbf;
    cv::Mat Ow = pKF->GetCameraCenter();
    const int nMPs = (int)vpMapPoints.size();
    WindowSet W(nMPs);
    std::vector<uint8_t> consider(nMPs, 0);
    for (int i = 0; i < nMPs; i++) consider[i] = vpMapPoints[i] != NULL;
    // the geometry of :855-888 on the device (none of it changes during the loop): z < 0 rejected, 1/z in float, x = xc*invz,
    // IsInImage, u_right = u - bf*invz, viewing angle
    orb_projection P = MakeProjection(fx, fy, cx, cy, bf, pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY, pKF->mnMaxY);
    SetPose(P.R1, P.t1, Rcw, tcw);
    for (int k = 0; k < 3; k++) P.Ow[k] = Ow.at<float>(k);
    P.reject_negative_z = 1; P.invz_rule = 1; P.uv_rule = 1; P.bounds_rule = 1; P.check_view_angle = 1;
    Projected G;
    Project(P, vpMapPoints, consider, G);
    for (int i = 0; i < nMPs; i++) {
        if (!G.ok[i]) continue;
        MapPoint *pMP = vpMapPoints[i];
        const int nPredictedLevel = pMP->PredictScale(G.dist[i], pKF->mfLogScaleFactor);
        if (!LevelOk(nPredictedLevel, pKF)) continue;
        W.Set(i, G.u[i], G.v[i], G.ur[i], nPredictedLevel, th * pKF->mvScaleFactors[nPredictedLevel], pMP);
    }
    W.Search(pKF, true);
    int nFused = 0;
    for (int i = 0; i < nMPs; i++) {                                   // the reference's loop order; state-dependent tests evaluated here
        MapPoint *pMP = vpMapPoints[i];
        if (!pMP) continue;
        if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;          // :851-852
        if (!W.valid[i]) continue;
        const int bestDist = W.bestDist[i], bestIdx = W.bestIdx[i];
        if (bestDist <= TH_LOW) {                                      // :954-973
            MapPoint *pMPinKF = pKF->GetMapPoint(bestIdx);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) {
                    if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
                    else pMPinKF->Replace(pMP);
                }
            } else {
                pMP->AddObservation(pKF, bestIdx);
                pKF->AddMapPoint(pMP, bestIdx);
            }
            nFused++;
        }
    }
    return nFused;
}

int Fuse(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, float th, std::vector<MapPoint *> &vpReplacePoint) {
    const float &fx = pKF->fx;
    const float &fy = pKF->fy;
    const float &cx = pKF->cx;
    const float &cy = pKF->cy;
    // decompose Scw, :989-994
    cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    cv::Mat Rcw = sRcw / scw;
    cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
    cv::Mat Ow = -Rcw.t() * tcw;
    const std::set<MapPoint *> spAlreadyFound = pKF->GetMapPoints();   // :997
    const int nPoints = (int)vpPoints.size();
    WindowSet W(nPoints);
    std::vector<uint8_t> consider(nPoints, 0);
    for (int iMP = 0; iMP < nPoints; iMP++)                            // :1009 (neither changes inside this function)
        consider[iMP] = !vpPoints[iMP]->isBad() && !spAlreadyFound.count(vpPoints[iMP]);
    // :1012-1043 on the device: z < 0 rejected, 1.0/z in double, x = xc*invz, IsInImage, viewing angle
    orb_projection P = MakeProjection(fx, fy, cx, cy, 0.f, pKF->mnMinX, pKF->mnMaxX, pKF->mnMinY, pKF->mnMaxY);
    SetPose(P.R1, P.t1, Rcw, tcw);
    for (int k = 0; k < 3; k++) P.Ow[k] = Ow.at<float>(k);
    P.reject_negative_z = 1; P.invz_rule = 0; P.uv_rule = 1; P.bounds_rule = 1; P.check_view_angle = 1;
    Projected G;
    Project(P, vpPoints, consider, G);
    for (int iMP = 0; iMP < nPoints; iMP++) {
        if (!G.ok[iMP]) continue;
        MapPoint *pMP = vpPoints[iMP];
        const int nPredictedLevel = pMP->PredictScale(G.dist[iMP], pKF->mfLogScaleFactor);
        if (!LevelOk(nPredictedLevel, pKF)) continue;
        W.Set(iMP, G.u[iMP], G.v[iMP], 0.f, nPredictedLevel, th * pKF->mvScaleFactors[nPredictedLevel], pMP);
    }
    W.Search(pKF, false);
    int nFused = 0;
    for (int iMP = 0; iMP < nPoints; iMP++) {                          // :1075-1090, in order: AddMapPoint changes GetMapPoint for later points
        if (!W.valid[iMP]) continue;
        MapPoint *pMP = vpPoints[iMP];
        const int bestDist = W.bestDist[iMP], bestIdx = W.bestIdx[iMP];
        if (bestDist <= TH_LOW) {
            MapPoint *pMPinKF = pKF->GetMapPoint(bestIdx);
            if (pMPinKF) {
                if (!pMPinKF->isBad()) vpReplacePoint[iMP] = pMPinKF;
            } else {
                pMP->AddObservation(pKF, bestIdx);
                pKF->AddMapPoint(pMP, bestIdx);
            }
            nFused++;
        }
    }
    return nFused;
}

int SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, const float &s12, const cv::Mat &R12,
                 const cv::Mat &t12, float th) {
    const float &fx = pKF1->fx;
    const float &fy = pKF1->fy;
    const float &cx = pKF1->cx;
    const float &cy = pKF1->cy;
    // camera 1 / 2 from world and the similarity between them, :1113-1123
    cv::Mat R1w = pKF1->GetRotation();
    cv::Mat t1w = pKF1->GetTranslation();
    cv::Mat R2w = pKF2->GetRotation();
    cv::Mat t2w = pKF2->GetTranslation();
    cv::Mat sR12 = s12 * R12;
    cv::Mat sR21 = (1.0 / s12) * R12.t();
    cv::Mat t21 = -sR21 * t12;
    const std::vector<MapPoint *> vpMapPoints1 = pKF1->GetMapPointMatches();
    const int N1 = (int)vpMapPoints1.size();
    const std::vector<MapPoint *> vpMapPoints2 = pKF2->GetMapPointMatches();
    const int N2 = (int)vpMapPoints2.size();
    std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
    for (int i = 0; i < N1; i++) {                                     // :1134-1144
        MapPoint *pMP = vpMatches12[i];
        if (pMP) {
            vbAlreadyMatched1[i] = true;
            int idx2 = pMP->GetIndexInKeyFrame(pKF2);
            if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
        }
    }
    // both directions share one shape: the map points of KF `a` into KF `b` (:1150-1226 and :1229-1305), on the device: camera a, then
    // the similarity into camera b, z < 0 rejected, 1.0/z in double, x = xc*invz, IsInImage, the distance measured in camera b
    struct Dir {
        static void Gather(const std::vector<MapPoint *> &pts, const std::vector<bool> &already, const cv::Mat &Raw, const cv::Mat &taw,
                           const cv::Mat &sRba, const cv::Mat &tba, KeyFrame *pKFb, float fx, float fy, float cx, float cy, float th,
                           WindowSet &W) {
            const int n = (int)pts.size();
            std::vector<uint8_t> consider(n, 0);
            for (int i = 0; i < n; i++) consider[i] = pts[i] && !already[i] && !pts[i]->isBad();
            orb_projection P = MakeProjection(fx, fy, cx, cy, 0.f, pKFb->mnMinX, pKFb->mnMaxX, pKFb->mnMinY, pKFb->mnMaxY);
            SetPose(P.R1, P.t1, Raw, taw);
            SetPose(P.R2, P.t2, sRba, tba);
            P.two_stage = 1; P.reject_negative_z = 1; P.invz_rule = 0; P.uv_rule = 1; P.bounds_rule = 1; P.dist_rule = 1;
            Projected G;
            Project(P, pts, consider, G);
            for (int i = 0; i < n; i++) {
                if (!G.ok[i]) continue;
                MapPoint *pMP = pts[i];
                const int nPredictedLevel = pMP->PredictScale(G.dist[i], pKFb->mfLogScaleFactor);
                if (!LevelOk(nPredictedLevel, pKFb)) continue;
                W.Set(i, G.u[i], G.v[i], 0.f, nPredictedLevel, th * pKFb->mvScaleFactors[nPredictedLevel], pMP);
            }
        }
    };
    WindowSet W1(N1), W2(N2);
    Dir::Gather(vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, pKF2, fx, fy, cx, cy, th, W1);
    Dir::Gather(vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, pKF1, fx, fy, cx, cy, th, W2);
    W1.Search(pKF2, false);
    W2.Search(pKF1, false);
    std::vector<int> vnMatch1(N1, -1), vnMatch2(N2, -1);
    for (int i1 = 0; i1 < N1; i1++)
        if (W1.valid[i1] && W1.bestDist[i1] <= TH_HIGH) vnMatch1[i1] = W1.bestIdx[i1];   // :1222-1225
    for (int i2 = 0; i2 < N2; i2++)
        if (W2.valid[i2] && W2.bestDist[i2] <= TH_HIGH) vnMatch2[i2] = W2.bestIdx[i2];   // :1301-1304
    int nFound = 0;                                                    // :1308-1326
    for (int i1 = 0; i1 < N1; i1++) {
        int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            int idx1 = vnMatch2[idx2];
            if (idx1 == i1) {
                vpMatches12[i1] = vpMapPoints2[idx2];
                nFound++;
            }
        }
    }
    return nFound;
}

}  // namespace orbgpu_shim
}  // namespace ORB_SLAM2
