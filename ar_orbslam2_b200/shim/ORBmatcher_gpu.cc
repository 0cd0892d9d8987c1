// See ORBmatcher_gpu.h.  Line references are to ORB_SLAM2/src/ORBmatcher.cc of the reference.
#include "ORBmatcher_gpu.h"

#include <cstring>
#include <stdexcept>
#include <string>

#include "orbgpu.h"

namespace ORB_SLAM2 {
namespace orbgpu_shim {

static void Check(int rc, const char *what) {
    if (rc != ORB_OK) throw std::runtime_error(std::string("liborbgpu: ") + what + ": " + orb_last_error());
}

// The part of a Frame the matchers read, uploaded as an orb_frame (mvKeysUn, mDescriptors, mvuRight, the image bounds that
// define the 64x48 grid, mvScaleFactors).  Built per call; a production integration caches it in the Frame right after
// AssignFeaturesToGrid() (ref Frame.cc:232).
struct GpuFrame {
    orb_frame *h = nullptr;
    std::vector<orb_keypoint> keys;
    std::vector<uint8_t> desc;
    template <class KeyVec>
    GpuFrame(const KeyVec &keysUn, const cv::Mat &descriptors, const std::vector<float> &uRight, const std::vector<float> &scaleFactors) {
        const int n = (int)keysUn.size();
        keys.resize(n);
        desc.resize((size_t)n * 32 + 32);
        for (int i = 0; i < n; i++) {
            const cv::KeyPoint &k = keysUn[i];
            keys[i].x = k.pt.x; keys[i].y = k.pt.y; keys[i].size = k.size; keys[i].angle = k.angle; keys[i].response = k.response;
            keys[i].octave = k.octave; keys[i].class_id = k.class_id;
            std::memcpy(&desc[(size_t)i * 32], descriptors.ptr(i), 32);
        }
        orb_frame_desc d;
        d.n = n; d.keys_un = keys.data(); d.descriptors = desc.data();
        d.u_right = ((int)uRight.size() == n && n > 0) ? uRight.data() : nullptr;
        d.min_x = Frame::mnMinX; d.max_x = Frame::mnMaxX; d.min_y = Frame::mnMinY; d.max_y = Frame::mnMaxY;
        d.nlevels = (int)scaleFactors.size(); d.scale_factors = scaleFactors.data();
        Check(orb_frame_create(&d, &h), "orb_frame_create");
    }
    ~GpuFrame() { orb_frame_destroy(h); }
    GpuFrame(const GpuFrame &) = delete;
    GpuFrame &operator=(const GpuFrame &) = delete;
};

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

}  // namespace orbgpu_shim
}  // namespace ORB_SLAM2
