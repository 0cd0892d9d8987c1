// ORB_SLAM2::ORBmatcher over liborbgpu: see ORBmatcher.h.  Replaces ORB_SLAM2/src/ORBmatcher.cc in the reference's
// CMakeLists.txt:44-66 together with ORBmatcher_gpu.cc (the bodies).
#include "ORBmatcher.h"

#include <stdint.h>

#include "ORBmatcher_gpu.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// One pair on the host: Frame::ComputeStereoMatches' reference body (Frame.cc:544) and MapPoint::ComputeDistinctiveDescriptors
// (MapPoint.cc:260) call this per pair; bulk distances go through orb_descriptor_distance / the matchers.
int ORBmatcher::DescriptorDistance(const cv::Mat &a, const cv::Mat &b) {
    const uint64_t *pa = a.ptr<uint64_t>(), *pb = b.ptr<uint64_t>();
    int dist = 0;
    for (int i = 0; i < 4; i++) dist += __builtin_popcountll(pa[i] ^ pb[i]);
    return dist;
}

int ORBmatcher::SearchByProjection(Frame &F, const std::vector<MapPoint *> &vpMapPoints, const float th) {
    return orbgpu_shim::SearchByProjection(F, vpMapPoints, th, mfNNratio);
}
int ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono) {
    return orbgpu_shim::SearchByProjection(CurrentFrame, LastFrame, th, bMono, mbCheckOrientation);
}
int ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const std::set<MapPoint *> &sAlreadyFound, const float th,
                                   const int ORBdist) {
    return orbgpu_shim::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist, mbCheckOrientation);
}
int ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints,
                                   std::vector<MapPoint *> &vpMatched, int th) {
    return orbgpu_shim::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th);
}
int ORBmatcher::SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint *> &vpMapPointMatches) {
    return orbgpu_shim::SearchByBoW(pKF, F, vpMapPointMatches, mfNNratio, mbCheckOrientation);
}
int ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12) {
    return orbgpu_shim::SearchByBoW(pKF1, pKF2, vpMatches12, mfNNratio, mbCheckOrientation);
}
int ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12,
                                        int windowSize) {
    return orbgpu_shim::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize, mfNNratio, mbCheckOrientation);
}
int ORBmatcher::SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> > &vMatchedPairs,
                                       const bool bOnlyStereo) {
    return orbgpu_shim::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, mbCheckOrientation);
}
int ORBmatcher::SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, const float &s12, const cv::Mat &R12,
                             const cv::Mat &t12, const float th) {
    return orbgpu_shim::SearchBySim3(pKF1, pKF2, vpMatches12, s12, R12, t12, th);
}
int ORBmatcher::Fuse(KeyFrame *pKF, const std::vector<MapPoint *> &vpMapPoints, const float th) {
    return orbgpu_shim::Fuse(pKF, vpMapPoints, th);
}
int ORBmatcher::Fuse(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, float th, std::vector<MapPoint *> &vpReplacePoint) {
    return orbgpu_shim::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint);
}

}  // namespace ORB_SLAM2
