// Drop-in replacement for ORB_SLAM2/include/ORBmatcher.h (ref :34-114): the class ORB_SLAM2::ORBmatcher with the reference's
// constructor, static DescriptorDistance, its eleven matcher entry points (same names, parameter types, defaults and return
// values) and the three public constants, so Tracking.cc, LocalMapping.cc, LoopClosing.cc, MapPoint.cc, Frame.cc and the
// AR-1.3 marker pipeline compile and link against it unchanged.  Put this directory in front of ORB_SLAM2/include on the
// include path and build ORBmatcher.cc + ORBmatcher_gpu.cc instead of ORB_SLAM2/src/ORBmatcher.cc.
//
// Every member forwards to the body in ORBmatcher_gpu.cc, which gathers flat arrays under the reference's getters and calls
// liborbgpu (include/orbgpu.h); no arithmetic lives here.  An ORBmatcher is a value type the callers construct on their stack
// from three threads (ref Tracking.cc:952, LocalMapping.cc:191, LoopClosing.cc:252): it holds two scalars and no GPU state;
// the device-side frame indices are cached per thread inside ORBmatcher_gpu.cc.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <set>
#include <utility>
#include <vector>

#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"

namespace ORB_SLAM2 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);                                   // ref ORBmatcher.h:37

    static int DescriptorDistance(const cv::Mat &a, const cv::Mat &b);                       // ref :40, ORBmatcher.cc:1650-1666

    int SearchByProjection(Frame &F, const std::vector<MapPoint *> &vpMapPoints, const float th = 3);                   // ref :44
    int SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono);              // ref :49
    int SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const std::set<MapPoint *> &sAlreadyFound, const float th,
                           const int ORBdist);                                                                          // ref :54
    int SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints,
                           std::vector<MapPoint *> &vpMatched, int th);                                                 // ref :60
    int SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint *> &vpMapPointMatches);                               // ref :67
    int SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12);                              // ref :69
    int SearchForInitialization(Frame &F1, Frame &F2, std::vector<cv::Point2f> &vbPrevMatched,
                                std::vector<int> &vnMatches12, int windowSize = 10);                                    // ref :73
    int SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12,
                               std::vector<std::pair<size_t, size_t> > &vMatchedPairs, const bool bOnlyStereo);         // ref :78
    int SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, const float &s12,
                     const cv::Mat &R12, const cv::Mat &t12, const float th);                                           // ref :84
    int Fuse(KeyFrame *pKF, const std::vector<MapPoint *> &vpMapPoints, const float th = 3.0);                           // ref :89
    int Fuse(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, float th,
             std::vector<MapPoint *> &vpReplacePoint);                                                                  // ref :93

public:
    static const int TH_LOW;         // 50   (ref ORBmatcher.cc:38)
    static const int TH_HIGH;        // 100  (ref :37)
    static const int HISTO_LENGTH;   // 30   (ref :39)

protected:
    // (the reference's protected helpers CheckDistEpipolarLine / RadiusByViewingCos / ComputeThreeMaxima are device code
    // now: csrc/orb_track.cu k_triangulate, orb_search_by_projection_points, three_maxima)
    float mfNNratio;
    bool mbCheckOrientation;
};

}  // namespace ORB_SLAM2

#endif  // ORBMATCHER_H
