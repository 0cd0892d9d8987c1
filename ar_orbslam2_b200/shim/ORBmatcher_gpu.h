// GPU bodies for the tracking-thread entry points of ORB_SLAM2::ORBmatcher (ref: ORB_SLAM2/include/ORBmatcher.h:42-94).
// Meant to be compiled INSIDE the reference tree (it includes the reference's Frame.h / KeyFrame.h / MapPoint.h): each
// function gathers the flat arrays liborbgpu wants under the reference's own getters, calls the C ABI and writes the
// results back exactly where the reference method writes them.  ORBmatcher.cc keeps its class; a maintainer replaces the
// body of the corresponding method by a one-line call.
#ifndef ORBMATCHER_GPU_H
#define ORBMATCHER_GPU_H

#include <set>
#include <utility>
#include <vector>

#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"

namespace ORB_SLAM2 {
namespace orbgpu_shim {

// Every thread that calls a matcher below keeps up to 48 device-resident frame indices (about 0.4 MB each) in a per-thread cache.
// Nothing is freed at thread or process exit (CUDA objects must not be destroyed from exit-time destructors); a thread that ends
// while the process continues -- System::Shutdown joining LocalMapping / LoopClosing, say -- calls this first.  No reference
// counterpart.
void ReleaseThreadFrameCache();

// ref ORBmatcher::SearchByProjection(Frame &F, const std::vector<MapPoint*> &vpMapPoints, const float th), ORBmatcher.cc:45-129
int SearchByProjection(Frame &F, const std::vector<MapPoint *> &vpMapPoints, float th, float mfNNratio);

// ref ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, const float th, const bool bMono), :1331-1473
int SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, float th, bool bMono, bool mbCheckOrientation);

// ref ORBmatcher::SearchForInitialization(Frame &F1, Frame &F2, vector<cv::Point2f>&, vector<int>&, int windowSize), :405-523
int SearchForInitialization(Frame &F1, Frame &F2, std::vector<cv::Point2f> &vbPrevMatched, std::vector<int> &vnMatches12,
                            int windowSize, float mfNNratio, bool mbCheckOrientation);

// ref ORBmatcher::SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint*> &vpMapPointMatches), :159-288
int SearchByBoW(KeyFrame *pKF, Frame &F, std::vector<MapPoint *> &vpMapPointMatches, float mfNNratio, bool mbCheckOrientation);

// ref ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12), :525-658
int SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, float mfNNratio, bool mbCheckOrientation);

// ref ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th, ORBdist), :1475-1602
int SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const std::set<MapPoint *> &sAlreadyFound, float th, int ORBdist,
                       bool mbCheckOrientation);

// ref ORBmatcher::SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, vector<MapPoint*> &vpMatched, int th), :290-403
int SearchByProjection(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, std::vector<MapPoint *> &vpMatched, int th);

// ---- LocalMapping / LoopClosing matchers (M9) ----
// ref ORBmatcher::SearchForTriangulation(KeyFrame*, KeyFrame*, cv::Mat F12, vector<pair<size_t,size_t>>&, bool bOnlyStereo), :660-826
int SearchForTriangulation(KeyFrame *pKF1, KeyFrame *pKF2, cv::Mat F12, std::vector<std::pair<size_t, size_t> > &vMatchedPairs,
                           bool bOnlyStereo, bool mbCheckOrientation);

// ref ORBmatcher::Fuse(KeyFrame *pKF, const vector<MapPoint*> &vpMapPoints, const float th), :828-978
int Fuse(KeyFrame *pKF, const std::vector<MapPoint *> &vpMapPoints, float th);

// ref ORBmatcher::Fuse(KeyFrame *pKF, cv::Mat Scw, const vector<MapPoint*> &vpPoints, float th, vector<MapPoint*> &vpReplacePoint), :980-1103
int Fuse(KeyFrame *pKF, cv::Mat Scw, const std::vector<MapPoint *> &vpPoints, float th, std::vector<MapPoint *> &vpReplacePoint);

// ref ORBmatcher::SearchBySim3(KeyFrame*, KeyFrame*, vector<MapPoint*> &vpMatches12, s12, R12, t12, th), :1105-1329
int SearchBySim3(KeyFrame *pKF1, KeyFrame *pKF2, std::vector<MapPoint *> &vpMatches12, const float &s12, const cv::Mat &R12,
                 const cv::Mat &t12, float th);

}  // namespace orbgpu_shim
}  // namespace ORB_SLAM2
#endif
