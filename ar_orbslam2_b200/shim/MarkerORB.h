// Drop-in pieces for the reference's marker pipeline (ORB_SLAM2/src/Marker.cc), which calls OpenCV 2.4 directly:
//   `ORB orb; orb(image, Mat(), keys, descriptors);`             Marker.cc:83-84 (setTargetImage), :107-108 (Match)
//   `BruteForceMatcher<HammingLUT> matcher; matcher.match(d1, d2, matches);`                       Marker.cc:111-121
//   `Marker::searchMatches(keys1, descp1, keys2, descp2, matches)` / `hamdist2`                    Marker.cc:314-358
// orbgpu_shim::ORB has cv::ORB's constructor and operator() (features2d.hpp of OpenCV 2.4); a maintainer writes
// `orbgpu_shim::ORB orb;` (or `using orbgpu_shim::ORB;` after `using namespace cv;` is narrowed) and nothing else changes.
// The arithmetic runs in liborbgpu (orb_cvorb_*, orb_hamming_top2, orb_search_matches_ratio) and is pinned to cv2 4.13's
// ORB: key points (fields and ORDER) and descriptors are bit-identical to it.  No CPU fallback.
#ifndef ORBGPU_MARKER_ORB_H
#define ORBGPU_MARKER_ORB_H

#include <vector>

#include "cv_compat.h"

namespace orbgpu_shim {

class ORB {
public:
    enum { kBytes = 32, HARRIS_SCORE = 0, FAST_SCORE = 1 };
    explicit ORB(int nfeatures = 500, float scaleFactor = 1.2f, int nlevels = 8, int edgeThreshold = 31, int firstLevel = 0,
                 int WTA_K = 2, int scoreType = HARRIS_SCORE, int patchSize = 31);
    // cv::ORB::operator()(image, mask, keypoints, descriptors, useProvidedKeypoints = false).  The marker path passes an empty
    // mask and lets the detector find the key points; a non-empty mask or useProvidedKeypoints throws std::invalid_argument.
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints, cv::OutputArray descriptors,
                    bool useProvidedKeypoints = false) const;
    int descriptorSize() const { return kBytes; }

private:
    int nfeatures, nlevels, edgeThreshold, scoreType;
    float scaleFactor;
};

// BruteForceMatcher<HammingLUT>::match(queryDescriptors, trainDescriptors, matches): for every query row the nearest train row
// (first one on ties), DMatch(queryIdx, trainIdx, 0, distance); no match entries when the train set is empty.
void BruteForceMatchHamming(const cv::Mat &queryDescriptors, const cv::Mat &trainDescriptors, std::vector<cv::DMatch> &matches);

// Marker::searchMatches (Marker.cc:314-349): descp1 = marker descriptors (train), descp2 = frame descriptors (query); pushes
// DMatch(i, trainIdx, 0, min_dist) for every accepted query i (ratio 0.8, distance <= 50).  Returns true as the reference does.
bool SearchMatches(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches);

// AR-1.3's copies (AR-1.3/src/ORBMatcher.cpp): naive_nn_search (:44-68, nearest train row, kept when distance <= 50) and
// naive_nn_search2 (:70-102, the ratio matcher above plus the file-scope `minD` / `maxD` bookkeeping over all pairs that
// ORBMatcher::Match filters with, :147-152): pass the reference's two globals.
void NaiveNNSearch(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches);
void NaiveNNSearch2(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches, unsigned int &minD, unsigned int &maxD);

// The ORB objects above share per-thread device handles that are NOT freed at thread or process exit (no CUDA call may run from
// an exit-time destructor).  A thread that ends while the process continues calls this first.  No reference counterpart.
void ReleaseThreadHandles();

}  // namespace orbgpu_shim
#endif
