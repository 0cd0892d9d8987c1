// Drop-in replacement for ORB_SLAM2/include/ORBextractor.h (ref :45-118): same class name, namespace,
// constructor, operator(), getters and public mvImagePyramid, so Frame.cc / Tracking.cc / the AR-1.3 copies
// compile unchanged.  The work happens in liborbgpu (include/orbgpu.h); this header holds no arithmetic.
//
// Differences a maintainer should know about:
//   * the extractor owns CUDA buffers sized for max_width x max_height (defaults fit 1920x1080; larger frames
//     re-create the handle transparently);
//   * mvImagePyramid is a LAZY host mirror of the device pyramid: nothing is copied until somebody indexes it; the first
//     read after an extraction downloads the eight levels (with the reference's 19-px border around each, ref :1052-1056).
//     Frame::ComputeStereoMatches, the reference's only reader (Frame.cc:476,566,578,583), is replaced by Frame_stereo.cc,
//     which reads the device pyramid in place, so a drop-in build never pays for the mirror; the reference's own
//     ComputeStereoMatches keeps working against it.  SetMirrorPyramid(true) restores the eager copy after every call;
//   * without a CUDA device the constructor throws std::runtime_error -- there is no CPU fallback.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>

#include "cv_compat.h"

struct orb_extractor;

namespace ORB_SLAM2 {

class ORBextractor;

// What readers of the reference's `std::vector<cv::Mat> mvImagePyramid` use of it -- indexing, size(), iteration, conversion
// to the vector -- with the download from the device deferred to the first such access after an extraction.
class PyramidMirror {
public:
    PyramidMirror() : mpOwner(nullptr), mbStale(false) {}
    const cv::Mat &operator[](size_t level) const { Fill(); return mvLevels[level]; }
    cv::Mat &operator[](size_t level) { Fill(); return mvLevels[level]; }
    size_t size() const { return mvLevels.size(); }
    bool empty() const { return mvLevels.empty(); }
    void resize(size_t n) { mvLevels.resize(n); }
    std::vector<cv::Mat>::const_iterator begin() const { Fill(); return mvLevels.begin(); }
    std::vector<cv::Mat>::const_iterator end() const { Fill(); return mvLevels.end(); }
    operator const std::vector<cv::Mat> &() const { Fill(); return mvLevels; }

private:
    friend class ORBextractor;
    void Fill() const;
    ORBextractor *mpOwner;
    mutable std::vector<cv::Mat> mvLevels;
    mutable bool mbStale;   // the device pyramid is newer than mvLevels
};

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ~ORBextractor();
    ORBextractor(const ORBextractor &) = delete;
    ORBextractor &operator=(const ORBextractor &) = delete;

    // Compute the ORB features and descriptors on an image.  Mask is ignored, as in the reference (ref :60).
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    PyramidMirror mvImagePyramid;   // ref :88 `std::vector<cv::Mat> mvImagePyramid`

    // --- additions (not in the reference)
    void SetMirrorPyramid(bool on) { mbMirrorPyramid = on; }   // true: copy the pyramid to the host after EVERY call (default: lazily)
    void DownloadPyramid();                                     // fills mvImagePyramid's levels from the device now
    orb_extractor *Handle() { return mpHandle; }   // for ORBGPU_ComputeStereoMatches
    int DefaultCapacity() const { return nfeatures + 4 * nlevels + 256; }   // output slots handed to the library (grown on demand)
    orb_extractor *StereoHandle(int width, int height);   // two-frame handle of ORBGPU_ExtractStereo, created on first use

protected:
    friend class PyramidMirror;
    void EnsureHandle(int width, int height);

    int nfeatures;
    double scaleFactor;   // double, as in the reference (ORBextractor.h:105)
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orb_extractor *mpHandle;
    orb_extractor *mpStereoHandle;
    int mStereoMaxWidth, mStereoMaxHeight;
    int mMaxWidth, mMaxHeight;
    bool mbMirrorPyramid;
    std::vector<unsigned char> mKpBuf, mDescBuf;
    std::vector<cv::Mat> mPlanes;   // padded planes backing mvImagePyramid
};

// ref: void Frame::ComputeStereoMatches(), Frame.cc:471-643, as a free function over the members it touches.
// mvuRight / mvDepth are resized to mvKeys.size().  Returns false when no match survived (the reference indexes an
// empty vector there, Frame.cc:630).
bool ORBGPU_ComputeStereoMatches(ORBextractor *left, ORBextractor *right, const std::vector<cv::KeyPoint> &mvKeys,
                                 const cv::Mat &mDescriptors, const std::vector<cv::KeyPoint> &mvKeysRight,
                                 const cv::Mat &mDescriptorsRight, float mbf, float mb, std::vector<float> &mvuRight,
                                 std::vector<float> &mvDepth);

// ref: the stereo Frame constructor's front end, Frame.cc:83-95 (ExtractORB on both images in two threads, then
// ComputeStereoMatches), as ONE library call on a two-frame handle owned by `left` (orb_stereo_frame): both extractions run
// as one batch, the association reads the device-resident keypoints, the host waits once.  Fills what the three reference
// calls fill: mvKeys / mDescriptors, mvKeysRight / mDescriptorsRight, mvuRight / mvDepth.  `left` supplies the extractor
// parameters (the reference builds both extractors with the same ones, Tracking.cc:455-461); mvImagePyramid is not mirrored
// on this path.  Returns false when no match survived.
bool ORBGPU_ExtractStereo(ORBextractor *left, const cv::Mat &imLeft, const cv::Mat &imRight, std::vector<cv::KeyPoint> &mvKeys,
                          cv::Mat &mDescriptors, std::vector<cv::KeyPoint> &mvKeysRight, cv::Mat &mDescriptorsRight, float mbf,
                          float mb, std::vector<float> &mvuRight, std::vector<float> &mvDepth);

}  // namespace ORB_SLAM2

#endif
