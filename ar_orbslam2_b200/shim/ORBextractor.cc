// Drop-in ORBextractor over liborbgpu; see ORBextractor.h.  ref: ORB_SLAM2/src/ORBextractor.cc.
#include "ORBextractor.h"

#include <cassert>
#include <cstring>
#include <stdexcept>
#include <string>

#include "../../include/orbgpu.h"

static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint must be the 28-byte POD liborbgpu writes");

namespace ORB_SLAM2 {

static void Check(int rc, const char *what) {
    if (rc != ORB_OK) throw std::runtime_error(std::string("liborbgpu: ") + what + ": " + orb_last_error());
}

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mpHandle(nullptr), mpStereoHandle(nullptr), mStereoMaxWidth(0), mStereoMaxHeight(0), mMaxWidth(0), mMaxHeight(0),
      mbMirrorPyramid(false) {
    mvImagePyramid.resize(nlevels);
    mvImagePyramid.mpOwner = this;
    EnsureHandle(1920, 1080);
}

ORBextractor::~ORBextractor() {
    orb_extractor_destroy(mpHandle);
    if (mpStereoHandle) orb_extractor_destroy(mpStereoHandle);
}

void ORBextractor::EnsureHandle(int width, int height) {
    if (mpHandle && width <= mMaxWidth && height <= mMaxHeight) return;
    if (mpHandle) orb_extractor_destroy(mpHandle);
    mpHandle = nullptr;
    mMaxWidth = width > mMaxWidth ? width : mMaxWidth;
    mMaxHeight = height > mMaxHeight ? height : mMaxHeight;
    orb_params p;
    std::memset(&p, 0, sizeof(p));
    p.nfeatures = nfeatures; p.scale_factor = (float)scaleFactor; p.nlevels = nlevels;
    p.ini_th_fast = iniThFAST; p.min_th_fast = minThFAST;
    p.max_width = mMaxWidth; p.max_height = mMaxHeight; p.max_batch = 1; p.device = -1; p.blur_taps = 0;
    Check(orb_extractor_create(&p, &mpHandle), "orb_extractor_create");
    // the tables the reference constructor builds (ref :408-437) come back from the library
    mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
    Check(orb_extractor_tables(mpHandle, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data(), mnFeaturesPerLevel.data()), "orb_extractor_tables");
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint> &_keypoints,
                              cv::OutputArray _descriptors) {
    if (_image.empty()) return;                       // ref :987-988
    cv::Mat image = _image.getMat();
    assert(image.type() == CV_8UC1);                  // ref :991
    EnsureHandle(image.cols, image.rows);

    int capacity = DefaultCapacity(), n = 0;
    for (;;) {
        mKpBuf.resize((size_t)capacity * sizeof(orb_keypoint));
        mDescBuf.resize((size_t)capacity * 32);
        const int rc = orb_extract(mpHandle, image.data, image.cols, image.rows, (int)image.step, ORB_HOST,
                                   (orb_keypoint *)mKpBuf.data(), mDescBuf.data(), capacity, &n);
        if (rc == ORB_ERR_CAPACITY && n > capacity) { capacity = n; continue; }   // n_out holds the required size
        Check(rc, "orb_extract");
        break;
    }
    if (n == 0) _descriptors.release();               // ref :1005-1006
    else {
        _descriptors.create(n, 32, CV_8U);            // ref :1008
        cv::Mat d = _descriptors.getMat();
        for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), mDescBuf.data() + (size_t)i * 32, 32);
    }
    _keypoints.resize(n);                             // ref :1012-1043
    if (n) std::memcpy((void *)_keypoints.data(), mKpBuf.data(), (size_t)n * sizeof(orb_keypoint));

    mvImagePyramid.mbStale = true;                    // the host mirror is refreshed when somebody reads it
    if (mbMirrorPyramid) DownloadPyramid();
}

void PyramidMirror::Fill() const {
    if (mbStale && mpOwner) mpOwner->DownloadPyramid();
}

void ORBextractor::DownloadPyramid() {
    mvImagePyramid.mbStale = false;
    mPlanes.resize(nlevels);
    for (int l = 0; l < nlevels; l++) {
        int w = 0, h = 0, pitch = 0;
        Check(orb_pyramid_level(mpHandle, 0, l, nullptr, &w, &h, &pitch), "orb_pyramid_level");
        // the reference's level is the interior ROI of a (w+38)x(h+38) plane (ref :1052-1056): keep that layout
        // so code that walks into the 19-px border (Frame.cc:583) finds the same bytes
        cv::Mat plane(h + 2 * ORB_EDGE_THRESHOLD, w + 2 * ORB_EDGE_THRESHOLD, CV_8UC1);
        Check(orb_pyramid_download(mpHandle, 0, l, 1, plane.data, (int)plane.step), "orb_pyramid_download");
        mPlanes[l] = plane;
        mvImagePyramid.mvLevels[l] = cv::Mat(h, w, CV_8UC1, plane.data + (size_t)ORB_EDGE_THRESHOLD * plane.step + ORB_EDGE_THRESHOLD,
                                             plane.step);
    }
}

bool ORBGPU_ComputeStereoMatches(ORBextractor *left, ORBextractor *right, const std::vector<cv::KeyPoint> &mvKeys,
                                 const cv::Mat &mDescriptors, const std::vector<cv::KeyPoint> &mvKeysRight,
                                 const cv::Mat &mDescriptorsRight, float mbf, float mb, std::vector<float> &mvuRight,
                                 std::vector<float> &mvDepth) {
    const int N = (int)mvKeys.size(), Nr = (int)mvKeysRight.size();
    mvuRight.assign(N, -1.0f);                         // ref :473-474
    mvDepth.assign(N, -1.0f);
    if (N == 0) return false;
    assert(mDescriptors.isContinuous() && (Nr == 0 || mDescriptorsRight.isContinuous()));
    int matched = 0;
    const int rc = orb_stereo_match(left->Handle(), right->Handle(), 0, (const orb_keypoint *)mvKeys.data(), mDescriptors.data, N,
                                    (const orb_keypoint *)mvKeysRight.data(), Nr ? mDescriptorsRight.data : nullptr, Nr, mbf, mb,
                                    mvuRight.data(), mvDepth.data(), &matched);
    if (rc == ORB_ERR_EMPTY) return false;
    Check(rc, "orb_stereo_match");
    return true;
}

orb_extractor *ORBextractor::StereoHandle(int width, int height) {
    if (mpStereoHandle && width <= mStereoMaxWidth && height <= mStereoMaxHeight) return mpStereoHandle;
    if (mpStereoHandle) orb_extractor_destroy(mpStereoHandle);
    mpStereoHandle = nullptr;
    mStereoMaxWidth = width > mStereoMaxWidth ? width : mStereoMaxWidth;
    mStereoMaxHeight = height > mStereoMaxHeight ? height : mStereoMaxHeight;
    orb_params p;
    std::memset(&p, 0, sizeof(p));
    p.nfeatures = nfeatures; p.scale_factor = (float)scaleFactor; p.nlevels = nlevels;
    p.ini_th_fast = iniThFAST; p.min_th_fast = minThFAST;
    p.max_width = mStereoMaxWidth; p.max_height = mStereoMaxHeight; p.max_batch = 2; p.device = -1; p.blur_taps = 0;
    Check(orb_extractor_create(&p, &mpStereoHandle), "orb_extractor_create");
    return mpStereoHandle;
}

bool ORBGPU_ExtractStereo(ORBextractor *left, const cv::Mat &imLeft, const cv::Mat &imRight, std::vector<cv::KeyPoint> &mvKeys,
                          cv::Mat &mDescriptors, std::vector<cv::KeyPoint> &mvKeysRight, cv::Mat &mDescriptorsRight, float mbf,
                          float mb, std::vector<float> &mvuRight, std::vector<float> &mvDepth) {
    assert(imLeft.type() == CV_8UC1 && imRight.type() == CV_8UC1 && imLeft.cols == imRight.cols && imLeft.rows == imRight.rows);
    orb_extractor *h = left->StereoHandle(imLeft.cols, imLeft.rows);
    int capacity = left->DefaultCapacity(), nL = 0, nR = 0, matched = 0, rc = 0;
    std::vector<unsigned char> kl, kr, dl, dr;
    for (;;) {
        kl.resize((size_t)capacity * sizeof(orb_keypoint)); kr.resize(kl.size());
        dl.resize((size_t)capacity * 32); dr.resize(dl.size());
        mvuRight.assign(capacity, -1.0f); mvDepth.assign(capacity, -1.0f);
        rc = orb_stereo_frame(h, imLeft.data, imRight.data, imLeft.cols, imLeft.rows, (int)imLeft.step, (int)imRight.step,
                              (orb_keypoint *)kl.data(), dl.data(), &nL, (orb_keypoint *)kr.data(), dr.data(), &nR, capacity, mbf, mb,
                              mvuRight.data(), mvDepth.data(), &matched);
        if (rc == ORB_ERR_CAPACITY && (nL > capacity || nR > capacity)) { capacity = nL > nR ? nL : nR; continue; }
        if (rc != ORB_ERR_EMPTY) Check(rc, "orb_stereo_frame");
        break;
    }
    auto fill = [](int n, const std::vector<unsigned char> &k, const std::vector<unsigned char> &d, std::vector<cv::KeyPoint> &keys,
                   cv::Mat &desc) {
        keys.resize(n);
        if (n) std::memcpy((void *)keys.data(), k.data(), (size_t)n * sizeof(orb_keypoint));
        if (n == 0) { desc.release(); return; }
        desc.create(n, 32, CV_8U);
        for (int i = 0; i < n; i++) std::memcpy(desc.ptr(i), d.data() + (size_t)i * 32, 32);
    };
    fill(nL, kl, dl, mvKeys, mDescriptors);
    fill(nR, kr, dr, mvKeysRight, mDescriptorsRight);
    mvuRight.resize(nL); mvDepth.resize(nL);           // ref :473-474: one entry per left keypoint
    return rc != ORB_ERR_EMPTY;
}

}  // namespace ORB_SLAM2
