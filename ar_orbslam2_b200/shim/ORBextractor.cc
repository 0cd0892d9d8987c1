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
      mpHandle(nullptr), mMaxWidth(0), mMaxHeight(0), mbMirrorPyramid(true) {
    mvImagePyramid.resize(nlevels);
    EnsureHandle(1920, 1080);
}

ORBextractor::~ORBextractor() { orb_extractor_destroy(mpHandle); }

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

    int capacity = nfeatures + 4 * nlevels + 256, n = 0;
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

    if (mbMirrorPyramid) {
        for (int l = 0; l < nlevels; l++) {
            int w = 0, h = 0, pitch = 0;
            Check(orb_pyramid_level(mpHandle, 0, l, nullptr, &w, &h, &pitch), "orb_pyramid_level");
            // the reference's level is the interior ROI of a (w+38)x(h+38) plane (ref :1052-1056): keep that layout
            // so code that walks into the 19-px border (Frame.cc:583) finds the same bytes
            cv::Mat plane(h + 2 * ORB_EDGE_THRESHOLD, w + 2 * ORB_EDGE_THRESHOLD, CV_8UC1);
            Check(orb_pyramid_download(mpHandle, 0, l, 1, plane.data, (int)plane.step), "orb_pyramid_download");
            mPlanes.resize(nlevels);
            mPlanes[l] = plane;
            mvImagePyramid[l] = cv::Mat(h, w, CV_8UC1, plane.data + (size_t)ORB_EDGE_THRESHOLD * plane.step + ORB_EDGE_THRESHOLD,
                                        plane.step);
        }
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

}  // namespace ORB_SLAM2
