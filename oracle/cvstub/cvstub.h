// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// A minimal stand-in for the slice of OpenCV that ORB_SLAM2/src/ORBextractor.cc uses, so that the REFERENCE'S OWN
// SOURCE FILE can be compiled unmodified into oracle/_ref/liborbslam_ref.so (see oracle/Makefile.ref).  The reference
// needs OpenCV 2.4 C++ (not installed, not vendored); this header supplies cv::Mat / KeyPoint / InputArray plumbing,
// and the arithmetic primitives (resize, copyMakeBorder, GaussianBlur, FAST, fastAtan2, cvRound) are the models of
// oracle/orbref_cv.h, which are pinned against cv2 4.13.0 by tests/golden/.  What this buys: the reference's control
// flow -- cell loop, threshold retry, DistributeOctTree with its std::list bookkeeping, orientation, descriptor,
// level scaling -- runs from the reference's own code, and tests compare the restated oracle against it.
#pragma once
#include "../orbref_cv.h"

#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <iterator>
#include <list>
#include <memory>
#include <vector>

typedef unsigned char uchar;
#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5

static inline int cvRound(double v) { return orbref::cv_round(v); }
static inline int cvFloor(double v) { return orbref::cv_floor(v); }
static inline int cvCeil(double v) { return orbref::cv_ceil(v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16, INTER_LINEAR = 1 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
static inline Point2f &operator*=(Point2f &a, float b) { a.x = a.x * b; a.y = a.y * b; return a; }

struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {} };

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

template <typename T> struct Point3_ { T x, y, z; Point3_() : x(0), y(0), z(0) {} Point3_(T a, T b, T c) : x(a), y(b), z(c) {} };
typedef Point3_<float> Point3f;

struct DMatch {   // features2d.hpp (OpenCV 2.4): Marker.cc:341 uses the 4-argument form (queryIdx, trainIdx, imgIdx, distance)
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.402823466e+38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
    DMatch(int q, int t, int i, float d) : queryIdx(q), trainIdx(t), imgIdx(i), distance(d) {}
};

struct HammingLUT {   // features2d.hpp (OpenCV 2.4): number of differing bits of two byte strings (Marker.cc:353-356)
    typedef unsigned char ValueType;
    typedef int ResultType;
    ResultType operator()(const unsigned char *a, const unsigned char *b, int size) const {
        int r = 0;
        for (int i = 0; i < size; i++) r += __builtin_popcount((unsigned)(a[i] ^ b[i]));
        return r;
    }
};

struct MatZeros { int rows, cols, type; };

class Mat {
public:
    struct Step {
        size_t v;
        operator size_t() const { return v; }
    };
    int rows, cols;
    uchar *data;
    Step step;
    int type_;
    size_t esz;
    std::shared_ptr<std::vector<uchar>> buf;

    Mat() : rows(0), cols(0), data(nullptr), step{0}, type_(CV_8UC1), esz(1) {}
    Mat(Size sz, int t) { alloc(sz.height, sz.width, t); }
    Mat(int r, int c, int t) { alloc(r, c, t); }
    Mat(int r, int c, int t, void *ext, size_t stp = 0) : rows(r), cols(c), data((uchar *)ext), type_(t), esz(t == CV_32F ? 4 : 1) {   // header over caller-owned pixels
        step.v = stp ? stp : (size_t)c * esz;
    }
    bool isContinuous() const { return step.v == (size_t)cols * esz; }
    Mat(const MatZeros &z) : rows(0), cols(0), data(nullptr), step{0}, type_(CV_8UC1), esz(1) { *this = z; }   // cv::Mat m = cv::Mat::zeros(..), MapPoint.cc:322
    void alloc(int r, int c, int t = CV_8UC1) {
        rows = r; cols = c; type_ = t; esz = (t == CV_32F) ? 4 : 1;
        buf = std::make_shared<std::vector<uchar>>((size_t)r * c * esz);
        data = buf->data();
        step.v = (size_t)c * esz;
    }
    void create(int r, int c, int t) { if (r != rows || c != cols || t != type_ || !data) alloc(r, c, t); }
    void release() { rows = cols = 0; data = nullptr; buf.reset(); step.v = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t step1() const { return step.v / esz; }
    Mat operator()(const Rect &r) const { Mat m = *this; m.data = data + (size_t)r.y * step.v + r.x * esz; m.rows = r.height; m.cols = r.width; return m; }
    Mat rowRange(int a, int b) const { Mat m = *this; m.data = data + (size_t)a * step.v; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { Mat m = *this; m.data = data + a * esz; m.cols = b - a; return m; }
    Mat row(int a) const { return rowRange(a, a + 1); }
    Mat col(int a) const { return colRange(a, a + 1); }
    template <typename T> T &at(int y, int x) { return *(T *)(data + (size_t)y * step.v + x * sizeof(T)); }
    template <typename T> const T &at(int y, int x) const { return *(const T *)(data + (size_t)y * step.v + x * sizeof(T)); }
    template <typename T> T &at(int i) { return cols == 1 ? at<T>(i, 0) : at<T>(0, i); }
    template <typename T> const T &at(int i) const { return cols == 1 ? at<T>(i, 0) : at<T>(0, i); }
    uchar *ptr(int y = 0) { return data + (size_t)y * step.v; }
    const uchar *ptr(int y = 0) const { return data + (size_t)y * step.v; }
    template <typename T> T *ptr(int y = 0) { return (T *)(data + (size_t)y * step.v); }
    template <typename T> const T *ptr(int y = 0) const { return (const T *)(data + (size_t)y * step.v); }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(m.ptr(y), ptr(y), cols * esz);
        return m;
    }
    void copyTo(Mat &m) const { m = clone(); }
    Mat reshape(int) const { return *this; }   // only on the undistortion path, which the zero-distortion configs never take
    void convertTo(Mat &dst, int t) const {    // 8U -> 32F (Frame.cc:567,584); dst may alias *this (a ROI header)
        assert(type_ == CV_8UC1 && t == CV_32F);
        Mat m(rows, cols, CV_32F);
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) m.at<float>(y, x) = (float)at<uchar>(y, x);
        dst = m;
    }
    static Mat ones(int r, int c, int t) {
        assert(t == CV_32F);
        Mat m(r, c, CV_32F);
        for (int y = 0; y < r; y++) for (int x = 0; x < c; x++) m.at<float>(y, x) = 1.0f;
        return m;
    }
    static MatZeros zeros(int r, int c, int t) { return MatZeros{r, c, t}; }
    // Mat = MatExpr(zeros): keeps the existing buffer when size and type match (cv::Mat::create semantics), which is
    // how ORBextractor.cc:978 zero-fills a row range of the output descriptor matrix in place.
    Mat &operator=(const MatZeros &z) {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) std::memset(ptr(y), 0, cols * esz);
        return *this;
    }
    // ---- CV_32F algebra used by ORBmatcher.cc (products accumulate in double like cv::gemm, results are float)
    Mat t() const {
        Mat m(cols, rows, CV_32F);
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    double dot(const Mat &o) const {
        double s = 0;
        for (int y = 0; y < rows; y++) for (int x = 0; x < cols; x++) s += (double)at<float>(y, x) * o.at<float>(y, x);
        return s;
    }
};
static inline Mat operator*(const Mat &a, const Mat &b) {
    assert(a.cols == b.rows && a.type() == CV_32F);
    Mat m(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            double s = 0;
            for (int k = 0; k < a.cols; k++) s += (double)a.at<float>(y, k) * b.at<float>(k, x);
            m.at<float>(y, x) = (float)s;
        }
    return m;
}
#define CVSTUB_EW(OP)                                                                                      \
    static inline Mat operator OP(const Mat &a, const Mat &b) {                                            \
        assert(a.rows == b.rows && a.cols == b.cols);                                                      \
        Mat m(a.rows, a.cols, CV_32F);                                                                     \
        for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) OP b.at<float>(y, x); \
        return m;                                                                                          \
    }
CVSTUB_EW(+)
CVSTUB_EW(-)
#undef CVSTUB_EW
static inline Mat operator*(const Mat &a, double s) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)(a.at<float>(y, x) * s);
    return m;
}
static inline Mat operator*(double s, const Mat &a) { return a * s; }
static inline Mat operator/(const Mat &a, double s) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = (float)(a.at<float>(y, x) / s);
    return m;
}
static inline Mat operator-(const Mat &a) { return a * -1.0; }
static inline double norm(const Mat &a) { return std::sqrt(a.dot(a)); }
enum { NORM_L1 = 2 };
static inline double norm(const Mat &a, const Mat &b, int normType) {   // cv::norm(IL, IR, NORM_L1), Frame.cc:587
    assert(normType == NORM_L1 && a.rows == b.rows && a.cols == b.cols);
    double s = 0;
    for (int y = 0; y < a.rows; y++) for (int x = 0; x < a.cols; x++) s += std::fabs((double)(a.at<float>(y, x) - b.at<float>(y, x)));
    return s;
}
template <typename T> class Mat_;
template <typename T> struct MatCommaInit {   // (cv::Mat_<float>(3,1) << x, y, z), Frame.cc:678
    Mat m;
    int i;
    MatCommaInit &operator,(T v) { m.at<T>(i / m.cols, i % m.cols) = v; i++; return *this; }
    operator Mat() const { return m; }
};
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
    MatCommaInit<T> operator<<(T v) { MatCommaInit<T> ci{*this, 0}; ci, v; return ci; }
};

class _InputArray {
public:
    const Mat *m;
    Mat tmp;
    _InputArray(const Mat &mat) : m(&mat) {}
    bool empty() const { return m->empty(); }
    Mat getMat() const { return *m; }
};
class _OutputArray {
public:
    Mat *m;
    _OutputArray(Mat &mat) : m(&mat) {}
    void create(int r, int c, int t) const { m->create(r, c, t); }
    void create(Size sz, int t) const { m->create(sz.height, sz.width, t); }
    void release() const { m->release(); }
    Mat getMat() const { return *m; }
};
typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;

static inline float fastAtan2(float y, float x) { return orbref::fast_atan2(y, x); }

static inline void FAST(InputArray image, std::vector<KeyPoint> &keypoints, int threshold, bool nonmaxSuppression = true) {
    (void)nonmaxSuppression;
    Mat m = image.getMat();
    std::vector<orbref::FastPoint> pts;
    orbref::fast9_16_nms(m.data, m.cols, m.rows, (int)m.step.v, threshold, pts);
    keypoints.clear();
    for (const orbref::FastPoint &p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1.f, (float)p.response));
}

static inline void resize(InputArray src, OutputArray dst, Size dsize, double = 0, double = 0, int = INTER_LINEAR) {
    Mat s = src.getMat();
    dst.create(dsize, s.type());
    Mat d = dst.getMat();
    orbref::resize_linear_u8(s.data, s.cols, s.rows, (int)s.step.v, d.data, d.cols, d.rows, (int)d.step.v);
}

static inline void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int) {
    Mat s = src.getMat();
    dst.create(s.rows + top + bottom, s.cols + left + right, s.type());
    Mat d = dst.getMat();
    // interior first (a no-op when src already is the interior of dst), then the reflected frame
    for (int y = 0; y < s.rows; y++)
        if (d.ptr(y + top) + left != s.ptr(y)) std::memmove(d.ptr(y + top) + left, s.ptr(y), s.cols);
    assert(top == bottom && left == right && top == left);
    orbref::make_border_reflect101(d.ptr(top) + left, s.cols, s.rows, (int)d.step.v, top);
}

static inline void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int = BORDER_REFLECT_101) {
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2);
    Mat s = src.getMat().clone();
    dst.create(s.rows, s.cols, s.type());
    Mat d = dst.getMat();
    orbref::gaussian_blur7_u8(s.data, s.cols, s.rows, (int)s.step.v, d.data, (int)d.step.v, 0);
}

static inline void undistortPoints(InputArray, OutputArray, InputArray, InputArray, InputArray, InputArray) {
    std::abort();   // Frame::UndistortKeyPoints / ComputeImageBounds with non-zero distortion: not on the tested path
}

struct KeyPointsFilter {   // only referenced by the dead ComputeKeyPointsOld (ORBextractor.cc:821-974)
    static void retainBest(std::vector<KeyPoint> &, int) { std::abort(); }
};

}  // namespace cv
