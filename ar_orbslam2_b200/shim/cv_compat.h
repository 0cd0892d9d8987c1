// cv_compat.h -- the handful of OpenCV types the ORB_SLAM2 front-end interfaces mention, for builds where the
// real OpenCV headers are absent (this repo's CI box).  Define ORBGPU_USE_OPENCV to use the real ones instead:
// the shims (ORBextractor.h / ORBmatcher_gpu.h) only need cv::Mat, cv::KeyPoint, cv::Point2f and
// cv::InputArray / cv::OutputArray, and use them identically in both modes.
//
// Nothing here computes anything: the arithmetic lives in liborbgpu (include/orbgpu.h).
#pragma once

#ifdef ORBGPU_USE_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#else

#include <cstddef>
#include <cstring>
#include <memory>
#include <vector>

typedef unsigned char uchar;
#ifndef CV_8U
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#endif

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct KeyPoint {   // 28 bytes, same field order as OpenCV's (and as orb_keypoint)
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

struct DMatch {   // features2d.hpp
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.402823466e+38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
    DMatch(int q, int t, int i, float d) : queryIdx(q), trainIdx(t), imgIdx(i), distance(d) {}
    bool operator<(const DMatch &m) const { return distance < m.distance; }
};

// Reference-counted byte matrix: enough of cv::Mat for images (8UC1) and descriptor tables.
class Mat {
public:
    int rows, cols;
    uchar *data;
    size_t step;
    Mat() : rows(0), cols(0), data(nullptr), step(0), elem_(1), type_(CV_8UC1) {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void *ext, size_t ext_step)   // header over caller memory
        : rows(r), cols(c), data((uchar *)ext), step(ext_step), elem_(type == CV_32F ? 4 : 1), type_(type) {}
    void create(int r, int c, int type) {
        const size_t e = type == CV_32F ? 4 : 1;
        if (data && r == rows && c == cols && type == type_ && buf_) return;
        rows = r; cols = c; elem_ = e; type_ = type; step = (size_t)c * e;
        buf_ = std::make_shared<std::vector<uchar>>((size_t)r * step);
        data = buf_->data();
    }
    void release() { rows = cols = 0; data = nullptr; step = 0; buf_.reset(); }
    bool empty() const { return !data || rows == 0 || cols == 0; }
    int type() const { return type_; }
    bool isContinuous() const { return step == (size_t)cols * elem_; }
    uchar *ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar *ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T *ptr(int y = 0) { return (T *)(data + (size_t)y * step); }
    template <typename T> const T *ptr(int y = 0) const { return (const T *)(data + (size_t)y * step); }
    template <typename T> T &at(int y, int x) { return ((T *)(data + (size_t)y * step))[x]; }
    template <typename T> const T &at(int y, int x) const { return ((const T *)(data + (size_t)y * step))[x]; }
    Mat row(int y) const { Mat m = *this; m.data = data + (size_t)y * step; m.rows = 1; return m; }
    Mat rowRange(int a, int b) const { Mat m = *this; m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(m.ptr(y), ptr(y), (size_t)cols * elem_);
        return m;
    }

private:
    size_t elem_;
    int type_;
    std::shared_ptr<std::vector<uchar>> buf_;
};

class _InputArray {
public:
    _InputArray(const Mat &m) : m_(&m) {}
    bool empty() const { return m_->empty(); }
    Mat getMat() const { return *m_; }

private:
    const Mat *m_;
};
class _OutputArray {
public:
    _OutputArray(Mat &m) : m_(&m) {}
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
    Mat getMat() const { return *m_; }

private:
    Mat *m_;
};
typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;

}  // namespace cv
#endif  // ORBGPU_USE_OPENCV
