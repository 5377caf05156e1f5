// cv_shim.h — the slice of the OpenCV C++ API that the reference's CPU extractor uses, so that its OWN source
// (app/src/main/cpp/src/ORBextractor_old.cc) compiles in this container without OpenCV headers.
//
// TEST INFRASTRUCTURE (oracle/): nothing under orbslam3lib_b200/ includes or links this.
//
// What is real and what is forwarded:
//   * the ORB-SLAM glue (IC_Angle, computeOrbDescriptor, DivideNode, compareNodes, DistributeOctTree,
//     ComputeKeyPointsOctTree, ComputePyramid, operator()) is the reference's own text, compiled from
//     /root/reference by make_ref_ext.py — none of it is restated here;
//   * cv::Mat / Point_ / KeyPoint / Size / Rect / InputArray below are containers with OpenCV's semantics for the
//     calls the glue makes (ROI views share the parent's buffer, at<>/ptr/step address bytes, implicit float -> int
//     argument conversions happen exactly where OpenCV's signatures make them happen);
//   * the five OpenCV ALGORITHMS the glue calls — cv::FAST, cv::GaussianBlur, cv::resize, cv::copyMakeBorder,
//     cv::fastAtan2 — forward to the oracle's primitives (oracle/cpp/orb_oracle.h), which tests/test_oracle_vs_cv2.py
//     and tests/golden/primitives.npz pin bit-for-bit against the genuine OpenCV build (cv2) of this container.
//   * cvRound/cvFloor/cvCeil follow opencv2/core/fast_math.hpp: round-half-to-even via the current rounding mode
//     (SSE2 cvtsd2si / ARM vcvtr == lrint), floor, ceil.
#ifndef ORBFE_ORACLE_CV_SHIM_H
#define ORBFE_ORACLE_CV_SHIM_H

#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <list>
#include <memory>
#include <stdexcept>
#include <vector>

#include "../cpp/orb_oracle.h"

typedef unsigned char uchar;

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum { BORDER_REFLECT_101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}    // cv::Point2i(float, int): the float converts to int here, as in OpenCV
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

// opencv2/core/types.hpp: a.x = saturate_cast<_Tp>(a.x * b); for _Tp = float that is the float product
static inline Point2f& operator*=(Point2f& a, float b) {
    a.x = a.x * b;
    a.y = a.y * b;
    return a;
}

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

class KeyPoint {
 public:
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
};

struct MatStep {
    size_t v;
    MatStep() : v(0) {}
    operator size_t() const { return v; }
};

// CV_8UC1 only. Header + shared buffer, like cv::Mat: copies and ROIs alias the same pixels.
class Mat {
 public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    MatStep step;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size sz, int type) { create(sz.height, sz.width, type); }
    // external data, not owned (cv::Mat(rows, cols, type, void* data, size_t step))
    Mat(int r, int c, int type, void* ext, size_t stp) : rows(r), cols(c), data((uchar*)ext) {
        assert(type == CV_8UC1);
        (void)type;
        step.v = stp;
    }

    void create(int r, int c, int type) {
        assert(type == CV_8UC1);
        (void)type;
        if (data && r == rows && c == cols) return;    // cv::Mat::create keeps a buffer of the right shape
        rows = r;
        cols = c;
        step.v = (size_t)c;
        buf.reset(new uchar[(size_t)r * c + 1], std::default_delete<uchar[]>());
        data = buf.get();
    }
    void release() {
        buf.reset();
        data = nullptr;
        rows = cols = 0;
        step.v = 0;
    }
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        std::memset(m.data, 0, (size_t)r * c);
        return m;
    }

    int type() const { return CV_8UC1; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step.v; }
    bool isContinuous() const { return step.v == (size_t)cols || rows == 1; }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const {
        return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T));
    }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }

    Mat roi(int x, int y, int w, int h) const {
        if (x < 0 || y < 0 || w < 0 || h < 0 || x + w > cols || y + h > rows) throw std::out_of_range("cv_shim: ROI");
        Mat m;
        m.rows = h;
        m.cols = w;
        m.step = step;
        m.buf = buf;
        m.data = data + (size_t)y * step.v + x;
        return m;
    }
    // the glue calls rowRange(float, float): the arguments convert to int at the call, as with cv::Mat
    Mat rowRange(int a, int b) const { return roi(0, a, cols, b - a); }
    Mat colRange(int a, int b) const { return roi(a, 0, b - a, rows); }
    Mat row(int r) const { return roi(0, r, cols, 1); }
    Mat operator()(const Rect& r) const { return roi(r.x, r.y, r.width, r.height); }

    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, CV_8UC1);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), (size_t)cols);
        return m;
    }
    // copyTo into an existing Mat of the same shape writes through (desc.row(i).copyTo(descriptors.row(k)))
    void copyTo(const Mat& dst) const {
        if (dst.rows != rows || dst.cols != cols) throw std::runtime_error("cv_shim: copyTo needs a preallocated dst");
        for (int r = 0; r < rows; ++r) std::memmove(const_cast<uchar*>(dst.ptr(r)), ptr(r), (size_t)cols);
    }

 private:
    std::shared_ptr<uchar> buf;
};

class _InputArray {
 public:
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& _m) : m(const_cast<Mat*>(&_m)) {}
    Mat getMat() const { return m ? *m : Mat(); }
    bool empty() const { return m == nullptr || m->empty(); }

 protected:
    Mat* m;
};
typedef const _InputArray& InputArray;

class _OutputArray : public _InputArray {
 public:
    _OutputArray(Mat& _m) : _InputArray(_m) {}
    void release() const { m->release(); }
    void create(int r, int c, int type) const { m->create(r, c, type); }
};
typedef const _OutputArray& OutputArray;

static inline float fastAtan2(float y, float x) { return orc_fast_atan2(y, x); }

// cv::FAST(image, keypoints, threshold, nonmaxSuppression) — FastFeatureDetector::TYPE_9_16.
// opencv/modules/features2d/src/fast.cpp: keypoints.push_back(KeyPoint((float)j, (float)(i-1), 7.f, -1, (float)score))
static inline void FAST(const Mat& image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
    if (!nonmaxSuppression) throw std::runtime_error("cv_shim: FAST without NMS is not on the path");
    keypoints.clear();
    if (image.empty()) return;
    int cap = image.rows * image.cols / 4 + 16;
    std::vector<int> xyr((size_t)cap * 3);
    int n = orc_fast9_nms(image.data, image.cols, image.rows, (int)image.step1(), threshold, xyr.data(), cap);
    if (n > cap) throw std::runtime_error("cv_shim: FAST capacity");
    keypoints.reserve(n);
    for (int i = 0; i < n; ++i)
        keypoints.push_back(KeyPoint((float)xyr[3 * i], (float)xyr[3 * i + 1], 7.f, -1, (float)xyr[3 * i + 2]));
}

static inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
    if (ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 || borderType != BORDER_REFLECT_101)
        throw std::runtime_error("cv_shim: only GaussianBlur 7x7 sigma 2 REFLECT_101 is on the path");
    Mat in = (src.data == dst.data) ? src.clone() : src;
    dst.create(src.rows, src.cols, CV_8UC1);
    orc_gaussian7x7_u8(in.data, in.cols, in.rows, (int)in.step1(), dst.data, (int)dst.step1());
}

static inline void resize(const Mat& src, Mat& dst, Size dsize, double fx, double fy, int interpolation) {
    if (fx != 0 || fy != 0 || interpolation != INTER_LINEAR) throw std::runtime_error("cv_shim: resize arguments");
    dst.create(dsize.height, dsize.width, CV_8UC1);    // no-op for the preallocated pyramid ROI, as in OpenCV
    orc_resize_linear_u8(src.data, src.cols, src.rows, (int)src.step1(), dst.data, dst.cols, dst.rows, (int)dst.step1());
}

// The glue calls it with equal borders, REFLECT_101 (+ ISOLATED for the in-place ROI case of levels >= 1, where src is
// the interior of dst).  The bordered image is built in a scratch buffer, then copied out.
static inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType) {
    if ((borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101 || top != bottom || top != left || top != right)
        throw std::runtime_error("cv_shim: copyMakeBorder arguments");
    const int W = src.cols + 2 * left, H = src.rows + 2 * top;
    std::vector<uchar> tmp((size_t)W * H);
    orc_border_reflect101_u8(src.data, src.cols, src.rows, (int)src.step1(), tmp.data(), W, top);
    dst.create(H, W, CV_8UC1);
    for (int r = 0; r < H; ++r) std::memcpy(dst.ptr(r), tmp.data() + (size_t)r * W, (size_t)W);
}

}    // namespace cv

#endif
