// ORACLE -- TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
//
// A minimal cv:: namespace -- just the types and functions the reference's ORB front-end touches -- so that
// /root/reference/src/ORBextractor.cc (unmodified) and the line ranges oracle/ref_extract.py lifts out of
// src/Frame.cc and src/ORBmatcher.cc compile here, where OpenCV's C++ headers and libraries do not exist.
// The image primitives forward to oracle/cvprim.cpp, each of which is pinned bit-for-bit against the genuine
// OpenCV of this image (cv2 4.13.0) in tests/test_oracle_cv2.py.  Container semantics follow OpenCV where the
// reference depends on them:
//   * Mat is a reference-counted view (ROI views share the buffer; create() on a matching Mat is a no-op, so
//     resize()/copyMakeBorder() write INTO mvImagePyramid's ROI of the padded buffer, src/ORBextractor.cc:1184-1197);
//   * `m = Mat::zeros(r, c, t)` assigns in place when m already has that shape (src/ORBextractor.cc:1106 clears
//     the caller's descriptor rows, it does not detach them);
//   * cvRound is round-half-to-even; KeyPoint is the 28-byte struct; Point is two ints (the pattern cast, :448).
// What is written here is the builder's; what is compiled against it is the reference's own text.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../cvprim.hpp"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_PI 3.1415926535897932384626433832795

typedef unsigned char uchar;

namespace cv {

inline int cvRound(double v) { return orc::cv_round(v); }
inline int cvRound(float v) { return orc::cv_round(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U>
    Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
};
template <typename T>
inline Point_<T>& operator*=(Point_<T>& a, float b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

template <typename T>
struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T>
struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Scalar { double v[4] = {0, 0, 0, 0}; };

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");
static_assert(sizeof(Point) == 8, "cv::Point layout");

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };

class Mat;
struct MatExpr {   // only what src/ORBextractor.cc:1106 and src/Frame.cc:997,1019 build
    enum Kind { FILL, SUB_FILL } kind;
    int rows, cols, type;
    double value;        // fill value
    const Mat* a;        // SUB_FILL: a - value
};

class Mat {
public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    size_t step = 0;

    Mat() {}
    Mat(int r, int c, int t) { create(r, c, t); }
    Mat(Size s, int t) { create(s.height, s.width, t); }
    Mat(int r, int c, int t, void* ext, size_t st) : rows(r), cols(c), data((uchar*)ext), step(st), type_(t) {}
    Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), type_(o.type_), cn_(o.cn_), blk_(o.blk_) { retain(); }
    Mat(const MatExpr& e) { *this = e; }
    ~Mat() { drop(); }
    Mat& operator=(const Mat& o) {
        if (this != &o) {
            o.retain();
            drop();
            rows = o.rows; cols = o.cols; data = o.data; step = o.step; type_ = o.type_; cn_ = o.cn_; blk_ = o.blk_;
        }
        return *this;
    }
    Mat& operator=(const MatExpr& e);

    void create(int r, int c, int t) {
        if (data && rows == r && cols == c && type_ == t) return;   // cv::Mat::create: matching shape -> untouched
        drop();
        rows = r; cols = c; type_ = t; cn_ = 1;
        step = (size_t)c * elemSize();
        const size_t bytes = step * (size_t)r;
        blk_ = (Block*)std::malloc(sizeof(Block) + bytes + 64);
        blk_->refs = 1;
        data = (uchar*)(blk_ + 1);
    }
    void release() { drop(); rows = cols = 0; data = nullptr; step = 0; }
    int type() const { return type_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t elemSize() const { return (type_ == CV_32F ? 4 : 1) * (size_t)cn_; }
    int channels() const { return cn_; }
    // cv::Mat::reshape(cn) with the row count kept: an N x 2 one-channel matrix <-> N x 1 two-channel points (src/Frame.cc:687,689)
    // on a continuous matrix; the view shares the buffer, so an in-place call through the reshaped header lands in the original
    Mat reshape(int cn) const {
        assert(step == (size_t)cols * elemSize() && (cols * cn_) % cn == 0);
        Mat m(*this);
        m.cols = cols * cn_ / cn; m.cn_ = cn;
        return m;
    }
    size_t step1() const { return step / elemSize(); }

    Mat operator()(const Rect& r) const { return view(r.y, r.y + r.height, r.x, r.x + r.width); }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat row(int y) const { return view(y, y + 1, 0, cols); }
    Mat col(int x) const { return view(0, rows, x, x + 1); }
    // CV_32F only, as the pose algebra of src/ORBmatcher.cc needs it (-Rcw.t() * tcw, sRcw.row(0).dot(sRcw.row(0)), ...)
    Mat t() const {
        assert(type_ == CV_32F);
        Mat m(cols, rows, CV_32F);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    double dot(const Mat& o) const {           // cv::Mat::dot on CV_32F accumulates in double
        assert(type_ == CV_32F && o.type_ == CV_32F && rows * cols == o.rows * o.cols);
        double s = 0;
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * (double)o.at<float>(rows == o.rows ? y : x, rows == o.rows ? x : y);
        return s;
    }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    template <typename T> T& at(int y, int x) { return ((T*)(data + (size_t)y * step))[x]; }
    template <typename T> const T& at(int y, int x) const { return ((const T*)(data + (ptrdiff_t)y * (ptrdiff_t)step))[x]; }
    // single-index form on a row or column vector (include/Frame.h:185,  src/Frame.cc:1226: Pc.at<float>(2))
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i / cols, i % cols); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i / cols, i % cols); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }

    // 8U -> 32F only (src/Frame.cc:996,1018); a type change always lands in a fresh buffer, like OpenCV
    void convertTo(Mat& dst, int t) const {
        assert(type_ == CV_8U && t == CV_32F);
        Mat out(rows, cols, t);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) out.at<float>(y, x) = (float)at<uchar>(y, x);
        dst = out;
    }
    static MatExpr zeros(int r, int c, int t) { return MatExpr{MatExpr::FILL, r, c, t, 0.0, nullptr}; }
    static MatExpr ones(int r, int c, int t) { return MatExpr{MatExpr::FILL, r, c, t, 1.0, nullptr}; }

private:
    struct Block { long refs; long pad; };
    int type_ = 0, cn_ = 1;
    Block* blk_ = nullptr;
    void retain() const { if (blk_) __atomic_add_fetch(&blk_->refs, 1, __ATOMIC_RELAXED); }
    void drop() {
        if (blk_ && __atomic_sub_fetch(&blk_->refs, 1, __ATOMIC_ACQ_REL) == 0) std::free(blk_);
        blk_ = nullptr;
    }
    Mat view(int y0, int y1, int x0, int x1) const {
        Mat m(*this);
        m.rows = y1 - y0; m.cols = x1 - x0;
        m.data = data + (ptrdiff_t)y0 * (ptrdiff_t)step + (ptrdiff_t)x0 * (ptrdiff_t)elemSize();
        return m;
    }
};

inline MatExpr operator*(double s, const MatExpr& e) { MatExpr r = e; assert(e.kind == MatExpr::FILL); r.value *= s; return r; }
inline MatExpr operator-(const Mat& a, const MatExpr& e) {
    assert(e.kind == MatExpr::FILL && e.rows == a.rows && e.cols == a.cols && e.type == a.type());
    return MatExpr{MatExpr::SUB_FILL, a.rows, a.cols, a.type(), e.value, &a};
}
inline Mat& Mat::operator=(const MatExpr& e) {
    if (e.kind == MatExpr::FILL) {
        create(e.rows, e.cols, e.type);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) {
                if (type_ == CV_32F) at<float>(y, x) = (float)e.value;
                else at<uchar>(y, x) = (uchar)e.value;
            }
    } else {   // element-wise float subtraction of a constant matrix (cv::subtract on CV_32F)
        assert(e.type == CV_32F);
        const Mat src = *e.a;                 // keeps the operand alive when it is *this
        const float s = (float)e.value;
        Mat out(e.rows, e.cols, CV_32F);
        for (int y = 0; y < e.rows; ++y)
            for (int x = 0; x < e.cols; ++x) out.at<float>(y, x) = src.at<float>(y, x) - s;
        *this = out;
    }
    return *this;
}

// CV_32F matrix product and sum (include/Frame.h:182: Pc = mRcw*Pw+mtcw).  cv::gemm on CV_32F accumulates each dot
// product in double and stores float (GEMMSingleMul<float,double>); the sum is element-wise float.
inline Mat operator*(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
    Mat out(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < b.cols; ++x) {
            double s = 0;
            for (int k = 0; k < a.cols; ++k) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
            out.at<float>(y, x) = (float)s;
        }
    return out;
}
inline Mat operator+(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat out(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) out.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    return out;
}

inline Mat operator-(const Mat& a) {
    assert(a.type() == CV_32F);
    Mat out(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) out.at<float>(y, x) = -a.at<float>(y, x);
    return out;
}
inline Mat operator-(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat out(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) out.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    return out;
}
// scalar * matrix and matrix / scalar on CV_32F: the element times (divided by) the double scalar, rounded to float
inline Mat operator*(double s, const Mat& a) {
    assert(a.type() == CV_32F);
    Mat out(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) out.at<float>(y, x) = (float)((double)a.at<float>(y, x) * s);
    return out;
}
inline Mat operator/(const Mat& a, double s) {
    assert(a.type() == CV_32F);
    Mat out(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) out.at<float>(y, x) = (float)((double)a.at<float>(y, x) / s);
    return out;
}
// cv::norm(a) = NORM_L2 on CV_32F: square root of the sum of squares accumulated in double
inline double norm(const Mat& a) {
    assert(a.type() == CV_32F);
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += (double)a.at<float>(y, x) * (double)a.at<float>(y, x);
    return std::sqrt(s);
}

// cv::norm(a, b, NORM_L1) on CV_32F: sum of |a - b| accumulated in double (src/Frame.cc:1021)
inline double norm(const Mat& a, const Mat& b, int normType) {
    assert(normType == NORM_L1 && a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    (void)normType;
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) s += std::fabs((double)(a.at<float>(y, x) - b.at<float>(y, x)));
    return s;
}

class _InputArray {
public:
    _InputArray() : m_(nullptr) {}
    _InputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    bool empty() const { return m_ == nullptr || m_->empty(); }
    Mat getMat() const { return m_ ? *m_ : Mat(); }
protected:
    Mat* m_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& m) : _InputArray(m) {}
    void create(int r, int c, int t) const { m_->create(r, c, t); }
    void release() const { if (m_) m_->release(); }
    void assign(const Mat& m) const { *m_ = m; }
    Mat* mat_ptr() const { return m_; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

// test hook: when set, every cv::FAST call appends its survivors in absolute coordinates of the buffer the
// cell view was cut from (identified by `origin`), so a test can read the cell loop's vToDistributeKeys
struct FastLogEntry { const uchar* origin; size_t step; int x, y, score; };
struct FastCall { const uchar* data; size_t step; size_t first; int threshold; };
struct FastLog { std::vector<FastLogEntry> pts; std::vector<FastCall> calls; };
extern thread_local FastLog* g_fast_log;

void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType,
                    const Scalar& value = Scalar());
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);
inline float fastAtan2(float y, float x) { return orc::fast_atan2(y, x); }
inline _InputArray noArray() { return _InputArray(); }
// cv::undistortPoints on N x 1 two-channel CV_32F points, CV_32F camera matrices (src/Frame.cc:688,710,745,795,858): forwards to the
// restated primitive of oracle/undistort.cpp, which tests/test_oracle_cv2.py holds bit-identical to the genuine cv2.undistortPoints.
// R empty -> identity; P 3x3 or 3x4 (left 3x3 used), empty -> identity; in place allowed.
void undistortPoints(InputArray src, OutputArray dst, InputArray cameraMatrix, InputArray distCoeffs,
                     InputArray R = _InputArray(), InputArray P = _InputArray());

}  // namespace cv
