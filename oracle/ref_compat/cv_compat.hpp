// cv_compat.hpp — the slice of the OpenCV C++ API that the reference's
// src/geometry/fextractor.cpp uses, so that the UNMODIFIED reference translation
// unit compiles in this image (which has no OpenCV C++ headers or libraries).
//
// *** TEST INFRASTRUCTURE ONLY (part of the oracle; see oracle/orb_oracle.cpp). ***
//
// What is here and what it stands for (OpenCV 4.x public API, opencv2/core/{types,mat}.hpp,
// opencv2/imgproc.hpp, opencv2/features2d.hpp):
//   types      Point_/Point/Point2i/Point2f, Size, Rect, KeyPoint (28-byte layout), Mat (CV_8UC1 only:
//              reference-counted buffer, ROI views via operator()(Rect)/rowRange/colRange/row, clone,
//              copyTo, at<>, ptr, step/step1, create/release, zeros), InputArray / OutputArray proxies,
//              KeyPointsFilter::retainBest
//   scalars    cvRound (round-half-to-even, what the SSE2 build of OpenCV does), cvFloor, cvCeil, CV_PI
//   primitives FAST, resize, copyMakeBorder, GaussianBlur, fastAtan2 — DECLARED here; cv_compat.cpp
//              forwards them to the restatements in oracle/orb_oracle.cpp, which tests/test_oracle_golden.py
//              pins bit-exactly against the real OpenCV (cv2 4.13 wheel) outputs.
// The control flow that runs on top of these — ComputePyramid, the FAST cell loop with its ini/min
// threshold fallback, DistributeOctTree/DivideNode, IC_Angle, computeOrbDescriptor, compute()'s output
// ordering — is the reference's own object code.
#pragma once

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_PI 3.1415926535897932384626433832795
#define CV_Assert(expr) do { if (!(expr)) throw std::runtime_error("CV_Assert failed: " #expr); } while (0)

// cvRound: OpenCV's x86-64 build uses cvtss2si / cvtsd2si, i.e. the current rounding mode
// (nearest-even).  lrintf/lrint do the same.
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum BorderTypes {
    BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3,
    BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16
};
enum InterpolationFlags { INTER_NEAREST = 0, INTER_LINEAR = 1 };

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U>
    Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
};
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, float b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, double b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
template <typename T> static inline Point_<T>& operator*=(Point_<T>& a, int b) { a.x = (T)(a.x * b); a.y = (T)(a.y * b); return a; }
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
template <typename T>
struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
};
struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

// opencv2/core/types.hpp: class KeyPoint — 28 bytes, defaults as in KeyPoint::KeyPoint()
struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct MatStep {
    size_t v;
    MatStep() : v(0) {}
    MatStep(size_t s) : v(s) {}
    operator size_t() const { return v; }
};

class _InputArray;
class _OutputArray;
class MatExpr;
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

// CV_8UC1 matrices (everything fextractor.cpp touches) and single-channel CV_32F ones (the 11 x 11 correlation patches of
// Frame::ComputeStereoMatches, oracle/ref_compat/ref_slices.cpp).  A Mat is a view (data, rows, cols, step in bytes) into
// a shared whole buffer.
class Mat {
public:
    struct Whole { uchar* base; int rows, cols; size_t step; bool owned; ~Whole() { if (owned) std::free(base); } };

    int rows, cols;
    uchar* data;
    MatStep step;
    std::shared_ptr<Whole> whole;
    int tp = CV_8UC1;
    size_t esz() const { return tp == CV_32F ? 4 : 1; }
    size_t elemSize() const { return esz(); }

    Mat() : rows(0), cols(0), data(nullptr) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr) { create(r, c, type); }
    Mat(Size s, int type) : rows(0), cols(0), data(nullptr) { create(s.height, s.width, type); }
    // header over caller memory (no copy), like cv::Mat(rows, cols, type, void* data, size_t step)
    Mat(int r, int c, int type, void* ptr, size_t stp = 0) : rows(r), cols(c), data((uchar*)ptr), tp(type) {
        assert(type == CV_8UC1 || type == CV_32F);
        step = stp ? stp : (size_t)c * esz();
        whole = std::shared_ptr<Whole>(new Whole{(uchar*)ptr, r, c, (size_t)step, false});
    }

    void create(int r, int c, int type) {
        assert(type == CV_8UC1 || type == CV_32F);
        if (data && rows == r && cols == c && tp == type) return;          // Mat::create keeps a matching buffer
        tp = type;
        rows = r; cols = c; step = (size_t)c * esz();
        size_t bytes = (size_t)r * c * esz();
        uchar* p = (uchar*)std::calloc(bytes ? bytes : 1, 1);
        whole = std::shared_ptr<Whole>(new Whole{p, r, c, (size_t)step, true});
        data = p;
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    void release() { rows = cols = 0; data = nullptr; step = 0; whole.reset(); }

    static Mat zeros(int r, int c, int type) { Mat m; m.create(r, c, type); std::memset(m.data, 0, (size_t)r * c * m.esz()); return m; }
    static Mat ones(int r, int c, int type) {
        Mat m; m.create(r, c, type);
        for (int y = 0; y < r; ++y)
            for (int x = 0; x < c; ++x) { if (type == CV_32F) m.at<float>(y, x) = 1.f; else m.at<uchar>(y, x) = 1; }
        return m;
    }

    int type() const { return tp; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    size_t step1(int = 0) const { return step; }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows == 1 || (size_t)step == (size_t)cols * esz(); }
    bool isSubmatrix() const { return whole && (rows != whole->rows || cols != whole->cols); }

    uchar* ptr(int i = 0) { return data + (size_t)i * step; }
    const uchar* ptr(int i = 0) const { return data + (size_t)i * step; }
    template <typename T> T* ptr(int i = 0) { return (T*)(data + (size_t)i * step); }
    template <typename T> const T* ptr(int i = 0) const { return (const T*)(data + (size_t)i * step); }
    template <typename T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    template <typename T> const T& at(int r, int c) const { return ((const T*)(data + (size_t)r * step))[c]; }
    // Mat::at(i0): element i0 of a row vector / a continuous matrix, row i0 of a column vector
    template <typename T> T& at(int i) { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    template <typename T> const T& at(int i) const { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    // A = <matrix expression>: OpenCV evaluates the expression INTO an already allocated destination of the right size and
    // type - which is how `A.row(0) = p1.x*Tcw1.row(2)-Tcw1.row(0)` (kannalabrandt8.cpp:489-492) fills a row of A
    inline Mat& operator=(const MatExpr& e);
    double dot(const Mat& b) const {                // Mat::dot on CV_32F: products accumulated in double
        assert(tp == CV_32F && b.tp == CV_32F && rows * cols == b.rows * b.cols);
        double s = 0;
        for (int i = 0; i < rows * cols; ++i) s += (double)at<float>(i) * (double)b.at<float>(i);
        return s;
    }
    // Mat::reshape(cn): the N x 2 one-channel <-> N x 1 two-channel views of frame.cpp:778-780 share one memory layout
    // (N rows of x, y); this stand-in keeps the N x 2 header and cv::undistortPoints below reads rows as points
    Mat reshape(int) const { return *this; }
    Mat t() const {
        assert(tp == CV_32F);
        Mat o; o.create(cols, rows, CV_32F);
        for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) o.at<float>(x, y) = at<float>(y, x);
        return o;
    }

    Mat operator()(const Rect& r) const {
        assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
        Mat m(*this);
        m.data = data + (size_t)r.y * step + (size_t)r.x * esz(); m.rows = r.height; m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat row(int i) const { return rowRange(i, i + 1); }
    Mat col(int i) const { return colRange(i, i + 1); }

    // Mat::locateROI / adjustROI, used by copyMakeBorder without BORDER_ISOLATED
    void locateROI(Size& wholeSize, Point& ofs) const {
        size_t d = (size_t)(data - whole->base);
        ofs.y = (int)(d / whole->step); ofs.x = (int)((d % whole->step) / esz());
        wholeSize = Size(whole->cols, whole->rows);
    }
    Mat& adjustROI(int dtop, int dbottom, int dleft, int dright) {
        Size ws; Point ofs; locateROI(ws, ofs);
        int r1 = std::max(ofs.y - dtop, 0), r2 = std::min(ofs.y + rows + dbottom, ws.height);
        int c1 = std::max(ofs.x - dleft, 0), c2 = std::min(ofs.x + cols + dright, ws.width);
        data = whole->base + (size_t)r1 * whole->step + (size_t)c1 * esz(); rows = r2 - r1; cols = c2 - c1;
        return *this;
    }

    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, tp);
        for (int y = 0; y < rows; ++y) std::memcpy(m.ptr(y), ptr(y), (size_t)cols * esz());
        return m;
    }
    inline void copyTo(OutputArray dst) const;
    // Mat::convertTo without scaling: 8U -> 32F (exact) or a same-type copy.  `dst` may be *this (the reference converts a
    // ROI view in place): a new buffer is made first, the source buffer is never written.
    void convertTo(Mat& dst, int rtype) const {
        Mat out;
        out.create(rows, cols, rtype);
        for (int y = 0; y < rows; ++y)
            for (int x = 0; x < cols; ++x) {
                const float v = tp == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
                if (rtype == CV_32F) out.at<float>(y, x) = v; else out.at<uchar>(y, x) = (uchar)v;
            }
        dst = out;
    }
};

// Result of a matrix expression: a value that converts to Mat, and that Mat::operator= copies into an existing destination.
class MatExpr {
public:
    Mat m;
    MatExpr() {}
    explicit MatExpr(const Mat& v) : m(v) {}
    operator Mat() const { return m; }
    template <typename T> const T& at(int i) const { return m.at<T>(i); }
    Mat t() const { return m.t(); }
};
inline Mat& Mat::operator=(const MatExpr& e) {
    if (data && rows == e.m.rows && cols == e.m.cols && tp == e.m.tp) {
        for (int y = 0; y < rows; ++y) std::memcpy(ptr(y), e.m.ptr(y), (size_t)cols * esz());
        return *this;
    }
    return *this = e.m;
}

// the matrix expressions of frame.cpp:921-937 ((float) * Mat::ones(...), Mat - Mat, norm(a, b, NORM_L1) on CV_32F patches), the
// pose algebra of fmatcher.cpp:2481-2505 (-A, A * B as gemm with float products accumulated in double per element, A + B)
// and kannalabrandt8.cpp:324-393, 485-498 (A / s, hconcat, dot, norm, Mat_<float>(r, c) << a, b, ...)
enum NormTypes { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };
static inline MatExpr operator*(float s, const Mat& m) {
    assert(m.type() == CV_32F);
    Mat o; o.create(m.rows, m.cols, CV_32F);
    for (int y = 0; y < m.rows; ++y) for (int x = 0; x < m.cols; ++x) o.at<float>(y, x) = s * m.at<float>(y, x);
    return MatExpr(o);
}
static inline MatExpr operator*(const Mat& m, float s) { return s * m; }
static inline MatExpr operator/(const Mat& m, float s) {
    assert(m.type() == CV_32F);
    Mat o; o.create(m.rows, m.cols, CV_32F);
    for (int y = 0; y < m.rows; ++y) for (int x = 0; x < m.cols; ++x) o.at<float>(y, x) = m.at<float>(y, x) / s;
    return MatExpr(o);
}
static inline MatExpr operator-(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat o; o.create(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) o.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    return MatExpr(o);
}
static inline MatExpr operator-(const Mat& a) { return -1.f * a; }
static inline MatExpr operator+(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat o; o.create(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) o.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    return MatExpr(o);
}
static inline MatExpr operator*(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
    Mat o; o.create(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < b.cols; ++x) {
            double s = 0;
            for (int k = 0; k < a.cols; ++k) s += (double)a.at<float>(y, k) * (double)b.at<float>(k, x);
            o.at<float>(y, x) = (float)s;
        }
    return MatExpr(o);
}
// MatExpr on the left: -R21*t12, (Rlw*twc)+tlw, p1.x*row - row ...
static inline MatExpr operator*(const MatExpr& a, const Mat& b) { return a.m * b; }
static inline MatExpr operator+(const MatExpr& a, const Mat& b) { return a.m + b; }
static inline MatExpr operator-(const MatExpr& a, const Mat& b) { return a.m - b; }
static inline MatExpr operator-(const Mat& a, const MatExpr& b) { return a - b.m; }
static inline MatExpr operator/(const MatExpr& a, float s) { return a.m / s; }
// cv::undistortPoints(src, dst, K, D, R = Mat(), P = K) on an N x 2 CV_32F matrix: forwarded to the oracle's restatement
// (orb_oracle.cpp), which tests/test_undistort.py pins bit for bit against the real cv2.undistortPoints
extern "C" void orc_undistort_points(const float* xy, int n, const float* cam, const float* dist, int ndist, float* out);
static inline void undistortPoints(const Mat& src, Mat& dst, const Mat& K, const Mat& D, const Mat& R, const Mat& P) {
    assert(src.type() == CV_32F && src.cols == 2 && R.empty() && P.data == K.data); (void)R; (void)P;
    const float cam[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    float dist[8];
    const int nd = D.rows * D.cols;
    for (int i = 0; i < nd && i < 8; ++i) dist[i] = D.at<float>(i);
    Mat in = src.clone(), out;
    out.create(src.rows, 2, CV_32F);
    orc_undistort_points(in.ptr<float>(0), src.rows, cam, dist, nd, out.ptr<float>(0));
    dst = out;
}
static inline double norm(const Mat& a) {            // L2 norm of a CV_32F vector, accumulated in double
    return std::sqrt(a.dot(a));
}
static inline void hconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows);
    Mat o; o.create(a.rows, a.cols + b.cols, CV_32F);
    for (int y = 0; y < a.rows; ++y) {
        for (int x = 0; x < a.cols; ++x) o.at<float>(y, x) = a.at<float>(y, x);
        for (int x = 0; x < b.cols; ++x) o.at<float>(y, a.cols + x) = b.at<float>(y, x);
    }
    dst = o;
}
// cv::Mat_<float>(r, c) << a, b, c ...
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
};
struct MatCommaInitializer {
    Mat m; int i;
    MatCommaInitializer& operator,(float v) { m.at<float>(i / m.cols, i % m.cols) = v; ++i; return *this; }
    operator Mat() const { return m; }
};
template <typename T> static inline MatCommaInitializer operator<<(const Mat_<T>& m, float v) {
    MatCommaInitializer c{m, 0};
    return c, v;
}
static inline double norm(const Mat& a, const Mat& b, int normType) {
    assert(normType == NORM_L1 && a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols); (void)normType;
    double s = 0;                       // cv::norm accumulates 32F L1 norms in double
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) s += std::fabs((double)a.at<float>(y, x) - (double)b.at<float>(y, x));
    return s;
}

// Proxies with the call shapes fextractor.cpp uses: image.empty()/getMat()/type(), descriptors.create()/
// release()/getMat().
class _InputArray {
public:
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& mat) : m(const_cast<Mat*>(&mat)) {}
    Mat getMat() const { return m ? *m : Mat(); }
    bool empty() const { return !m || m->empty(); }
    int type() const { return CV_8UC1; }
    Size size() const { return m ? m->size() : Size(); }
protected:
    Mat* m;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& mat) : _InputArray(mat) {}
    _OutputArray(const Mat& mat) : _InputArray(mat) {}   // OpenCV also accepts a const (temporary) header
    void create(int r, int c, int type) const { assert(m); m->create(r, c, type); }
    void create(Size s, int type) const { create(s.height, s.width, type); }
    void release() const { if (m) m->release(); }
    Mat& getMatRef() const { return *m; }
};
static inline InputArray noArray() { static _OutputArray none; return none; }

inline void Mat::copyTo(OutputArray dst) const {
    dst.create(rows, cols, tp);
    Mat d = dst.getMat();
    for (int y = 0; y < rows; ++y)
        if (d.ptr(y) != ptr(y)) std::memmove(d.ptr(y), ptr(y), (size_t)cols * esz());
}

// opencv2/core/types.hpp: class DMatch; opencv2/core/mat.hpp: Mat1b = Mat_<uchar>
struct DMatch {
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.402823466e+38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
    DMatch(int q, int t, int i, float d) : queryIdx(q), trainIdx(t), imgIdx(i), distance(d) {}
};
class Mat1b : public Mat {
public:
    Mat1b() {}
    Mat1b(const Mat& m) : Mat(m) {}
    Mat1b(int r, int c) : Mat(r, c, CV_8UC1) {}
};

// opencv2/features2d.hpp: KeyPointsFilter::retainBest — keep the n_points strongest, plus every
// keypoint tied with the weakest kept one (only ComputeKeyPointsOld, which compute() never calls, uses it).
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& keypoints, int n_points);
};

// ---- primitives (cv_compat.cpp; arithmetic = the cv2-pinned restatements of oracle/orb_oracle.cpp) ----
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);
float fastAtan2(float y, float x);
int borderInterpolate(int p, int len, int borderType);

}  // namespace cv

// ---- trace of the FAST calls the reference made (cv_compat.cpp), read by ref_driver.cpp ----
namespace cv_compat {
struct FastCall {
    const uchar* origin;     // first pixel of the ROI the reference passed
    size_t step;
    int w, h, threshold;
    std::vector<cv::KeyPoint> out;
};
std::vector<FastCall>& fast_trace();       // thread-local
void fast_trace_enable(bool on);
}
