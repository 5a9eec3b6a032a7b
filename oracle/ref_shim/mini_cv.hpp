// mini_cv.hpp — a small stand-in for the OpenCV C++ API surface that the reference's front-end sources touch.
// TEST INFRASTRUCTURE ONLY (used to build oracle/_ref/libmcsref.so).
//
// Purpose: /root/reference/MCSlam/src/ORBextractor.cpp (whole file) and line ranges of MultiCameraFrame.cpp / FrontEnd.cpp
// are compiled UNMODIFIED against these headers (recipe: oracle/ref_build.py), so that the control flow, containers,
// float arithmetic, std::sort / std::list order and tie rules of the checker are the reference's own object code.
// OpenCV itself is absent from this image; the OpenCV *primitives* those sources call (cv::FAST, cv::resize,
// cv::GaussianBlur, cv::fastAtan2, BFMatcher::knnMatch) forward to the cv2-pinned restatements in oracle/mcs_oracle.cpp
// (orc_* symbols of liboracle.so, bit-exact against cv2 4.13.0 goldens: tests/test_oracle_golden.py); copyMakeBorder,
// Mat algebra, SVD and the opencv_contrib sfm routines are written here from their published behaviour.
//
// Only what the compiled sources use is present: single-channel Mat of depth 8U / 32F / 64F, reference-counted ROI views,
// eager MatExpr, Mat_<T>, comma initialiser, Point_/Size/Rect/Range/Vec, KeyPoint, DMatch, InputArray/OutputArray.
//
// Documented differences from real OpenCV (also listed in BASELINE.md "deviations"):
//  * Mat(rows, cols, type) zero-fills (OpenCV leaves the buffer uninitialised) so that runs are deterministic.
//  * `view = MatExpr` with a different depth but the same size converts into the view's buffer.  Real OpenCV
//    re-allocates the temporary header, so MultiCameraFrame::triangulate (MultiCameraFrame.cpp:493-498, A is CV_32F, the
//    projection rows are CV_64F) leaves A unwritten there; the shim implements the evident intent.
//  * copyMakeBorder always behaves as BORDER_ISOLATED (a sub-matrix input never borrows its parent's pixels); only the
//    19 px frame of level 0 of a ROI input could differ, and nothing on the path reads that frame.
#pragma once
#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_8U
#define CV_32SC1 CV_32S
#define CV_32FC1 CV_32F
#define CV_64FC1 CV_64F
#define CV_PI 3.1415926535897932384626433832795

namespace cv {

[[noreturn]] inline void shim_fail(const char* what, const char* file, int line) {
    throw std::runtime_error(std::string("mini_cv assertion failed: ") + what + " at " + file + ":" + std::to_string(line));
}
}  // namespace cv
#define CV_Assert(expr) do { if (!(expr)) cv::shim_fail(#expr, __FILE__, __LINE__); } while (0)

// cvRound: OpenCV uses cvtsd2si / lrint = round half to even in the default rounding mode.
inline int cvRound(double v) { return (int)std::lrint(v); }
inline int cvRound(float v) { return (int)std::lrintf(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }

namespace cv {

using std::string;
template <typename T> using Ptr = std::shared_ptr<T>;

template <typename T> inline T saturate_cast(double v) { return (T)v; }
template <> inline int saturate_cast<int>(double v) { return cvRound(v); }
template <> inline uchar saturate_cast<uchar>(double v) { int i = cvRound(v); return (uchar)(i < 0 ? 0 : i > 255 ? 255 : i); }
template <> inline float saturate_cast<float>(double v) { return (float)v; }

enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_REFLECT101 = 4,
       BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };

float fastAtan2(float y, float x);  // OpenCV's degree polynomial; forwards to the cv2-pinned orc_fast_atan2

// ---------------------------------------------------------------------------------------------------------------------
template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U>& p) : x(saturate_cast<T>(p.x)), y(saturate_cast<T>(p.y)) {}
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, float b) { a.x = saturate_cast<T>(a.x * b); a.y = saturate_cast<T>(a.y * b); return a; }
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, double b) { a.x = saturate_cast<T>(a.x * b); a.y = saturate_cast<T>(a.y * b); return a; }
template <typename T> inline Point_<T>& operator*=(Point_<T>& a, int b) { a.x = saturate_cast<T>(a.x * b); a.y = saturate_cast<T>(a.y * b); return a; }
template <typename T> inline std::ostream& operator<<(std::ostream& o, const Point_<T>& p) { return o << "[" << p.x << ", " << p.y << "]"; }

template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;
typedef Point3_<double> Point3d;

template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
};
typedef Size_<int> Size;

template <typename T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

struct Range {
    int start, end;
    Range() : start(0), end(0) {}
    Range(int s, int e) : start(s), end(e) {}
    static Range all() { return Range(INT_MIN, INT_MAX); }
};

template <typename T, int N> struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; i++) val[i] = T(0); }
    Vec(T a, T b, T c) { static_assert(N == 3, ""); val[0] = a; val[1] = b; val[2] = c; }
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
};
typedef Vec<double, 3> Vec3d;
typedef Vec<float, 3> Vec3f;

struct KeyPoint {  // field order and defaults of cv::KeyPoint (28 bytes)
    Point2f pt;
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
    KeyPoint(Point2f pt_, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(pt_), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "KeyPoint layout");

struct DMatch {
    int queryIdx, trainIdx, imgIdx;
    float distance;
    DMatch() : queryIdx(-1), trainIdx(-1), imgIdx(-1), distance(3.402823466e+38f) {}
    DMatch(int q, int t, float d) : queryIdx(q), trainIdx(t), imgIdx(-1), distance(d) {}
    bool operator<(const DMatch& m) const { return distance < m.distance; }
};

// ---------------------------------------------------------------------------------------------------------------------
class Mat;
class MatExpr;
class _InputArray;
class _OutputArray;
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline size_t shim_elem_size(int type) {
    switch (type) {
        case CV_8U: case CV_8S: return 1;
        case CV_16U: case CV_16S: return 2;
        case CV_32S: case CV_32F: return 4;
        case CV_64F: return 8;
    }
    shim_fail("unsupported Mat type", __FILE__, __LINE__);
}

class Mat {
public:
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    size_t step = 0;  // bytes per row (OpenCV: MatStep, convertible to size_t)

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size sz, int type) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0) : rows(r), cols(c), data((uchar*)ext), type_(type) {  // caller-owned buffer
        step = ext_step ? ext_step : (size_t)c * shim_elem_size(type);
    }
    Mat(const Mat&) = default;
    Mat(const MatExpr& e);
    Mat& operator=(const Mat&) = default;
    Mat& operator=(const MatExpr& e);  // create(size, type) + write: an equally sized view is written in place

    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == type_ && cn_ == 1) return;
        rows = r; cols = c; type_ = type; cn_ = 1;
        step = (size_t)c * shim_elem_size(type);
        size_t bytes = step * (size_t)std::max(r, 0);
        buf_ = std::shared_ptr<uchar>(new uchar[bytes + 16](), std::default_delete<uchar[]>());
        data = buf_.get();
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }

    int type() const { return type_; }
    int depth() const { return type_; }
    int channels() const { return cn_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t elemSize() const { return shim_elem_size(type_) * cn_; }
    size_t step1() const { return step / shim_elem_size(type_); }
    // reshape(cn): same rows, the row's scalars regrouped into cn-channel elements (UndistortKeyPoints, MultiCameraFrame.cpp:332-334)
    Mat reshape(int cn) const {
        Mat m(*this);
        CV_Assert(cn >= 1 && (cols * cn_) % cn == 0);
        m.cols = cols * cn_ / cn;
        m.cn_ = cn;
        return m;
    }
    size_t total() const { return (size_t)rows * cols; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows == 1; }

    Mat rowRange(int a, int b) const { Mat m(*this); m.rows = b - a; m.data = data + (size_t)a * step; return m; }
    Mat colRange(int a, int b) const { Mat m(*this); m.cols = b - a; m.data = data + (size_t)a * elemSize(); return m; }
    Mat row(int i) const { return rowRange(i, i + 1); }
    Mat col(int j) const { return colRange(j, j + 1); }
    Mat operator()(const Rect& r) const { return rowRange(r.y, r.y + r.height).colRange(r.x, r.x + r.width); }
    Mat operator()(Range rr, Range cr) const {
        if (rr.start == INT_MIN) rr = Range(0, rows);
        if (cr.start == INT_MIN) cr = Range(0, cols);
        return rowRange(rr.start, rr.end).colRange(cr.start, cr.end);
    }

    template <typename T> T& at(int i, int j) { return *(T*)(data + (size_t)i * step + (size_t)j * sizeof(T)); }
    template <typename T> const T& at(int i, int j) const { return *(const T*)(data + (size_t)i * step + (size_t)j * sizeof(T)); }
    template <typename T> T& at(int i) { return (isContinuous() || rows == 1) ? ((T*)data)[i] : *(T*)(data + (size_t)i * step); }
    template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }
    template <typename T> T& at(Point p) { return at<T>(p.y, p.x); }
    template <typename T> const T& at(Point p) const { return at<T>(p.y, p.x); }
    uchar* ptr(int i = 0) { return data + (size_t)i * step; }
    const uchar* ptr(int i = 0) const { return data + (size_t)i * step; }
    template <typename T> T* ptr(int i = 0) { return (T*)(data + (size_t)i * step); }
    template <typename T> const T* ptr(int i = 0) const { return (const T*)(data + (size_t)i * step); }

    Mat clone() const;
    void copyTo(OutputArray dst) const;
    void convertTo(OutputArray dst, int rtype) const;
    MatExpr t() const;
    MatExpr inv() const;

    static MatExpr zeros(int r, int c, int type);
    static MatExpr ones(int r, int c, int type);
    static MatExpr eye(int r, int c, int type);

    double getd(int i, int j) const {  // element as double, any depth
        switch (type_) {
            case CV_8U: return at<uchar>(i, j);
            case CV_32S: return at<int>(i, j);
            case CV_32F: return at<float>(i, j);
            case CV_64F: return at<double>(i, j);
        }
        shim_fail("getd: unsupported depth", __FILE__, __LINE__);
    }
    void setd(int i, int j, double v) {
        switch (type_) {
            case CV_8U: at<uchar>(i, j) = saturate_cast<uchar>(v); return;
            case CV_32S: at<int>(i, j) = saturate_cast<int>(v); return;
            case CV_32F: at<float>(i, j) = (float)v; return;
            case CV_64F: at<double>(i, j) = v; return;
        }
        shim_fail("setd: unsupported depth", __FILE__, __LINE__);
    }

protected:
    int type_ = CV_8U;
    int cn_ = 1;  // channel count: only reshape() produces anything but 1, and only undistortPoints consumes it
    std::shared_ptr<uchar> buf_;
};

// Eagerly evaluated expression result.  Deriving from Mat lets every operator take `const Mat&`.
class MatExpr : public Mat {
public:
    MatExpr() {}
    explicit MatExpr(const Mat& m) { static_cast<Mat&>(*this) = m; }
    MatExpr t() const { return Mat::t(); }
    MatExpr inv() const { return Mat::inv(); }
};

inline Mat::Mat(const MatExpr& e) { *this = static_cast<const Mat&>(e); }
inline Mat& Mat::operator=(const MatExpr& e) {
    const Mat& s = e;
    if (data && rows == s.rows && cols == s.cols) {  // same size: written into the existing buffer (converted if the depth differs)
        for (int i = 0; i < rows; i++) {
            if (type_ == s.type()) std::memcpy(ptr(i), s.ptr(i), (size_t)cols * elemSize());
            else for (int j = 0; j < cols; j++) setd(i, j, s.getd(i, j));
        }
        return *this;
    }
    return *this = s;
}

template <typename T> struct DataType;
template <> struct DataType<uchar> { enum { type = CV_8U }; };
template <> struct DataType<int> { enum { type = CV_32S }; };
template <> struct DataType<float> { enum { type = CV_32F }; };
template <> struct DataType<double> { enum { type = CV_64F }; };

template <typename T> class MatCommaInitializer_;

template <typename T> class Mat_ : public Mat {
public:
    Mat_() { type_ = DataType<T>::type; }
    Mat_(int r, int c) : Mat(r, c, DataType<T>::type) {}
    Mat_(const Mat& m) { assign(m); }
    Mat_(const MatExpr& m) { assign(m); }
    Mat_(const Mat_&) = default;
    Mat_& operator=(const Mat_&) = default;
    Mat_& operator=(const Mat& m) { assign(m); return *this; }
    T& operator()(int i, int j) { return this->template at<T>(i, j); }
    const T& operator()(int i, int j) const { return this->template at<T>(i, j); }
    T& operator()(int i) { return this->template at<T>(i); }
    Mat_ clone() const { return Mat_(Mat::clone()); }
    static MatExpr zeros(int r, int c) { return Mat::zeros(r, c, DataType<T>::type); }
    static MatExpr ones(int r, int c) { return Mat::ones(r, c, DataType<T>::type); }
    static MatExpr eye(int r, int c) { return Mat::eye(r, c, DataType<T>::type); }

private:
    void assign(const Mat& m) {
        if (m.type() == DataType<T>::type || m.empty()) { static_cast<Mat&>(*this) = m; type_ = DataType<T>::type; }
        else m.convertTo(*this, DataType<T>::type);
    }
};

template <typename T> class MatCommaInitializer_ {
public:
    explicit MatCommaInitializer_(const Mat_<T>& m) : m_(m), k_(0) {}
    template <typename S> MatCommaInitializer_& operator,(S v) {
        CV_Assert(k_ < (int)m_.total());
        m_(k_ / m_.cols, k_ % m_.cols) = (T)v;
        k_++;
        return *this;
    }
    operator Mat_<T>() const { return m_; }
    operator Mat() const { return m_; }

private:
    Mat_<T> m_;
    int k_;
};
template <typename T, typename S> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, S v) {
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}

// ---------------------------------------------------------------------------------------------------------------------
class _InputArray {
public:
    _InputArray() {}
    _InputArray(const Mat& m) : m_(m) {}
    template <typename T> _InputArray(const Mat_<T>& m) : m_(m) {}
    _InputArray(const MatExpr& m) : m_(m) {}
    Mat getMat() const { return m_; }
    bool empty() const { return m_.empty(); }

protected:
    Mat m_;
};

class _OutputArray {
public:
    _OutputArray() : p_(nullptr) {}
    _OutputArray(Mat& m) : p_(&m) {}
    _OutputArray(const Mat& m) : p_(const_cast<Mat*>(&m)) {}  // OpenCV also lets temporaries (views) bind as outputs
    template <typename T> _OutputArray(Mat_<T>& m) : p_(&m) {}
    void create(int r, int c, int type) const { if (p_) p_->create(r, c, type); }
    void create(Size sz, int type) const { create(sz.height, sz.width, type); }
    void release() const { if (p_) p_->release(); }
    Mat getMat() const { return p_ ? *p_ : Mat(); }
    bool needed() const { return p_ != nullptr; }

private:
    Mat* p_;
};
inline const _InputArray& noArray() { static _InputArray none; return none; }

// ---------------------------------------------------------------------------------------------------------------------
inline Mat Mat::clone() const {
    Mat m;
    if (empty()) return m;
    m.create(rows, cols, type_);
    for (int i = 0; i < rows; i++) std::memcpy(m.ptr(i), ptr(i), (size_t)cols * elemSize());
    return m;
}
inline void Mat::copyTo(OutputArray dst) const {
    dst.create(rows, cols, type_);
    Mat d = dst.getMat();
    for (int i = 0; i < rows; i++)
        if (d.ptr(i) != ptr(i)) std::memmove(d.ptr(i), ptr(i), (size_t)cols * elemSize());
}
inline void Mat::convertTo(OutputArray dst, int rtype) const {
    if (rtype < 0) rtype = type_;
    Mat src = *this;  // dst may alias this
    Mat out;
    out.create(rows, cols, rtype);
    for (int i = 0; i < rows; i++)
        for (int j = 0; j < cols; j++) out.setd(i, j, src.getd(i, j));
    dst.create(rows, cols, rtype);
    Mat d = dst.getMat();
    for (int i = 0; i < rows; i++) std::memcpy(d.ptr(i), out.ptr(i), (size_t)cols * out.elemSize());
}
inline MatExpr Mat::zeros(int r, int c, int type) { return MatExpr(Mat(r, c, type)); }
inline MatExpr Mat::ones(int r, int c, int type) {
    Mat m(r, c, type);
    for (int i = 0; i < r; i++) for (int j = 0; j < c; j++) m.setd(i, j, 1.0);
    return MatExpr(m);
}
inline MatExpr Mat::eye(int r, int c, int type) {
    Mat m(r, c, type);
    for (int i = 0; i < std::min(r, c); i++) m.setd(i, i, 1.0);
    return MatExpr(m);
}
inline MatExpr Mat::t() const {
    Mat m(cols, rows, type_);
    for (int i = 0; i < rows; i++) for (int j = 0; j < cols; j++) m.setd(j, i, getd(i, j));
    return MatExpr(m);
}

inline int shim_float_type(const Mat& a) { CV_Assert(a.type() == CV_32F || a.type() == CV_64F); return a.type(); }
template <typename F> inline MatExpr shim_binary(const Mat& a, const Mat& b, F f) {
    CV_Assert(a.rows == b.rows && a.cols == b.cols && a.type() == b.type());
    Mat m(a.rows, a.cols, shim_float_type(a));
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) {
            if (a.type() == CV_32F) m.at<float>(i, j) = f(a.at<float>(i, j), b.at<float>(i, j));
            else m.at<double>(i, j) = f(a.at<double>(i, j), b.at<double>(i, j));
        }
    return MatExpr(m);
}
inline MatExpr operator+(const Mat& a, const Mat& b) { return shim_binary(a, b, [](auto x, auto y) { return x + y; }); }
inline MatExpr operator-(const Mat& a, const Mat& b) { return shim_binary(a, b, [](auto x, auto y) { return x - y; }); }
inline MatExpr shim_scale(const Mat& a, double s, bool divide) {
    Mat m(a.rows, a.cols, shim_float_type(a));
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) {
            // OpenCV's convertScale / divide compute in double for 64F and in float (from a double scale) for 32F
            if (a.type() == CV_32F) m.at<float>(i, j) = divide ? (float)(a.at<float>(i, j) / s) : (float)(a.at<float>(i, j) * s);
            else m.at<double>(i, j) = divide ? a.at<double>(i, j) / s : a.at<double>(i, j) * s;
        }
    return MatExpr(m);
}
inline MatExpr operator*(const Mat& a, double s) { return shim_scale(a, s, false); }
inline MatExpr operator*(double s, const Mat& a) { return shim_scale(a, s, false); }
inline MatExpr operator/(const Mat& a, double s) { return shim_scale(a, s, true); }
inline MatExpr operator-(const Mat& a) { return shim_scale(a, -1.0, false); }
inline MatExpr operator*(const Mat& a, const Mat& b) {  // gemm; accumulates in double in k order
    CV_Assert(a.cols == b.rows && a.type() == b.type());
    Mat m(a.rows, b.cols, shim_float_type(a));
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < b.cols; j++) {
            double s = 0;
            for (int k = 0; k < a.cols; k++) s += a.getd(i, k) * b.getd(k, j);
            m.setd(i, j, s);
        }
    return MatExpr(m);
}
std::ostream& operator<<(std::ostream& o, const Mat& m);

// cv::invert(DECOMP_LU): closed forms up to 3x3 (as OpenCV), Gauss-Jordan with partial pivoting above; zero matrix if singular.
MatExpr shim_invert(const Mat& a);
inline MatExpr Mat::inv() const { return shim_invert(*this); }
inline double norm(const Mat& a) {  // NORM_L2 over all elements, accumulated in double in element order
    double s = 0;
    for (int i = 0; i < a.rows; i++)
        for (int j = 0; j < a.cols; j++) s += a.getd(i, j) * a.getd(i, j);
    return std::sqrt(s);
}

// ---------------------------------------------------------------------------------------------------------------------
struct SVD {
    enum { MODIFY_A = 1, NO_UV = 2, FULL_UV = 4 };
    // singular values descending in w, rows of vt = right singular vectors (one-sided Jacobi, like OpenCV's JacobiSVD)
    static void compute(InputArray A, OutputArray w, OutputArray u, OutputArray vt, int flags = 0);
    static void solveZ(InputArray A, OutputArray z);  // right singular vector of the smallest singular value
};

// imgproc / features2d primitives: forwarded to liboracle's cv2-pinned restatements (mini_cv.cpp)
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);

// calib3d: cv::undistortPoints(src, dst, K, dist, R = noArray(), P): src / dst N x 1 two-channel CV_32F (may alias)
void undistortPoints(InputArray src, OutputArray dst, InputArray cameraMatrix, InputArray distCoeffs, InputArray R, InputArray P);

struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);  // only ComputeKeyPointsOld (dead code) uses it
};

class DescriptorMatcher {
public:
    static Ptr<DescriptorMatcher> create(const std::string& name);
    // BruteForce-Hamming: per query the k nearest train rows ordered by (distance, trainIdx); rows hold min(k, ntrain) entries
    void knnMatch(InputArray query, InputArray train, std::vector<std::vector<DMatch>>& matches, int k);
};

namespace sfm {
// opencv_contrib sfm: conversions.cpp / triangulation.cpp
void homogeneousToEuclidean(InputArray src, OutputArray dst);
void homogeneousToEuclidean(InputArray src, Vec3d& dst);
// points2d: one 2xN matrix per view; projection_matrices: one 3x4 per view; points3d: 3xN CV_64F.
// 2 views: 4x4 DLT (triangulateDLT), more: (3n)x(4+n) design (triangulateNViews); both SVD::solveZ.
void triangulatePoints(const std::vector<Mat_<double>>& points2d, const std::vector<Mat>& projection_matrices, OutputArray points3d);
}  // namespace sfm

}  // namespace cv
