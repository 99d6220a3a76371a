// oracle/shim/opencv2/opencv.hpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A header-only stand-in for the subset of the OpenCV C++ API that the reference's hot-path
// translation units use (src/math/se3.cpp, src/math/gaussian.cpp, src/core/transform.cpp,
// src/core/convert.cpp, src/track/optimize.cpp, src/track/tracker.cpp, src/map/implement.cpp,
// src/map/mapper.cpp, src/system/frame.cpp and the headers they include).  OpenCV C++ is an
// un-vendored, un-pinned dependency of the reference ("OpenCV >=3", README.md:8) and is not
// installed in this image, so oracle/Makefile compiles the reference's OWN sources, where they lie
// under /root/reference, against this header (-Ishim) to obtain oracle/_ref/libdvo_ref.so.
//
// Only single-channel float matrices exist in the reference's hot path, so cv::Mat here IS a float
// matrix.  What is restated from OpenCV's published behaviour (each piece is checked against
// opencv-python 4.13 known answers in tests/test_oracle_known_answers.py where Python exposes it):
//   * MatExpr is LAZY like OpenCV's: scale factors fold into a double alpha/beta and are applied at
//     assignment (convertTo / scaleAdd / addWeighted / gemm semantics), see "MatExpr" below;
//   * small cv::gemm (2..4 inner length): fp32 dot products left to right, no FMA, then
//     (float)(t*(float)alpha + c*(float)beta) evaluated in double;
//   * Point_<float> -> Point_<int> = saturate_cast = cvRound (round half to even, SSE cvtss2si);
//   * cv::norm / cv::trace / Mat::dot accumulate in double, sequentially;
//   * cv::Rodrigues, cv::solve(DECOMP_SVD) (JacobiSVDImpl_<float> + SVBkSb), cv::randn (MWC +
//     Ziggurat) follow modules/core/src/{lapack,rand}.cpp and modules/calib3d/src/calibration.cpp.
//   * forEach runs serially in row-major order; -DSHIM_PARALLEL_FOREACH parallelises
//     it over rows with OpenMP for the timing-only build (stand-in for cv::parallel_for_).
#ifndef DVO_SHIM_OPENCV_HPP
#define DVO_SHIM_OPENCV_HPP

#include <algorithm>
#include <array>
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <iostream>
#include <memory>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#define CV_32F 5
#define CV_64F 6
#define CV_32FC1 5

namespace cv
{
typedef uint64_t uint64;
typedef unsigned char uchar;

enum { DECOMP_LU = 0, DECOMP_SVD = 1, DECOMP_EIG = 2, DECOMP_CHOLESKY = 3, DECOMP_QR = 4, DECOMP_NORMAL = 16 };
enum { WINDOW_NORMAL = 0, WINDOW_AUTOSIZE = 1 };

// ---------------------------------------------------------------------------------------------
// saturate_cast (core/saturate.hpp, core/fast_math.hpp): float -> int is cvRound = cvtss2si
// ---------------------------------------------------------------------------------------------
inline int cvRound(float v)
{
    if (!(v > -2147483648.0f && v < 2147483648.0f)) return INT_MIN; // cvtss2si "integer indefinite"
    return (int)lrintf(v);
}
inline int cvRound(double v)
{
    if (!(v > -2147483649.0 && v < 2147483648.0)) return INT_MIN;
    return (int)lrint(v);
}
inline int cvRound(int v) { return v; }
template <typename T> inline T saturate_cast(float v) { return T(v); }
template <typename T> inline T saturate_cast(double v) { return T(v); }
template <typename T> inline T saturate_cast(int v) { return T(v); }
template <> inline int saturate_cast<int>(float v) { return cvRound(v); }
template <> inline int saturate_cast<int>(double v) { return cvRound(v); }

// ---------------------------------------------------------------------------------------------
// Size_, Point_, Point3_, Scalar (core/types.hpp)
// ---------------------------------------------------------------------------------------------
template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
    bool operator==(const Size_& o) const { return width == o.width && height == o.height; }
    bool operator!=(const Size_& o) const { return !(*this == o); }
    T area() const { return width * height; }
};
template <typename T> inline Size_<T> operator/(const Size_<T>& a, T b) { return Size_<T>(a.width / b, a.height / b); }
template <typename T> inline std::ostream& operator<<(std::ostream& os, const Size_<T>& s) { return os << "[" << s.width << " x " << s.height << "]"; }
typedef Size_<int> Size2i;
typedef Size2i Size;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    // converting constructor-by-operator, like OpenCV's `template<typename _Tp2> operator Point_<_Tp2>() const`
    template <typename U> operator Point_<U>() const { return Point_<U>(saturate_cast<U>(x), saturate_cast<U>(y)); }
    Point_& operator+=(const Point_& b) { x += b.x; y += b.y; return *this; }
    Point_& operator-=(const Point_& b) { x -= b.x; y -= b.y; return *this; }
};
template <typename T> inline Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(saturate_cast<T>(a.x + b.x), saturate_cast<T>(a.y + b.y)); }
template <typename T> inline Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(saturate_cast<T>(a.x - b.x), saturate_cast<T>(a.y - b.y)); }
template <typename T> inline Point_<T> operator-(const Point_<T>& a) { return Point_<T>(saturate_cast<T>(-a.x), saturate_cast<T>(-a.y)); }
template <typename T> inline Point_<T> operator*(const Point_<T>& a, int b) { return Point_<T>(saturate_cast<T>(a.x * b), saturate_cast<T>(a.y * b)); }
template <typename T> inline Point_<T> operator*(int a, const Point_<T>& b) { return Point_<T>(saturate_cast<T>(b.x * a), saturate_cast<T>(b.y * a)); }
template <typename T> inline Point_<T> operator*(const Point_<T>& a, float b) { return Point_<T>(saturate_cast<T>(a.x * b), saturate_cast<T>(a.y * b)); }
template <typename T> inline Point_<T> operator*(float a, const Point_<T>& b) { return Point_<T>(saturate_cast<T>(b.x * a), saturate_cast<T>(b.y * a)); }
template <typename T> inline Point_<T> operator*(const Point_<T>& a, double b) { return Point_<T>(saturate_cast<T>(a.x * b), saturate_cast<T>(a.y * b)); }
template <typename T> inline Point_<T> operator*(double a, const Point_<T>& b) { return Point_<T>(saturate_cast<T>(b.x * a), saturate_cast<T>(b.y * a)); }
// types.hpp: operator/= divides each coordinate (no reciprocal)
template <typename T> inline Point_<T> operator/(const Point_<T>& a, int b) { return Point_<T>(saturate_cast<T>(a.x / b), saturate_cast<T>(a.y / b)); }
template <typename T> inline Point_<T> operator/(const Point_<T>& a, float b) { return Point_<T>(saturate_cast<T>(a.x / b), saturate_cast<T>(a.y / b)); }
template <typename T> inline Point_<T> operator/(const Point_<T>& a, double b) { return Point_<T>(saturate_cast<T>(a.x / b), saturate_cast<T>(a.y / b)); }
template <typename T> inline bool operator==(const Point_<T>& a, const Point_<T>& b) { return a.x == b.x && a.y == b.y; }
template <typename T> inline std::ostream& operator<<(std::ostream& os, const Point_<T>& p) { return os << "[" << p.x << ", " << p.y << "]"; }
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
typedef Point2i Point;

// cv::norm(Point_<T>): std::sqrt((double)pt.x*pt.x + (double)pt.y*pt.y)   (types.hpp)
template <typename T> inline double norm(const Point_<T>& pt) { return std::sqrt((double)pt.x * pt.x + (double)pt.y * pt.y); }

template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

struct Scalar {
    double val[4];
    Scalar(double v0 = 0, double v1 = 0, double v2 = 0, double v3 = 0) { val[0] = v0; val[1] = v1; val[2] = v2; val[3] = v3; }
    double operator[](int i) const { return val[i]; }
};

class Mat;
class MatExpr;
template <typename T> class Mat_;
template <typename T> class MatCommaInitializer_;

// ---------------------------------------------------------------------------------------------
// Mat: ref-counted single-channel float matrix with row/column views (core/mat.hpp)
// ---------------------------------------------------------------------------------------------
class Mat
{
public:
    int rows = 0, cols = 0;
    size_t step = 0; // elements per row of the underlying buffer
    float* data = nullptr;
    std::shared_ptr<float> buf;

    Mat() {}
    Mat(int r, int c) { create(r, c); }
    Mat(int r, int c, int /*type*/) { create(r, c); }
    Mat(int r, int c, float v) { create(r, c); setTo(v); }
    explicit Mat(Size s) { create(s.height, s.width); }
    Mat(Size s, int /*type*/) { create(s.height, s.width); }
    Mat(Size s, float v) { create(s.height, s.width); setTo(v); }
    template <typename T> explicit Mat(const Point3_<T>& p) { create(3, 1); data[0] = (float)p.x; data[step] = (float)p.y; data[2 * step] = (float)p.z; }
    Mat(const MatExpr& e);            // defined after MatExpr
    Mat& operator=(const MatExpr& e); // defined after MatExpr
    Mat& operator=(float v) { setTo(v); return *this; }

    void create(int r, int c)
    {
        if (r == rows && c == cols && data && step == (size_t)cols && buf && buf.get() == data) return;
        rows = r; cols = c; step = (size_t)c;
        const size_t n = (size_t)r * c;
        buf = std::shared_ptr<float>(n ? new float[n] : nullptr, std::default_delete<float[]>());
        data = buf.get();
    }
    void setTo(float v)
    {
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) data[y * step + x] = v;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    int type() const { return CV_32FC1; }
    int channels() const { return 1; }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return step == (size_t)cols || rows <= 1; }

    float* ptr(int r = 0) { return data + (size_t)r * step; }
    const float* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T& at(int r, int c) { return data[(size_t)r * step + c]; }
    template <typename T> const T& at(int r, int c) const { return data[(size_t)r * step + c]; }
    template <typename T> T& at(Point p) { return data[(size_t)p.y * step + p.x]; }
    template <typename T> const T& at(Point p) const { return data[(size_t)p.y * step + p.x]; }
    template <typename T> T& at(int i) { return elem(i); }
    template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->elem(i); }
    float& elem(int i) // Mat::at(int i0)
    {
        if (rows == 1) return data[i];
        if (cols == 1) return data[(size_t)i * step];
        return data[(size_t)(i / cols) * step + (i % cols)];
    }

    Mat view(int r0, int r1, int c0, int c1) const
    {
        Mat m;
        m.rows = r1 - r0; m.cols = c1 - c0; m.step = step; m.buf = buf;
        m.data = data + (size_t)r0 * step + c0;
        return m;
    }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat row(int i) const { return view(i, i + 1, 0, cols); }
    Mat col(int i) const { return view(0, rows, i, i + 1); }

    Mat clone() const
    {
        Mat m(rows, cols);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, sizeof(float) * cols);
        return m;
    }
    // copyTo(OutputArray): OutputArray binds to temporaries (views) in OpenCV, hence const Mat&
    void copyTo(const Mat& dst_) const
    {
        Mat& dst = const_cast<Mat&>(dst_);
        if (dst.rows != rows || dst.cols != cols || !dst.data) dst.create(rows, cols);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) dst.data[(size_t)y * dst.step + x] = data[(size_t)y * step + x];
    }
    void convertTo(Mat& dst, int /*rtype*/, double alpha = 1, double beta = 0) const;
    MatExpr t() const;
    // Mat::dot: dotProd_32f — products and sum in double for lengths below the SIMD block
    double dot(const Mat& m) const
    {
        double r = 0;
        assert(total() == m.total());
        const int n = (int)total();
        for (int i = 0; i < n; i++) r += (double)const_cast<Mat*>(this)->elem(i) * (double)const_cast<Mat&>(m).elem(i);
        return r;
    }

    template <typename F> void forEach(const F& f) const
    {
        Mat* self = const_cast<Mat*>(this);
#ifdef SHIM_PARALLEL_FOREACH
#pragma omp parallel for schedule(dynamic, 4)
#endif
        for (int y = 0; y < rows; y++) {
            int pos[2];
            pos[0] = y;
            for (int x = 0; x < cols; x++) {
                pos[1] = x;
                f(self->data[(size_t)y * step + x], pos);
            }
        }
    }

    static MatExpr zeros(int r, int c, int type);
    static MatExpr zeros(Size s, int type);
    static MatExpr ones(int r, int c, int type);
    static MatExpr ones(Size s, int type);
    static MatExpr eye(int r, int c, int type);

    template <typename T> operator Point3_<T>() const
    {
        Mat* s = const_cast<Mat*>(this);
        return Point3_<T>((T)s->elem(0), (T)s->elem(1), (T)s->elem(2));
    }
};

template <typename T> class Mat_ : public Mat
{
public:
    typedef T value_type;
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c) {}
    Mat_(int r, int c, const T& v) : Mat(r, c, (float)v) {}
    explicit Mat_(Size s) : Mat(s) {}
    Mat_(Size s, const T& v) : Mat(s, (float)v) {}
    Mat_(const Mat& m) : Mat(m) {}
    Mat_(const Mat_& m) : Mat(m) {}
    Mat_(const MatExpr& e) : Mat(e) {}
    explicit Mat_(const Point3_<T>& p) : Mat(p) {}
    Mat_(const MatCommaInitializer_<T>& ci);
    Mat_& operator=(const Mat& m) { Mat::operator=(m); return *this; }
    Mat_& operator=(const Mat_& m) { Mat::operator=(m); return *this; }
    Mat_& operator=(const MatExpr& e) { Mat::operator=(e); return *this; }
    Mat_& operator=(const T& v) { setTo((float)v); return *this; }

    T& operator()(int r, int c) { return data[(size_t)r * step + c]; }
    const T& operator()(int r, int c) const { return data[(size_t)r * step + c]; }
    T& operator()(int i) { return elem(i); }
    const T& operator()(int i) const { return const_cast<Mat_*>(this)->elem(i); }
    T& operator()(Point p) { return data[(size_t)p.y * step + p.x]; }
    const T& operator()(Point p) const { return data[(size_t)p.y * step + p.x]; }

    Mat_ rowRange(int a, int b) const { return Mat_(Mat::rowRange(a, b)); }
    Mat_ colRange(int a, int b) const { return Mat_(Mat::colRange(a, b)); }
    Mat_ row(int i) const { return Mat_(Mat::row(i)); }
    Mat_ col(int i) const { return Mat_(Mat::col(i)); }
    Mat_ clone() const { return Mat_(Mat::clone()); }

    static MatExpr zeros(int r, int c);
    static MatExpr zeros(Size s);
    static MatExpr ones(int r, int c);
    static MatExpr ones(Size s);
    static MatExpr eye(int r, int c);
};
typedef Mat_<float> Mat1f;

// `(cv::Mat1f(3, 3) << a, b, c, ...)`
template <typename T> class MatCommaInitializer_
{
public:
    Mat_<T> m;
    int idx = 0;
    MatCommaInitializer_(const Mat_<T>& m_) : m(m_) {}
    template <typename U> MatCommaInitializer_& operator,(U v)
    {
        m.elem(idx++) = (T)v;
        return *this;
    }
    operator Mat() const { return m; }
};
template <typename T, typename U> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v)
{
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}
template <typename T> inline Mat_<T>::Mat_(const MatCommaInitializer_<T>& ci) : Mat(ci.m) {}

// ---------------------------------------------------------------------------------------------
// MatExpr — lazy matrix expressions (modules/core/src/matop.cpp).  Restated behaviour:
//   IDENT   a
//   ADDEX   a*alpha + b*beta + s     (b optional)    -> convertTo / add / subtract / scaleAdd / addWeighted
//   GEMM    alpha*a*b + beta*c       (c optional)    -> cv::gemm
//   INIT    zeros/ones/eye * alpha
//   TRANSP  alpha * a^T
//   BIN     max/min(a, scalar)
// Scalar multiplication/division folds into alpha (division multiplies by 1.0/s, matop.cpp
// "operator / (const MatExpr& e, double s)"), sums of products fold into gemm's C operand exactly
// as MatOp_GEMM::add / ::subtract and MatOp::add / ::subtract do.
// ---------------------------------------------------------------------------------------------
namespace shim
{
// cv::gemm for float (modules/core/src/matmul.dispatch.cpp): inner lengths 2..4 take the unrolled
// path: float t = a0*b0 + a1*b1 + ... (fp32, left to right), d = (float)(t*alpha + c*beta) with
// alpha, beta rounded to float and that line evaluated in double (bit-exact against cv2.gemm 4.13,
// tests/test_oracle_known_answers.py).  Other shapes: GEMMSingleMul<float,double> (double accumulation).
inline Mat gemm(const Mat& A, const Mat& B, double alpha_, const Mat* C, double beta)
{
    assert(A.cols == B.rows);
    const int M = A.rows, N = B.cols, len = A.cols;
    Mat D(M, N);
    const bool small = (len >= 2 && len <= 4 && (len == N || len == M));
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) {
            const double c = C ? (double)C->data[(size_t)i * C->step + j] : 0.0;
            const double be = C ? (double)(float)beta : 0.0;
            const double alpha = (double)(float)alpha_; // alpha/beta are rounded to float first (checked against cv2.gemm)
            if (small) {
                float t = A.data[(size_t)i * A.step] * B.data[j];
                for (int k = 1; k < len; k++) t = t + A.data[(size_t)i * A.step + k] * B.data[(size_t)k * B.step + j];
                D.data[(size_t)i * D.step + j] = (float)((double)t * alpha + c * be);
            } else {
                double t = 0;
                for (int k = 0; k < len; k++) t += (double)A.data[(size_t)i * A.step + k] * (double)B.data[(size_t)k * B.step + j];
                D.data[(size_t)i * D.step + j] = (float)(t * alpha + c * be);
            }
        }
    return D;
}
} // namespace shim

class MatExpr
{
public:
    enum Kind { IDENT, ADDEX, GEMM, INIT, TRANSP, BINMAX, BINMIN };
    Kind kind = IDENT;
    Mat a, b, c;
    double alpha = 1, beta = 1, s = 0;
    int init_kind = 0, irows = 0, icols = 0; // INIT: '0', '1', 'I'

    MatExpr() {}
    MatExpr(const Mat& m) : kind(IDENT), a(m) {}

    static MatExpr addex(const Mat& a, const Mat& b, double alpha, double beta, double s = 0)
    {
        MatExpr e; e.kind = ADDEX; e.a = a; e.b = b; e.alpha = alpha; e.beta = beta; e.s = s;
        return e;
    }
    static MatExpr gemm(const Mat& a, const Mat& b, double alpha = 1, const Mat& c = Mat(), double beta = 1)
    {
        MatExpr e; e.kind = GEMM; e.a = a; e.b = b; e.c = c; e.alpha = alpha; e.beta = beta;
        return e;
    }
    static MatExpr init(int k, int r, int c_, double alpha = 1)
    {
        MatExpr e; e.kind = INIT; e.init_kind = k; e.irows = r; e.icols = c_; e.alpha = alpha;
        return e;
    }
    // predicates of matop.cpp
    bool isIdentity() const { return kind == IDENT; }
    bool isAddEx() const { return kind == ADDEX; }
    bool isScaled() const { return kind == ADDEX && (b.empty() || beta == 0) && s == 0; }
    bool isUnaryAddEx() const { return kind == ADDEX && (b.empty() || beta == 0); }
    bool isMatProd() const { return kind == GEMM && (c.empty() || beta == 0); }
    bool isT() const { return kind == TRANSP; }

    Size size() const
    {
        switch (kind) {
        case GEMM: return Size(b.cols, a.rows);
        case INIT: return Size(icols, irows);
        case TRANSP: return Size(a.rows, a.cols);
        default: return a.size();
        }
    }

    Mat eval() const
    {
        switch (kind) {
        case IDENT: return a;
        case INIT: {
            Mat m(irows, icols);
            m.setTo(init_kind == '1' ? (float)alpha : 0.f);
            if (init_kind == 'I')
                for (int i = 0; i < std::min(irows, icols); i++) m.data[(size_t)i * m.step + i] = (float)alpha;
            return m;
        }
        case TRANSP: {
            Mat m(a.cols, a.rows);
            const float al = (float)alpha;
            for (int y = 0; y < a.rows; y++)
                for (int x = 0; x < a.cols; x++) {
                    const float v = a.data[(size_t)y * a.step + x];
                    m.data[(size_t)x * m.step + y] = alpha == 1 ? v : v * al;
                }
            return m;
        }
        case GEMM: return shim::gemm(a, b, alpha, (c.empty() || beta == 0) ? nullptr : &c, beta);
        case BINMAX:
        case BINMIN: {
            Mat m(a.rows, a.cols);
            const float sv = (float)s;
            for (int y = 0; y < a.rows; y++)
                for (int x = 0; x < a.cols; x++) {
                    const float v = a.data[(size_t)y * a.step + x];
                    m.data[(size_t)y * m.step + x] = kind == BINMAX ? std::max(v, sv) : std::min(v, sv);
                }
            return m;
        }
        case ADDEX: break;
        }
        // MatOp_AddEx::assign
        Mat m(a.rows, a.cols);
        const float sf = (float)s;
        if (b.empty() || beta == 0) {
            // a.convertTo(m, type, alpha) (+ s): cvtScale32f -> src*(float)alpha + (float)beta
            const float al = (float)alpha;
            for (int y = 0; y < a.rows; y++)
                for (int x = 0; x < a.cols; x++) {
                    const float v = a.data[(size_t)y * a.step + x];
                    float r = alpha == 1 ? v : v * al;
                    if (s != 0) r = r + sf;
                    m.data[(size_t)y * m.step + x] = r;
                }
            return m;
        }
        assert(a.rows == b.rows && a.cols == b.cols);
        for (int y = 0; y < a.rows; y++)
            for (int x = 0; x < a.cols; x++) {
                const float va = a.data[(size_t)y * a.step + x], vb = b.data[(size_t)y * b.step + x];
                float r;
                if (alpha == 1 && beta == 1) r = va + vb;                   // cv::add
                else if (alpha == 1 && beta == -1) r = va - vb;             // cv::subtract
                // cv::scaleAdd is v_muladd(src1, (float)alpha, src2): fused on FMA3 hardware (as measured
                // with cv2 4.13 here).  In the reference's hot path its operands have disjoint non-zeros
                // (eye + alpha*hat(w), se3.cpp:85,114), where fused and unfused results coincide.
                else if (alpha == 1) r = std::fmaf(vb, (float)beta, va);    // cv::scaleAdd(b, beta, a)
                else if (beta == 1 && alpha == -1) r = vb - va;             // cv::subtract(b, a)
                else if (beta == 1) r = std::fmaf(va, (float)alpha, vb);    // cv::scaleAdd(a, alpha, b)
                else r = (float)((double)va * alpha + (double)vb * beta);   // cv::addWeighted (gamma 0): double
                if (s != 0) r = r + sf;
                m.data[(size_t)y * m.step + x] = r;
            }
        return m;
    }
    operator Mat() const { return eval(); }

    MatExpr t() const;
    // conveniences used on expressions in the reference
    Mat colRange(int a_, int b_) const { return eval().colRange(a_, b_); }
    Mat rowRange(int a_, int b_) const { return eval().rowRange(a_, b_); }
    double dot(const Mat& m) const { return eval().dot(m); }
};

inline Mat::Mat(const MatExpr& e) { *this = e.eval(); }
inline Mat& Mat::operator=(const MatExpr& e)
{
    Mat r = e.eval();
    if (data && rows == r.rows && cols == r.cols && r.data != data) r.copyTo(*this); // Mat::create reuses the buffer
    else { rows = r.rows; cols = r.cols; step = r.step; data = r.data; buf = r.buf; }
    return *this;
}
inline void Mat::convertTo(Mat& dst, int, double alpha, double beta) const
{
    Mat r = MatExpr::addex(*this, Mat(), alpha, 0, beta).eval();
    dst = r;
}
inline MatExpr Mat::t() const { MatExpr e; e.kind = MatExpr::TRANSP; e.a = *this; e.alpha = 1; return e; }
inline MatExpr MatExpr::t() const
{
    if (kind == TRANSP) return MatExpr::addex(a, Mat(), alpha, 0);
    return eval().t();
}
inline MatExpr Mat::zeros(int r, int c, int) { return MatExpr::init('0', r, c); }
inline MatExpr Mat::zeros(Size s, int) { return MatExpr::init('0', s.height, s.width); }
inline MatExpr Mat::ones(int r, int c, int) { return MatExpr::init('1', r, c); }
inline MatExpr Mat::ones(Size s, int) { return MatExpr::init('1', s.height, s.width); }
inline MatExpr Mat::eye(int r, int c, int) { return MatExpr::init('I', r, c); }
template <typename T> inline MatExpr Mat_<T>::zeros(int r, int c) { return MatExpr::init('0', r, c); }
template <typename T> inline MatExpr Mat_<T>::zeros(Size s) { return MatExpr::init('0', s.height, s.width); }
template <typename T> inline MatExpr Mat_<T>::ones(int r, int c) { return MatExpr::init('1', r, c); }
template <typename T> inline MatExpr Mat_<T>::ones(Size s) { return MatExpr::init('1', s.height, s.width); }
template <typename T> inline MatExpr Mat_<T>::eye(int r, int c) { return MatExpr::init('I', r, c); }

// ---- operator dispatch (matop.cpp: MatOp::add/subtract/matmul/multiply, MatOp_GEMM::add/subtract) ----
namespace shim
{
inline bool same_op(const MatExpr& x, const MatExpr& y)
{
    auto cls = [](MatExpr::Kind k) { return k == MatExpr::BINMIN ? MatExpr::BINMAX : k; };
    return cls(x.kind) == cls(y.kind);
}
inline Mat transposed(const Mat& m) { return m.t().eval(); }
// generic branch of MatOp::add / MatOp::subtract (sign = +1 / -1)
inline MatExpr addsub_generic(const MatExpr& e1, const MatExpr& e2, double sign)
{
    double alpha = 1, beta = sign, s = 0;
    Mat m1, m2;
    if (e1.isUnaryAddEx()) { m1 = e1.a; alpha = e1.alpha; s = e1.s; }
    else m1 = e1.eval();
    if (e2.isUnaryAddEx()) { m2 = e2.a; beta = sign * e2.alpha; s += sign * e2.s; }
    else m2 = e2.eval();
    return MatExpr::addex(m1, m2, alpha, beta, s);
}
inline MatExpr addsub(const MatExpr& e1, const MatExpr& e2, double sign)
{
    if (e1.kind == MatExpr::GEMM || e2.kind == MatExpr::GEMM) { // MatOp_GEMM::add / ::subtract
        const bool i1 = e1.isIdentity(), i2 = e2.isIdentity();
        const double alpha1 = i1 ? 1 : e1.alpha, alpha2 = i2 ? 1 : e2.alpha;
        if (e1.isMatProd() && (i2 || e2.isScaled() || e2.isT()))
            return MatExpr::gemm(e1.a, e1.b, alpha1, e2.isT() ? transposed(e2.a) : e2.a, sign * alpha2);
        if (e2.isMatProd() && (i1 || e1.isScaled() || e1.isT()))
            return MatExpr::gemm(e2.a, e2.b, sign * alpha2, e1.isT() ? transposed(e1.a) : e1.a, alpha1);
    }
    return addsub_generic(e1, e2, sign);
}
inline MatExpr scale(const MatExpr& e, double s) // op->multiply(e, s)
{
    MatExpr r = e;
    switch (e.kind) {
    case MatExpr::ADDEX: r.alpha *= s; r.beta *= s; r.s *= s; return r;
    case MatExpr::GEMM: r.alpha *= s; r.beta *= s; return r;
    case MatExpr::INIT: r.alpha *= s; return r;
    case MatExpr::TRANSP: r.alpha *= s; return r;
    default: return MatExpr::addex(e.eval(), Mat(), s, 0);
    }
}
inline MatExpr negate(const MatExpr& e) // op->subtract(Scalar(0), e)
{
    if (e.kind == MatExpr::ADDEX) { MatExpr r = e; r.alpha = -r.alpha; r.beta = -r.beta; r.s = 0 - r.s; return r; }
    return MatExpr::addex(e.eval(), Mat(), -1, 0);
}
inline MatExpr matmul(const MatExpr& e1, const MatExpr& e2) // MatOp::matmul
{
    double sc = 1;
    Mat m1, m2;
    if (e1.isT()) { sc = e1.alpha; m1 = transposed(e1.a); }
    else if (e1.isScaled()) { sc = e1.alpha; m1 = e1.a; }
    else m1 = e1.eval();
    if (e2.isT()) { sc *= e2.alpha; m2 = transposed(e2.a); }
    else if (e2.isScaled()) { sc *= e2.alpha; m2 = e2.a; }
    else m2 = e2.eval();
    return MatExpr::gemm(m1, m2, sc);
}
} // namespace shim

inline MatExpr operator+(const Mat& a, const Mat& b) { return MatExpr::addex(a, b, 1, 1); }
inline MatExpr operator+(const MatExpr& e, const Mat& m) { return shim::addsub(e, MatExpr(m), 1); }
inline MatExpr operator+(const Mat& m, const MatExpr& e) { return shim::addsub(e, MatExpr(m), 1); }
inline MatExpr operator+(const MatExpr& e1, const MatExpr& e2) { return shim::addsub(e1, e2, 1); }
inline MatExpr operator-(const Mat& a, const Mat& b) { return MatExpr::addex(a, b, 1, -1); }
inline MatExpr operator-(const MatExpr& e, const Mat& m) { return shim::addsub(e, MatExpr(m), -1); }
inline MatExpr operator-(const Mat& m, const MatExpr& e) { return shim::addsub(MatExpr(m), e, -1); }
inline MatExpr operator-(const MatExpr& e1, const MatExpr& e2) { return shim::addsub(e1, e2, -1); }
inline MatExpr operator-(const Mat& m) { return MatExpr::addex(m, Mat(), -1, 0); }
inline MatExpr operator-(const MatExpr& e) { return shim::negate(e); }
inline MatExpr operator*(const Mat& a, const Mat& b) { return MatExpr::gemm(a, b); }
inline MatExpr operator*(const MatExpr& e, const Mat& m) { return shim::matmul(e, MatExpr(m)); }
inline MatExpr operator*(const Mat& m, const MatExpr& e) { return shim::matmul(MatExpr(m), e); }
inline MatExpr operator*(const MatExpr& e1, const MatExpr& e2) { return shim::matmul(e1, e2); }
inline MatExpr operator*(const Mat& a, double s) { return MatExpr::addex(a, Mat(), s, 0); }
inline MatExpr operator*(double s, const Mat& a) { return MatExpr::addex(a, Mat(), s, 0); }
inline MatExpr operator*(const MatExpr& e, double s) { return shim::scale(e, s); }
inline MatExpr operator*(double s, const MatExpr& e) { return shim::scale(e, s); }
inline MatExpr operator/(const Mat& a, double s) { return MatExpr::addex(a, Mat(), 1. / s, 0); }
inline MatExpr operator/(const MatExpr& e, double s) { return shim::scale(e, 1. / s); }
inline MatExpr max(const Mat& a, double s) { MatExpr e; e.kind = MatExpr::BINMAX; e.a = a; e.s = s; return e; }
inline MatExpr min(const Mat& a, double s) { MatExpr e; e.kind = MatExpr::BINMIN; e.a = a; e.s = s; return e; }

inline std::ostream& operator<<(std::ostream& os, const Mat& m)
{
    os << "[";
    for (int y = 0; y < m.rows; y++) {
        for (int x = 0; x < m.cols; x++) os << m.data[(size_t)y * m.step + x] << (x + 1 < m.cols ? ", " : "");
        if (y + 1 < m.rows) os << ";\n ";
    }
    return os << "]";
}
inline std::ostream& operator<<(std::ostream& os, const MatExpr& e) { return os << e.eval(); }

// ---------------------------------------------------------------------------------------------
// reductions: cv::norm (NORM_L2), cv::trace, cv::minMaxIdx
// ---------------------------------------------------------------------------------------------
inline double norm(const Mat& m) // normL2_32f: squares and sum in double, sequential
{
    double s = 0;
    for (int y = 0; y < m.rows; y++)
        for (int x = 0; x < m.cols; x++) {
            const double v = m.data[(size_t)y * m.step + x];
            s += v * v;
        }
    return std::sqrt(s);
}
inline double norm(const MatExpr& e) { return norm(e.eval()); }
inline Scalar trace(const Mat& m) // double accumulation over the diagonal
{
    double s = 0;
    for (int i = 0; i < std::min(m.rows, m.cols); i++) s += m.data[(size_t)i * m.step + i];
    return Scalar(s);
}
inline void minMaxIdx(const Mat& m, double* minVal, double* maxVal, int* = nullptr, int* = nullptr)
{
    float lo = FLT_MAX, hi = -FLT_MAX;
    for (int y = 0; y < m.rows; y++)
        for (int x = 0; x < m.cols; x++) {
            const float v = m.data[(size_t)y * m.step + x];
            lo = std::min(lo, v); hi = std::max(hi, v);
        }
    if (minVal) *minVal = lo;
    if (maxVal) *maxVal = hi;
}

// ---------------------------------------------------------------------------------------------
// cv::Rodrigues, rotation vector -> matrix (modules/calib3d/src/calibration.cpp cvRodrigues2):
// double arithmetic, R = c*I + (1-c)*r*r^T + s*[r]x, converted to the input depth (float).
// ---------------------------------------------------------------------------------------------
inline void Rodrigues(const Mat& src, Mat& dst)
{
    assert(src.total() == 3);
    Mat& sm = const_cast<Mat&>(src);
    double rx = sm.elem(0), ry = sm.elem(1), rz = sm.elem(2);
    const double theta = std::sqrt(rx * rx + ry * ry + rz * rz);
    Mat R(3, 3);
    if (theta < DBL_EPSILON) {
        R.setTo(0.f);
        R.data[0] = R.data[4] = R.data[8] = 1.f;
    } else {
        const double c = std::cos(theta), s = std::sin(theta), c1 = 1. - c, itheta = 1. / theta;
        rx *= itheta; ry *= itheta; rz *= itheta;
        const double rrt[9] = {rx * rx, rx * ry, rx * rz, rx * ry, ry * ry, ry * rz, rx * rz, ry * rz, rz * rz};
        const double r_x[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
        const double eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        for (int k = 0; k < 9; k++) R.data[k] = (float)((c * eye[k] + c1 * rrt[k]) + s * r_x[k]);
    }
    dst = R;
}

// ---------------------------------------------------------------------------------------------
// cv::solve(A, B, X, DECOMP_SVD) for float, m >= n (modules/core/src/lapack.cpp): A is transposed
// into At (n x m), JacobiSVDImpl_<float> (one-sided Jacobi, double accumulators, float rotations,
// eps = 2*FLT_EPSILON, minval = FLT_MIN), then SVBkSbImpl_ with threshold 2*FLT_EPSILON*sum(w).
// A build of OpenCV with LAPACK (e.g. the opencv-python wheels) routes through sgesdd instead and
// agrees to ~1e-5 relative only; the restated path is the HAL-less one.
// ---------------------------------------------------------------------------------------------
inline bool solve(const Mat& A_, const Mat& B_, Mat& X, int method)
{
    assert((method & ~DECOMP_NORMAL) == DECOMP_SVD);
    (void)method;
    const Mat A = A_, B = B_;
    const int m = A.rows, n = A.cols, nb = B.cols;
    assert(m >= n && B.rows == m);
    std::vector<float> At((size_t)n * m), Vt((size_t)n * n), Wf(n);
    std::vector<double> W(n);
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) At[(size_t)j * m + i] = A.data[(size_t)i * A.step + j];
    const float eps = FLT_EPSILON * 2;
    const double minval = FLT_MIN;
    const int max_iter = std::max(m, 30);
    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) { const float t = At[(size_t)i * m + k]; sd += (double)t * t; }
        W[i] = sd;
        for (int k = 0; k < n; k++) Vt[(size_t)i * n + k] = 0;
        Vt[(size_t)i * n + i] = 1;
    }
    for (int iter = 0; iter < max_iter; iter++) {
        bool changed = false;
        for (int i = 0; i < n - 1; i++)
            for (int j = i + 1; j < n; j++) {
                float *Ai = &At[(size_t)i * m], *Aj = &At[(size_t)j * m];
                double a = W[i], p = 0, b = W[j];
                for (int k = 0; k < m; k++) p += (double)Ai[k] * Aj[k];
                if (std::abs(p) <= eps * std::sqrt((double)a * b)) continue;
                p *= 2;
                const double beta = a - b, gamma = hypot((double)p, beta);
                float c, s;
                if (beta < 0) {
                    const double delta = (gamma - beta) * 0.5;
                    s = (float)std::sqrt(delta / gamma);
                    c = (float)(p / (gamma * s * 2));
                } else {
                    c = (float)std::sqrt((gamma + beta) / (gamma * 2));
                    s = (float)(p / (gamma * c * 2));
                }
                a = b = 0;
                for (int k = 0; k < m; k++) {
                    const float t0 = c * Ai[k] + s * Aj[k];
                    const float t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0; Aj[k] = t1;
                    a += (double)t0 * t0; b += (double)t1 * t1;
                }
                W[i] = a; W[j] = b;
                changed = true;
                float *Vi = &Vt[(size_t)i * n], *Vj = &Vt[(size_t)j * n];
                for (int k = 0; k < n; k++) {
                    const float t0 = c * Vi[k] + s * Vj[k];
                    const float t1 = -s * Vi[k] + c * Vj[k];
                    Vi[k] = t0; Vj[k] = t1;
                }
            }
        if (!changed) break;
    }
    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) { const float t = At[(size_t)i * m + k]; sd += (double)t * t; }
        W[i] = std::sqrt(sd);
    }
    for (int i = 0; i < n - 1; i++) {
        int j = i;
        for (int k = i + 1; k < n; k++)
            if (W[j] < W[k]) j = k;
        if (i != j) {
            std::swap(W[i], W[j]);
            for (int k = 0; k < m; k++) std::swap(At[(size_t)i * m + k], At[(size_t)j * m + k]);
            for (int k = 0; k < n; k++) std::swap(Vt[(size_t)i * n + k], Vt[(size_t)j * n + k]);
        }
    }
    for (int i = 0; i < n; i++) Wf[i] = (float)W[i];
    // U rows: At_i / w_i.  (Null singular values get a random completion in OpenCV; those rows are
    // dropped by the threshold below and cannot reach X.)
    for (int i = 0; i < n; i++) {
        const double sd = W[i];
        const float s = (float)(sd > minval ? 1 / sd : 0.);
        for (int k = 0; k < m; k++) At[(size_t)i * m + k] *= s;
    }
    // SVBkSbImpl_: x = sum_i v_i * (u_i . b) / w_i over w_i > threshold
    double threshold = 0;
    for (int i = 0; i < n; i++) threshold += Wf[i];
    threshold *= (double)(float)(FLT_EPSILON * 2);
    X.create(n, nb);
    X.setTo(0.f);
    for (int col = 0; col < nb; col++)
        for (int i = 0; i < n; i++) {
            double wi = Wf[i];
            if (std::abs(wi) <= threshold) continue;
            wi = 1 / wi;
            double s = 0;
            for (int j = 0; j < m; j++) s += At[(size_t)i * m + j] * B.data[(size_t)j * B.step + col];
            s *= wi;
            for (int j = 0; j < n; j++) {
                float& xj = X.data[(size_t)j * X.step + col];
                xj = (float)(xj + s * Vt[(size_t)i * n + j]);
            }
        }
    return true;
}
inline bool solve(const Mat& A, const MatExpr& B, Mat& X, int method) { return solve(A, B.eval(), X, method); }

// ---------------------------------------------------------------------------------------------
// cv::RNG / cv::randn (modules/core/src/rand.cpp): multiply-with-carry generator, Ziggurat normal
// deviates (randn_0_1_32f), then dst = src*stddev + mean in float (randnScale_).
// shim::randn_source lets the test harness substitute the next randn() fill (the depth
// prior of the first frame is an INPUT of the parity runs).
// ---------------------------------------------------------------------------------------------
struct RNG {
    uint64 state;
    RNG() : state(0xffffffff) {}
    explicit RNG(uint64 s) : state(s ? s : 0xffffffff) {}
};
inline RNG& theRNG() { static RNG r; return r; }
inline void setRNGSeed(int seed) { theRNG() = RNG((uint64)seed); }
namespace shim
{
struct RandnSource { const float* data = nullptr; size_t n = 0; };
inline RandnSource& randn_source() { static RandnSource s; return s; }
inline void randn_0_1_32f(float* arr, int len, uint64* state)
{
    const float r = 3.442620f;
    const float rng_flt = 2.3283064365386962890625e-10f;
    static unsigned kn[128];
    static float wn[128], fn[128];
    static bool initialized = false;
    uint64 temp = *state;
    auto next = [](uint64 x) { return (uint64)(unsigned)x * 4164903690U + (x >> 32); };
    if (!initialized) {
        const double m1 = 2147483648.0;
        double dn = 3.442619855899, tn = dn, vn = 9.91256303526217e-3;
        const double q = vn / std::exp(-.5 * dn * dn);
        kn[0] = (unsigned)((dn / q) * m1);
        kn[1] = 0;
        wn[0] = (float)(q / m1);
        wn[127] = (float)(dn / m1);
        fn[0] = 1.f;
        fn[127] = (float)std::exp(-.5 * dn * dn);
        for (int i = 126; i >= 1; i--) {
            dn = std::sqrt(-2. * std::log(vn / dn + std::exp(-.5 * dn * dn)));
            kn[i + 1] = (unsigned)((dn / tn) * m1);
            tn = dn;
            fn[i] = (float)std::exp(-.5 * dn * dn);
            wn[i] = (float)(dn / m1);
        }
        initialized = true;
    }
    for (int i = 0; i < len; i++) {
        float x, y;
        for (;;) {
            const int hz = (int)temp;
            temp = next(temp);
            const int iz = hz & 127;
            x = hz * wn[iz];
            if ((unsigned)std::abs(hz) < kn[iz]) break;
            if (iz == 0) {
                do {
                    x = (unsigned)temp * rng_flt; temp = next(temp);
                    y = (unsigned)temp * rng_flt; temp = next(temp);
                    x = (float)(-std::log(x + FLT_MIN) * 0.2904764);
                    y = (float)-std::log(y + FLT_MIN);
                } while (y + y < x * x);
                x = hz > 0 ? r + x : -r - x;
                break;
            }
            y = (unsigned)temp * rng_flt; temp = next(temp);
            if (fn[iz] + y * (fn[iz - 1] - fn[iz]) < std::exp(-.5 * x * x)) break;
        }
        arr[i] = x;
    }
    *state = temp;
}
} // namespace shim
inline void randn(Mat& dst, double mean, double stddev)
{
    shim::RandnSource& src = shim::randn_source();
    const size_t total = dst.total();
    if (src.data && src.n == total) {
        size_t k = 0;
        for (int y = 0; y < dst.rows; y++)
            for (int x = 0; x < dst.cols; x++) dst.data[(size_t)y * dst.step + x] = src.data[k++];
        src.data = nullptr; src.n = 0;
        return;
    }
    const float b = (float)mean, a = (float)stddev;
    std::vector<float> tmp(total);
    const size_t block = 1024;
    uint64 st = theRNG().state;
    for (size_t j = 0; j < total; j += block) {
        const int len = (int)std::min(block, total - j);
        shim::randn_0_1_32f(tmp.data() + j, len, &st);
    }
    theRNG().state = st;
    size_t k = 0;
    for (int y = 0; y < dst.rows; y++)
        for (int x = 0; x < dst.cols; x++) dst.data[(size_t)y * dst.step + x] = tmp[k++] * a + b;
}

// ---------------------------------------------------------------------------------------------
// highgui / display: no graphics stack on this box — no-ops
// ---------------------------------------------------------------------------------------------
inline void namedWindow(const std::string&, int = 0) {}
inline void resizeWindow(const std::string&, int, int) {}
inline void imshow(const std::string&, const Mat&) {}
inline int waitKey(int = 0) { return -1; }
inline void vconcat(const Mat& a, const Mat&, Mat& dst) { dst = a; }
inline void hconcat(const Mat& a, const Mat&, Mat& dst) { dst = a; }

} // namespace cv
#endif // DVO_SHIM_OPENCV_HPP
