// dvo/mat.hpp — the image type of the drop-in C++ layer.
//
// The reference passes every image as a ref-counted cv::Mat1f header.  The classes in
// include/dvo/ are written against the small subset of that API listed below, so they compile
//   * stand-alone, with the owned dvo::Mat1f defined here (default), or
//   * inside an OpenCV project: define DVO_WITH_OPENCV before including and dvo::Mat1f IS cv::Mat1f
//     (row-major continuous float32), so the reference's call sites compile unchanged.
// Subset used: Mat1f(rows, cols), rows, cols, empty(), clone(), isContinuous(), ptr<float>(r),
// operator()(r, c), operator()(i), static zeros/ones(rows, cols), total().
#pragma once
#include <cstddef>
#include <cstring>
#include <memory>
#include <stdexcept>

#ifdef DVO_WITH_OPENCV
#include <opencv2/opencv.hpp>
namespace dvo {
using Mat1f = cv::Mat1f;
using Point2i = cv::Point2i;
inline float* data(Mat1f& m) { return reinterpret_cast<float*>(m.data); }
inline const float* data(const Mat1f& m) { return reinterpret_cast<const float*>(m.data); }
}  // namespace dvo
#else
namespace dvo {

struct Point2i {
    int x = 0, y = 0;
    Point2i() = default;
    Point2i(int x_, int y_) : x(x_), y(y_) {}
};

// Shallow-copy, ref-counted, row-major continuous float32 image (the cv::Mat1f semantics the
// reference relies on: copies of a header share pixels, clone() does not).
class Mat1f
{
public:
    int rows = 0, cols = 0;

    Mat1f() = default;
    Mat1f(int rows_, int cols_) : rows(rows_), cols(cols_), m_data(alloc((size_t)rows_ * cols_)) {}
    Mat1f(int rows_, int cols_, float value) : Mat1f(rows_, cols_)
    {
        for (size_t i = 0, n = total(); i < n; i++) m_data.get()[i] = value;
    }
    // copies `src` (rows*cols floats)
    Mat1f(int rows_, int cols_, const float* src) : Mat1f(rows_, cols_) { std::memcpy(m_data.get(), src, sizeof(float) * total()); }

    static Mat1f zeros(int rows, int cols) { return Mat1f(rows, cols, 0.0f); }
    static Mat1f ones(int rows, int cols) { return Mat1f(rows, cols, 1.0f); }

    bool empty() const { return rows == 0 || cols == 0 || !m_data; }
    bool isContinuous() const { return true; }
    size_t total() const { return (size_t)rows * cols; }
    Mat1f clone() const { return empty() ? Mat1f() : Mat1f(rows, cols, m_data.get()); }

    template <typename T = float>
    T* ptr(int r = 0) { return reinterpret_cast<T*>(m_data.get() + (size_t)r * cols); }
    template <typename T = float>
    const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(m_data.get() + (size_t)r * cols); }

    float& operator()(int r, int c) { return m_data.get()[(size_t)r * cols + c]; }
    const float& operator()(int r, int c) const { return m_data.get()[(size_t)r * cols + c]; }
    float& operator()(int i) { return m_data.get()[i]; }
    const float& operator()(int i) const { return m_data.get()[i]; }
    float& operator()(const Point2i& p) { return (*this)(p.y, p.x); }
    const float& operator()(const Point2i& p) const { return (*this)(p.y, p.x); }

private:
    static std::shared_ptr<float> alloc(size_t n) { return std::shared_ptr<float>(new float[n ? n : 1], std::default_delete<float[]>()); }
    std::shared_ptr<float> m_data;
};

inline float* data(Mat1f& m) { return m.ptr<float>(0); }
inline const float* data(const Mat1f& m) { return m.ptr<float>(0); }

}  // namespace dvo
#endif

namespace dvo {

// 6x1 twist / 3x3 intrinsics / 4x4 pose helpers over Mat1f
inline Mat1f make_xi(const float* v = nullptr)
{
    Mat1f xi = Mat1f::zeros(6, 1);
    if (v)
        for (int i = 0; i < 6; i++) xi(i) = v[i];
    return xi;
}
inline void require(bool ok, const char* what)
{
    if (!ok) throw std::invalid_argument(what);
}
// continuous float pointer of an image (clones a non-continuous cv::Mat view)
inline const float* pixels(const Mat1f& m, Mat1f& keep)
{
    if (m.isContinuous()) return data(m);
    keep = m.clone();
    return data(keep);
}

}  // namespace dvo
