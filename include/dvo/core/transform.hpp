// dvo/core/transform.hpp — namespace Transform of the reference (include/core/transform.hpp:8-28, src/core/transform.cpp:7-78).
// The per-image functions (warpImage, mapDepthtoGray) run on the device and take raw images like the reference
// (test/warp.cpp:42, src/core/loader.cpp:116-119).  The scalar ones (transform, project, backProject, warp) are host arithmetic
// inside libdvo_cuda.so in the reference's fp32 order (dvo_transform_point, ...), so a caller's compiler flags cannot change a
// rounding; a 6x1 twist is exponentiated by the device code the tracker itself uses (math::se3::exp), as transform.cpp:13-14 does.
#pragma once
#include "../device.hpp"
#include "../math/se3.hpp"

#include <utility>

namespace Transform
{
using dvo::Mat1f;

struct Point2f {
    float x = 0, y = 0;
    Point2f() = default;
    Point2f(float x_, float y_) : x(x_), y(y_) {}
    Point2f(const dvo::Point2i& p) : x((float)p.x), y((float)p.y) {}  // cv::Point2i -> cv::Point2f
};
struct Point3f {
    float x = 0, y = 0, z = 0;
    Point3f() = default;
    Point3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};

// transform.cpp:7-18: T (4x4) or xi (6x1), x -> R x + t
inline Point3f transform(const Mat1f& T, const Point3f& x)
{
    if (T.rows == 6 && T.cols == 1) return transform(math::se3::exp(T), x);
    dvo::require(T.rows == 4 && T.cols == 4, "invalid transform");  // the reference aborts here
    float t16[16], in[3] = {x.x, x.y, x.z}, out[3];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) t16[r * 4 + c] = T(r, c);
    dvo_transform_point(t16, in, out);
    return Point3f(out[0], out[1], out[2]);
}
// transform.cpp:20-23: x_c -> x_i
inline Point2f project(const Mat1f& K, const Point3f& point)
{
    float k9[9], in[3] = {point.x, point.y, point.z}, out[2];
    for (int i = 0; i < 9; i++) k9[i] = K(i);
    dvo_project_point(k9, in, out);
    return Point2f(out[0], out[1]);
}
// transform.cpp:25-28: x_i -> x_c
inline Point3f backProject(const Mat1f& K, const Point2f& point, float depth)
{
    float k9[9], in[2] = {point.x, point.y}, out[3];
    for (int i = 0; i < 9; i++) k9[i] = K(i);
    dvo_back_project_point(k9, in, depth, out);
    return Point3f(out[0], out[1], out[2]);
}
// transform.cpp:30-33
inline Point2f warp(const Mat1f& xi, const Point2f& x_i, const float depth, const Mat1f& K)
{
    return project(K, transform(xi, backProject(K, x_i, depth)));
}

// transform.cpp:53-78: <mapped, sigma> - the colour image registered onto the depth camera (sigma 0.1 where mapped, 1.0 elsewhere)
inline std::pair<Mat1f, Mat1f> mapDepthtoGray(const Mat1f& depth_image, const Mat1f& gray_image, const Mat1f& rgb_K, const Mat1f& depth_K,
                                              const Mat1f& invT)
{
    dvo::require(rgb_K.total() == 9 && depth_K.total() == 9 && invT.rows == 4 && invT.cols == 4, "mapDepthtoGray: bad calibration");
    float rk[9], dk[9], t16[16];
    for (int i = 0; i < 9; i++) { rk[i] = rgb_K(i); dk[i] = depth_K(i); }
    for (int i = 0; i < 16; i++) t16[i] = invT(i);
    Mat1f mapped(depth_image.rows, depth_image.cols), sigma(depth_image.rows, depth_image.cols), k1, k2;
    const int rc = dvo_map_depth_to_gray(dvo::pixels(depth_image, k1), depth_image.cols, depth_image.rows, dvo::pixels(gray_image, k2),
                                         gray_image.cols, gray_image.rows, rk, dk, t16, dvo::data(mapped), dvo::data(sigma));
    if (rc) throw dvo::Error(rc, "dvo_map_depth_to_gray failed (no CUDA device?)");
    return {mapped, sigma};
}

// out(x) = INVALID if |depth(x)| < 1e-6 else getSubpixel(gray, warp(-xi, x, depth(x), K))
inline Mat1f warpImage(const Mat1f& xi, const Mat1f& gray_image, const Mat1f& depth_image, const Mat1f& K)
{
    dvo::require(xi.total() == 6, "xi must have 6 elements");
    dvo::require(gray_image.rows == depth_image.rows && gray_image.cols == depth_image.cols, "gray/depth size mismatch");
    auto ctx = dvo::Device::get(gray_image.cols, gray_image.rows, 1, 0, K);
    dvo::Slot slot(ctx, gray_image);
    Mat1f keep;
    ctx->check(dvo_frame_set_depth(ctx->h, slot.id, dvo::pixels(depth_image, keep), nullptr, nullptr, 0));
    float tw[6];
    for (int i = 0; i < 6; i++) tw[i] = xi(i);
    Mat1f out(gray_image.rows, gray_image.cols);
    ctx->check(dvo_warp_image(ctx->h, slot.id, 0, tw, dvo::data(out)));
    return out;
}
}  // namespace Transform
