// dvo/math/se3.hpp — math::se3 of the reference (include/math/se3.hpp, src/math/se3.cpp:76-131)
// forwarded to the device code the tracker itself uses (dvo_se3_* in include/dvo_cuda.h), so host-side
// pose bookkeeping (Frame::updateXi, Mapper::needNewFrame) agrees with the kernels.
#pragma once
#include "../device.hpp"

#include <array>
#include <cmath>

#include "util.hpp"

namespace math
{
namespace se3
{
using dvo::Mat1f;

inline Mat1f xi() { return Mat1f::zeros(6, 1); }  // se3.hpp: zero twist
inline Mat1f xi(std::initializer_list<float> v)
{
    dvo::require(v.size() == 6, "xi needs 6 values");
    Mat1f out(6, 1);
    int i = 0;
    for (float f : v) out(i++) = f;
    return out;
}
inline Mat1f xi(const std::array<float, 6>& v)  // test/warp.cpp:40
{
    Mat1f out(6, 1);
    for (int i = 0; i < 6; i++) out(i) = v[i];
    return out;
}
inline Mat1f T()  // 4x4 identity
{
    Mat1f I = Mat1f::zeros(4, 4);
    for (int i = 0; i < 4; i++) I(i, i) = 1.0f;
    return I;
}

// se3::exp (src/math/se3.cpp:76-99): 6x1 twist -> 4x4
inline Mat1f exp(const Mat1f& twist)
{
    dvo::require(twist.total() == 6, "twist must have 6 elements");
    auto ctx = dvo::Device::scratch();
    float in[6];
    for (int i = 0; i < 6; i++) in[i] = twist(i);
    Mat1f out(4, 4);
    ctx->check(dvo_se3_exp(ctx->h, in, dvo::data(out), 1));
    return out;
}
// se3::log (src/math/se3.cpp:102-124): 4x4 -> 6x1
inline Mat1f log(const Mat1f& pose)
{
    dvo::require(pose.rows == 4 && pose.cols == 4, "pose must be 4x4");
    auto ctx = dvo::Device::scratch();
    float in[16];
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) in[r * 4 + c] = pose(r, c);
    Mat1f out(6, 1);
    ctx->check(dvo_se3_log(ctx->h, in, dvo::data(out), 1));
    return out;
}
// se3::concatenate (src/math/se3.cpp:127-131): log(exp(a) * exp(b))
inline Mat1f concatenate(const Mat1f& a, const Mat1f& b)
{
    dvo::require(a.total() == 6 && b.total() == 6, "twists must have 6 elements");
    auto ctx = dvo::Device::scratch();
    float x[6], y[6];
    for (int i = 0; i < 6; i++) { x[i] = a(i); y[i] = b(i); }
    Mat1f out(6, 1);
    ctx->check(dvo_se3_concatenate(ctx->h, x, y, dvo::data(out), 1));
    return out;
}
}  // namespace se3
}  // namespace math
