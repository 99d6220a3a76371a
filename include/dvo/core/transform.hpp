// dvo/core/transform.hpp — Transform::warpImage (include/core/transform.hpp:28, src/core/transform.cpp:35-51)
// on the device.  Takes raw images like the reference (test/warp.cpp:42 calls it that way): a one-level
// context of the images' size is used, the result is downloaded.
#pragma once
#include "../device.hpp"

namespace Transform
{
using dvo::Mat1f;

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
