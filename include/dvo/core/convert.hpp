// dvo/core/convert.hpp — Convert::{gradiate, cullImage, cullIntrinsic} (include/core/convert.hpp:31-40, src/core/convert.cpp:7-29,
// 41-75) with the reference's signatures on raw images (test/update.cpp:55-56, test/warp.cpp:30-32).  The per-image functions run
// on the device: a one-level context of the image's size builds the frame (Convert::cullImage is what Frame's constructor does)
// and the requested plane is downloaded - the same kernels every Frame goes through.  Convert::inversePose lives in core/draw.hpp.
#pragma once
#include "../device.hpp"

namespace Convert
{
using dvo::Mat1f;

namespace detail
{
inline Mat1f identityK()
{
    Mat1f K = Mat1f::zeros(3, 3);
    K(0, 0) = K(1, 1) = K(2, 2) = 1.0f;
    return K;
}
}  // namespace detail

// convert.cpp:7-20: keep the top-left pixel of every 2^times block (through Convert::getPixel: values <= INVALID become INVALID)
inline Mat1f cullImage(const Mat1f& src_image, int times = 1)
{
    if (times == 0) return src_image;  // the reference returns a shallow copy of the header
    dvo::require(times > 0 && (src_image.cols >> times) > 0 && (src_image.rows >> times) > 0, "cullImage: image too small");
    auto ctx = dvo::Device::get(src_image.cols, src_image.rows, 1, times, detail::identityK());
    dvo::Slot slot(ctx, src_image);
    return slot.download(0, DVO_IMG_GRAY);
}

// convert.cpp:22-29: K / 2^times with K(2,2) = 1 (`/ s` is a multiplication by 1.0 / s in OpenCV's MatExpr; exact for powers of two)
inline Mat1f cullIntrinsic(const Mat1f& intrinsic, int times = 1)
{
    dvo::require(intrinsic.rows == 3 && intrinsic.cols == 3, "K must be 3x3");
    if (times == 0) return intrinsic;
    Mat1f K(3, 3);
    const float s = (float)(1.0 / (double)(1 << times));
    for (int i = 0; i < 9; i++) K(i) = intrinsic(i) * s;
    K(2, 2) = 1.0f;
    return K;
}

// convert.cpp:41-75: I(x+1) - I(x-1) (no 1/2), INVALID on the border column / row and next to an invalid tap
inline Mat1f gradiate(const Mat1f& gray_image, bool x)
{
    auto ctx = dvo::Device::get(gray_image.cols, gray_image.rows, 1, 0, detail::identityK());
    dvo::Slot slot(ctx, gray_image);
    return slot.download(0, x ? DVO_IMG_GRADX : DVO_IMG_GRADY);
}
}  // namespace Convert
