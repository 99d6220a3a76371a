// dvo/map/implement.hpp — Map::Implement::{update, propagate, regularize} (include/map/implement.hpp:10-28,
// src/map/implement.cpp:156-256) on the device.  These take raw images like the reference (test/update.cpp,
// test/propagate.cpp, test/regularize.cpp call them that way): a one-level context of the images' size is
// used and the results are downloaded.  Mapper (mapper.hpp) works on resident frames instead.
#pragma once
#include "../device.hpp"

#include <tuple>
#include <utility>

namespace Map
{
namespace Implement
{
using dvo::Mat1f;

// Batched form of `update` below: n points of obj_gray searched in ref_gray.  Uploads the two images once.
inline void update(const Mat1f& obj_gray, const Mat1f& ref_gray, const Mat1f& relative_xi, const Mat1f& K, int n, const int* px,
                   const int* py, const float* depth, const float* sigma, float* out_depth, float* out_sigma)
{
    dvo::require(obj_gray.rows == ref_gray.rows && obj_gray.cols == ref_gray.cols, "obj/ref size mismatch");
    dvo::require(relative_xi.total() == 6, "relative_xi must have 6 elements");
    auto ctx = dvo::Device::get(obj_gray.cols, obj_gray.rows, 1, 0, K);
    dvo::Slot obj(ctx, obj_gray), born(ctx, ref_gray);
    float tw[6];
    for (int i = 0; i < 6; i++) tw[i] = relative_xi(i);
    ctx->check(dvo_map_update_points(ctx->h, obj.id, born.id, tw, n, px, py, depth, sigma, out_depth, out_sigma));
}

// return: depth, sigma; (-1, -1) if failed (implement.cpp:182-214).  The gradient images are accepted for
// signature compatibility; the device recomputes Convert::gradiate(ref_gray) where it samples.
inline std::pair<float, float> update(const Mat1f& obj_gray, const Mat1f& ref_gray, const Mat1f& /*ref_gradx*/, const Mat1f& /*ref_grady*/,
                                      const Mat1f& relative_xi, const Mat1f& K, const dvo::Point2i& x_i, float depth, float sigma)
{
    float od = -1.f, os = -1.f;
    update(obj_gray, ref_gray, relative_xi, K, 1, &x_i.x, &x_i.y, &depth, &sigma, &od, &os);
    return {od, os};
}

// implement.cpp:217-256: forward scatter of (depth, sigma, age) by xi; last source in row-major order wins
inline std::tuple<Mat1f, Mat1f, Mat1f> propagate(const Mat1f& ref_depth, const Mat1f& ref_sigma, const Mat1f& ref_age, const Mat1f& xi,
                                                  const Mat1f& K)
{
    const int h = ref_depth.rows, w = ref_depth.cols;
    dvo::require(ref_sigma.rows == h && ref_sigma.cols == w && ref_age.rows == h && ref_age.cols == w, "map size mismatch");
    dvo::require(xi.total() == 6, "xi must have 6 elements");
    auto ctx = dvo::Device::get(w, h, 1, 0, K);
    const Mat1f blank = Mat1f::zeros(h, w);
    dvo::Slot ref(ctx, blank), dst(ctx, blank);
    Mat1f k0, k1, k2;
    ctx->check(dvo_frame_set_depth(ctx->h, ref.id, dvo::pixels(ref_depth, k0), dvo::pixels(ref_sigma, k1), dvo::pixels(ref_age, k2), 0));
    float tw[6];
    for (int i = 0; i < 6; i++) tw[i] = xi(i);
    ctx->check(dvo_map_propagate(ctx->h, ref.id, dst.id, tw));
    return {dst.download(0, DVO_IMG_DEPTH), dst.download(0, DVO_IMG_SIGMA), dst.download(0, DVO_IMG_AGE)};
}

// implement.cpp:156-180: 4-neighbour Gaussian fusion from the input maps, clamped to 6.0; sigma is not returned
inline Mat1f regularize(const Mat1f& depth, const Mat1f& sigma)
{
    const int h = depth.rows, w = depth.cols;
    dvo::require(sigma.rows == h && sigma.cols == w, "map size mismatch");
    Mat1f K = Mat1f::zeros(3, 3);
    K(0, 0) = K(1, 1) = K(2, 2) = 1.0f;
    auto ctx = dvo::Device::get(w, h, 1, 0, K);
    dvo::Slot s(ctx, Mat1f::zeros(h, w));
    Mat1f k0, k1;
    ctx->check(dvo_frame_set_depth(ctx->h, s.id, dvo::pixels(depth, k0), dvo::pixels(sigma, k1), nullptr, 0));
    ctx->check(dvo_map_regularize(ctx->h, s.id));
    return s.download(0, DVO_IMG_DEPTH);
}

}  // namespace Implement
}  // namespace Map
