// dvo/math/gaussian.hpp — math::Gaussian (include/math/gaussian.hpp:8-17, src/math/gaussian.cpp:12-50): the per-pixel depth
// hypothesis the reference's callers fuse by hand (test/update.cpp:101-106).  The arithmetic lives in libdvo_cuda.so
// (dvo_gaussian_fuse / dvo_gaussian_update, host code compiled without FMA contraction) so that it is bit-identical to what the
// mapping kernels do per pixel, whatever flags the caller compiles with.  update()'s reset draws from a process-wide
// default-seeded std::mt19937 like the reference's file-scope engine (gaussian.cpp:8-9).
#pragma once
#include "../../dvo_cuda.h"
#include "util.hpp"

namespace math
{
struct Gaussian {
    Gaussian(float depth, float sigma) : depth(depth), sigma(sigma) {}
    float depth;
    float sigma;

    // gaussian.cpp:12-31: fuse, or reset to (min(dist(engine), 4), 0.5) when the observation is too far; returns whether it fused
    bool update(float d, float s) { return dvo_gaussian_update(&depth, &sigma, d, s) == 1; }
    // gaussian.cpp:33-50: fuse, or leave the hypothesis untouched
    bool operator()(float d, float s) { return dvo_gaussian_fuse(&depth, &sigma, d, s) == 1; }
};
}  // namespace math
