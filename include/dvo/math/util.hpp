// dvo/math/util.hpp — the scalar helpers of include/math/util.hpp:6-55 the reference's callers use (test/track.cpp:55 math::testXi,
// the validity predicates everywhere).  Host code; nothing here rounds.
#pragma once
#include "../mat.hpp"

#include <cmath>

namespace math
{
#ifndef DVO_MATH_CONSTANTS
#define DVO_MATH_CONSTANTS
constexpr float EPSILON = 1e-6f;   // include/math/util.hpp:6
constexpr float INVALID = -2.0f;   // include/math/util.hpp:7
#endif

inline bool isValid(float num) { return INVALID < num; }      // util.hpp:9
inline bool isInvalid(float num) { return num <= INVALID; }   // util.hpp:10
inline bool isEpsilon(float num) { return std::abs(num) < EPSILON; }  // util.hpp:29-32

template <typename T>
inline T square(T num) { return num * num; }  // util.hpp:12-16

// util.hpp:34-44: false when the 6x1 twist holds a NaN (Tracker::track keeps the previous pose then, tracker.cpp:47-51)
inline bool testXi(const dvo::Mat1f& xi)
{
    dvo::require(xi.total() == 6, "xi must have 6 elements");
    for (int i = 0; i < 6; i++)
        if (std::isnan(xi(i))) return false;
    return true;
}

// util.hpp:46-55
inline bool inRange(const dvo::Point2i& pt, int width, int height)
{
    if (pt.x < 0 || width <= pt.x) return false;
    if (pt.y < 0 || height <= pt.y) return false;
    return true;
}
}  // namespace math
