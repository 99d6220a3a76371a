// dvo/track/optimize.hpp — Track::{Stuff, Outcome, optimize} (include/track/optimize.hpp:10-65,
// src/track/optimize.cpp:10-99).  One call = one fused Gauss-Newton step on the device (dvo_gn_step):
// warp + bilinear intensity/gradient + residual + weight + 1x6 Jacobian + normal-equation reduction +
// 6x6 solve.  `Stuff::levels` is the pyramid INDEX (the reference's misnomer, tracker.cpp:40).
#pragma once
#include "../core/transform.hpp"
#include "../system/frame.hpp"

namespace Track
{
using dvo::Mat1f;

struct Stuff {
    Stuff(std::shared_ptr<System::Scene> obj, std::shared_ptr<System::Scene> ref, const Mat1f& xi, const int levels)
        : obj(obj), ref(ref), xi(xi.clone()), K(ref->K()), cols(ref->cols), rows(ref->rows), levels(levels)
    {
        dvo::require(obj->slot()->ctx == ref->slot()->ctx, "obj and ref scenes must share one geometry");
        dvo::require(obj->level() == ref->level(), "obj and ref scenes must be the same pyramid level");
    }

    void update(const Mat1f _xi) { xi = _xi.clone(); }  // the warped image is recomputed inside the fused step

    // the images the reference's Stuff aliases, as snapshots
    Mat1f obj_gray() const { return obj->gray(); }
    Mat1f ref_gray() const { return ref->gray(); }
    Mat1f ref_depth() const { return ref->depth(); }
    Mat1f ref_sigma() const { return ref->sigma(); }
    Mat1f grad_x() const { return ref->gradX(); }
    Mat1f grad_y() const { return ref->gradY(); }
    Mat1f warped_gray() const
    {
        float tw[6];
        for (int i = 0; i < 6; i++) tw[i] = xi(i);
        Mat1f out(rows, cols);
        dvo::Context& c = *ref->slot()->ctx;
        c.check(dvo_warp_image(c.h, ref->slot()->id, ref->level(), tw, dvo::data(out)));
        return out;
    }

    const std::shared_ptr<System::Scene> obj;
    const std::shared_ptr<System::Scene> ref;
    Mat1f xi;
    const Mat1f K;
    const int cols;
    const int rows;
    const int levels;
};

struct Outcome {
    Outcome(const Mat1f& xi_update, float residual, int valid_pixels) : xi_update(xi_update), residual(residual), valid_pixels(valid_pixels) {}
    const Mat1f xi_update;
    const float residual;   // sum r^2 / valid, -1 when no pixel contributed (optimize.cpp:92-93)
    const int valid_pixels;
};

inline Outcome optimize(const Stuff& stuff)
{
    dvo::Context& c = *stuff.ref->slot()->ctx;
    dvo::require(stuff.levels == stuff.ref->level(), "Stuff::levels must be the scenes' pyramid index");
    float tw[6], upd[6], residual = -1.0f;
    int valid = 0;
    for (int i = 0; i < 6; i++) tw[i] = stuff.xi(i);
    c.check(dvo_gn_step(c.h, stuff.obj->slot()->id, stuff.ref->slot()->id, stuff.ref->level(), tw, upd, &residual, &valid, nullptr));
    return Outcome(dvo::make_xi(upd), residual, valid);
}

}  // namespace Track
