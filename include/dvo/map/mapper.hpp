// dvo/map/mapper.hpp — Map::Mapper (include/map/mapper.hpp:16-42, src/map/mapper.cpp:16-144) over
// device-resident frames: propagate = dvo_map_propagate, update = dvo_map_update (epipolar search +
// Gaussian fusion + mt19937-ordered resets), regularize = dvo_map_regularize.
#pragma once
#include "../system/frame.hpp"
#include "implement.hpp"

#include <cmath>

namespace Map
{
using Frame = System::Frame;
using FrameHistory = System::FrameHistory;
using pFrame = std::shared_ptr<System::Frame>;
using pScene = std::shared_ptr<System::Scene>;

class Mapper
{
public:
    Mapper() {}

    // mapper.cpp:16-33
    void estimate(FrameHistory& frame_history, pFrame frame)
    {
        if (needNewFrame(frame)) {
            propagate(frame);
            frame_history.setRefFrame(frame);
        } else {
            update(frame_history, frame);
        }
        regularize(frame_history.getRefFrame());
    }

    // mapper.cpp:45-60: ||t_rel|| > 0.02 (cv::norm: double accumulation) or 6 frames since the keyframe
    bool needNewFrame(const pFrame frame)
    {
        const dvo_params& P = frame->ctx().P;
        double s = 0;
        for (int i = 0; i < 3; i++) s += (double)frame->m_relative_xi(i) * (double)frame->m_relative_xi(i);
        if (std::sqrt(s) > (double)P.min_movement) return true;
        if (frame->id - frame->m_ref_frame->id >= P.max_forward) return true;
        return false;
    }

    // mapper.cpp:62-74: the keyframe's maps warped into `frame` (Implement::propagate + updateDepthSigmaAge)
    void propagate(pFrame frame)
    {
        const pFrame& ref = frame->m_ref_frame;
        dvo::require(ref != nullptr, "frame has no reference frame");
        dvo::Context& c = frame->ctx();
        float tw[6];
        for (int i = 0; i < 6; i++) tw[i] = frame->m_relative_xi(i);
        c.check(dvo_map_propagate(c.h, ref->slot()->id, frame->slot()->id, tw));
    }

    // mapper.cpp:76-137: stereo update of the keyframe's depth/sigma/age from `frame`
    void update(const FrameHistory& frame_history, pFrame frame)
    {
        const pFrame& ref = frame->m_ref_frame;
        dvo::require(ref != nullptr, "frame has no reference frame");
        dvo::Context& c = frame->ctx();
        std::vector<int> hist(frame_history.size());
        for (int a = 0; a < frame_history.size(); a++) hist[a] = frame_history[a]->slot()->id;  // by age, newest first
        c.check(dvo_map_update(c.h, frame->slot()->id, ref->slot()->id, hist.data(), (int)hist.size(), &m_valid_update, nullptr));
    }

    // mapper.cpp:139-144
    void regularize(pFrame frame)
    {
        dvo::Context& c = frame->ctx();
        c.check(dvo_map_regularize(c.h, frame->slot()->id));
    }

    int lastValidUpdate() const { return m_valid_update; }  // "valid update" count the reference prints (mapper.cpp:136)

private:
    int m_valid_update = 0;
};
}  // namespace Map
