// dvo/track/tracker.hpp — Track::Tracker (include/track/tracker.hpp:9-27, src/track/tracker.cpp:22-85).
// track() is ONE persistent kernel launch: all pyramid levels x <= 15 iterations, 6x6 solve, se3
// composition and the stop rules stay on the device.  No highgui window is opened (tracker.hpp:14-18);
// the 200 ms wall-clock budget (tracker.cpp:70) is treated as never reached.
#pragma once
#include "optimize.hpp"

#include <vector>

namespace Track
{
class Tracker
{
public:
    using pScene = std::shared_ptr<System::Scene>;
    using pFrame = std::shared_ptr<System::Frame>;
    Tracker() {}

    // returns the 6x1 twist of obj relative to ref
    Mat1f track(const pFrame obj_frame, const pFrame ref_frame)
    {
        dvo::require(obj_frame && ref_frame, "null frame");
        dvo::require(obj_frame->slot()->ctx == ref_frame->slot()->ctx, "frames must share one geometry");
        dvo::Context& c = obj_frame->ctx();
        float xi[6];
        const int cap = c.P.levels * c.P.max_iterations;
        m_log.resize(cap);
        int n = 0;
        c.check(dvo_track(c.h, obj_frame->slot()->id, ref_frame->slot()->id, xi, m_log.data(), cap, &n, -1.0, -1.0f));
        m_log.resize(n < cap ? n : cap);
        return dvo::make_xi(xi);
    }

    // what Tracker::track prints per iteration (tracker.cpp:56-61), for the last call
    const std::vector<dvo_iter_log>& iterations() const { return m_log; }

private:
    std::vector<dvo_iter_log> m_log;
};
}  // namespace Track
