// dvo/system/system.hpp — System::VisualOdometry (include/system/system.hpp:12-104).
//
// odometrize(gray) is the whole per-frame hot path (Frame ctor -> Tracker::track -> Frame::updateXi ->
// Mapper::estimate) and runs device-resident through dvo_vo_push: pyramid, one persistent tracking
// launch, keyframe decision, propagate | update, regularize, with no host round trip in between.  Each
// VisualOdometry owns its own dvo_ctx (keyframe pool + mt19937 replay position are per context).
// odometrizeUsingDepth (tracking only, system.hpp:77-93) uses Frame + Tracker like the reference.
#pragma once
#include "../map/mapper.hpp"
#include "../track/tracker.hpp"
#include "frame.hpp"

namespace System
{

class VisualOdometry
{
public:
    explicit VisualOdometry(const Mat1f& K) : K(K.clone()) {}

    // with an initial depth / sigma (system.hpp:24-32): the first keyframe is Frame(gray, depth, sigma, K, 4, 1)
    VisualOdometry(const Mat1f& gray, const Mat1f& depth, const Mat1f& sigma, const Mat1f& K) : K(K.clone())
    {
        std::shared_ptr<Frame> frame = std::make_shared<Frame>(gray, depth, sigma, K, 4, 1);
        m_history.setRefFrame(frame);
    }

    // Frame-0 prior of odometrize (Scene's cv::randn prior, frame.hpp:18-21), finest-level sized
    // (input size >> 2).  Optional: without it a std::mt19937 normal prior max(N(1.5, 0.5), 0.5) / 0.5 is drawn.
    void setInitialDepth(const Mat1f& depth, const Mat1f& sigma)
    {
        m_init_depth = depth.clone();
        m_init_sigma = sigma.clone();
    }

    // system.hpp:44-74.  Returns se3::exp(frame->m_xi) (identity for the first frame).
    Mat1f odometrize(const Mat1f& gray_image)
    {
        const bool first = (m_vo == nullptr);
        if (first) {
            // like the reference, odometrize tracks against its own keyframes only: a keyframe given to the
            // second ctor has Frame(., 4, 1) geometry and cannot be tracked by Frame(., 3, 2) frames
            dvo_params p = dvo::make_params(gray_image.cols, gray_image.rows, 3, 2, K);  // Frame(gray, K, 3, 2), system.hpp:47
            p.max_frames = 2;
            p.max_keyframes = 4096;  // FrameHistory is never pruned; pool entries are allocated on first use
            m_vo = std::make_shared<dvo::Context>(p);
        }
        dvo::Context& c = *m_vo;
        dvo::require(gray_image.rows == c.P.height && gray_image.cols == c.P.width, "frame size changed");
        Mat1f T(4, 4), keep;
        const float* d = nullptr;
        const float* s = nullptr;
        if (first) {
            const int w = c.levelCols(2), h = c.levelRows(2);
            if (m_init_depth.empty()) {
                m_init_depth = Mat1f(h, w);
                m_init_sigma = Mat1f(h, w, 0.5f);
                std::normal_distribution<float> dist(1.5f, 0.5f);
                for (int i = 0; i < w * h; i++) {
                    const float v = dist(Frame::priorEngine());
                    m_init_depth(i) = v < 0.5f ? 0.5f : v;
                }
            }
            dvo::require(m_init_depth.rows == h && m_init_depth.cols == w && m_init_sigma.rows == h && m_init_sigma.cols == w,
                         "initial depth/sigma must have the finest level's size");
            d = dvo::data(m_init_depth);
            s = dvo::data(m_init_sigma);
        }
        c.check(dvo_vo_push(c.h, dvo::pixels(gray_image, keep), d, s, dvo::data(T)));
        c.check(dvo_vo_error(c.h));
        return T;
    }

    // NOTE: tracking-only mode, system.hpp:77-93
    Mat1f odometrizeUsingDepth(const Mat1f& gray, const Mat1f& depth, const Mat1f& sigma)
    {
        std::shared_ptr<Frame> frame = std::make_shared<Frame>(gray, depth, sigma, K, 4, 1);
        if (m_ref_frame == nullptr) {
            m_ref_frame = frame;
            return math::se3::T();
        }
        Mat1f relative_xi = m_tracker.track(frame, m_ref_frame);
        frame->updateXi(relative_xi, m_ref_frame);
        m_ref_frame = frame;
        return math::se3::exp(relative_xi);
    }

    // ---- state of the last odometrize (what the reference prints / shows) ----
    struct State {
        Mat1f xi, relative_xi;
        bool was_keyframe;
        int n_keyframes, valid_updates, iterations;
    };
    State state() const
    {
        dvo::require(m_vo != nullptr, "no frame pushed");
        State st;
        st.xi = Mat1f(6, 1); st.relative_xi = Mat1f(6, 1);
        int kf = 0;
        m_vo->check(dvo_vo_state(m_vo->h, dvo::data(st.xi), dvo::data(st.relative_xi), &kf, &st.n_keyframes, &st.valid_updates, &st.iterations));
        st.was_keyframe = kf != 0;
        return st;
    }
    // depth / sigma / age / gray (dvo_image_kind) of the current keyframe at pyramid level `level`
    Mat1f keyframeImage(int kind, int level = 2) const
    {
        dvo::require(m_vo != nullptr, "no frame pushed");
        int slot = -1;
        m_vo->check(dvo_vo_ref_slot(m_vo->h, 0, &slot));
        Mat1f out(m_vo->levelRows(level), m_vo->levelCols(level));
        m_vo->check(dvo_frame_download(m_vo->h, slot, level, kind, dvo::data(out)));
        return out;
    }

private:
    const Mat1f K;

    Track::Tracker m_tracker;
    Map::Mapper m_mapper;
    FrameHistory m_history;
    std::shared_ptr<dvo::Context> m_vo;
    Mat1f m_init_depth, m_init_sigma;

    // NOTE: tracking-only mode
    std::shared_ptr<Frame> m_ref_frame = nullptr;
};
}  // namespace System
