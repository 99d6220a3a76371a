// dvo/system/frame.hpp — System::{Scene, Frame, FrameHistory} with the reference's names and
// signatures (include/system/frame.hpp:9-188, src/system/frame.cpp), backed by a device frame slot.
//
// What differs from the reference, by design:
//  * Pixels live in HBM.  A Frame owns one slot of the geometry's dvo_ctx holding the finest level only;
//    coarser levels and gradients are views computed where sampled.  Scene accessors
//    (gray(), depth(), sigma(), gradX(), gradY()) therefore return SNAPSHOTS downloaded from the device,
//    not writable aliases; writes go through Frame::updateDepth* as in src/system/frame.cpp:39-61.
//  * The 4-argument Frame ctor's random prior (cv::randn(1.5, 0.5) from OpenCV's global RNG,
//    frame.hpp:18-21) comes from a std::mt19937-driven normal distribution; supply depth/sigma through
//    the 6-argument ctor (as System::VisualOdometry's second ctor does) when a specific prior matters.
#pragma once
#include "../device.hpp"
#include "../math/se3.hpp"

#include <iostream>
#include <random>

namespace System
{
using dvo::Mat1f;

class Frame;

class Scene
{
public:
    Scene(std::shared_ptr<dvo::Slot> slot, int level)
        : cols(slot->ctx->levelCols(level)), rows(slot->ctx->levelRows(level)), m_slot(std::move(slot)), m_level(level) {}

    const int cols;
    const int rows;

    Mat1f gray() const { return m_slot->download(m_level, DVO_IMG_GRAY); }
    Mat1f depth() const { return m_slot->download(m_level, DVO_IMG_DEPTH); }
    Mat1f sigma() const { return m_slot->download(m_level, DVO_IMG_SIGMA); }
    Mat1f gradX() const { return m_slot->download(m_level, DVO_IMG_GRADX); }  // Convert::gradiate(gray, true)
    Mat1f gradY() const { return m_slot->download(m_level, DVO_IMG_GRADY); }  // Convert::gradiate(gray, false)
    Mat1f K() const { return m_slot->ctx->levelK(m_level); }                  // Convert::cullIntrinsic

    const std::shared_ptr<dvo::Slot>& slot() const { return m_slot; }
    int level() const { return m_level; }

private:
    std::shared_ptr<dvo::Slot> m_slot;
    int m_level;
};

class Frame
{
public:
    static int& latestId()
    {
        static int latest_id = -1;  // frame.cpp:5
        return latest_id;
    }

    const int id;
    const int cols;
    const int rows;
    const int levels;
    const int culls;

    std::vector<std::shared_ptr<Scene>> m_scenes;

    Mat1f m_xi;
    Mat1f m_relative_xi;
    std::shared_ptr<Frame> m_ref_frame;

    // Frame(gray, depth, sigma, K, levels, culls), frame.hpp:91-106: depth/sigma are input-sized and culled
    Frame(const Mat1f& gray_image, const Mat1f& depth_image, const Mat1f& sigma_image, const Mat1f& K, int levels, int culls)
        : id(++latestId()), cols(gray_image.cols), rows(gray_image.rows), levels(levels), culls(culls),
          m_xi(math::se3::xi()), m_relative_xi(math::se3::xi()), m_ref_frame(nullptr)
    {
        init(gray_image, K);
        dvo::require(depth_image.rows == rows && depth_image.cols == cols && sigma_image.rows == rows && sigma_image.cols == cols,
                     "depth/sigma must have the gray image's size");
        Mat1f k0, k1;
        ctx().check(dvo_frame_set_depth(ctx().h, m_slot->id, dvo::pixels(depth_image, k0), dvo::pixels(sigma_image, k1), nullptr, 1));
    }

    // Frame(gray, K, levels, culls), frame.hpp:108-117: random prior max(N(1.5, 0.5), 0.5), sigma 0.5
    Frame(const Mat1f& gray_image, const Mat1f& K, int levels, int culls)
        : id(++latestId()), cols(gray_image.cols), rows(gray_image.rows), levels(levels), culls(culls),
          m_xi(math::se3::xi()), m_relative_xi(math::se3::xi()), m_ref_frame(nullptr)
    {
        init(gray_image, K);
        const int w = top()->cols, h = top()->rows;
        Mat1f depth(h, w), sigma(h, w, 0.5f);
        std::normal_distribution<float> dist(1.5f, 0.5f);
        for (int i = 0; i < w * h; i++) {
            const float d = dist(priorEngine());
            depth(i) = d < 0.5f ? 0.5f : d;
        }
        ctx().check(dvo_frame_set_depth(ctx().h, m_slot->id, dvo::data(depth), dvo::data(sigma), nullptr, 0));
    }

    static std::mt19937& priorEngine()
    {
        static std::mt19937 e(0x0D50);
        return e;
    }

    std::shared_ptr<Scene> at(int level) { return m_scenes.at(level); }
    std::shared_ptr<Scene> top() { return m_scenes.at(levels - 1); }
    std::shared_ptr<const Scene> top() const { return m_scenes.at(levels - 1); }

    Mat1f gray() const { return top()->gray(); }
    Mat1f depth() const { return top()->depth(); }
    Mat1f sigma() const { return top()->sigma(); }
    Mat1f gradX() const { return top()->gradX(); }
    Mat1f gradY() const { return top()->gradY(); }
    Mat1f K() const { return top()->K(); }
    Mat1f age() const { return m_slot->download(levels - 1, DVO_IMG_AGE); }

    // frame.cpp:7-14
    void updateXi(const Mat1f& relative_xi, std::shared_ptr<Frame> ref_frame)
    {
        if (ref_frame == nullptr) return;
        m_relative_xi = relative_xi.clone();
        m_ref_frame = ref_frame;
        m_xi = math::se3::concatenate(ref_frame->m_xi, relative_xi);
        pushXi();
    }
    // frame.cpp:39-61: finest-level sized arrays; the coarser levels are views, so "re-culling" is implicit
    void updateDepthSigmaAge(const Mat1f& depth_image, const Mat1f& sigma_image, const Mat1f& age_image)
    {
        setTop(&depth_image, &sigma_image, &age_image);
    }
    void updateDepthSigma(const Mat1f& depth_image, const Mat1f& sigma_image) { setTop(&depth_image, &sigma_image, nullptr); }
    void updateDepth(const Mat1f& depth_image) { setTop(&depth_image, nullptr, nullptr); }

    // device handles for Track:: / Map::
    dvo::Context& ctx() const { return *m_slot->ctx; }
    const std::shared_ptr<dvo::Slot>& slot() const { return m_slot; }

private:
    void init(const Mat1f& gray_image, const Mat1f& K)
    {
        auto c = dvo::Device::get(cols, rows, levels, culls, K);
        m_slot = std::make_shared<dvo::Slot>(c, gray_image);
        for (int i = 0; i < levels; i++) m_scenes.push_back(std::make_shared<Scene>(m_slot, i));  // at(0) coarsest
        pushXi();
    }
    void pushXi()
    {
        float a[6], b[6];
        for (int i = 0; i < 6; i++) { a[i] = m_xi(i); b[i] = m_relative_xi(i); }
        ctx().check(dvo_frame_set_xi(ctx().h, m_slot->id, a, b));
    }
    void setTop(const Mat1f* d, const Mat1f* s, const Mat1f* a)
    {
        const int w = m_scenes.back()->cols, h = m_scenes.back()->rows;
        Mat1f k0, k1, k2;
        for (const Mat1f* m : {d, s, a})
            if (m) dvo::require(m->rows == h && m->cols == w, "map must have the finest level's size");
        ctx().check(dvo_frame_set_depth(ctx().h, m_slot->id, d ? dvo::pixels(*d, k0) : nullptr, s ? dvo::pixels(*s, k1) : nullptr,
                                        a ? dvo::pixels(*a, k2) : nullptr, 0));
    }

    std::shared_ptr<dvo::Slot> m_slot;
};

// include/system/frame.hpp:146-188
class FrameHistory
{
public:
    FrameHistory() {}

    void setRefFrame(std::shared_ptr<Frame> frame) { m_history.push_back(frame); }

    std::shared_ptr<Frame> getRefFrame()
    {
        if (m_history.empty()) return nullptr;
        return m_history.back();
    }
    const std::shared_ptr<Frame> getRefFrame() const
    {
        if (m_history.empty()) return nullptr;
        return m_history.at(0);  // sic, frame.hpp:168-173
    }

    int size() const { return static_cast<int>(m_history.size()); }

    // i keyframes ago (0 = newest), frame.hpp:176
    const std::shared_ptr<Frame> operator[](size_t i) const { return m_history.at(size() - i - 1); }

private:
    std::vector<std::shared_ptr<Frame>> m_history;
};

}  // namespace System
