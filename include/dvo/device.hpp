// dvo/device.hpp — C++ owner of the C-ABI contexts (include/dvo_cuda.h) behind the drop-in classes.
//
// The reference's classes carry no device notion: a System::Frame is built from (gray, K, levels,
// culls) and everything else follows.  Here every distinct geometry (width, height, levels, culls, K)
// maps to one dvo_ctx (n_streams = 1) kept in a process-wide registry; frames own a slot of it.
// There is no CPU fallback: without a CUDA device the first call throws dvo::Error.
#pragma once
#include "../dvo_cuda.h"
#include "mat.hpp"

#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace dvo {

struct Error : std::runtime_error {
    int code;
    Error(int code_, const std::string& what) : std::runtime_error(what), code(code_) {}
};

class Context
{
public:
    explicit Context(const dvo_params& p) : P(p)
    {
        const int rc = dvo_create(&h, &P);
        if (rc) {
            const std::string msg = h ? dvo_last_error(h) : "dvo_create failed";
            if (h) dvo_destroy(h);
            h = nullptr;
            throw Error(rc, "dvo_create: " + msg);
        }
    }
    ~Context() { if (h) dvo_destroy(h); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;

    void check(int rc) const
    {
        if (rc) throw Error(rc, std::string("libdvo_cuda: ") + dvo_last_error(h));
    }
    int levelCols(int level) const { int w, hh; check(dvo_level_size(h, level, &w, &hh)); return w; }
    int levelRows(int level) const { int w, hh; check(dvo_level_size(h, level, &w, &hh)); return hh; }
    Mat1f levelK(int level) const
    {
        Mat1f K(3, 3);
        check(dvo_level_K(h, level, dvo::data(K)));
        return K;
    }

    dvo_ctx* h = nullptr;
    dvo_params P;
};

struct Geometry {
    int w, h, levels, culls;
    float K[9];
    bool operator<(const Geometry& o) const { return std::memcmp(this, &o, sizeof(Geometry)) < 0; }
};

inline dvo_params make_params(int w, int h, int levels, int culls, const Mat1f& K)
{
    require(!K.empty() && K.rows == 3 && K.cols == 3, "K must be 3x3");
    dvo_params p;
    dvo_params_default(&p);
    p.n_streams = 1; p.width = w; p.height = h; p.levels = levels; p.culls = culls;
    p.max_frames = 1024;  // FrameHistory is never pruned (frame.hpp:179-185); slots are allocated on first use
    p.max_keyframes = p.max_frames;  // Map::Mapper::update is handed the whole FrameHistory: as many keyframes as there can be frames
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) p.K[r * 3 + c] = K(r, c);
    return p;
}

// Process-wide registry: one context per geometry, created on first use, destroyed when the last
// frame of that geometry and the registry entry are gone (Device::reset()).
class Device
{
public:
    static std::shared_ptr<Context> get(int w, int h, int levels, int culls, const Mat1f& K)
    {
        Geometry g;
        std::memset(&g, 0, sizeof g);
        g.w = w; g.h = h; g.levels = levels; g.culls = culls;
        require(!K.empty() && K.rows == 3 && K.cols == 3, "K must be 3x3");
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) g.K[r * 3 + c] = K(r, c);
        std::lock_guard<std::mutex> lock(mutex());
        auto& reg = registry();
        auto it = reg.find(g);
        if (it != reg.end()) return it->second;
        auto ctx = std::make_shared<Context>(make_params(w, h, levels, culls, K));
        reg[g] = ctx;
        return ctx;
    }
    // context for geometry-free calls (math::se3)
    static std::shared_ptr<Context> scratch()
    {
        Mat1f K = Mat1f::zeros(3, 3);
        K(0, 0) = K(1, 1) = K(2, 2) = 1.0f;
        return get(8, 8, 1, 0, K);
    }
    static void reset()
    {
        std::lock_guard<std::mutex> lock(mutex());
        registry().clear();
    }

private:
    static std::map<Geometry, std::shared_ptr<Context>>& registry()
    {
        static std::map<Geometry, std::shared_ptr<Context>> r;
        return r;
    }
    static std::mutex& mutex()
    {
        static std::mutex m;
        return m;
    }
};

// RAII frame slot
class Slot
{
public:
    Slot(std::shared_ptr<Context> ctx_, const Mat1f& gray) : ctx(std::move(ctx_))
    {
        require(gray.rows == ctx->P.height && gray.cols == ctx->P.width, "gray size does not match the context");
        Mat1f keep;
        ctx->check(dvo_frame_create(ctx->h, &id, pixels(gray, keep)));
        ctx->check(dvo_sync(ctx->h)); // `gray` may be pageable and released by the caller
    }
    ~Slot() { if (id >= 0) dvo_frame_release(ctx->h, id); }
    Slot(const Slot&) = delete;
    Slot& operator=(const Slot&) = delete;

    Mat1f download(int level, int kind) const
    {
        Mat1f out(ctx->levelRows(level), ctx->levelCols(level));
        ctx->check(dvo_frame_download(ctx->h, id, level, kind, dvo::data(out)));
        return out;
    }

    std::shared_ptr<Context> ctx;
    int id = -1;
};

}  // namespace dvo
