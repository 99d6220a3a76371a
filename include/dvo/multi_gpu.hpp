// dvo/multi_gpu.hpp — dvo::MultiGpu: batched independent sequences over every GPU of the node from ONE host process
// (SURVEY §8-e, §7 step 8): one host thread + one dvo_ctx + one CUDA stream per device, sequence ids sharded `id mod nGPU`,
// NO data-path collective - sequences never exchange a pixel, so there is nothing to reduce or gather on the device side; the
// host collects each rank's poses.  The reference has no counterpart (it runs one sequence on one CPU); this is the C++ form of
// direct-visual-odometry_b200/dvo_b200/shard.py, which bench.py uses under torchrun.
//
//   dvo::MultiGpu mg(n_sequences, params);            // params.n_streams is overwritten per device
//   mg.push(gray, depth0, sigma0, T);                 // frame 0 of EVERY sequence: [n_sequences][h][w], prior [n_sequences][h_top][w_top]
//   mg.push(gray, nullptr, nullptr, T);               // T: [n_sequences][16], in sequence-id order
//
// Sharding itself (ids, owner, local index) is host-only arithmetic and is tested without a GPU.
#pragma once
#include "device.hpp"

#include <cstdint>
#include <thread>
#include <vector>

namespace dvo {

struct Sharding {
    int n_sequences, world;
    Sharding(int n, int w) : n_sequences(n), world(w) { require(n >= 1 && w >= 1, "bad sharding"); }
    int owner(int id) const { return id % world; }                       // id mod nGPU
    int local_index(int id) const { return id / world; }                 // position of `id` among its owner's sequences
    int local_count(int rank) const { return rank < world ? (n_sequences - rank + world - 1) / world : 0; }
    int global_id(int rank, int local) const { return rank + local * world; }
    std::vector<int> ids(int rank) const
    {
        std::vector<int> v;
        for (int id = rank; id < n_sequences; id += world) v.push_back(id);
        return v;
    }
};

class MultiGpu
{
public:
    // n_gpus <= 0: every device of the node (never more than there are sequences)
    MultiGpu(int n_sequences, dvo_params params, int n_gpus = 0) : shard_(n_sequences, pick_world(n_sequences, n_gpus)), P(params)
    {
        ctx_.resize(shard_.world);
        std::vector<std::string> err(shard_.world);
        run([&](int r) {
            if (dvo_set_device(r) != DVO_OK) { err[r] = "dvo_set_device failed"; return; }
            dvo_params p = P;
            p.n_streams = shard_.local_count(r);
            try { ctx_[r] = std::make_shared<Context>(p); } catch (const std::exception& e) { err[r] = e.what(); }
        });
        for (auto& e : err)
            if (!e.empty()) throw Error(DVO_ERR_CUDA, "MultiGpu: " + e);
        const int ntop_w = ctx_[0]->levelCols(P.levels - 1), ntop_h = ctx_[0]->levelRows(P.levels - 1);
        n_in_ = (size_t)P.width * P.height;
        n_top_ = (size_t)ntop_w * ntop_h;
        for (int r = 0; r < shard_.world; r++) {
            const size_t s = (size_t)shard_.local_count(r);
            gray_.emplace_back(s * n_in_); d0_.emplace_back(s * n_top_); s0_.emplace_back(s * n_top_); T_.emplace_back(s * 16);
        }
    }
    int world() const { return shard_.world; }
    const Sharding& sharding() const { return shard_; }

    // System::VisualOdometry::odometrize for every sequence: gray [n_sequences][height][width]; init_depth / init_sigma
    // [n_sequences][h_top][w_top] (first frame only, nullptr afterwards); T_out [n_sequences][16] in sequence-id order.
    void push(const float* gray, const float* init_depth, const float* init_sigma, float* T_out)
    {
        require(gray && T_out, "null argument");
        std::vector<int> rc(shard_.world, 0);
        run([&](int r) {
            dvo_set_device(r);
            const int n = shard_.local_count(r);
            for (int k = 0; k < n; k++) {   // gather this rank's sequences (id = r + k world) into its stream-major batch
                const size_t id = (size_t)shard_.global_id(r, k);
                std::memcpy(&gray_[r][k * n_in_], gray + id * n_in_, sizeof(float) * n_in_);
                if (init_depth) std::memcpy(&d0_[r][k * n_top_], init_depth + id * n_top_, sizeof(float) * n_top_);
                if (init_sigma) std::memcpy(&s0_[r][k * n_top_], init_sigma + id * n_top_, sizeof(float) * n_top_);
            }
            rc[r] = dvo_vo_push(ctx_[r]->h, gray_[r].data(), init_depth ? d0_[r].data() : nullptr, init_sigma ? s0_[r].data() : nullptr,
                                T_[r].data());
            for (int k = 0; k < n && rc[r] == 0; k++)
                std::memcpy(T_out + (size_t)shard_.global_id(r, k) * 16, &T_[r][(size_t)k * 16], sizeof(float) * 16);
        });
        for (int r = 0; r < shard_.world; r++) ctx_[r]->check(rc[r]);
    }
    Context& context(int rank) { return *ctx_[rank]; }

private:
    static int pick_world(int n_sequences, int n_gpus)
    {
        int have = dvo_device_count();
        if (have <= 0) throw Error(DVO_ERR_CUDA, "MultiGpu: no CUDA device (libdvo_cuda has no CPU fallback)");
        int w = n_gpus > 0 ? (n_gpus < have ? n_gpus : have) : have;
        return w < n_sequences ? w : n_sequences;
    }
    template <typename F>
    void run(F f)
    {
        std::vector<std::thread> th;
        for (int r = 1; r < shard_.world; r++) th.emplace_back(f, r);
        f(0);
        for (auto& t : th) t.join();
    }
    Sharding shard_;
    dvo_params P;
    std::vector<std::shared_ptr<Context>> ctx_;
    std::vector<std::vector<float>> gray_, d0_, s0_, T_;
    size_t n_in_ = 0, n_top_ = 0;
};

}  // namespace dvo
