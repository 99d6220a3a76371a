// dvo_cuda.cu — C-ABI (include/dvo_cuda.h) over the sm_100a kernels in dvo_kernels.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo -O3 -shared
// No CPU fallback: every entry point needs a CUDA device.
#include "dvo_kernels.cuh"
#include "dvo_synth.cuh"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

using namespace dvo;

struct SlotHost {
    void* base = nullptr;
    bool in_use = false;
    bool quad_valid = false; // granular slots: dev.quad matches dev.gray (VO pool slots are kept valid by the kernels themselves)
    SlotDev dev{};
};

struct dvo_ctx {
    dvo_params P{};
    Consts c{};
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    std::vector<SlotHost> slots;
    size_t slot_bytes = 0;
    SlotDev* d_slots = nullptr;
    TrackState* d_ts = nullptr;
    MapState* d_ms = nullptr;
    VoState* d_vo = nullptr;
    int* d_hist = nullptr;
    PoseTab* d_tab = nullptr;
    double* d_partials = nullptr;
    int trk_B = 1;          // CTAs per stream of the tracking kernel (1: no inter-CTA barrier)
    int trk_capacity = 0;   // co-resident tracking CTAs on this device
    MapRoi roi{};           // Mapper::update's ROI clipped to the finest level
    dvo_iter_log* d_log = nullptr;
    int log_cap = 0;
    uint8_t* d_flags = nullptr;
    uint8_t* d_mask = nullptr;
    unsigned int* d_blockcnt = nullptr;
    int nblk_top = 0;
    int* d_winner = nullptr;
    int* d_remap = nullptr;        // [n_top] source pixel of every finest-level pixel (bgr8 ingest), built on first use
    // Kinect ingest (dvo_kinect_configure)
    bool kin_ready = false;
    KinectDev kin{};
    int* d_kin_cidx = nullptr;     // [color_h][color_w]
    int* d_kin_didx = nullptr;     // [n_top]
    uint8_t* d_kin_color = nullptr; // [S][color_h][color_w][channels]
    uint16_t* d_kin_depth = nullptr; // [S][in_h][in_w]
    float* d_tmp = nullptr;
    uint32_t* d_rng = nullptr;
    uint64_t rng_bound = 0;        // upper bound of any stream's replay position (grows by roi.n per Mapper::update; tightened on read-back)
    ReadyQ* d_rq = nullptr;        // tracking -> Mapper::update hand-over of one frame
    int* d_status = nullptr;       // first device-side error of any stream
    int* h_status = nullptr;       // pinned: [0], [1] = status read back with pipeline buffer k's poses; [2] = synchronous pushes
    unsigned int* d_noalloc = nullptr; // device address of g_noalloc_counts on this context's device (dvo_mask_compact_noalloc)
    bool overlap = true;           // k_map_update as a programmatic dependent of k_track (DVO_OVERLAP=0 disables)
    bool fuse_prop = true;         // Implement::propagate in k_track's epilogue (DVO_FUSE_PROP=0: the two wide kernels)
    // keyframe-pool growth: max over streams of the keyframe count, read back asynchronously (d_status[1])
    int kf_known = 0, kf_known_frame = -1;
    uint64_t rng_known = 0;        // furthest replay position of any stream after frame kf_known_frame
    int kf_rb_frame[2] = {-1, -1};
    bool kf_rb_pending[2] = {false, false};
    cudaEvent_t ev_kf[2] = {nullptr, nullptr};
    unsigned kf_gen = 0;
    int status_pending[2] = {0, 0};
    float* d_stage_in = nullptr;   // [S][in_h][in_w]
    float* d_stage[3] = {nullptr, nullptr, nullptr}; // depth/sigma/age uploads, download staging
    float* d_small = nullptr;      // [S][64] small vectors (xi, poses)
    float* d_forced = nullptr;     // [S][6] forced relative pose for the next dvo_vo_push
    bool forced_pending = false;
    int* d_ismall = nullptr;       // small int scratch
    void* h_pinned = nullptr;      // pinned staging for small readbacks
    size_t h_pinned_bytes = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t launches = 0;
    // per-kernel device timing (CUDA events around every launch of the timed kernels)
    bool ktiming = false;
    std::vector<cudaEvent_t> k_ev;
    std::vector<int> k_ev_id;
    size_t k_ev_used = 0;
    std::vector<cudaEvent_t> user_ev; // dvo_event_record / dvo_event_elapsed
    bool rows_only_staged = false; // d_stage_in holds only the sampled input rows (host path)
    // pipelined submission (dvo_vo_submit / dvo_vo_collect)
    cudaStream_t copy_stream = nullptr;
    float* d_stage_pp[2] = {nullptr, nullptr};
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    uint64_t pp_head = 0;
    int pp_in_flight = 0;
    // VO host mirror
    int vo_frame_id = -1;
    std::mt19937 engine; // default seed 5489, src/math/gaussian.cpp:8
};

// Scratch of dvo_mask_compact_noalloc: one count per 256-byte tile of the mask.  Host code never passes this name to a kernel (there it
// is a host-side handle, not a device address): dvo_create resolves the address per device with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned int g_noalloc_counts[DVO_NOALLOC_MAX_N / 256];

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e__);                        \
            return DVO_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)
#define KCHECK()                                                                                   \
    do {                                                                                           \
        ctx->launches++;                                                                           \
        cudaError_t e__ = cudaGetLastError();                                                      \
        if (e__ != cudaSuccess) {                                                                  \
            ctx->err = std::string("kernel launch: ") + cudaGetErrorString(e__);                   \
            return DVO_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)
#define ARG(cond, msg)                                                                             \
    do {                                                                                           \
        if (!(cond)) {                                                                             \
            ctx->err = msg;                                                                        \
            return DVO_ERR_ARG;                                                                    \
        }                                                                                          \
    } while (0)

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

extern "C" int dvo_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
extern "C" int dvo_set_device(int device) { return cudaSetDevice(device) == cudaSuccess ? DVO_OK : DVO_ERR_CUDA; }

extern "C" void dvo_params_default(dvo_params* p)
{
    memset(p, 0, sizeof(*p));
    p->n_streams = 1;
    p->width = 640; p->height = 480;
    p->levels = 3; p->culls = 2; // system.hpp:47
    const float K[9] = {7.93852356e+02f, 0.f, 3.77845917e+02f, 0.f, 7.96643677e+02f, 2.20205841e+02f, 0.f, 0.f, 1.f};
    memcpy(p->K, K, sizeof K); // external/camera-calibration/data/logicool_00/config.yaml
    p->max_frames = 64;
    p->max_keyframes = 64;
    p->min_residual = 5e-3f; p->min_update = 5e-4f; p->max_iterations = 15;
    p->roi_level = 2; p->roi_x0 = 20; p->roi_x1 = 140; p->roi_y0 = 20; p->roi_y1 = 100;
    p->min_depth = 0.20; p->sigma_lo = 0.01f; p->sigma_hi = 0.5f;
    p->step_default = 2.0f; p->step_level1 = 1.5f; p->step_level2 = 1.0f;
    p->map_roi_x0 = 16; p->map_roi_x1 = 144; p->map_roi_y0 = 12; p->map_roi_y1 = 108;
    p->min_movement = 0.02f; p->max_forward = 6;
    p->accept_depth_lo = 0.2f; p->accept_depth_hi = 6.0f; p->accept_sigma_hi = 0.5f;
    p->rng_table = 1 << 22;
}

static int alloc_slot(dvo_ctx* ctx, int idx)
{
    SlotHost& sh = ctx->slots[idx];
    if (sh.base) {
        // A recycled slot must not hand its previous occupant's depth / sigma / age / pose to the new Frame: the reference's
        // Frame ctor starts from m_age = zeros and a fresh prior (frame.hpp:108-117).  Same state as a first allocation: all zero.
        CK(cudaMemsetAsync(sh.base, 0, ctx->slot_bytes, ctx->stream));
        sh.quad_valid = false;
        return DVO_OK;
    }
    CK(cudaMalloc(&sh.base, ctx->slot_bytes));
    CK(cudaMemsetAsync(sh.base, 0, ctx->slot_bytes, ctx->stream));
    char* p = (char*)sh.base;
    const Consts& c = ctx->c;
    const size_t ntop = (size_t)c.lv[c.L - 1].n;
    auto take = [&](size_t bytes) { char* r = p; p += (bytes + 255) & ~(size_t)255; return r; };
    sh.dev.quad = (float4*)take((size_t)c.S * ntop * sizeof(float4));
    sh.dev.ds = (float2*)take((size_t)c.S * ntop * sizeof(float2));
    sh.dev.gray = (float*)take((size_t)c.S * ntop * sizeof(float));
    sh.dev.age = (float*)take((size_t)c.S * ntop * sizeof(float));
    sh.dev.xi = (float*)take((size_t)c.S * 6 * sizeof(float));
    sh.dev.rel_xi = (float*)take((size_t)c.S * 6 * sizeof(float));
    CK(cudaMemcpyAsync(ctx->d_slots + idx, &sh.dev, sizeof(SlotDev), cudaMemcpyHostToDevice, ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_create(dvo_ctx** out, const dvo_params* p)
{
    if (!out || !p) return DVO_ERR_ARG;
    *out = nullptr;
    dvo_ctx* ctx = new dvo_ctx();
    *out = ctx; // returned even on failure so that dvo_last_error can be read
    ctx->P = *p;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        ctx->err = "no CUDA device: libdvo_cuda has no CPU fallback";
        return DVO_ERR_CUDA;
    }
    CK(cudaGetDevice(&ctx->device));
    // dvo_mask_compact_noalloc's scratch: in host code `g_noalloc_counts` is only a handle, the array's address on THIS device (the
    // current one) comes from the runtime.  Once per context = once per device the caller works on; allocates nothing.
    CK(cudaGetSymbolAddress((void**)&ctx->d_noalloc, g_noalloc_counts));
    ARG(p->n_streams >= 1 && p->width > 0 && p->height > 0, "bad geometry");
    ARG(p->levels >= 1 && p->levels <= MAXL && p->culls >= 0 && p->culls < 16, "bad levels/culls");
    ARG(p->max_frames >= 2 && p->max_keyframes >= 1, "bad capacities");
    Consts& c = ctx->c;
    c.S = p->n_streams; c.L = p->levels; c.in_w = p->width; c.in_h = p->height; c.culls = p->culls;
    const int bw = p->width >> p->culls, bh = p->height >> p->culls;
    ARG((bw >> (p->levels - 1)) >= 1 && (bh >> (p->levels - 1)) >= 1, "image too small for the pyramid");
    for (int l = 0; l < c.L; l++) {
        const int t = c.L - 1 - l;
        LevelGeom& g = c.lv[l];
        g.w = bw >> t; g.h = bh >> t; g.n = g.w * g.h;
        // Convert::cullIntrinsic (convert.cpp:22-29): K * (1/2^times) from the base level's K
        const float s0 = (float)(1.0 / (double)(1 << p->culls));
        float Kb[4] = {p->K[0], p->K[4], p->K[2], p->K[5]};
        if (p->culls) for (float& v : Kb) v = v * s0;
        const float s1 = (float)(1.0 / (double)(1 << t));
        if (t) for (float& v : Kb) v = v * s1;
        g.fx = Kb[0]; g.fy = Kb[1]; g.cx = Kb[2]; g.cy = Kb[3];
    }
    c.max_kf = p->max_keyframes;
    c.kf_stride = (p->max_keyframes + 31) & ~31;
    c.vo_base = p->max_frames;
    c.max_iter = p->max_iterations;
    c.roi_level = p->roi_level; c.roi_x0 = p->roi_x0; c.roi_x1 = p->roi_x1; c.roi_y0 = p->roi_y0; c.roi_y1 = p->roi_y1;
    {   // (double)d < min_depth  <=>  d < smallest float >= min_depth
        float f = (float)p->min_depth;
        if ((double)f < p->min_depth) f = nextafterf(f, INFINITY);
        c.min_depth_f = f;
    }
    c.sigma_lo = p->sigma_lo; c.sigma_hi = p->sigma_hi;
    c.step_default = p->step_default; c.step_level1 = p->step_level1; c.step_level2 = p->step_level2;
    c.huber_delta = p->huber_delta > 0.0f ? p->huber_delta : 0.0f;
    c.min_grad2 = p->min_gradient > 0.0f ? p->min_gradient * p->min_gradient : 0.0f;
    c.map_x0 = p->map_roi_x0; c.map_x1 = p->map_roi_x1; c.map_y0 = p->map_roi_y0; c.map_y1 = p->map_roi_y1;
    c.min_movement = p->min_movement; c.max_forward = p->max_forward;
    c.acc_d_lo = p->accept_depth_lo; c.acc_d_hi = p->accept_depth_hi; c.acc_s_hi = p->accept_sigma_hi;
    c.rng_table = (unsigned)(p->rng_table > 0 ? p->rng_table : (1 << 22));

    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&ctx->ev0));
    CK(cudaEventCreate(&ctx->ev1));
    const int S = c.S, ntop = c.lv[c.L - 1].n;
    size_t sb = 0;
    auto rnd = [](size_t b) { return (b + 255) & ~(size_t)255; };
    sb += rnd((size_t)S * ntop * 16) + rnd((size_t)S * ntop * 8) + 2 * rnd((size_t)S * ntop * 4) + 2 * rnd((size_t)S * 24);
    ctx->slot_bytes = sb;
    // slot ids [0, max_frames): granular API; [max_frames, max_frames + max_keyframes]: the VO pool
    // (entry k = keyframe k of every stream, entry n_kf = the stream's current frame)
    const size_t n_slots = (size_t)p->max_frames + p->max_keyframes + 1;
    ctx->slots.resize(n_slots);
    CK(cudaMalloc(&ctx->d_slots, sizeof(SlotDev) * n_slots));
    CK(cudaMemset(ctx->d_slots, 0, sizeof(SlotDev) * n_slots));
    CK(cudaMalloc(&ctx->d_ts, sizeof(TrackState) * S));
    CK(cudaMemset(ctx->d_ts, 0, sizeof(TrackState) * S));
    CK(cudaMalloc(&ctx->d_ms, sizeof(MapState) * S));
    CK(cudaMemset(ctx->d_ms, 0, sizeof(MapState) * S));
    CK(cudaMalloc(&ctx->d_vo, sizeof(VoState) * S));
    CK(cudaMalloc(&ctx->d_hist, sizeof(int) * (size_t)S * c.kf_stride));
    CK(cudaMalloc(&ctx->d_tab, sizeof(PoseTab) * (size_t)S * c.kf_stride));
    CK(cudaMalloc(&ctx->d_rq, sizeof(ReadyQ) + sizeof(int) * (size_t)S));
    CK(cudaMemset(ctx->d_rq, 0, sizeof(ReadyQ) + sizeof(int) * (size_t)S));
    CK(cudaMalloc(&ctx->d_status, sizeof(int) * 4));   // [0] first device-side error, [1] max over streams of the keyframe count,
    CK(cudaMemset(ctx->d_status, 0, sizeof(int) * 4));  // [2] max over streams of the mt19937 replay position
    CK(cudaMallocHost(&ctx->h_status, sizeof(int) * 8)); // [0..2] error read-backs, [4..5] keyframe-count read-backs
    memset(ctx->h_status, 0, sizeof(int) * 8);
    if (const char* e = getenv("DVO_OVERLAP")) ctx->overlap = atoi(e) != 0;
    if (const char* e = getenv("DVO_FUSE_PROP")) ctx->fuse_prop = atoi(e) != 0;
    for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&ctx->ev_kf[k], cudaEventDisableTiming));
    for (int k = 0; k < 2; k++) { // pipeline events exist from the start: buffer 0 is d_stage_in, which the plain entry points use too
        CK(cudaEventCreateWithFlags(&ctx->ev_copied[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&ctx->ev_consumed[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&ctx->ev_done[k], cudaEventDisableTiming));
    }
    // Tracking launch shape: one CTA per stream when the streams alone fill the GPU; otherwise a
    // stream's pixels are split over B co-resident CTAs (cooperative launch, global barrier).
    {
        int per_sm = 0, n_sm = 0;
        // the split form runs k_track_staged (coarse reference levels in shared memory): its occupancy bounds the co-resident grid
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_track_staged, TRK_THREADS, 0));
        CK(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, ctx->device));
        ctx->trk_capacity = per_sm * n_sm;
        int coop = 0;
        CK(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, ctx->device));
        int B = coop ? ctx->trk_capacity / S : 1;
        const int useful = cdiv(ntop, TRK_THREADS * 4); // >= 4 pixels per thread at the finest level
        if (B > useful) B = useful;
        if (B < 1) B = 1;
        // tests: force the one-CTA-per-stream form (the one 256 streams take: fused propagate, overlapped Mapper::update) on few streams
        if (const char* e = getenv("DVO_TRK_B")) { const int f = atoi(e); if (f >= 1 && f <= B) B = f; }
        ctx->trk_B = B;
    }
    CK(cudaMalloc(&ctx->d_partials, sizeof(double) * 32 * (size_t)S * ctx->trk_B));
    ctx->log_cap = c.L * c.max_iter;
    CK(cudaMalloc(&ctx->d_log, sizeof(dvo_iter_log) * (size_t)S * ctx->log_cap));
    CK(cudaMemset(ctx->d_log, 0, sizeof(dvo_iter_log) * (size_t)S * ctx->log_cap));
    size_t nmax = 0;
    for (int l = 0; l < c.L; l++) nmax = nmax > (size_t)c.lv[l].n ? nmax : (size_t)c.lv[l].n;
    CK(cudaMalloc(&ctx->d_flags, (size_t)S * ntop));
    CK(cudaMalloc(&ctx->d_mask, (size_t)S * nmax));
    CK(cudaMemset(ctx->d_flags, 0, (size_t)S * ntop));
    {   // Mapper::update's ROI (mapper.cpp:90) clipped to the finest level
        const LevelGeom& gt = c.lv[c.L - 1];
        MapRoi& r = ctx->roi;
        r.x0 = c.map_x0 > 0 ? c.map_x0 : 0; r.y0 = c.map_y0 > 0 ? c.map_y0 : 0;
        const int x1 = c.map_x1 < gt.w - 1 ? c.map_x1 : gt.w - 1, y1 = c.map_y1 < gt.h - 1 ? c.map_y1 : gt.h - 1;
        r.w = x1 - r.x0 + 1 > 0 ? x1 - r.x0 + 1 : 0; r.h = y1 - r.y0 + 1 > 0 ? y1 - r.y0 + 1 : 0;
        r.n = r.w * r.h;
        r.nblk = cdiv(r.n > 0 ? r.n : 1, MAPU_THREADS);
        if (r.w == 0) r.w = 1;
    }
    ctx->nblk_top = ctx->roi.nblk;
    // The warp-cooperative epipolar search keeps 4.6 KB of checkpoints per warp (9.2 KB per block, 16 blocks per
    // SM); the rest of the 256 KB stays L1 for the gathers.  Measured on B200 (640x480 x 256 streams, k_map_update):
    // carve-out 50 % 822 us, 60-66 % 761 us, 80 % 1416 us.  DVO_MAP_CARVEOUT overrides for experiments.
    {
        int carve = 66;
        if (const char* e = getenv("DVO_MAP_CARVEOUT")) carve = atoi(e);
        CK(cudaFuncSetAttribute(k_map_update, cudaFuncAttributePreferredSharedMemoryCarveout, carve));
    }
    CK(cudaMalloc(&ctx->d_blockcnt, sizeof(unsigned) * (size_t)S * ctx->nblk_top));
    CK(cudaMalloc(&ctx->d_winner, sizeof(int) * (size_t)S * ntop));
    CK(cudaMemset(ctx->d_winner, 0xFF, sizeof(int) * (size_t)S * ntop));
    CK(cudaMalloc(&ctx->d_tmp, sizeof(float) * (size_t)S * ntop));
    CK(cudaMalloc(&ctx->d_stage_in, sizeof(float) * (size_t)S * c.in_w * c.in_h));
    for (int k = 0; k < 3; k++) CK(cudaMalloc(&ctx->d_stage[k], sizeof(float) * (size_t)S * c.in_w * c.in_h));
    CK(cudaMalloc(&ctx->d_small, sizeof(float) * (size_t)S * 64));
    CK(cudaMalloc(&ctx->d_forced, sizeof(float) * (size_t)S * 6));
    CK(cudaMalloc(&ctx->d_ismall, sizeof(int) * (size_t)(S > 4096 ? S : 4096)));
    ctx->h_pinned_bytes = sizeof(float) * (size_t)S * 64 + sizeof(dvo_iter_log) * (size_t)S * ctx->log_cap + sizeof(TrackState) * (size_t)S;
    CK(cudaMallocHost(&ctx->h_pinned, ctx->h_pinned_bytes));
    {   // mt19937 replay table (src/math/gaussian.cpp:8): raw 32-bit outputs, default seed
        std::vector<uint32_t> raw(c.rng_table);
        for (auto& v : raw) v = (uint32_t)ctx->engine();
        CK(cudaMalloc(&ctx->d_rng, sizeof(uint32_t) * (size_t)c.rng_table));
        CK(cudaMemcpy(ctx->d_rng, raw.data(), sizeof(uint32_t) * (size_t)c.rng_table, cudaMemcpyHostToDevice));
    }
    int rc = dvo_vo_reset(ctx);
    if (rc) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" void dvo_destroy(dvo_ctx* ctx)
{
    if (!ctx) return;
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (auto& s : ctx->slots) if (s.base) cudaFree(s.base);
    cudaFree(ctx->d_slots); cudaFree(ctx->d_ts); cudaFree(ctx->d_ms); cudaFree(ctx->d_vo); cudaFree(ctx->d_hist);
    cudaFree(ctx->d_tab); cudaFree(ctx->d_partials); cudaFree(ctx->d_log); cudaFree(ctx->d_flags); cudaFree(ctx->d_mask);
    cudaFree(ctx->d_blockcnt); cudaFree(ctx->d_winner); cudaFree(ctx->d_remap);
    cudaFree(ctx->d_kin_cidx); cudaFree(ctx->d_kin_didx); cudaFree(ctx->d_kin_color); cudaFree(ctx->d_kin_depth); cudaFree(ctx->d_tmp); cudaFree(ctx->d_rng); cudaFree(ctx->d_stage_in);
    for (int k = 0; k < 3; k++) cudaFree(ctx->d_stage[k]);
    cudaFree(ctx->d_small); cudaFree(ctx->d_ismall); cudaFree(ctx->d_forced); cudaFree(ctx->d_rq); cudaFree(ctx->d_status);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    if (ctx->h_status) cudaFreeHost(ctx->h_status);
    if (ctx->copy_stream) {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaFree(ctx->d_stage_pp[1]);
        cudaStreamDestroy(ctx->copy_stream);
    }
    for (int k = 0; k < 2; k++)
        if (ctx->ev_copied[k]) { cudaEventDestroy(ctx->ev_copied[k]); cudaEventDestroy(ctx->ev_consumed[k]); cudaEventDestroy(ctx->ev_done[k]); }
    for (auto e : ctx->user_ev) if (e) cudaEventDestroy(e);
    for (int k = 0; k < 2; k++) if (ctx->ev_kf[k]) cudaEventDestroy(ctx->ev_kf[k]);
    for (auto e : ctx->k_ev) cudaEventDestroy(e);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char* dvo_last_error(const dvo_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }
extern "C" int dvo_sync(dvo_ctx* ctx) { CK(cudaStreamSynchronize(ctx->stream)); return DVO_OK; }
extern "C" uint64_t dvo_launch_count(const dvo_ctx* ctx) { return ctx->launches; }

extern "C" int dvo_level_size(const dvo_ctx* ctx, int level, int* w, int* h)
{
    if (!ctx || level < 0 || level >= ctx->c.L) return DVO_ERR_ARG;
    *w = ctx->c.lv[level].w; *h = ctx->c.lv[level].h;
    return DVO_OK;
}
extern "C" int dvo_level_K(const dvo_ctx* ctx, int level, float K[9])
{
    if (!ctx || level < 0 || level >= ctx->c.L) return DVO_ERR_ARG;
    const LevelGeom& g = ctx->c.lv[level];
    const float k[9] = {g.fx, 0, g.cx, 0, g.fy, g.cy, 0, 0, 1};
    memcpy(K, k, sizeof k);
    return DVO_OK;
}
extern "C" void* dvo_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void dvo_host_free(void* p) { if (p) cudaFreeHost(p); }

// ---- frames ---------------------------------------------------------------------------------
static int find_free_slot(dvo_ctx* ctx, int* slot)
{
    for (size_t i = 0; i < (size_t)ctx->P.max_frames; i++)
        if (!ctx->slots[i].in_use) {
            int rc = alloc_slot(ctx, (int)i);
            if (rc) return rc;
            ctx->slots[i].in_use = true;
            *slot = (int)i;
            return DVO_OK;
        }
    ctx->err = "frame slots exhausted (dvo_params.max_frames)";
    return DVO_ERR_SLOT;
}
static int check_slot(dvo_ctx* ctx, int slot)
{
    if (slot < 0 || slot >= (int)ctx->slots.size() || !ctx->slots[slot].in_use) {
        ctx->err = "bad frame slot";
        return DVO_ERR_SLOT;
    }
    return DVO_OK;
}

// ---- per-kernel device timing --------------------------------------------------------------------
enum KernelId { K_CULL = 0, K_TRACK, K_PROP_SCATTER, K_PROP_GATHER, K_MAP_UPDATE, K_MAP_APPLY, K_REGULARIZE, K_REG_COMMIT, K_COUNT };
static const char* const kernel_names[K_COUNT] = {"k_cull_gray", "k_track", "k_propagate_scatter", "k_propagate_gather", "k_map_update",
                                                  "k_map_update_apply", "k_regularize", "k_regularize_commit"};
struct KTimer { // records an event pair around one launch when timing is on
    dvo_ctx* ctx;
    cudaEvent_t e1 = nullptr;
    KTimer(dvo_ctx* c, int id) : ctx(c)
    {
        if (!ctx->ktiming) return;
        if (ctx->k_ev_used + 2 > ctx->k_ev.size())
            for (int k = 0; k < 2; k++) {
                cudaEvent_t e;
                if (cudaEventCreate(&e) != cudaSuccess) return;
                ctx->k_ev.push_back(e);
            }
        ctx->k_ev_id.resize(ctx->k_ev.size() / 2);
        ctx->k_ev_id[ctx->k_ev_used / 2] = id;
        cudaEvent_t e0 = ctx->k_ev[ctx->k_ev_used++];
        e1 = ctx->k_ev[ctx->k_ev_used++];
        cudaEventRecord(e0, ctx->stream);
    }
    ~KTimer() { if (e1) cudaEventRecord(e1, ctx->stream); }
};

// slot >= 0: that slot for every stream; slot < 0: each stream's current VO slot.
// rows_only: gray_dev holds only the input rows that Convert::cullImage samples (host staging path).
// kind: 0 float gray, 1 8-bit gray (aux = invalid_code), 2 8-bit BGR / BGRA (aux = channels; whole frames, through the remap index)
static int build_frame(dvo_ctx* ctx, int slot, const float* gray_dev, bool rows_only, int kind = 0, int aux = -1)
{
    const Consts& c = ctx->c;
    const VoState* vo = slot < 0 ? ctx->d_vo : nullptr;
    const LevelGeom& gt = c.lv[c.L - 1];
    dim3 grid(cdiv(gt.n, 1024), c.S);
    if (slot >= 0) ctx->slots[slot].quad_valid = false;
    ReadyQ* rq = vo ? ctx->d_rq : nullptr; // the VO frame's first kernel clears the tracking -> mapping hand-over queue
    {
        KTimer t(ctx, K_CULL);
        if (kind == 2)
            k_cull_bgr8<<<dim3(cdiv(gt.n, 256), c.S), 256, 0, ctx->stream>>>(c, (const uint8_t*)gray_dev, aux, ctx->d_remap, ctx->d_slots, slot, vo, rq);
        else if (kind == 1)
            k_cull_gray8<<<grid, 256, 0, ctx->stream>>>(c, (const uint8_t*)gray_dev, rows_only ? gt.h : c.in_h, rows_only ? 0 : c.culls,
                                                        aux, ctx->d_slots, slot, vo, rq);
        else
            k_cull_gray<<<grid, 256, 0, ctx->stream>>>(c, gray_dev, rows_only ? gt.h : c.in_h, rows_only ? 0 : c.culls, ctx->d_slots, slot, vo, rq);
        KCHECK();
    }
    // d_stage_in doubles as pipeline buffer 0: whoever filled it (dvo_frame_create*, dvo_vo_push*), a later dvo_vo_submit must not
    // overwrite it on the copy stream before this kernel has read it
    if ((const void*)gray_dev == (const void*)ctx->d_stage_in) CK(cudaEventRecord(ctx->ev_consumed[0], ctx->stream));
    return DVO_OK;
}

// Host frame -> device staging.  Convert::cullImage only samples rows that are multiples of 2^culls,
// so when the geometry allows one uniform pitch only those rows cross PCIe (1/2^culls of the bytes).
static int stage_gray_to(dvo_ctx* ctx, float* dst, cudaStream_t st, const void* gray, bool* rows_only, size_t elem = sizeof(float))
{
    const Consts& c = ctx->c;
    const int step = 1 << c.culls;
    const LevelGeom& gt = c.lv[c.L - 1];
    if (c.culls > 0 && c.in_h % step == 0 && gt.h == c.in_h / step) {
        const size_t row_bytes = elem * (size_t)c.in_w;
        // one strided copy for all streams: the copy engine keeps the link full even on 640-byte rows (53.6 GB/s measured,
        // scripts/micro/h2d_rows.cu; one call per image drops to 15 GB/s, zero-copy reads by the SMs to 13 GB/s)
        CK(cudaMemcpy2DAsync(dst, row_bytes, gray, row_bytes * step, row_bytes, (size_t)c.S * gt.h, cudaMemcpyHostToDevice, st));
        *rows_only = true;
    } else {
        CK(cudaMemcpyAsync(dst, gray, elem * (size_t)c.S * c.in_w * c.in_h, cudaMemcpyHostToDevice, st));
        *rows_only = false;
    }
    return DVO_OK;
}
static int stage_gray(dvo_ctx* ctx, const float* gray, bool* rows_only) { return stage_gray_to(ctx, ctx->d_stage_in, ctx->stream, gray, rows_only); }

extern "C" int dvo_frame_create_dev(dvo_ctx* ctx, int* slot, const float* gray_dev)
{
    ARG(slot && gray_dev, "null argument");
    int rc = find_free_slot(ctx, slot);
    if (rc) return rc;
    return build_frame(ctx, *slot, gray_dev, false);
}

extern "C" int dvo_frame_create(dvo_ctx* ctx, int* slot, const float* gray)
{
    ARG(slot && gray, "null argument");
    bool rows_only = false;
    int rc = stage_gray(ctx, gray, &rows_only);
    if (rc) return rc;
    rc = find_free_slot(ctx, slot);
    if (rc) return rc;
    return build_frame(ctx, *slot, ctx->d_stage_in, rows_only);
}

extern "C" int dvo_frame_create_gray8(dvo_ctx* ctx, int* slot, const uint8_t* gray8, int invalid_code)
{
    ARG(slot && gray8 && invalid_code >= -1 && invalid_code <= 255, "bad argument");
    bool rows_only = false;
    int rc = stage_gray_to(ctx, ctx->d_stage_in, ctx->stream, gray8, &rows_only, 1);
    if (rc) return rc;
    rc = find_free_slot(ctx, slot);
    if (rc) return rc;
    return build_frame(ctx, *slot, ctx->d_stage_in, rows_only, 1, invalid_code);
}

// ---- 8-bit BGR ingest: Loader::getNormalizedUndistortedImages on the device ------------------------
static int remap_build(dvo_ctx* ctx, const float* map_x, const float* map_y)
{
    const Consts& c = ctx->c;
    const LevelGeom& gt = c.lv[c.L - 1];
    if (!ctx->d_remap) CK(cudaMalloc(&ctx->d_remap, sizeof(int) * (size_t)gt.n));
    float *dx = nullptr, *dy = nullptr;
    if (map_x) { // the two staging planes hold [S][in_h][in_w] floats each: room for one map
        const size_t b = sizeof(float) * (size_t)c.in_w * c.in_h;
        dx = ctx->d_stage[0]; dy = ctx->d_stage[1];
        CK(cudaMemcpyAsync(dx, map_x, b, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(dy, map_y, b, cudaMemcpyHostToDevice, ctx->stream));
    }
    k_build_remap<<<cdiv(gt.n, 256), 256, 0, ctx->stream>>>(c, dx, dy, ctx->d_remap);
    KCHECK();
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_set_undistort_map(dvo_ctx* ctx, const float* map_x, const float* map_y)
{
    ARG((map_x == nullptr) == (map_y == nullptr), "give both maps or neither");
    return remap_build(ctx, map_x, map_y);
}

static int stage_bgr(dvo_ctx* ctx, float* dst, cudaStream_t st, const uint8_t* bgr8, int channels)
{
    const Consts& c = ctx->c;
    ARG(bgr8 && (channels == 3 || channels == 4), "bgr8: null frame or channels not 3 / 4");
    if (!ctx->d_remap) { // no map given: identity (Loader::getNormalizedImages)
        int rc = remap_build(ctx, nullptr, nullptr);
        if (rc) return rc;
    }
    CK(cudaMemcpyAsync(dst, bgr8, (size_t)c.S * c.in_w * c.in_h * channels, cudaMemcpyHostToDevice, st));
    return DVO_OK;
}

extern "C" int dvo_frame_create_bgr8(dvo_ctx* ctx, int* slot, const uint8_t* bgr8, int channels)
{
    ARG(slot, "null argument");
    int rc = stage_bgr(ctx, ctx->d_stage_in, ctx->stream, bgr8, channels);
    if (rc) return rc;
    rc = find_free_slot(ctx, slot);
    if (rc) return rc;
    return build_frame(ctx, *slot, ctx->d_stage_in, false, 2, channels);
}

// ---- Kinect ingest: KinectLoader::getMappedImages + mapDepthtoGray + Frame(gray, depth, sigma, ...) ------------------
extern "C" int dvo_kinect_configure(dvo_ctx* ctx, const dvo_kinect_params* kp, const float* rgb_map_x, const float* rgb_map_y,
                                    const float* depth_map_x, const float* depth_map_y)
{
    const Consts& c = ctx->c;
    ARG(kp && rgb_map_x && rgb_map_y && depth_map_x && depth_map_y, "null argument");
    ARG(kp->color_width > 0 && kp->color_height > 0 && kp->color_width < 32768 && kp->color_height < 32768 &&
        (kp->channels == 3 || kp->channels == 4), "bad colour image geometry");
    const LevelGeom& gt = c.lv[c.L - 1];
    const size_t cn = (size_t)kp->color_width * kp->color_height;
    cudaFree(ctx->d_kin_cidx); cudaFree(ctx->d_kin_didx); cudaFree(ctx->d_kin_color); cudaFree(ctx->d_kin_depth);
    ctx->d_kin_cidx = ctx->d_kin_didx = nullptr; ctx->d_kin_color = nullptr; ctx->d_kin_depth = nullptr;
    ctx->kin_ready = false;
    CK(cudaMalloc(&ctx->d_kin_cidx, sizeof(int) * cn));
    CK(cudaMalloc(&ctx->d_kin_didx, sizeof(int) * (size_t)gt.n));
    CK(cudaMalloc(&ctx->d_kin_color, (size_t)c.S * cn * kp->channels));
    CK(cudaMalloc(&ctx->d_kin_depth, sizeof(uint16_t) * (size_t)c.S * c.in_w * c.in_h));
    float *mx = nullptr, *my = nullptr; // temporary device copies of the colour maps
    CK(cudaMalloc(&mx, sizeof(float) * cn));
    CK(cudaMalloc(&my, sizeof(float) * cn));
    CK(cudaMemcpyAsync(mx, rgb_map_x, sizeof(float) * cn, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(my, rgb_map_y, sizeof(float) * cn, cudaMemcpyHostToDevice, ctx->stream));
    k_build_remap_full<<<cdiv((int)cn, 256), 256, 0, ctx->stream>>>(mx, my, kp->color_width, kp->color_height, ctx->d_kin_cidx);
    KCHECK();
    // depth maps: only the pixels Convert::cullImage keeps (the generic builder, into the Kinect's own index)
    const size_t db = sizeof(float) * (size_t)c.in_w * c.in_h;
    CK(cudaMemcpyAsync(ctx->d_stage[0], depth_map_x, db, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->d_stage[1], depth_map_y, db, cudaMemcpyHostToDevice, ctx->stream));
    k_build_remap<<<cdiv(gt.n, 256), 256, 0, ctx->stream>>>(c, ctx->d_stage[0], ctx->d_stage[1], ctx->d_kin_didx);
    KCHECK();
    CK(cudaStreamSynchronize(ctx->stream));
    cudaFree(mx); cudaFree(my);
    KinectDev& k = ctx->kin;
    k.cw = kp->color_width; k.ch = kp->color_height; k.channels = kp->channels;
    k.fx = kp->rgb_K[0]; k.fy = kp->rgb_K[4]; k.cx = kp->rgb_K[2]; k.cy = kp->rgb_K[5];
    k.dfx = ctx->P.K[0]; k.dfy = ctx->P.K[4]; k.dcx = ctx->P.K[2]; k.dcy = ctx->P.K[5];
    for (int r = 0; r < 3; r++)
        for (int q = 0; q < 4; q++) k.T[r * 4 + q] = kp->invT[r * 4 + q];
    k.cidx = ctx->d_kin_cidx; k.didx = ctx->d_kin_didx;
    ctx->kin_ready = true;
    return DVO_OK;
}

extern "C" int dvo_frame_create_kinect(dvo_ctx* ctx, int* slot, const uint8_t* bgr8, const uint16_t* depth16)
{
    const Consts& c = ctx->c;
    ARG(slot && bgr8 && depth16, "null argument");
    if (!ctx->kin_ready) {
        ctx->err = "dvo_frame_create_kinect: call dvo_kinect_configure first";
        return DVO_ERR_STATE;
    }
    const KinectDev& k = ctx->kin;
    CK(cudaMemcpyAsync(ctx->d_kin_color, bgr8, (size_t)c.S * k.cw * k.ch * k.channels, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->d_kin_depth, depth16, sizeof(uint16_t) * (size_t)c.S * c.in_w * c.in_h, cudaMemcpyHostToDevice, ctx->stream));
    int rc = find_free_slot(ctx, slot);
    if (rc) return rc;
    const LevelGeom& gt = c.lv[c.L - 1];
    k_kinect_frame<<<dim3(cdiv(gt.n, 256), c.S), 256, 0, ctx->stream>>>(c, k, ctx->d_kin_color, ctx->d_kin_depth, ctx->slots[*slot].dev);
    KCHECK();
    ctx->slots[*slot].quad_valid = false;
    return DVO_OK;
}

extern "C" int dvo_frame_release(dvo_ctx* ctx, int slot)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    ctx->slots[slot].in_use = false;
    return DVO_OK;
}

static int set_depth_impl(dvo_ctx* ctx, int slot, const float* depth, const float* sigma, const float* age, int full_res)
{
    const Consts& c = ctx->c;
    const LevelGeom& gt = c.lv[c.L - 1];
    const int sw = full_res ? c.in_w : gt.w, sh = full_res ? c.in_h : gt.h;
    const float* src[3] = {depth, sigma, age};
    const float* dev[3] = {nullptr, nullptr, nullptr};
    for (int k = 0; k < 3; k++)
        if (src[k]) {
            CK(cudaMemcpyAsync(ctx->d_stage[k], src[k], sizeof(float) * (size_t)c.S * sw * sh, cudaMemcpyHostToDevice, ctx->stream));
            dev[k] = ctx->d_stage[k];
        }
    dim3 grid(cdiv(gt.n, 256), c.S);
    k_set_top<<<grid, 256, 0, ctx->stream>>>(c, ctx->d_slots, slot, slot < 0 ? ctx->d_vo : nullptr, dev[0], dev[1], dev[2], sw, sh,
                                              full_res ? c.culls : 0);
    KCHECK();
    CK(cudaStreamSynchronize(ctx->stream)); // host buffers may be pageable: do not return before the copies finish
    return DVO_OK;
}

extern "C" int dvo_frame_set_depth(dvo_ctx* ctx, int slot, const float* depth, const float* sigma, const float* age, int full_res)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    return set_depth_impl(ctx, slot, depth, sigma, age, full_res);
}

extern "C" int dvo_frame_download(dvo_ctx* ctx, int slot, int level, int kind, float* out)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(level >= 0 && level < c.L && out, "bad level");
    ARG(kind >= 0 && kind <= 5, "bad image kind");
    ARG(kind != DVO_IMG_AGE || level == c.L - 1, "age exists on the finest level only");
    const LevelGeom& g = c.lv[level];
    dim3 grid(cdiv(g.n, 256), c.S);
    k_extract<<<grid, 256, 0, ctx->stream>>>(c, ctx->slots[slot].dev, level, kind, ctx->d_stage[0]);
    KCHECK();
    CK(cudaMemcpyAsync(out, ctx->d_stage[0], sizeof(float) * (size_t)c.S * g.n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_visualize(dvo_ctx* ctx, int slot, int level, int kind, uint8_t* bgr_out)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(level >= 0 && level < c.L && bgr_out, "bad level");
    ARG(kind >= DVO_VIS_GRAY && kind <= DVO_VIS_AGE, "bad visualisation kind");
    ARG(kind != DVO_VIS_AGE || level == c.L - 1, "age exists on the finest level only");
    const LevelGeom& g = c.lv[level];
    uint8_t* d_out = reinterpret_cast<uint8_t*>(ctx->d_stage[0]); // 3 B per pixel of a level fit the 4 B per input pixel staged there
    dim3 grid(cdiv(g.n, 256), c.S);
    k_visualize<<<grid, 256, 0, ctx->stream>>>(c, ctx->slots[slot].dev, level, kind, d_out);
    KCHECK();
    CK(cudaMemcpyAsync(bgr_out, d_out, (size_t)3 * c.S * g.n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

// Draw::visualize*(image) on raw host images, any size: no context needed (temporary device buffers on `device`'s default stream)
extern "C" int dvo_visualize_image(int kind, const float* src, const float* aux, int width, int height, uint8_t* bgr_out)
{
    if (!src || !bgr_out || width <= 0 || height <= 0 || kind < DVO_VIS_GRAY || kind > DVO_VIS_AGE) return DVO_ERR_ARG;
    if (kind == DVO_VIS_DEPTH_SIGMA && !aux) return DVO_ERR_ARG;
    const size_t n = (size_t)width * height;
    float *d_src = nullptr, *d_aux = nullptr;
    uint8_t* d_out = nullptr;
    bool ok = cudaMalloc(&d_src, n * 4) == cudaSuccess && cudaMalloc(&d_out, n * 3) == cudaSuccess;
    const bool use_aux = (kind == DVO_VIS_DEPTH_SIGMA);
    if (ok && use_aux) ok = cudaMalloc(&d_aux, n * 4) == cudaSuccess;
    if (ok) ok = cudaMemcpy(d_src, src, n * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok && use_aux) ok = cudaMemcpy(d_aux, aux, n * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok) {
        k_visualize_image<<<(unsigned)((n + 255) / 256), 256>>>(kind, d_src, d_aux, n, d_out);
        ok = cudaGetLastError() == cudaSuccess && cudaMemcpy(bgr_out, d_out, n * 3, cudaMemcpyDeviceToHost) == cudaSuccess;
    }
    cudaFree(d_src); cudaFree(d_aux); cudaFree(d_out);
    return ok ? DVO_OK : DVO_ERR_CUDA;
}

// ---- scalar geometry and the depth filter of the reference's boundary, host arithmetic ---------------------------------------
// Compiled here once (x86-64 baseline: no FMA contraction) so that callers' own compiler flags cannot change the rounding.
extern "C" int dvo_transform_point(const float T[16], const float x[3], float out[3])
{   // Transform::transform, transform.cpp:7-12: R x + t, fp32 left to right
    if (!T || !x || !out) return DVO_ERR_ARG;
    const volatile float X = x[0], Y = x[1], Z = x[2];
    for (int r = 0; r < 3; r++) {
        volatile float a = T[r * 4 + 0] * X; volatile float b = T[r * 4 + 1] * Y; volatile float c2 = T[r * 4 + 2] * Z;
        volatile float s = a + b; s = s + c2;
        out[r] = s + T[r * 4 + 3];
    }
    return DVO_OK;
}
extern "C" int dvo_project_point(const float K[9], const float p[3], float out[2])
{   // Transform::project, transform.cpp:20-23
    if (!K || !p || !out) return DVO_ERR_ARG;
    volatile float a = p[0] * K[0]; a = a / p[2]; out[0] = a + K[2];
    volatile float b = p[1] * K[4]; b = b / p[2]; out[1] = b + K[5];
    return DVO_OK;
}
extern "C" int dvo_back_project_point(const float K[9], const float xy[2], float depth, float out[3])
{   // Transform::backProject, transform.cpp:25-28
    if (!K || !xy || !out) return DVO_ERR_ARG;
    volatile float a = xy[0] - K[2]; a = depth * a; out[0] = a / K[0];
    volatile float b = xy[1] - K[5]; b = depth * b; out[1] = b / K[4];
    out[2] = depth;
    return DVO_OK;
}
namespace {
std::mt19937 g_gauss_engine; // the file-scope engine of src/math/gaussian.cpp:8 for the host-side math::Gaussian
bool gaussian_gate(float depth, float sigma, float d, float s, float& v1, float& v2, float& v)
{
    volatile float a = sigma * sigma, b = s * s;
    v1 = a; v2 = b;
    volatile float sum = a + b;
    v = sum;
    const float diff = std::fabs(d - depth);
    const float m = d < diff ? d : diff;
    volatile float q = m / 0.8f; q = q * 0.5f;
    const float gain = ((double)m < 0.8) ? 0.5f + q : 1.0f;
    const float mx = sigma < s ? s : sigma;
    volatile float lim = gain * mx;
    return !(diff > lim);
}
void gaussian_merge(float& depth, float& sigma, float d, float v1, float v2, float v)
{
    volatile float a = v2 * depth, b = v1 * d;
    volatile float num = a + b;
    depth = num / v;
    volatile float p = v1 * v2; p = p / v;
    sigma = std::sqrt(p);
}
} // namespace
extern "C" int dvo_gaussian_fuse(float* depth, float* sigma, float d, float s)
{   // Gaussian::operator(), gaussian.cpp:33-50: 1 = fused, 0 = refused (state untouched)
    if (!depth || !sigma) return DVO_ERR_ARG;
    float v1, v2, v;
    if (!gaussian_gate(*depth, *sigma, d, s, v1, v2, v)) return 0;
    gaussian_merge(*depth, *sigma, d, v1, v2, v);
    return 1;
}
extern "C" int dvo_gaussian_update(float* depth, float* sigma, float d, float s)
{   // Gaussian::update, gaussian.cpp:12-31: a refused observation resets the hypothesis from the process-wide mt19937
    if (!depth || !sigma) return DVO_ERR_ARG;
    float v1, v2, v;
    if (!gaussian_gate(*depth, *sigma, d, s, v1, v2, v)) {
        // libstdc++ uniform_real_distribution<float>(2.0, 0.5): one 32-bit draw, (raw * 2^-32) * (0.5 - 2.0) + 2.0
        volatile float u = (float)(uint32_t)g_gauss_engine() / 4294967296.0f;
        if (u >= 1.0f) u = 0.99999994f;
        volatile float r = u * (0.5f - 2.0f); r = r + 2.0f;
        *depth = 4.0f < r ? 4.0f : r;
        *sigma = 0.5f;
        return 0;
    }
    gaussian_merge(*depth, *sigma, d, v1, v2, v);
    return 1;
}
extern "C" void dvo_gaussian_reseed(void) { g_gauss_engine = std::mt19937(); }

// Transform::mapDepthtoGray on raw host images (no context: temporary device buffers)
extern "C" int dvo_map_depth_to_gray(const float* depth, int w, int h, const float* gray, int gw, int gh, const float rgb_K[9],
                                     const float depth_K[9], const float invT[16], float* mapped, float* sigma)
{
    if (!depth || !gray || !rgb_K || !depth_K || !invT || !mapped || !sigma || w <= 0 || h <= 0 || gw <= 0 || gh <= 0) return DVO_ERR_ARG;
    MapD2G P;
    P.w = w; P.h = h; P.gw = gw; P.gh = gh;
    P.dfx = depth_K[0]; P.dfy = depth_K[4]; P.dcx = depth_K[2]; P.dcy = depth_K[5];
    P.fx = rgb_K[0]; P.fy = rgb_K[4]; P.cx = rgb_K[2]; P.cy = rgb_K[5];
    for (int r = 0; r < 3; r++)
        for (int q = 0; q < 4; q++) P.T[r * 4 + q] = invT[r * 4 + q];
    const size_t n = (size_t)w * h, gn = (size_t)gw * gh;
    float *d_d = nullptr, *d_g = nullptr, *d_m = nullptr, *d_s = nullptr;
    bool ok = cudaMalloc(&d_d, n * 4) == cudaSuccess && cudaMalloc(&d_g, gn * 4) == cudaSuccess && cudaMalloc(&d_m, n * 4) == cudaSuccess &&
              cudaMalloc(&d_s, n * 4) == cudaSuccess;
    if (ok) ok = cudaMemcpy(d_d, depth, n * 4, cudaMemcpyHostToDevice) == cudaSuccess && cudaMemcpy(d_g, gray, gn * 4, cudaMemcpyHostToDevice) == cudaSuccess;
    if (ok) {
        k_map_depth_to_gray<<<(unsigned)((n + 255) / 256), 256>>>(P, d_d, d_g, d_m, d_s);
        ok = cudaGetLastError() == cudaSuccess && cudaMemcpy(mapped, d_m, n * 4, cudaMemcpyDeviceToHost) == cudaSuccess &&
             cudaMemcpy(sigma, d_s, n * 4, cudaMemcpyDeviceToHost) == cudaSuccess;
    }
    cudaFree(d_d); cudaFree(d_g); cudaFree(d_m); cudaFree(d_s);
    return ok ? DVO_OK : DVO_ERR_CUDA;
}

// ---- headless outputs: host side ----------------------------------------------------------------
extern "C" int dvo_inverse_pose(const float T[16], float out[16])
{
    if (!T || !out) return DVO_ERR_ARG;
    float r[16] = {0};
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) r[i * 4 + j] = T[j * 4 + i];
        r[i * 4 + 3] = -T[i * 4 + 3];
    }
    memcpy(out, r, sizeof(r));
    return DVO_OK;
}

extern "C" int dvo_write_trajectory_csv(const char* path, const float* T, int n, int first_frame, int append)
{
    if (!path || (!T && n > 0) || n < 0) return DVO_ERR_ARG;
    FILE* f = fopen(path, append ? "a" : "w");
    if (!f) return DVO_ERR_ARG;
    for (int k = 0; k < n; k++) {
        fprintf(f, "%d", first_frame + k);
        for (int q = 0; q < 16; q++) fprintf(f, ",%.9g", (double)T[(size_t)k * 16 + q]);
        fputc('\n', f);
    }
    return fclose(f) == 0 ? DVO_OK : DVO_ERR_ARG;
}

namespace {
struct PngOut {
    FILE* f;
    uint32_t crc_table[256];
    PngOut(FILE* f_) : f(f_)
    {
        for (uint32_t n = 0; n < 256; n++) {
            uint32_t c = n;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            crc_table[n] = c;
        }
    }
    uint32_t crc(uint32_t c, const uint8_t* p, size_t n) const
    {
        for (size_t i = 0; i < n; i++) c = crc_table[(c ^ p[i]) & 0xff] ^ (c >> 8);
        return c;
    }
    static void be32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }
    bool chunk(const char type[4], const std::vector<uint8_t>& data) const
    {
        uint8_t hd[8], tl[4];
        be32(hd, (uint32_t)data.size());
        memcpy(hd + 4, type, 4);
        uint32_t c = crc(0xffffffffu, hd + 4, 4);
        c = crc(c, data.data(), data.size()) ^ 0xffffffffu;
        be32(tl, c);
        return fwrite(hd, 1, 8, f) == 8 && (data.empty() || fwrite(data.data(), 1, data.size(), f) == data.size()) && fwrite(tl, 1, 4, f) == 4;
    }
};
} // namespace

extern "C" int dvo_write_png(const char* path, const uint8_t* bgr, int width, int height)
{
    if (!path || !bgr || width <= 0 || height <= 0) return DVO_ERR_ARG;
    // raw scanlines: filter byte 0 + RGB
    const size_t row = 1 + (size_t)3 * width;
    std::vector<uint8_t> raw(row * height);
    for (int y = 0; y < height; y++) {
        uint8_t* d = raw.data() + row * y;
        *d++ = 0;
        const uint8_t* s = bgr + (size_t)3 * width * y;
        for (int x = 0; x < width; x++) { d[0] = s[2]; d[1] = s[1]; d[2] = s[0]; d += 3; s += 3; }
    }
    // zlib stream of stored blocks
    std::vector<uint8_t> z;
    z.reserve(raw.size() + raw.size() / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    for (size_t pos = 0; pos < raw.size();) {
        const size_t n = raw.size() - pos < 65535 ? raw.size() - pos : 65535;
        z.push_back(pos + n == raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 0xff)); z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xff)); z.push_back((uint8_t)((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = 0; i < n; i++) { a = (a + raw[pos + i]) % 65521u; b = (b + a) % 65521u; }
        pos += n;
    }
    uint8_t ad[4];
    PngOut::be32(ad, (b << 16) | a);
    z.insert(z.end(), ad, ad + 4);
    FILE* f = fopen(path, "wb");
    if (!f) return DVO_ERR_ARG;
    PngOut png(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::vector<uint8_t> ihdr(13);
    PngOut::be32(ihdr.data(), (uint32_t)width); PngOut::be32(ihdr.data() + 4, (uint32_t)height);
    ihdr[8] = 8; ihdr[9] = 2; ihdr[10] = 0; ihdr[11] = 0; ihdr[12] = 0; // 8-bit RGB, deflate, no filter, no interlace
    bool ok = fwrite(sig, 1, 8, f) == 8 && png.chunk("IHDR", ihdr) && png.chunk("IDAT", z) && png.chunk("IEND", std::vector<uint8_t>());
    ok = (fclose(f) == 0) && ok;
    return ok ? DVO_OK : DVO_ERR_ARG;
}

extern "C" int dvo_frame_set_xi(dvo_ctx* ctx, int slot, const float* xi, const float* rel_xi)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const size_t b = sizeof(float) * 6 * (size_t)ctx->c.S;
    if (xi) CK(cudaMemcpyAsync(ctx->slots[slot].dev.xi, xi, b, cudaMemcpyHostToDevice, ctx->stream));
    if (rel_xi) CK(cudaMemcpyAsync(ctx->slots[slot].dev.rel_xi, rel_xi, b, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_frame_get_xi(dvo_ctx* ctx, int slot, float* xi, float* rel_xi)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const size_t b = sizeof(float) * 6 * (size_t)ctx->c.S;
    if (xi) CK(cudaMemcpyAsync(xi, ctx->slots[slot].dev.xi, b, cudaMemcpyDeviceToHost, ctx->stream));
    if (rel_xi) CK(cudaMemcpyAsync(rel_xi, ctx->slots[slot].dev.rel_xi, b, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

// ---- tracking -------------------------------------------------------------------------------
extern "C" int dvo_warp_image(dvo_ctx* ctx, int ref_slot, int level, const float* xi, float* out)
{
    int rc = check_slot(ctx, ref_slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(level >= 0 && level < c.L && xi && out, "bad argument");
    const LevelGeom& g = c.lv[level];
    CK(cudaMemcpyAsync(ctx->d_small, xi, sizeof(float) * 6 * (size_t)c.S, cudaMemcpyHostToDevice, ctx->stream));
    dim3 grid(cdiv(g.n, 256), c.S);
    k_warp_image<<<grid, 256, 0, ctx->stream>>>(c, ctx->slots[ref_slot].dev, level, ctx->d_small, ctx->d_stage[0]);
    KCHECK();
    CK(cudaMemcpyAsync(out, ctx->d_stage[0], sizeof(float) * (size_t)c.S * g.n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

// One k_track launch = a whole Tracker::track (or one Track::optimize step) for every stream.
static int launch_track(dvo_ctx* ctx, TrackArgs a)
{
    const Consts& c = ctx->c;
    a.st = ctx->d_ts; a.partials = ctx->d_partials; a.B = ctx->trk_B;
    a.log = ctx->d_log; a.log_cap = ctx->log_cap;
    a.slots = ctx->d_slots;
    a.status = ctx->d_status;
    dim3 grid(ctx->trk_B, c.S);
    KTimer t(ctx, K_TRACK);
    if (ctx->trk_B > 1) { // CTAs of a stream meet at a global barrier: they must be co-resident
        Consts cc = c;
        void* params[] = {(void*)&cc, (void*)&a};
        CK(cudaLaunchCooperativeKernel((const void*)k_track_staged, grid, dim3(TRK_THREADS), params, 0, ctx->stream));
        ctx->launches++;
    } else {
        k_track<<<grid, TRK_THREADS, 0, ctx->stream>>>(c, a);
        KCHECK();
    }
    return DVO_OK;
}

extern "C" int dvo_kernel_timing(dvo_ctx* ctx, int enable, dvo_kernel_time* out, int cap, int* n_out)
{
    if (enable) {
        ctx->ktiming = true;
        ctx->k_ev_used = 0;
        return DVO_OK;
    }
    ctx->ktiming = false;
    CK(cudaStreamSynchronize(ctx->stream));
    float tot[K_COUNT] = {0};
    int cnt[K_COUNT] = {0};
    for (size_t k = 0; k + 1 < ctx->k_ev_used; k += 2) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, ctx->k_ev[k], ctx->k_ev[k + 1]));
        const int id = ctx->k_ev_id[k / 2];
        tot[id] += ms; cnt[id]++;
    }
    int n = 0;
    for (int id = 0; id < K_COUNT && out && n < cap; id++) {
        if (!cnt[id]) continue;
        memset(&out[n], 0, sizeof(out[n]));
        strncpy(out[n].name, kernel_names[id], sizeof(out[n].name) - 1);
        out[n].total_ms = tot[id]; out[n].launches = cnt[id];
        n++;
    }
    if (n_out) *n_out = n;
    ctx->k_ev_used = 0;
    return DVO_OK;
}

#ifdef DVO_TRK_PROF
// experiment builds only (scripts/trk_phases.py): per-phase nanoseconds of stream 0 inside k_track
extern "C" int dvo_debug_trk_prof(dvo_ctx* ctx, unsigned long long* out, int reset)
{
    CK(cudaStreamSynchronize(ctx->stream));
    if (out) CK(cudaMemcpyFromSymbol(out, g_trk_prof, sizeof(unsigned long long) * 8));
    if (reset) { unsigned long long z[8] = {0}; CK(cudaMemcpyToSymbol(g_trk_prof, z, sizeof(z))); }
    return DVO_OK;
}
extern "C" int dvo_debug_trk_streams(dvo_ctx* ctx, unsigned int* out, int n)
{
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpyFromSymbol(out, g_trk_stream_ns, sizeof(unsigned int) * (size_t)(n < 1024 ? n : 1024)));
    return DVO_OK;
}
#endif

extern "C" int dvo_gn_timing(dvo_ctx* ctx, int enable, float* total_ms, int* launches)
{
    if (enable) return dvo_kernel_timing(ctx, 1, nullptr, 0, nullptr);
    dvo_kernel_time kt[K_COUNT];
    int n = 0;
    int rc = dvo_kernel_timing(ctx, 0, kt, K_COUNT, &n);
    if (rc) return rc;
    if (total_ms) *total_ms = 0;
    if (launches) *launches = 0;
    for (int k = 0; k < n && k < K_COUNT; k++)
        if (!strncmp(kt[k].name, "k_track", sizeof(kt[k].name))) {
            if (total_ms) *total_ms = kt[k].total_ms;
            if (launches) *launches = kt[k].launches;
        }
    return DVO_OK;
}

extern "C" int dvo_gn_step(dvo_ctx* ctx, int obj_slot, int ref_slot, int level, const float* xi, float* upd, float* residual,
                           int* valid, uint8_t* mask)
{
    int rc = check_slot(ctx, obj_slot);
    if (rc) return rc;
    rc = check_slot(ctx, ref_slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(level >= 0 && level < c.L && xi, "bad argument");
    const LevelGeom& g = c.lv[level];
    CK(cudaMemcpyAsync(ctx->d_small, xi, sizeof(float) * 6 * (size_t)c.S, cudaMemcpyHostToDevice, ctx->stream));
    if (mask) CK(cudaMemsetAsync(ctx->d_mask, 0, (size_t)c.S * g.n, ctx->stream));
    TrackArgs a{};
    a.mask = mask ? ctx->d_mask : nullptr;
    a.obj_slot = obj_slot; a.ref_slot = ref_slot; a.mode = 1; a.level0 = level; a.xi0 = ctx->d_small;
    rc = launch_track(ctx, a);
    if (rc) return rc;
    dvo_iter_log* hl = (dvo_iter_log*)ctx->h_pinned;
    CK(cudaMemcpy2DAsync(hl, sizeof(dvo_iter_log), ctx->d_log, sizeof(dvo_iter_log) * ctx->log_cap, sizeof(dvo_iter_log), c.S,
                         cudaMemcpyDeviceToHost, ctx->stream));
    if (mask) CK(cudaMemcpyAsync(mask, ctx->d_mask, (size_t)c.S * g.n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) {
        if (upd) memcpy(upd + 6 * s, hl[s].update, sizeof(float) * 6);
        if (residual) residual[s] = hl[s].residual;
        if (valid) valid[s] = hl[s].valid_pixels;
    }
    return DVO_OK;
}

static int enqueue_track(dvo_ctx* ctx, int obj_slot, int ref_slot, bool use_vo, double stop_update, float stop_residual, int frame_id,
                         const float* forced_rel, float* T_out_dev = nullptr, bool overlapped = false, bool fused_propagate = false)
{
    if (stop_update < 0) stop_update = (double)ctx->P.min_update; // cv::norm(...) < MINIMUM_UPDATE: float promoted
    if (stop_residual < 0) stop_residual = ctx->P.min_residual;
    TrackArgs a{};
    a.obj_slot = obj_slot; a.ref_slot = ref_slot; a.mode = 0; a.level0 = 0; a.xi0 = nullptr;
    a.stop_update = stop_update; a.stop_residual = stop_residual;
    if (use_vo) {
        a.vo = ctx->d_vo; a.ms = ctx->d_ms; a.hist = ctx->d_hist; a.tab = ctx->d_tab;
        a.frame_id = frame_id; a.forced_rel = forced_rel;
        a.T_out = T_out_dev;
        a.rq = overlapped ? ctx->d_rq : nullptr;
        a.winner = fused_propagate ? ctx->d_winner : nullptr;
    }
    return launch_track(ctx, a);
}

extern "C" int dvo_track(dvo_ctx* ctx, int obj_slot, int ref_slot, float* xi_out, dvo_iter_log* log, int log_cap, int* n_iters,
                         double stop_update, float stop_residual)
{
    int rc = check_slot(ctx, obj_slot);
    if (rc) return rc;
    rc = check_slot(ctx, ref_slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    rc = enqueue_track(ctx, obj_slot, ref_slot, false, stop_update, stop_residual, 0, nullptr);
    if (rc) return rc;
    TrackState* hts = (TrackState*)((char*)ctx->h_pinned + sizeof(dvo_iter_log) * (size_t)c.S * ctx->log_cap);
    CK(cudaMemcpyAsync(hts, ctx->d_ts, sizeof(TrackState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    dvo_iter_log* hl = (dvo_iter_log*)ctx->h_pinned;
    if (log) CK(cudaMemcpyAsync(hl, ctx->d_log, sizeof(dvo_iter_log) * (size_t)c.S * ctx->log_cap, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) {
        if (xi_out) memcpy(xi_out + 6 * s, hts[s].xi, sizeof(float) * 6);
        if (n_iters) n_iters[s] = hts[s].n_total;
        if (log)
            for (int k = 0; k < hts[s].n_total && k < log_cap && k < ctx->log_cap; k++) log[(size_t)s * log_cap + k] = hl[(size_t)s * ctx->log_cap + k];
    }
    return DVO_OK;
}

// ---- mapping --------------------------------------------------------------------------------
__global__ void k_map_config(Consts c, MapState* mst, int ref_slot, int obj_slot, int new_slot, int n_hist, int do_p, int do_u, int do_r,
                             int reg_slot)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    MapState* m = mst + s;
    m->ref_slot = ref_slot; m->obj_slot = obj_slot; m->new_slot = new_slot; m->n_hist = n_hist;
    m->do_propagate = do_p; m->do_update = do_u; m->do_regularize = do_r; m->reg_slot = reg_slot;
}
__global__ void k_hist_fill(Consts c, int* hist, const int* list, int n)
{
    const int s = blockIdx.x;
    for (int k = threadIdx.x; k < n; k += blockDim.x) hist[(size_t)s * c.kf_stride + k] = list[k];
}

// Levels up to this many pixels take the fused forms: Implement::propagate in k_track's epilogue (one CTA per stream) and
// Frame::updateDepth in k_regularize's last CTA (one CTA copies the stream's whole level).  Larger levels (1080p, 4K single
// streams) keep the wide two-kernel forms.
static const int FUSE_MAX_PIXELS = 64 * 1024;
static bool small_level(const dvo_ctx* ctx) { return ctx->c.lv[ctx->c.L - 1].n <= FUSE_MAX_PIXELS; }

static int enqueue_propagate(dvo_ctx* ctx)
{
    const Consts& c = ctx->c;
    const int ntop = c.lv[c.L - 1].n;
    dim3 grid(cdiv(ntop, 256), c.S);
    {
        KTimer t(ctx, K_PROP_SCATTER);
        k_propagate_scatter<<<grid, 256, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_ms, ctx->d_winner);
        KCHECK();
    }
    {
        KTimer t(ctx, K_PROP_GATHER);
        k_propagate_gather<<<grid, 256, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_ms, ctx->d_winner);
        KCHECK();
    }
    return DVO_OK;
}

// mt19937 replay table (src/math/gaussian.cpp:8-9): the reference's engine is unbounded, so the table grows on demand.  Every
// Mapper::update can consume at most roi.n draws per stream; `rng_bound` bounds every stream's position from above.
static int rng_reserve(dvo_ctx* ctx, uint64_t need)
{
    Consts& c = ctx->c;
    if (need <= (uint64_t)c.rng_table) return DVO_OK;
    uint64_t len = (uint64_t)c.rng_table * 4; // growing costs a device synchronisation: in big steps
    while (len < need) len *= 2;
    if (len > 0xffffffffull && need <= 0xffffffffull) len = 0xffffffffull;
    if (len > 0xffffffffull) {
        ctx->err = "mt19937 replay table would exceed 2^32 draws";
        return DVO_ERR_RNG;
    }
    std::vector<uint32_t> tail((size_t)(len - c.rng_table));
    for (auto& v : tail) v = (uint32_t)ctx->engine(); // the host engine continues where the table ends
    uint32_t* d_new = nullptr;
    CK(cudaMalloc(&d_new, sizeof(uint32_t) * (size_t)len));
    CK(cudaStreamSynchronize(ctx->stream)); // kernels in flight still read the old table
    CK(cudaMemcpy(d_new, ctx->d_rng, sizeof(uint32_t) * (size_t)c.rng_table, cudaMemcpyDeviceToDevice));
    CK(cudaMemcpy(d_new + c.rng_table, tail.data(), sizeof(uint32_t) * tail.size(), cudaMemcpyHostToDevice));
    cudaFree(ctx->d_rng);
    ctx->d_rng = d_new;
    c.rng_table = (unsigned int)len;
    return DVO_OK;
}

// tables: the granular entry point computes Mapper::update's pose table here; in the VO loop the
// tracking kernel's epilogue already did.  overlapped: launched as a programmatic dependent of the k_track launch right before
// it, so that its CTAs start on whatever the tracking kernel leaves free and serve streams in tracking-completion order.
static int enqueue_update(dvo_ctx* ctx, bool tables, bool overlapped)
{
    const Consts& c = ctx->c;
    if (tables) {
        dim3 tgrid(cdiv(c.max_kf, 64), c.S);
        k_pose_tables<<<tgrid, 64, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_ms, ctx->d_hist, ctx->d_tab);
        KCHECK();
    }
    if (ctx->roi.n <= 0) return DVO_OK;
    dim3 grid(ctx->roi.nblk, c.S);
    const int fold = ctx->roi.nblk <= 4096 ? 1 : 0; // the stream's last phase-A CTA scans the reject counts itself
    {
        KTimer t(ctx, K_MAP_UPDATE);
        if (overlapped) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = grid; cfg.blockDim = dim3(MAPU_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = ctx->stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            CK(cudaLaunchKernelEx(&cfg, k_map_update, c, ctx->roi, (const SlotDev*)ctx->d_slots, ctx->d_ms, (const int*)ctx->d_hist,
                                  (const PoseTab*)ctx->d_tab, ctx->d_flags, ctx->d_blockcnt, ctx->d_rq, ctx->d_status, fold));
            ctx->launches++;
        } else {
            k_map_update<<<grid, MAPU_THREADS, 0, ctx->stream>>>(c, ctx->roi, ctx->d_slots, ctx->d_ms, ctx->d_hist, ctx->d_tab, ctx->d_flags,
                                                                ctx->d_blockcnt, nullptr, ctx->d_status, fold);
            KCHECK();
        }
    }
    {
        KTimer t(ctx, K_MAP_APPLY);
        if (!fold) {
            k_map_scan<<<c.S, 1024, 0, ctx->stream>>>(c, ctx->roi, ctx->d_ms, ctx->d_blockcnt);
            KCHECK();
        }
        dim3 agrid(cdiv(ctx->roi.nblk, APPLY_SUB), c.S);
        k_map_update_apply<<<agrid, MAPU_THREADS * APPLY_SUB, 0, ctx->stream>>>(c, ctx->roi, ctx->d_slots, ctx->d_ms, ctx->d_flags,
                                                                                 ctx->d_blockcnt, ctx->d_rng, ctx->d_status);
        KCHECK();
    }
    return DVO_OK;
}
static int enqueue_regularize(dvo_ctx* ctx, bool vo)
{
    const Consts& c = ctx->c;
    const int ntop = c.lv[c.L - 1].n;
    {
        KTimer t(ctx, K_REGULARIZE);
        k_regularize<<<dim3(cdiv(ntop, 256), c.S), 256, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_ms, ctx->d_tmp, vo ? 1 : 0);
        KCHECK();
    }
    {
        KTimer t(ctx, K_REG_COMMIT);
        k_regularize_commit<<<dim3(cdiv(ntop, 512), c.S), 256, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_ms, ctx->d_tmp);
        KCHECK();
    }
    return DVO_OK;
}

// the 2x2-footprint image of a granular slot that is about to serve as a born keyframe (VO pool slots are kept valid on the device)
static int ensure_quad(dvo_ctx* ctx, int slot)
{
    const Consts& c = ctx->c;
    SlotHost& sh = ctx->slots[slot];
    if (slot >= c.vo_base || sh.quad_valid) return DVO_OK;
    k_build_quad<<<dim3(cdiv(c.lv[c.L - 1].n, 256), c.S), 256, 0, ctx->stream>>>(c, ctx->d_slots, slot, nullptr);
    KCHECK();
    sh.quad_valid = true;
    return DVO_OK;
}

static int map_error(dvo_ctx* ctx)
{   // first error flagged by any stream
    const Consts& c = ctx->c;
    MapState* hm = (MapState*)ctx->h_pinned;
    if (sizeof(MapState) * (size_t)c.S > ctx->h_pinned_bytes) return DVO_OK;
    CK(cudaMemcpyAsync(hm, ctx->d_ms, sizeof(MapState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    {   // nothing is in flight: the replay-table bound becomes the furthest actual position
        uint64_t mx = 0;
        for (int s = 0; s < c.S; s++) mx = hm[s].rng_pos > mx ? hm[s].rng_pos : mx;
        ctx->rng_bound = mx;
    }
    for (int s = 0; s < c.S; s++)
        if (hm[s].err) {
            ctx->err = hm[s].err == DVO_ERR_HISTORY ? "age >= history size (FrameHistory::operator[] would throw)" : "mt19937 replay table exhausted";
            return hm[s].err;
        }
    return DVO_OK;
}

extern "C" int dvo_map_propagate(dvo_ctx* ctx, int ref_slot, int new_slot, const float* rel_xi)
{
    int rc = check_slot(ctx, ref_slot);
    if (rc) return rc;
    rc = check_slot(ctx, new_slot);
    if (rc) return rc;
    ARG(rel_xi && ref_slot != new_slot, "bad argument");
    const Consts& c = ctx->c;
    CK(cudaMemcpyAsync(ctx->d_small, rel_xi, sizeof(float) * 6 * (size_t)c.S, cudaMemcpyHostToDevice, ctx->stream));
    k_map_config<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_ms, ref_slot, new_slot, new_slot, 0, 1, 0, 0, new_slot);
    KCHECK();
    k_map_prepare<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_ms, ctx->d_small);
    KCHECK();
    rc = enqueue_propagate(ctx);
    if (rc) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_map_update(dvo_ctx* ctx, int obj_slot, int ref_slot, const int* history_slots, int n_hist, int* valid_updates,
                              uint8_t* mask)
{
    int rc = check_slot(ctx, obj_slot);
    if (rc) return rc;
    rc = check_slot(ctx, ref_slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(history_slots && n_hist >= 1 && n_hist <= c.max_kf, "bad history");
    for (int k = 0; k < n_hist; k++) {
        rc = check_slot(ctx, history_slots[k]);
        if (rc) return rc;
        rc = ensure_quad(ctx, history_slots[k]);
        if (rc) return rc;
    }
    // hist is stored in push order (oldest first); history_slots is indexed by age (newest first)
    std::vector<int> order(n_hist);
    for (int a = 0; a < n_hist; a++) order[n_hist - 1 - a] = history_slots[a];
    CK(cudaMemcpyAsync(ctx->d_ismall, order.data(), sizeof(int) * n_hist, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    k_hist_fill<<<c.S, 64, 0, ctx->stream>>>(c, ctx->d_hist, ctx->d_ismall, n_hist);
    KCHECK();
    k_map_config<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_ms, ref_slot, obj_slot, obj_slot, n_hist, 0, 1, 0, ref_slot);
    KCHECK();
    k_map_prepare<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_ms, ctx->slots[obj_slot].dev.rel_xi);
    KCHECK();
    ctx->rng_bound += (uint64_t)(ctx->roi.n > 0 ? ctx->roi.n : 0);
    rc = rng_reserve(ctx, ctx->rng_bound);
    if (rc) return rc;
    rc = enqueue_update(ctx, true, false); // ref->updateDepthSigma (mapper.cpp:135) is implicit: coarser levels are views
    if (rc) return rc;
    const int ntop = c.lv[c.L - 1].n;
    if (mask) CK(cudaMemcpyAsync(mask, ctx->d_flags, (size_t)c.S * ntop, cudaMemcpyDeviceToHost, ctx->stream));
    rc = map_error(ctx);
    if (rc) return rc;
    if (valid_updates) {
        MapState* hm = (MapState*)ctx->h_pinned;
        for (int s = 0; s < c.S; s++) valid_updates[s] = hm[s].valid_update;
    }
    return DVO_OK;
}

extern "C" int dvo_map_regularize(dvo_ctx* ctx, int slot)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    k_map_config<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_ms, slot, slot, slot, 0, 0, 0, 1, slot);
    KCHECK();
    rc = enqueue_regularize(ctx, false);
    if (rc) return rc;
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_rng_position(dvo_ctx* ctx, uint64_t* pos)
{
    const Consts& c = ctx->c;
    MapState* hm = (MapState*)ctx->h_pinned;
    CK(cudaMemcpyAsync(hm, ctx->d_ms, sizeof(MapState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) pos[s] = hm[s].rng_pos;
    return DVO_OK;
}

// Scratch device buffer of the granular entry points below (freed on every exit path).
struct DevScratch {
    void* p = nullptr;
    ~DevScratch() { if (p) cudaFree(p); }
};

extern "C" int dvo_map_update_points(dvo_ctx* ctx, int obj_slot, int born_slot, const float* relative_xi, int n, const int* px,
                                     const int* py, const float* depth, const float* sigma, float* out_depth, float* out_sigma)
{
    int rc = check_slot(ctx, obj_slot);
    if (rc) return rc;
    rc = check_slot(ctx, born_slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(relative_xi && n > 0 && px && py && depth && sigma && out_depth && out_sigma, "bad argument");
    rc = ensure_quad(ctx, born_slot);
    if (rc) return rc;
    const size_t N = (size_t)c.S * n;
    DevScratch sc;
    CK(cudaMalloc(&sc.p, N * 24 + sizeof(float) * 6 * (size_t)c.S));
    int* d_px = (int*)sc.p; int* d_py = d_px + N;
    float* d_d = (float*)(d_py + N); float* d_s = d_d + N; float* d_od = d_s + N; float* d_os = d_od + N; float* d_xi = d_os + N;
    CK(cudaMemcpyAsync(d_px, px, N * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_py, py, N * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_d, depth, N * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_s, sigma, N * 4, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(d_xi, relative_xi, sizeof(float) * 6 * (size_t)c.S, cudaMemcpyHostToDevice, ctx->stream));
    k_update_points<<<dim3(cdiv(n, 64), c.S), 64, 0, ctx->stream>>>(c, ctx->slots[obj_slot].dev, ctx->slots[born_slot].dev, d_xi, n, d_px, d_py,
                                                                   d_d, d_s, d_od, d_os);
    KCHECK();
    CK(cudaMemcpyAsync(out_depth, d_od, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(out_sigma, d_os, N * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

// ---- se3 algebra ----------------------------------------------------------------------------
static int se3_op(dvo_ctx* ctx, int op, const float* a, int na, const float* b, float* out, int nout, int n)
{
    ARG(a && out && n > 0 && (op != 2 || b), "bad argument");
    DevScratch sc;
    const size_t ba = sizeof(float) * (size_t)na * n, bo = sizeof(float) * (size_t)nout * n;
    CK(cudaMalloc(&sc.p, 2 * ba + bo));
    float* d_a = (float*)sc.p; float* d_b = d_a + (size_t)na * n; float* d_o = d_b + (size_t)na * n;
    CK(cudaMemcpyAsync(d_a, a, ba, cudaMemcpyHostToDevice, ctx->stream));
    if (b) CK(cudaMemcpyAsync(d_b, b, ba, cudaMemcpyHostToDevice, ctx->stream));
    k_se3_ops<<<cdiv(n, 64), 64, 0, ctx->stream>>>(op, d_a, d_b, d_o, n);
    KCHECK();
    CK(cudaMemcpyAsync(out, d_o, bo, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_se3_exp(dvo_ctx* ctx, const float* xi, float* T, int n) { return se3_op(ctx, 0, xi, 6, nullptr, T, 16, n); }
extern "C" int dvo_se3_log(dvo_ctx* ctx, const float* T, float* xi, int n) { return se3_op(ctx, 1, T, 16, nullptr, xi, 6, n); }
extern "C" int dvo_se3_concatenate(dvo_ctx* ctx, const float* a, const float* b, float* out, int n) { return se3_op(ctx, 2, a, 6, b, out, 6, n); }

// ---- mask compaction ------------------------------------------------------------------------
// mask [S][n] on the device -> idx [S][n], count [S]; scratch from the context (d_blockcnt is sized for roi.nblk CTAs of 64, so the
// compaction keeps its own)
static int compact_device(dvo_ctx* ctx, const uint8_t* d_mask, int S, int n, int32_t* d_idx, int* d_total)
{
    const int nb = cdiv(n, 256);
    DevScratch sc;
    CK(cudaMalloc(&sc.p, sizeof(unsigned) * (size_t)S * nb));
    unsigned int* d_c = (unsigned int*)sc.p;
    k_compact_count<<<dim3(nb, S), 256, 0, ctx->stream>>>(d_mask, n, d_c);
    KCHECK();
    k_compact_scan<<<S, 256, 0, ctx->stream>>>(nb, d_c, d_total);
    KCHECK();
    k_compact_scatter<<<dim3(nb, S), 256, 0, ctx->stream>>>(d_mask, n, d_c, d_idx);
    KCHECK();
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_mask_compact(dvo_ctx* ctx, const uint8_t* mask, int n, int32_t* idx, int* count)
{
    ARG(mask && idx && count && n > 0, "bad argument");
    DevScratch m, i, t;
    CK(cudaMalloc(&m.p, n));
    CK(cudaMalloc(&i.p, sizeof(int32_t) * (size_t)n));
    CK(cudaMalloc(&t.p, sizeof(int)));
    CK(cudaMemcpyAsync(m.p, mask, n, cudaMemcpyHostToDevice, ctx->stream));
    int rc = compact_device(ctx, (const uint8_t*)m.p, 1, n, (int32_t*)i.p, (int*)t.p);
    if (rc) return rc;
    int total = 0;
    CK(cudaMemcpy(&total, t.p, sizeof(int), cudaMemcpyDeviceToHost));
    if (total) CK(cudaMemcpy(idx, i.p, sizeof(int32_t) * (size_t)total, cudaMemcpyDeviceToHost));
    *count = total;
    return DVO_OK;
}

// The same compaction for callers that may not allocate (code under CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): mask, idx and count are DEVICE buffers of the caller, and the per-CTA counts - the one piece of scratch the three
// kernels share - live in a file-scope device array instead of a cudaMalloc'ed buffer.  Every GPU that loads the library holds
// these 128 KiB until exit.  Nothing in the array has to start at zero: k_compact_count overwrites every count a call reads, so the
// previous call's contents (which stay there) are harmless.  All calls share the array: two calls that overlap on the GPU (different
// contexts or streams on one device) race; calls on one context are ordered by its stream.
extern "C" int dvo_mask_compact_noalloc(dvo_ctx* ctx, const uint8_t* mask_dev, int n, int32_t* idx_dev, int* count_dev)
{
    ARG(mask_dev && idx_dev && count_dev && n > 0, "bad argument");
    if (n > DVO_NOALLOC_MAX_N) {
        ctx->err = "dvo_mask_compact_noalloc: n = " + std::to_string(n) + " exceeds DVO_NOALLOC_MAX_N = " + std::to_string(DVO_NOALLOC_MAX_N) +
                   " (the static scratch holds one count per 256 mask bytes); use dvo_mask_compact";
        return DVO_ERR_ARG;
    }
    if (!ctx->d_noalloc) {
        ctx->err = "dvo_mask_compact_noalloc: the scratch array's device address was not resolved (dvo_create failed early)";
        return DVO_ERR_STATE;
    }
    const int nb = cdiv(n, 256);
    k_compact_count<<<dim3(nb, 1), 256, 0, ctx->stream>>>(mask_dev, n, ctx->d_noalloc);
    KCHECK();
    k_compact_scan<<<1, 256, 0, ctx->stream>>>(nb, ctx->d_noalloc, count_dev);
    KCHECK();
    k_compact_scatter<<<dim3(nb, 1), 256, 0, ctx->stream>>>(mask_dev, n, ctx->d_noalloc, idx_dev);
    KCHECK();
    return DVO_OK; // asynchronous on the context's stream: dvo_sync (or any synchronising call) before the results are read
}

extern "C" int dvo_select_pixels(dvo_ctx* ctx, int slot, int level, float min_gradient, int32_t* idx, int* count)
{
    int rc = check_slot(ctx, slot);
    if (rc) return rc;
    const Consts& c = ctx->c;
    ARG(level >= 0 && level < c.L && idx && count && min_gradient >= 0.0f, "bad argument");
    const LevelGeom& g = c.lv[level];
    DevScratch i, t;
    CK(cudaMalloc(&i.p, sizeof(int32_t) * (size_t)c.S * g.n));
    CK(cudaMalloc(&t.p, sizeof(int) * (size_t)c.S));
    k_select_mask<<<dim3(cdiv(g.n, 256), c.S), 256, 0, ctx->stream>>>(c, ctx->slots[slot].dev, level, min_gradient * min_gradient, ctx->d_mask);
    KCHECK();
    rc = compact_device(ctx, ctx->d_mask, c.S, g.n, (int32_t*)i.p, (int*)t.p);
    if (rc) return rc;
    CK(cudaMemcpy(count, t.p, sizeof(int) * (size_t)c.S, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(idx, i.p, sizeof(int32_t) * (size_t)c.S * g.n, cudaMemcpyDeviceToHost));
    return DVO_OK;
}

// ---- full VO --------------------------------------------------------------------------------
__global__ void k_vo_reset(Consts c, VoState* vo, MapState* mst)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    vo[s].kf_slot = -1; vo[s].kf_frame_id = -1; vo[s].n_kf = 0; vo[s].need_new = 0; vo[s].was_keyframe = 0; vo[s].n_iters = 0; vo[s].cur_slot = c.vo_base;
    vo[s].gn_iters = 0; vo[s].gn_pixels = 0; vo[s].n_propagate = 0; vo[s].n_update = 0; vo[s].map_matched = 0;
    mst[s].rng_pos = 0; mst[s].err = 0; mst[s].valid_update = 0; mst[s].ticket = 0; mst[s].ticket_a = 0; mst[s].ticket_r = 0;
}

extern "C" int dvo_vo_reset(dvo_ctx* ctx)
{
    const Consts& c = ctx->c;
    for (auto& s : ctx->slots) s.in_use = false;
    // a new sequence starts from fresh frames: no map, age or pose of the previous run may survive in the keyframe pool
    for (size_t k = (size_t)c.vo_base; k < ctx->slots.size(); k++)
        if (ctx->slots[k].base) CK(cudaMemsetAsync(ctx->slots[k].base, 0, ctx->slot_bytes, ctx->stream));
    ctx->vo_frame_id = -1;
    ctx->forced_pending = false;
    ctx->rng_bound = 0;
    CK(cudaStreamSynchronize(ctx->stream)); // read-backs of the previous sequence must not land after this point
    ctx->kf_known = 0; ctx->kf_known_frame = -1; ctx->rng_known = 0;
    ctx->kf_rb_pending[0] = ctx->kf_rb_pending[1] = false;
    CK(cudaMemsetAsync(ctx->d_status, 0, sizeof(int) * 4, ctx->stream));
    memset(ctx->h_status, 0, sizeof(int) * 8);
    k_vo_reset<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_vo, ctx->d_ms);
    KCHECK();
    return DVO_OK;
}

extern "C" int dvo_vo_reserve(dvo_ctx* ctx, int n_keyframes)
{
    const Consts& c = ctx->c;
    ARG(n_keyframes >= 0, "bad argument");
    const int need = n_keyframes < c.max_kf ? n_keyframes : c.max_kf;
    for (int k = 0; k <= need; k++) {
        if (!ctx->slots[c.vo_base + k].base) {
            int rc = alloc_slot(ctx, c.vo_base + k);
            if (rc) return rc;
        }
    }
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

extern "C" int dvo_vo_force_pose(dvo_ctx* ctx, const float* rel_xi)
{
    if (!rel_xi) {
        ctx->forced_pending = false;
        return DVO_OK;
    }
    CK(cudaMemcpyAsync(ctx->d_forced, rel_xi, sizeof(float) * 6 * (size_t)ctx->c.S, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->forced_pending = true;
    return DVO_OK;
}

// One odometrize for every stream, enqueued on the context's stream.  after_cull (nullable) is recorded
// once the staging buffer has been consumed; sync_out = false leaves the pose read-back asynchronous.
static int status_error(dvo_ctx* ctx, int code)
{
    ctx->err = code == DVO_ERR_HISTORY ? "a stream ran out of keyframe capacity (dvo_params.max_keyframes) or met an age >= history size: "
                                         "it is tracked but no longer mapped"
                                       : "device-side error";
    return code;
}

static int vo_push_impl(dvo_ctx* ctx, const float* gray_dev, const float* init_depth, const float* init_sigma, float* T_out,
                        cudaEvent_t after_cull, bool sync_out, int kind = 0, int aux = -1, int status_slot = 2)
{
    const Consts& c = ctx->c;
    ARG(gray_dev, "null gray");
    const int frame_id = ++ctx->vo_frame_id; // Frame::latest_id, frame.cpp:5
    // Keyframe pool: entry k holds keyframe k of every stream, entry n_kf the frame being processed, so this push can touch entries
    // 0 .. M where M = the largest keyframe count of any stream after the previous frame.  M is read back asynchronously
    // (d_status[1], written by the tracking epilogue); between read-backs it grows by at most one per frame.  Entries are allocated
    // on first need and never freed (FrameHistory is never pruned, frame.hpp:179-185).
    {
        for (int k = 0; k < 2; k++)
            if (ctx->kf_rb_pending[k] && cudaEventQuery(ctx->ev_kf[k]) == cudaSuccess) {
                ctx->kf_rb_pending[k] = false;
                if (ctx->kf_rb_frame[k] > ctx->kf_known_frame) {
                    ctx->kf_known = ctx->h_status[4 + 2 * k]; ctx->rng_known = (uint64_t)(unsigned)ctx->h_status[5 + 2 * k];
                    ctx->kf_known_frame = ctx->kf_rb_frame[k];
                }
            }
        int need = ctx->kf_known + (frame_id - 1 - ctx->kf_known_frame);
        if (need > frame_id) need = frame_id;
        if (need > c.max_kf) need = c.max_kf;
        for (int k = 0; k <= need; k++) {
            SlotHost& sh = ctx->slots[c.vo_base + k];
            if (!sh.base) {
                int rc = alloc_slot(ctx, c.vo_base + k);
                if (rc) return rc;
            }
            sh.in_use = true;
        }
    }
    if (frame_id == 0 && !init_depth) {
        // Scene's ctor draws the first keyframe's depth from cv::randn and sets sigma = 0.5 (frame.hpp:18-21); that prior is an INPUT
        // here.  Without it every pixel fails optimize's depth gate and tracking would silently return identities.
        ctx->vo_frame_id = -1;
        ctx->err = "the first frame of a sequence needs init_depth (use dvo_vo_push* for frame 0; dvo_vo_submit* carries no prior)";
        return DVO_ERR_STATE;
    }
    int rc = build_frame(ctx, -1, gray_dev, ctx->rows_only_staged, kind, aux); // make_shared<Frame>(gray, K, 3, 2), system.hpp:47
    ctx->rows_only_staged = false;
    if (rc) return rc;
    if (after_cull) CK(cudaEventRecord(after_cull, ctx->stream));
    if (init_depth || init_sigma) {
        rc = set_depth_impl(ctx, -1, init_depth, init_sigma, nullptr, 0);
        if (rc) return rc;
    }
    const bool fused = ctx->fuse_prop && small_level(ctx) && ctx->trk_B == 1;   // propagate in the tracking kernel's epilogue
    const bool overlapped = ctx->overlap && ctx->trk_B == 1 && !ctx->ktiming;   // per-kernel timing needs the launches apart
    if (frame_id > 0) {
        // before the launches (Mapper::update must follow the tracking kernel back to back): every frame since the last read-back
        // can have consumed at most roi.n draws per stream
        const uint64_t per = (uint64_t)(ctx->roi.n > 0 ? ctx->roi.n : 0);
        uint64_t bound = ctx->rng_known + per * (uint64_t)(frame_id - ctx->kf_known_frame);
        if (bound > (uint64_t)c.rng_table && ctx->kf_known_frame < frame_id - 1) {
            // The bound is pessimistic by roi.n draws for every frame the host has run ahead of the last read-back (a caller that
            // never reads poses enqueues hundreds of frames).  Before paying for a bigger table, wait for the device and read the
            // actual position: one drain of the stream instead of a host-side mt19937 run of 10^8..10^9 draws.
            int now[2] = {0, 0};
            CK(cudaStreamSynchronize(ctx->stream));
            CK(cudaMemcpy(now, ctx->d_status + 1, sizeof(int) * 2, cudaMemcpyDeviceToHost));
            ctx->kf_known = now[0]; ctx->rng_known = (uint64_t)(unsigned)now[1];
            ctx->kf_known_frame = frame_id - 1;
            ctx->kf_rb_pending[0] = ctx->kf_rb_pending[1] = false; // both have landed; they are older than what was just read
            bound = ctx->rng_known + per;
        }
        if (bound > ctx->rng_bound) ctx->rng_bound = bound; // dvo_map_update on the same context continues from here
        rc = rng_reserve(ctx, bound);
        if (rc) return rc;
    }
    // Tracking (system.hpp:57) against each stream's current keyframe (no-op on the first frame), then
    // Frame::updateXi + Mapper::needNewFrame + setRefFrame (+ propagate, + odometrize's return value) in the same kernel's epilogue
    rc = enqueue_track(ctx, -1, -1, true, -1.0, -1.0f, frame_id, ctx->forced_pending ? ctx->d_forced : nullptr, T_out ? ctx->d_small : nullptr,
                       overlapped && frame_id > 0, fused);
    if (rc) return rc;
    ctx->forced_pending = false;
    // Mapping (system.hpp:67, mapper.cpp:16-33): per stream either propagate or update, then regularize
    if (frame_id > 0) {
        rc = enqueue_update(ctx, false, overlapped);
        if (rc) return rc;
        if (!fused) {
            rc = enqueue_propagate(ctx); // update and propagate touch disjoint streams: their order is free
            if (rc) return rc;
        }
        rc = enqueue_regularize(ctx, true);
        if (rc) return rc;
    } else {
        // frame 0 is the first keyframe: its 2x2-footprint image (later ones are written by k_regularize when they are created)
        k_build_quad<<<dim3(cdiv(c.lv[c.L - 1].n, 256), c.S), 256, 0, ctx->stream>>>(c, ctx->d_slots, -1, ctx->d_vo);
        KCHECK();
    }
    {   // keyframe-count read-back for the pool (see above): one of two slots, whichever is free
        const int k = ctx->kf_rb_pending[0] ? (ctx->kf_rb_pending[1] ? -1 : 1) : 0;
        if (k >= 0) {
            CK(cudaMemcpyAsync(ctx->h_status + 4 + 2 * k, ctx->d_status + 1, sizeof(int) * 2, cudaMemcpyDeviceToHost, ctx->stream));
            CK(cudaEventRecord(ctx->ev_kf[k], ctx->stream));
            ctx->kf_rb_pending[k] = true; ctx->kf_rb_frame[k] = frame_id;
        }
    }
    if (T_out) {
        // device-side errors (history overflow) come back with the poses: checked here when synchronous, in dvo_vo_collect otherwise
        CK(cudaMemcpyAsync(ctx->h_status + status_slot, ctx->d_status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaMemcpyAsync(T_out, ctx->d_small, sizeof(float) * 16 * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
        if (sync_out) {
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->h_status[status_slot]) return status_error(ctx, ctx->h_status[status_slot]);
        }
    }
    return DVO_OK;
}

extern "C" int dvo_vo_push_dev(dvo_ctx* ctx, const float* gray_dev, const float* init_depth, const float* init_sigma, float* T_out)
{
    return vo_push_impl(ctx, gray_dev, init_depth, init_sigma, T_out, nullptr, true);
}

extern "C" int dvo_vo_push_gray8_dev(dvo_ctx* ctx, const uint8_t* gray8_dev, int invalid_code, const float* init_depth,
                                     const float* init_sigma, float* T_out)
{
    ARG(gray8_dev && invalid_code >= -1 && invalid_code <= 255, "bad argument");
    return vo_push_impl(ctx, (const float*)gray8_dev, init_depth, init_sigma, T_out, nullptr, true, 1, invalid_code);
}

extern "C" int dvo_vo_push(dvo_ctx* ctx, const float* gray, const float* init_depth, const float* init_sigma, float* T_out)
{
    ARG(gray, "null gray");
    bool rows_only = false;
    int rc = stage_gray(ctx, gray, &rows_only);
    if (rc) return rc;
    ctx->rows_only_staged = rows_only;
    return vo_push_impl(ctx, ctx->d_stage_in, init_depth, init_sigma, T_out, nullptr, true);
}

extern "C" int dvo_vo_push_gray8(dvo_ctx* ctx, const uint8_t* gray8, int invalid_code, const float* init_depth, const float* init_sigma,
                                 float* T_out)
{
    ARG(gray8 && invalid_code >= -1 && invalid_code <= 255, "bad argument");
    bool rows_only = false;
    int rc = stage_gray_to(ctx, ctx->d_stage_in, ctx->stream, gray8, &rows_only, 1);
    if (rc) return rc;
    ctx->rows_only_staged = rows_only;
    return vo_push_impl(ctx, ctx->d_stage_in, init_depth, init_sigma, T_out, nullptr, true, 1, invalid_code);
}

extern "C" int dvo_vo_push_bgr8(dvo_ctx* ctx, const uint8_t* bgr8, int channels, const float* init_depth, const float* init_sigma,
                                float* T_out)
{
    int rc = stage_bgr(ctx, ctx->d_stage_in, ctx->stream, bgr8, channels);
    if (rc) return rc;
    ctx->rows_only_staged = false;
    return vo_push_impl(ctx, ctx->d_stage_in, init_depth, init_sigma, T_out, nullptr, true, 2, channels);
}

// Pipelined submission: two staging buffers filled on a copy stream, so frame n+1 crosses PCIe while
// frame n's kernels run.
static int pipe_init(dvo_ctx* ctx)
{
    if (ctx->copy_stream) return DVO_OK;
    const Consts& c = ctx->c;
    {   // highest priority: the staging kernel's few CTAs must not queue behind the previous frame's 50 000-CTA grids
        int lo = 0, hi = 0;
        CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CK(cudaStreamCreateWithPriority(&ctx->copy_stream, cudaStreamNonBlocking, hi));
    }
    ctx->d_stage_pp[0] = ctx->d_stage_in;
    CK(cudaMalloc(&ctx->d_stage_pp[1], sizeof(float) * (size_t)c.S * c.in_w * c.in_h));
    return DVO_OK;
}

static int vo_submit_impl(dvo_ctx* ctx, const void* gray, float* T_out, int kind, int aux)
{
    ARG(gray, "null gray");
    if (ctx->pp_in_flight >= 2) {
        ctx->err = "dvo_vo_submit: two frames already in flight (call dvo_vo_collect)";
        return DVO_ERR_STATE;
    }
    int rc = pipe_init(ctx);
    if (rc) return rc;
    const int k = (int)(ctx->pp_head & 1);
    // the buffer's previous contents must have been consumed, and d_stage_in (buffer 0) may also have been
    // filled by a plain dvo_vo_push on the compute stream before
    CK(cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_consumed[k], 0));
    bool rows_only = false;
    if (kind == 2) {
        if (!ctx->d_remap) { // before the copy stream is involved: remap_build works on the compute stream
            rc = remap_build(ctx, nullptr, nullptr);
            if (rc) return rc;
        }
        rc = stage_bgr(ctx, ctx->d_stage_pp[k], ctx->copy_stream, (const uint8_t*)gray, aux);
    } else {
        rc = stage_gray_to(ctx, ctx->d_stage_pp[k], ctx->copy_stream, gray, &rows_only, kind == 1 ? 1 : sizeof(float));
    }
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev_copied[k], ctx->copy_stream));
    CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_copied[k], 0));
    ctx->rows_only_staged = rows_only;
    rc = vo_push_impl(ctx, ctx->d_stage_pp[k], nullptr, nullptr, T_out, k == 0 ? nullptr : ctx->ev_consumed[k], false, kind, aux, k);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev_done[k], ctx->stream));
    ctx->pp_head++;
    ctx->pp_in_flight++;
    return DVO_OK;
}

extern "C" int dvo_vo_submit(dvo_ctx* ctx, const float* gray, float* T_out) { return vo_submit_impl(ctx, gray, T_out, 0, -1); }
extern "C" int dvo_vo_submit_bgr8(dvo_ctx* ctx, const uint8_t* bgr8, int channels, float* T_out)
{
    ARG(channels == 3 || channels == 4, "channels not 3 / 4");
    return vo_submit_impl(ctx, bgr8, T_out, 2, channels);
}
extern "C" int dvo_vo_submit_gray8(dvo_ctx* ctx, const uint8_t* gray8, int invalid_code, float* T_out)
{
    ARG(invalid_code >= -1 && invalid_code <= 255, "bad invalid_code");
    return vo_submit_impl(ctx, gray8, T_out, 1, invalid_code);
}

extern "C" int dvo_vo_collect(dvo_ctx* ctx)
{
    if (ctx->pp_in_flight <= 0) {
        ctx->err = "dvo_vo_collect: nothing in flight";
        return DVO_ERR_STATE;
    }
    const int k = (int)((ctx->pp_head - ctx->pp_in_flight) & 1);
    CK(cudaEventSynchronize(ctx->ev_done[k]));
    ctx->pp_in_flight--;
    if (ctx->h_status[k]) return status_error(ctx, ctx->h_status[k]); // read back with this frame's poses (0 when T_out was NULL)
    return DVO_OK;
}

extern "C" int dvo_vo_state(dvo_ctx* ctx, float* xi, float* rel_xi, int* was_keyframe, int* n_keyframes, int* valid_updates, int* n_iters)
{
    const Consts& c = ctx->c;
    ARG(ctx->vo_frame_id >= 0, "no frame pushed");
    std::vector<VoState> hv(c.S);
    std::vector<MapState> hm(c.S);
    std::vector<float> hx((size_t)c.S * 12);
    k_vo_gather_xi<<<cdiv(c.S, 64), 64, 0, ctx->stream>>>(c, ctx->d_slots, ctx->d_vo, ctx->d_small);
    KCHECK();
    CK(cudaMemcpyAsync(hv.data(), ctx->d_vo, sizeof(VoState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(hm.data(), ctx->d_ms, sizeof(MapState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(hx.data(), ctx->d_small, sizeof(float) * 12 * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) {
        if (xi) memcpy(xi + 6 * s, &hx[12 * s], sizeof(float) * 6);
        if (rel_xi) memcpy(rel_xi + 6 * s, &hx[12 * s + 6], sizeof(float) * 6);
        if (was_keyframe) was_keyframe[s] = hv[s].was_keyframe;
        if (n_keyframes) n_keyframes[s] = hv[s].n_kf;
        if (valid_updates) valid_updates[s] = hm[s].valid_update;
        if (n_iters) n_iters[s] = hv[s].n_iters;
    }
    return DVO_OK;
}

extern "C" int dvo_vo_counters(dvo_ctx* ctx, uint64_t* gn_iters, uint64_t* gn_pixels, uint64_t* n_propagate, uint64_t* n_update)
{
    const Consts& c = ctx->c;
    std::vector<VoState> hv(c.S);
    CK(cudaMemcpyAsync(hv.data(), ctx->d_vo, sizeof(VoState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) {
        if (gn_iters) gn_iters[s] = hv[s].gn_iters;
        if (gn_pixels) gn_pixels[s] = hv[s].gn_pixels;
        if (n_propagate) n_propagate[s] = hv[s].n_propagate;
        if (n_update) n_update[s] = hv[s].n_update;
    }
    return DVO_OK;
}

extern "C" int dvo_vo_matched(dvo_ctx* ctx, uint64_t* matched)
{
    const Consts& c = ctx->c;
    ARG(matched, "bad argument");
    std::vector<VoState> hv(c.S);
    std::vector<MapState> hm(c.S);
    CK(cudaMemcpyAsync(hv.data(), ctx->d_vo, sizeof(VoState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(hm.data(), ctx->d_ms, sizeof(MapState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) matched[s] = hv[s].map_matched + (uint64_t)(unsigned int)hm[s].valid_update;
    return DVO_OK;
}

extern "C" int dvo_vo_last_log(dvo_ctx* ctx, dvo_iter_log* log, int log_cap, int* n_iters)
{
    const Consts& c = ctx->c;
    ARG(log && log_cap > 0, "bad argument");
    TrackState* hts = (TrackState*)((char*)ctx->h_pinned + sizeof(dvo_iter_log) * (size_t)c.S * ctx->log_cap);
    dvo_iter_log* hl = (dvo_iter_log*)ctx->h_pinned;
    CK(cudaMemcpyAsync(hts, ctx->d_ts, sizeof(TrackState) * (size_t)c.S, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(hl, ctx->d_log, sizeof(dvo_iter_log) * (size_t)c.S * ctx->log_cap, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int s = 0; s < c.S; s++) {
        if (n_iters) n_iters[s] = hts[s].n_total;
        for (int k = 0; k < hts[s].n_total && k < log_cap && k < ctx->log_cap; k++) log[(size_t)s * log_cap + k] = hl[(size_t)s * ctx->log_cap + k];
    }
    return DVO_OK;
}

extern "C" int dvo_vo_ref_slot(dvo_ctx* ctx, int stream, int* slot)
{
    const Consts& c = ctx->c;
    ARG(stream >= 0 && stream < c.S && slot, "bad stream");
    VoState v;
    CK(cudaMemcpyAsync(&v, ctx->d_vo + stream, sizeof(VoState), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *slot = v.kf_slot;
    return DVO_OK;
}
extern "C" int dvo_vo_last_slot(dvo_ctx* ctx, int stream, int* slot)
{
    const Consts& c = ctx->c;
    ARG(stream >= 0 && stream < c.S && slot, "bad stream");
    VoState v;
    CK(cudaMemcpyAsync(&v, ctx->d_vo + stream, sizeof(VoState), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *slot = v.cur_slot;
    return DVO_OK;
}
extern "C" int dvo_vo_error(dvo_ctx* ctx) { return map_error(ctx); }

// ---- synthetic inputs -----------------------------------------------------------------------
static int synth_render(dvo_ctx* ctx, float* gray_dev, float* depth_dev, uint8_t* gray8_dev, int frame, uint64_t seed0, uint64_t seed_stride,
                        float trans_step, float rot_step_deg)
{
    ARG((gray_dev || gray8_dev) && frame >= 0, "bad argument");
    const Consts& c = ctx->c;
    SynthParams P;
    P.w = c.in_w; P.h = c.in_h;
    P.fx = ctx->P.K[0]; P.fy = ctx->P.K[4]; P.cx = ctx->P.K[2]; P.cy = ctx->P.K[5];
    P.seed0 = seed0; P.seed_stride = seed_stride; P.trans_step = trans_step; P.rot_step = rot_step_deg * 0.017453292519943295f; P.frame = frame;
    dim3 grid(cdiv(P.w * P.h, 256), c.S);
    k_synth_render<<<grid, 256, 0, ctx->stream>>>(P, gray_dev, depth_dev, gray8_dev);
    KCHECK();
    return DVO_OK;
}
extern "C" int dvo_synth_render_dev(dvo_ctx* ctx, float* gray_dev, float* depth_dev, int frame, uint64_t seed0, float trans_step,
                                    float rot_step_deg)
{
    return synth_render(ctx, gray_dev, depth_dev, nullptr, frame, seed0, 1, trans_step, rot_step_deg);
}
extern "C" int dvo_synth_render_gray8_dev(dvo_ctx* ctx, uint8_t* gray8_dev, float* depth_dev, int frame, uint64_t seed0, uint64_t seed_stride,
                                          float trans_step, float rot_step_deg)
{
    return synth_render(ctx, nullptr, depth_dev, gray8_dev, frame, seed0, seed_stride, trans_step, rot_step_deg);
}
static int synth_info(dvo_ctx* ctx, int frame, uint64_t seed, float trans_step, float rot_step_deg, float* host, int first, int count)
{
    float* d = ctx->d_small; // [S][64] floats: room for SY_INFO_FLOATS when S >= 3; otherwise a scratch allocation
    DevScratch sc;
    if ((size_t)ctx->c.S * 64 < (size_t)SY_INFO_FLOATS) {
        CK(cudaMalloc(&sc.p, sizeof(float) * SY_INFO_FLOATS));
        d = (float*)sc.p;
    }
    k_synth_info<<<1, 64, 0, ctx->stream>>>((unsigned long long)seed, frame, trans_step, rot_step_deg * 0.017453292519943295f, d);
    KCHECK();
    CK(cudaMemcpyAsync(host, d + first, sizeof(float) * count, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_synth_pose(dvo_ctx* ctx, int frame, uint64_t seed, float trans_step, float rot_step_deg, float* T_wc)
{
    ARG(T_wc && frame >= 0, "bad argument");
    int rc = synth_info(ctx, frame, seed, trans_step, rot_step_deg, T_wc, 0, 12);
    if (rc) return rc;
    T_wc[12] = T_wc[13] = T_wc[14] = 0.f; T_wc[15] = 1.f;
    return DVO_OK;
}
extern "C" int dvo_synth_scene(dvo_ctx* ctx, uint64_t seed, float* out, int cap)
{
    ARG(out && cap >= SY_INFO_FLOATS - 12, "bad argument");
    return synth_info(ctx, 0, seed, 0.f, 0.f, out, 12, SY_INFO_FLOATS - 12);
}
extern "C" void* dvo_dev_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void dvo_dev_free(void* p) { if (p) cudaFree(p); }
extern "C" int dvo_copy_to_host(dvo_ctx* ctx, void* host, const void* dev, size_t bytes)
{
    CK(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_copy_to_dev(dvo_ctx* ctx, void* dev, const void* host, size_t bytes)
{
    CK(cudaMemcpyAsync(dev, host, bytes, cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return DVO_OK;
}

// ---- measurement ----------------------------------------------------------------------------
extern "C" int dvo_timer_start(dvo_ctx* ctx)
{
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_timer_stop(dvo_ctx* ctx, float* ms)
{
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    CK(cudaEventSynchronize(ctx->ev1));
    CK(cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return DVO_OK;
}

extern "C" int dvo_event_record(dvo_ctx* ctx, int id)
{
    ARG(id >= 0 && id < 65536, "bad event id");
    if ((size_t)id >= ctx->user_ev.size()) ctx->user_ev.resize((size_t)id + 1, nullptr);
    if (!ctx->user_ev[id]) CK(cudaEventCreate(&ctx->user_ev[id]));
    CK(cudaEventRecord(ctx->user_ev[id], ctx->stream));
    return DVO_OK;
}
extern "C" int dvo_event_elapsed(dvo_ctx* ctx, int id0, int id1, float* ms)
{
    ARG(ms && id0 >= 0 && id1 >= 0 && (size_t)id0 < ctx->user_ev.size() && (size_t)id1 < ctx->user_ev.size() && ctx->user_ev[id0] &&
        ctx->user_ev[id1], "bad event id");
    CK(cudaEventSynchronize(ctx->user_ev[id1]));
    CK(cudaEventElapsedTime(ms, ctx->user_ev[id0], ctx->user_ev[id1]));
    return DVO_OK;
}

#ifdef DVO_MAP_STATS
extern "C" int dvo_debug_map_stats(unsigned long long* st, unsigned int* hist, int reset)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(st, g_map_stats, sizeof(unsigned long long) * 8);
    cudaMemcpyFromSymbol(hist, g_map_hist, sizeof(unsigned int) * 128);
    if (reset) {
        unsigned long long z[8] = {0}; unsigned int zh[128] = {0};
        cudaMemcpyToSymbol(g_map_stats, z, sizeof z); cudaMemcpyToSymbol(g_map_hist, zh, sizeof zh);
    }
    return 0;
}
#endif
