// dvo_internal.cuh — device-visible data model shared by the kernels and the C-ABI layer.
//
// HBM layout.  A frame slot keeps ONLY the finest pyramid level of every stream:
//   gray [S][n_top] float, ds [S][n_top] float2 {depth, sigma}, age [S][n_top] float, xi/rel_xi [S][6],
//   quad [S][n_top] float4 (the gray image as 2x2 bilinear footprints, keyframes only: what the epipolar search samples).
// Coarser levels are never materialised: Convert::cullImage is top-left point sampling through
// Convert::getPixel (convert.cpp:7-20,107-125), and the reference re-culls depth/sigma from the
// finest map after every map step (frame.cpp:39-61), so level l pixel (x, y) is exactly
// getPixel(top[(y << t) * W + (x << t)]) with t = L-1-l.  Gradients (Convert::gradiate,
// convert.cpp:41-75) are two-tap differences of that gray and are recomputed where sampled.
#pragma once
#include "../../include/dvo_cuda.h"
#include "dvo_device.cuh"

namespace dvo {

constexpr int MAXL = DVO_MAX_LEVELS;
#ifndef DVO_TRK_THREADS
#define DVO_TRK_THREADS 512
#endif
constexpr int TRK_THREADS = DVO_TRK_THREADS;   // threads per tracking CTA
constexpr int TRK_STAGE_PIXELS = 80 * 60; // k_track_staged: reference levels up to 80x60 (19.2 KB) as a shared-memory tile
constexpr int TRK_MINB = 1024 / TRK_THREADS;   // <= 64 registers per thread: 1024 tracking threads per SM
constexpr int NACC = 29;         // 21 J^T J + 6 J^T (w r) + sum r^2 + valid count

// Geometry + constants, passed to kernels by value (__grid_constant__).
struct Consts {
    int S, L;
    int in_w, in_h, culls;
    LevelGeom lv[MAXL];
    int max_kf;
    int kf_stride; // row pitch of the per-stream history / pose tables: max_kf rounded up to 32 entries, so that no 128-byte line
                   // holds entries of two streams (Mapper::update may run while OTHER streams' tables are still being written)
    int vo_base; // first slot id of the VO keyframe pool (= dvo_params.max_frames)
    // tracking
    int max_iter;
    int roi_level, roi_x0, roi_x1, roi_y0, roi_y1;
    float min_depth_f;  // smallest float >= (double)min_depth: (double)d < min_depth <=> d < min_depth_f
    float sigma_lo, sigma_hi;
    float step_default, step_level1, step_level2;
    float huber_delta, min_grad2; // extras (dvo_params.huber_delta, min_gradient^2); 0 = off = the reference
    // mapping
    int map_x0, map_x1, map_y0, map_y1;
    float min_movement;
    int max_forward;
    float acc_d_lo, acc_d_hi, acc_s_hi;
    unsigned int rng_table;
};

// One frame slot as the device sees it.  All images are stream-major: base + s * n_top.
struct SlotDev {
    float4* quad;  // finest-level gray as 2x2 footprints  [S][n_top]: {I(x,y), I(x+1,y), I(x,y+1), I(x+1,y+1)} with the border rule
                   // of Convert::getSubpixelFromDense applied (convert.cpp:77-105: a tap outside the image takes I(x,y)), so that one
                   // bilinear tap of the epipolar search is ONE 16-byte load.  Valid for keyframes only (built when a frame becomes one).
    float* gray;   // finest-level gray                    [S][n_top]
    float2* ds;    // finest-level {depth, sigma}          [S][n_top]
    float* age;    // finest level                         [S][n_top]
    float* xi;     // Frame::m_xi                          [S][6]
    float* rel_xi; // Frame::m_relative_xi                 [S][6]
};

// A pyramid level seen through the finest-level image: pixel (x, y) -> top[(y << t) * W + (x << t)].
struct LevelView {
    int w, h;     // level size
    int t;        // shift to the finest level
    int W;        // finest-level row pitch (elements)
};

// Per-stream Gauss-Newton state (Tracker::track, src/track/tracker.cpp:22-85).
struct TrackState {
    int level, iter, done, n_total;
    int mode;            // 0: full coarse-to-fine track; 1: single step (Track::optimize)
    float xi[6];
    float Tn[12];        // exp(-xi), rows of [R|t]
    double stop_update;
    float stop_residual;
    int px_total;        // pixels swept by this track (sum of level sizes over iterations)
    const float* obj_gray;  // finest level, pre-offset by stream
    const float* ref_gray;
    const float2* ref_ds;
    // inter-CTA barrier of the stream's CTA group (only used when a stream is split over > 1 CTA)
    unsigned int arrive; // monotonic: CTAs that finished accumulating, summed over iterations
    unsigned int gen;    // iterations whose tail has been published
};

// Per-stream mapping work descriptor (filled by the host for the granular API, by the tracking
// kernel's epilogue for the device-resident VO loop).
struct MapState {
    int ref_slot, obj_slot, new_slot;
    int n_hist;
    int do_propagate, do_update, do_regularize;
    int reg_slot;
    float Trel[12];   // exp(rel_xi)
    float rel_z;      // rel_xi[2]
    int valid_update;
    unsigned int n_reject;      // rejects of this update
    unsigned int ticket;        // phase-B CTAs that have finished (the last one advances rng_pos)
    unsigned long long rng_pos; // draws consumed so far
    int err;
    unsigned int ticket_a;      // phase-A CTAs that have finished (the last one turns the per-CTA reject counts into offsets)
    unsigned int ticket_r;      // regularize CTAs that have finished (the last one copies the result back, Frame::updateDepth)
    int pad[3];
};

// Hand-over from the tracking kernel to Mapper::update inside one frame (DESIGN.md §4 "overlap").  Streams are independent, so a
// stream whose tracking has finished can be mapped while slower streams still iterate: the tracking CTA appends its stream to
// `list` (completion order) when Mapper::needNewFrame chose update, and every finished stream counts in n_finished.  k_map_update
// CTAs take a ticket and serve (list position, tile) = (ticket / nblk, ticket % nblk).  Reset by the frame's ingest kernel.
struct ReadyQ {
    int n_reserved;        // list entries handed out
    int n_finished;        // streams whose tracking epilogue is complete (updating or not)
    unsigned int ticket;   // k_map_update CTAs started
    int pad;
    int list[1];           // [S] stream ids in completion order, -1 = not yet written
};

// Per-stream VO state (System::VisualOdometry + FrameHistory).
struct VoState {
    int kf_slot, kf_frame_id, n_kf;
    int need_new, was_keyframe;
    int n_iters;
    int cur_slot; // slot holding this stream's current frame: vo_base + min(n_kf, max_kf)
    int pad;
    unsigned long long gn_iters;  // Gauss-Newton iterations since dvo_vo_reset
    unsigned long long gn_pixels; // pixels swept by those iterations
    unsigned long long n_propagate, n_update; // frames mapped by propagate / update
    unsigned long long map_matched;           // pixels Mapper::update matched and fused (valid_update, mapper.cpp:122-128), frames before the last
};

struct PoseTab { // per (stream, age): exp(-r_xi) and -r_xi[0:3] (src/map/mapper.cpp:107, implement.cpp:56-58,193)
    float T[12];
    float t[3];
    float pad;
};

// slot holding stream s's current VO frame (entry k of the pool = keyframe k of every stream;
// the frame being processed lives in the entry after the stream's newest keyframe)
__host__ __device__ __forceinline__ int vo_cur_slot(const Consts& c, int n_kf) { return c.vo_base + (n_kf < c.max_kf ? n_kf : c.max_kf); }

} // namespace dvo
