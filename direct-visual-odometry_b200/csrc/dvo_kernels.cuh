// dvo_kernels.cuh — hand-written sm_100a kernels for the tracking + mapping hot path.
// Included once by dvo_cuda.cu.  Compiled with -fmad=false (see dvo_device.cuh).
#pragma once
#include "dvo_internal.cuh"

namespace dvo {

// =============================================================================================
// Frame construction: Convert::cullImage (convert.cpp:7-20), Convert::gradiate (convert.cpp:41-75),
// Frame::createScenePyramid (frame.cpp:30-37).  grid = (ceil(n_top/256), S, L)
// =============================================================================================
// Slot of stream s for the frame-construction kernels: a fixed slot (granular API) or, in the VO
// loop, the stream's own current slot (streams take keyframes at different times, so the current
// frame of each stream lives in the pool entry after its newest keyframe).
__device__ __forceinline__ const SlotDev& slot_of(const SlotDev* __restrict__ slots, int fixed, const VoState* __restrict__ vo, int s)
{
    return slots[vo ? vo[s].cur_slot : fixed];
}

__global__ void __launch_bounds__(256) k_cull_gray(const __grid_constant__ Consts c, const float* __restrict__ in,
                                                   const SlotDev* __restrict__ slots, int fixed, const VoState* __restrict__ vo)
{
    const int l = blockIdx.z, s = blockIdx.y;
    const SlotDev& slot = slot_of(slots, fixed, vo, s);
    const LevelGeom g = c.lv[l];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int sh = c.culls + (c.L - 1 - l);
    const int y = i / g.w, x = i - y * g.w;
    const float v = __ldg(in + (size_t)s * c.in_w * c.in_h + (size_t)(y << sh) * c.in_w + (x << sh));
    // cull of a cull == direct sample: getPixel maps every value <= -2 to exactly -2 (idempotent)
    slot.gray[l][(size_t)s * g.n + i] = (sh == 0) ? v : get_pixel_value(v);
}

__global__ void __launch_bounds__(256) k_build_tex(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, int fixed,
                                                   const VoState* __restrict__ vo)
{
    const int l = blockIdx.z, s = blockIdx.y;
    const SlotDev& slot = slot_of(slots, fixed, vo, s);
    const LevelGeom g = c.lv[l];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const float* gr = slot.gray[l] + (size_t)s * g.n;
    const float v = gr[i];
    float gx = DVO_INV, gy = DVO_INV;
    if (!(x - 1 <= -1 || x + 1 >= g.w)) {
        float a = get_pixel_value(gr[i - 1]), b = get_pixel_value(gr[i + 1]);
        if (!(is_invalid(a) || is_invalid(b))) gx = b - a;
    }
    if (!(y - 1 <= -1 || y + 1 >= g.h)) {
        float a = get_pixel_value(gr[i - g.w]), b = get_pixel_value(gr[i + g.w]);
        if (!(is_invalid(a) || is_invalid(b))) gy = b - a;
    }
    slot.tex[l][(size_t)s * g.n + i] = make_float4(v, gx, gy, 0.f);
}

// depth / sigma / age upload (Frame 6-arg ctor frame.hpp:91-106, updateDepthSigmaAge frame.cpp:47-54).
// src arrays are [S][sh][sw]; shift = culls when full-res, 0 otherwise.  grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_set_top(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, int fixed,
                                                 const VoState* __restrict__ vo, const float* __restrict__ depth,
                                                 const float* __restrict__ sigma, const float* __restrict__ age, int sw, int sh_,
                                                 int shift)
{
    const int s = blockIdx.y;
    const SlotDev& slot = slot_of(slots, fixed, vo, s);
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t src = (size_t)s * sw * sh_ + (size_t)(y << shift) * sw + (x << shift);
    float2* ds = slot.ds[c.L - 1] + (size_t)s * g.n;
    float2 v = ds[i];
    if (depth) { float d = depth[src]; v.x = shift ? get_pixel_value(d) : d; }
    if (sigma) { float d = sigma[src]; v.y = shift ? get_pixel_value(d) : d; }
    ds[i] = v;
    if (age) { float d = age[src]; slot.age[(size_t)s * g.n + i] = shift ? get_pixel_value(d) : d; }
}

// Re-cull the coarser depth/sigma levels from the finest (frame.cpp:39-61).
// grid = (ceil(n_{L-2}/256), S, L-1)
__global__ void __launch_bounds__(256) k_cull_ds(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, int fixed,
                                                 const VoState* __restrict__ vo)
{
    const int l = blockIdx.z, s = blockIdx.y;
    const SlotDev& slot = slot_of(slots, fixed, vo, s);
    const LevelGeom g = c.lv[l];
    const LevelGeom gt = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int t = c.L - 1 - l;
    const int y = i / g.w, x = i - y * g.w;
    const float2 v = slot.ds[c.L - 1][(size_t)s * gt.n + (size_t)(y << t) * gt.w + (x << t)];
    slot.ds[l][(size_t)s * g.n + i] = make_float2(get_pixel_value(v.x), get_pixel_value(v.y));
}

// unpack one image component for download.  grid = (ceil(n_l/256), S)
__global__ void __launch_bounds__(256) k_extract(const __grid_constant__ Consts c, SlotDev slot, int l, int kind, float* __restrict__ out)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[l];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const size_t o = (size_t)s * g.n + i;
    float v;
    switch (kind) {
    case DVO_IMG_GRAY: v = slot.gray[l][o]; break;
    case DVO_IMG_DEPTH: v = slot.ds[l][o].x; break;
    case DVO_IMG_SIGMA: v = slot.ds[l][o].y; break;
    case DVO_IMG_GRADX: v = slot.tex[l][o].y; break;
    case DVO_IMG_GRADY: v = slot.tex[l][o].z; break;
    default: v = slot.age[o]; break;
    }
    out[o] = v;
}

// =============================================================================================
// Sampling of the {gray, gx, gy} texel array at one warped point: Convert::getSubpixel for the
// gray channel (with the repair automaton) and Convert::getSubpixelFromDense for the gradients
// share x0,y0,h,v and the four taps (src/core/convert.cpp:77-105,128-177).
// =============================================================================================
struct Sample3 {
    float I, gx, gy;
    bool in; // (x0,y0) in range
};
__device__ __forceinline__ Sample3 sample_tex(const float4* __restrict__ tex, int w, int h, float px, float py)
{
    Sample3 r;
    r.I = r.gx = r.gy = DVO_INV;
    const int x0 = trunc_x86(px), y0 = trunc_x86(py);
    r.in = in_range(x0, y0, w, h);
    if (!r.in) return r;
    const float hh = px - (float)x0, v = py - (float)y0;
    const bool xin = (x0 + 1) < w, yin = (y0 + 1) < h;
    const float4* p = tex + (size_t)y0 * w + x0;
    const float4 t0 = __ldg(p);
    const float4 t1 = xin ? __ldg(p + 1) : t0;
    const float4 t2 = yin ? __ldg(p + w) : t0;
    const float4 t3 = (xin && yin) ? __ldg(p + w + 1) : t0;
    r.gx = bilerp(t0.y, t1.y, t2.y, t3.y, hh, v);
    r.gy = bilerp(t0.z, t1.z, t2.z, t3.z, hh, v);
    float g[4] = {t0.x, t1.x, t2.x, t3.x};
    if (!(is_valid(g[0]) && is_valid(g[1]) && is_valid(g[2]) && is_valid(g[3]))) {
        if (!repair_taps(g)) return r; // I stays INVALID
    }
    r.I = bilerp(g[0], g[1], g[2], g[3], hh, v);
    return r;
}

// Transform::warpImage (src/core/transform.cpp:35-51).  grid = (ceil(n/256), S)
__global__ void __launch_bounds__(256) k_warp_image(const __grid_constant__ Consts c, SlotDev ref, int l, const float* __restrict__ xi,
                                                    float* __restrict__ out)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[l];
    __shared__ float T[12];
    if (threadIdx.x == 0) se3_exp_neg12(xi + 6 * s, T);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const float d = ref.ds[l][(size_t)s * g.n + i].x;
    float p = DVO_INV;
    if (!(fabsf(d) < DVO_EPS)) {
        float wx, wy;
        warp_point(T, g, (float)x, (float)y, d, wx, wy);
        p = sample_tex(ref.tex[l] + (size_t)s * g.n, g.w, g.h, wx, wy).I;
    }
    out[(size_t)s * g.n + i] = p;
}

// =============================================================================================
// 6x6 solve of the normal equations (replaces cv::solve(A,-B,DECOMP_SVD), optimize.cpp:97):
// Cholesky in fp64; rank-deficient systems fall back to a Jacobi eigen-decomposition with the
// SVD back-substitution threshold (w_i <= 2*FLT_EPSILON*sum(w) dropped) so the minimum-norm
// solution is returned like OpenCV's SVBkSb.
// =============================================================================================
__device__ inline void solve6_eigen(const double* H, const double* b, double* x)
{
    double A[36], V[36];
    for (int i = 0; i < 36; i++) { A[i] = H[i]; V[i] = 0; }
    for (int i = 0; i < 6; i++) V[i * 6 + i] = 1;
    for (int sweep = 0; sweep < 30; sweep++) {
        double off = 0;
        for (int i = 0; i < 6; i++)
            for (int j = i + 1; j < 6; j++) off += A[i * 6 + j] * A[i * 6 + j];
        double diag = 0;
        for (int i = 0; i < 6; i++) diag += A[i * 6 + i] * A[i * 6 + i];
        if (off <= 1e-30 * diag || off == 0) break;
        for (int p = 0; p < 5; p++)
            for (int q = p + 1; q < 6; q++) {
                double apq = A[p * 6 + q];
                if (apq == 0) continue;
                double theta = (A[q * 6 + q] - A[p * 6 + p]) / (2 * apq);
                double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1));
                double cs = 1 / sqrt(t * t + 1), sn = t * cs;
                for (int k = 0; k < 6; k++) {
                    double akp = A[k * 6 + p], akq = A[k * 6 + q];
                    A[k * 6 + p] = cs * akp - sn * akq;
                    A[k * 6 + q] = sn * akp + cs * akq;
                }
                for (int k = 0; k < 6; k++) {
                    double apk = A[p * 6 + k], aqk = A[q * 6 + k];
                    A[p * 6 + k] = cs * apk - sn * aqk;
                    A[q * 6 + k] = sn * apk + cs * aqk;
                }
                for (int k = 0; k < 6; k++) {
                    double vkp = V[k * 6 + p], vkq = V[k * 6 + q];
                    V[k * 6 + p] = cs * vkp - sn * vkq;
                    V[k * 6 + q] = sn * vkp + cs * vkq;
                }
            }
    }
    double w[6], wsum = 0;
    for (int i = 0; i < 6; i++) { double l = A[i * 6 + i]; w[i] = l > 0 ? sqrt(l) : 0; wsum += w[i]; }
    const double thr = wsum * (double)(1.1920928955078125e-07f * 2);
    for (int i = 0; i < 6; i++) x[i] = 0;
    for (int k = 0; k < 6; k++) {
        if (w[k] <= thr) continue;
        double proj = 0;
        for (int i = 0; i < 6; i++) proj += V[i * 6 + k] * b[i];
        proj /= (w[k] * w[k]);
        for (int i = 0; i < 6; i++) x[i] += proj * V[i * 6 + k];
    }
}

__device__ inline void solve6(const double* H, const double* b, double* x)
{
    double Lm[36];
    double dmax = 0;
    for (int i = 0; i < 6; i++) dmax = fmax(dmax, H[i * 6 + i]);
    bool ok = dmax > 0;
    for (int j = 0; j < 6 && ok; j++) {
        double sum = H[j * 6 + j];
        for (int k = 0; k < j; k++) sum -= Lm[j * 6 + k] * Lm[j * 6 + k];
        if (!(sum > dmax * 1e-11)) { ok = false; break; }
        double ljj = sqrt(sum);
        Lm[j * 6 + j] = ljj;
        for (int i = j + 1; i < 6; i++) {
            double s2 = H[i * 6 + j];
            for (int k = 0; k < j; k++) s2 -= Lm[i * 6 + k] * Lm[j * 6 + k];
            Lm[i * 6 + j] = s2 / ljj;
        }
    }
    if (!ok) {
        solve6_eigen(H, b, x);
        return;
    }
    double yv[6];
    for (int i = 0; i < 6; i++) {
        double s2 = b[i];
        for (int k = 0; k < i; k++) s2 -= Lm[i * 6 + k] * yv[k];
        yv[i] = s2 / Lm[i * 6 + i];
    }
    for (int i = 5; i >= 0; i--) {
        double s2 = yv[i];
        for (int k = i + 1; k < 6; k++) s2 -= Lm[k * 6 + i] * x[k];
        x[i] = s2 / Lm[i * 6 + i];
    }
}

// =============================================================================================
// The fused Gauss-Newton iteration: Transform::warpImage + Track::optimize + the iteration tail
// of Tracker::track (src/core/transform.cpp:35-51, src/track/optimize.cpp:10-99,
// src/track/tracker.cpp:42-73) for every stream, one launch per iteration.
//   grid = (blocks_x, S), 256 threads, PPT pixels per thread (contiguous 256*PPT chunk per block).
// Per pixel: gates -> warp -> 4 texel taps -> residual, weight, 1x6 Jacobian -> 29 fp64 sums.
// Block partials are written to global memory; the last block of each stream (ticket) reduces
// them in a fixed order (deterministic), solves the 6x6 system, composes the pose, applies the
// stop rules and advances the per-stream state machine — no host round trip between iterations.
// =============================================================================================
struct GnArgs {
    TrackState* st;       // [S]
    double* partials;     // [S][max_blocks][32]
    int max_blocks;
    dvo_iter_log* log;    // [S][log_cap]
    int log_cap;
    uint8_t* mask;        // optional [S][n_level] (single-step mode)
    int ppt;
};

template <int PPT>
__global__ void __launch_bounds__(GN_THREADS) k_gn_step(const __grid_constant__ Consts c, const GnArgs a)
{
    const int s = blockIdx.y;
    TrackState* __restrict__ st = a.st + s;
    if (st->done) return;
    const int level = st->level;
    const LevelGeom g = c.lv[level];
    const int nb = (g.n + GN_THREADS * PPT - 1) / (GN_THREADS * PPT);
    if ((int)blockIdx.x >= nb) return;

    __shared__ float sT[12];
    __shared__ double sred[GN_THREADS / 32][NACC + 1];
    __shared__ double stot[NACC + 3];
    __shared__ int s_last;
    if (threadIdx.x < 12) sT[threadIdx.x] = st->Tn[threadIdx.x];
    const float* __restrict__ obj_gray = st->obj_gray[level];
    const float4* __restrict__ tex = st->ref_tex[level];
    const float2* __restrict__ ds = st->ref_ds[level];
    const bool roi = (level == c.roi_level);
    float step = c.step_default; // optimize.cpp:22-26
    if (level == 1) step = c.step_level1;
    if (level == 2) step = c.step_level2;
    __syncthreads();

    double acc[NACC];
#pragma unroll
    for (int k = 0; k < NACC; k++) acc[k] = 0.0;

#pragma unroll
    for (int k = 0; k < PPT; k++) {
        const int i = (blockIdx.x * PPT + k) * GN_THREADS + threadIdx.x;
        if (i >= g.n) break;
        const int y = i / g.w, x = i - y * g.w;
        bool ok = true;
        if (roi) ok = !(x < c.roi_x0 || x > c.roi_x1 || y < c.roi_y0 || y > c.roi_y1); // optimize.cpp:33-36
        float2 dsv = make_float2(0.f, 0.f);
        float I1 = DVO_INV;
        if (ok) {
            dsv = __ldg(ds + i);
            I1 = __ldg(obj_gray + i);
            ok = !(dsv.x < c.min_depth_f) && !is_invalid(I1); // optimize.cpp:39,46
        }
        if (ok) {
            const float depth = dsv.x;
            float X, Y, Z;
            back_project(g, (float)x, (float)y, depth, X, Y, Z);
            const float tx = (sT[0] * X + sT[1] * Y + sT[2] * Z) + sT[3];
            const float ty = (sT[4] * X + sT[5] * Y + sT[6] * Z) + sT[7];
            const float tz = (sT[8] * X + sT[9] * Y + sT[10] * Z) + sT[11];
            const float wx = tx * g.fx / tz + g.cx;
            const float wy = ty * g.fy / tz + g.cy;
            const Sample3 sm = sample_tex(tex, g.w, g.h, wx, wy);
            // I_2 valid (optimize.cpp:46); warped point inside the image (:52-56); gradients valid (:61)
            ok = sm.in && !is_invalid(sm.I) && !(wx < 0 || wy < 0 || (float)g.w <= wx || (float)g.h <= wy)
                 && !(is_invalid(sm.gx) || is_invalid(sm.gy));
            if (ok) {
                const float fgx = g.fx * sm.gx, fgy = g.fy * sm.gy; // optimize.cpp:67-77
                const float xz = X / Z, yz = Y / Z;
                float J[6];
                J[0] = fgx / Z;
                J[1] = fgy / Z;
                J[2] = -(fgx * X + fgy * Y) / Z / Z;
                J[3] = -fgx * xz * yz - fgy * (1.0f + yz * yz);
                J[4] = fgx * (1.0f + xz * xz) + fgy * xz * yz;
                J[5] = (-fgx * yz + fgy * xz);
                const float r = sm.I - I1;
                const float sg = (dsv.y < c.sigma_lo) ? c.sigma_lo : (c.sigma_hi < dsv.y) ? c.sigma_hi : dsv.y; // std::clamp :83
                const float bw = r * (step / sg);
                double Jd[6];
#pragma unroll
                for (int q = 0; q < 6; q++) Jd[q] = (double)J[q];
                int idx = 0;
#pragma unroll
                for (int p = 0; p < 6; p++)
#pragma unroll
                    for (int q = p; q < 6; q++) { acc[idx] = fma(Jd[p], Jd[q], acc[idx]); idx++; }
                const double bd = (double)bw;
#pragma unroll
                for (int q = 0; q < 6; q++) acc[21 + q] = fma(Jd[q], bd, acc[21 + q]);
                acc[27] += (double)(r * r);
                acc[28] += 1.0;
            }
        }
        if (a.mask) a.mask[(size_t)s * g.n + i] = ok ? 1 : 0;
    }

    // ---- block reduction (fixed order => deterministic) ----
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NACC; k++) {
        double v = acc[k];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
        if (lane == 0) sred[warp][k] = v;
    }
    __syncthreads();
    double* mypart = a.partials + ((size_t)s * a.max_blocks + blockIdx.x) * 32;
    if (threadIdx.x < NACC) {
        double v = 0;
#pragma unroll
        for (int wv = 0; wv < GN_THREADS / 32; wv++) v += sred[wv][threadIdx.x];
        mypart[threadIdx.x] = v;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = atomicAdd(&st->ticket, 1u);
        s_last = (t == (unsigned)(nb - 1));
    }
    __syncthreads();
    if (!s_last) return;

    // ---- tail: last block of this stream ----
    __threadfence();
    if (threadIdx.x < NACC) {
        const double* base = a.partials + (size_t)s * a.max_blocks * 32 + threadIdx.x;
        double v = 0;
        for (int b = 0; b < nb; b++) v += __ldcg(base + (size_t)b * 32);
        stot[threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;

    const int valid = (int)stot[28];
    float upd[6] = {0, 0, 0, 0, 0, 0};
    float residual = -1.0f; // optimize.cpp:92-93
    if (valid > 0) {
        double H[36], bv[6], xs[6];
        int idx = 0;
        for (int p = 0; p < 6; p++)
            for (int q = p; q < 6; q++) { H[p * 6 + q] = stot[idx]; H[q * 6 + p] = stot[idx]; idx++; }
        for (int q = 0; q < 6; q++) bv[q] = stot[21 + q];
        solve6(H, bv, xs);
        for (int q = 0; q < 6; q++) upd[q] = (float)xs[q];
        residual = (float)stot[27] / (float)valid; // optimize.cpp:98
    }
    float xi[6], xin[6];
    for (int q = 0; q < 6; q++) xi[q] = st->xi[q];
    const int n_total = st->n_total;
    bool advance_pose = (st->mode == 0);
    if (advance_pose) {
        se3_concatenate(xi, upd, xin); // tracker.cpp:46
        bool nan = false;
        for (int q = 0; q < 6; q++) nan |= isnan(xin[q]);
        if (!nan) // tracker.cpp:47-51
            for (int q = 0; q < 6; q++) xi[q] = xin[q];
    }
    if (a.log && n_total < a.log_cap) {
        dvo_iter_log* L = a.log + (size_t)s * a.log_cap + n_total;
        L->level = level; L->iteration = st->iter; L->valid_pixels = valid; L->residual = residual;
        for (int q = 0; q < 6; q++) { L->update[q] = upd[q]; L->xi_after[q] = xi[q]; }
    }
    st->n_total = n_total + 1;
    st->px_total += g.n;
    st->ticket = 0;
    if (!advance_pose) {
        st->done = 1;
        return;
    }
    for (int q = 0; q < 6; q++) st->xi[q] = xi[q];
    float Tn[12];
    se3_exp_neg12(xi, Tn); // stuff.update(xi): the next iteration warps with exp(-xi)
    for (int q = 0; q < 12; q++) st->Tn[q] = Tn[q];
    // stop rules, tracker.cpp:68-70 (time budget never reached)
    const bool stop = (norm_n(upd, 6) < st->stop_update) || (residual < st->stop_residual);
    int iter = st->iter + 1;
    if (stop || iter >= c.max_iter) {
        iter = 0;
        const int nl = level + 1;
        st->level = nl;
        if (nl >= c.L) st->done = 1;
    }
    st->iter = iter;
}

// per-stream tracker reset (xi = 0, level 0) — Tracker::track prologue, tracker.cpp:28-40.
// obj_slot / ref_slot: if ref_per_stream != nullptr the keyframe slot is taken from VoState.
__global__ void k_track_init(const __grid_constant__ Consts c, TrackState* st, const SlotDev* __restrict__ slots, int obj_slot,
                             int ref_slot, const VoState* __restrict__ vo, int mode, int level, const float* __restrict__ xi0,
                             double stop_update, float stop_residual)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    TrackState* t = st + s;
    const int rs = vo ? vo[s].kf_slot : ref_slot;
    if (vo) obj_slot = vo[s].cur_slot;
    const SlotDev o = slots[obj_slot];
    const SlotDev r = slots[rs < 0 ? obj_slot : rs];
    for (int l = 0; l < c.L; l++) {
        const size_t off = (size_t)s * c.lv[l].n;
        t->obj_gray[l] = o.gray[l] + off;
        t->ref_tex[l] = r.tex[l] + off;
        t->ref_ds[l] = r.ds[l] + off;
    }
    float xi[6];
    for (int q = 0; q < 6; q++) xi[q] = xi0 ? xi0[6 * s + q] : 0.f;
    for (int q = 0; q < 6; q++) t->xi[q] = xi[q];
    float Tn[12];
    se3_exp_neg12(xi, Tn);
    for (int q = 0; q < 12; q++) t->Tn[q] = Tn[q];
    t->level = level; t->iter = 0; t->n_total = 0; t->px_total = 0; t->mode = mode; t->ticket = 0;
    t->done = (vo && rs < 0) ? 1 : 0;
    t->stop_update = stop_update; t->stop_residual = stop_residual;
}

// =============================================================================================
// Mapping
// =============================================================================================

// Implement::update (src/map/implement.cpp:182-214): epipolar segment (:23-47), search (:106-152),
// depth (:49-71) and sigma (:73-104) estimates.  T = exp(-r_xi) (12 floats), tneg = -r_xi[0:3].
__device__ __forceinline__ double point_norm(float x, float y)
{
    return sqrt(__dadd_rn(__dmul_rn((double)x, (double)x), __dmul_rn((double)y, (double)y)));
}

__device__ inline void implement_update(const float* __restrict__ born_gray, const float4* __restrict__ born_tex, const LevelGeom& g,
                                        const float* __restrict__ T, const float* __restrict__ tneg, const float* __restrict__ Kc,
                                        int px, int py, float og, float depth, float sigma, float& new_depth, float& new_sigma)
{
    const int w = g.w, h = g.h;
    const float dmin = (depth - sigma) < 0.10f ? 0.10f : (depth - sigma); // :30
    const float dmax = depth + sigma;
    float sx, sy, ex, ey;
    warp_point(T, g, (float)px, (float)py, dmax, sx, sy); // start :31
    warp_point(T, g, (float)px, (float)py, dmin, ex, ey); // end   :32
    const float length = (float)point_norm(sx - ex, sy - ey); // :33

    // doMatching :106-152
    const float dirx = (ex - sx) / length, diry = (ey - sy) / length;
    float ptx = sx, pty = sy, bestx = sx, besty = sy;
    float min_ssd = 2.0f * 3;
    int count = 0;
    const double len_d = (double)length;
    while (point_norm(ptx - sx, pty - sy) < len_d) {
        float ssd = 0;
        ptx += dirx;
        pty += diry;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const float k = (float)(i - 1);
            const float tx = ptx + dirx * k, ty = pty + diry * k;
            const float sub = subpixel_dense(born_gray, w, h, tx, ty);
            if (is_invalid(sub)) {
                ssd = 2.0f * 3;
                break;
            }
            const float diff = sub - og;
            const double wk = (i == 0) ? (1.0 * 1 / 3) : (i == 1) ? (1.0 * 2 / 3) : (1.0 * 3 / 3); // :134
            ssd = (float)__dadd_rn((double)ssd, __dmul_rn(wk, (double)(diff * diff)));
        }
        if (ssd < min_ssd) {
            bestx = ptx; besty = pty;
            min_ssd = ssd;
        }
        if (count++ > 100) break;
    }
    float mx = bestx, my = besty;
    if ((double)min_ssd > 3 * 0.1) { mx = -1; my = -1; } // :145
    if (mx < 0 || my < 0 || mx > (float)w || my > (float)h) { // :196-200
        new_depth = -1; new_sigma = -1;
        return;
    }
    // depthEstimate :49-71
    {
        const float xq0 = 1.0f * ((float)px - g.cx) / g.fx, xq1 = 1.0f * ((float)py - g.cy) / g.fy, xq2 = 1.0f;
        const float R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
        const float xi3[3] = {mx, my, 1.0f};
        double r3q = 0;
        r3q = __dadd_rn(r3q, __dmul_rn((double)R[6], (double)xq0));
        r3q = __dadd_rn(r3q, __dmul_rn((double)R[7], (double)xq1));
        r3q = __dadd_rn(r3q, __dmul_rn((double)R[8], (double)xq2));
        float KR[9], KRq[3], Kt[3];
        const float xq[3] = {xq0, xq1, xq2};
        mat3_mul(Kc, R, KR);
        mat3_vec(KR, xq, KRq);
        mat3_vec(Kc, tneg, Kt);
        double ab = 0, aa = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            // `s*x_i - (K*R)*x_q`, `t(2)*x_i - K*t`: folded by MatOp_GEMM::subtract into
            // cv::gemm(A, B, -1, x_i, s) = (float)(t*(-1) + c*(float)s) evaluated in double
            const float av = (float)__dadd_rn(__dmul_rn((double)KRq[i], -1.0), __dmul_rn((double)xi3[i], (double)(float)r3q));
            const float bv = (float)__dadd_rn(__dmul_rn((double)Kt[i], -1.0), __dmul_rn((double)xi3[i], (double)tneg[2]));
            ab = __dadd_rn(ab, __dmul_rn((double)av, (double)bv));
            aa = __dadd_rn(aa, __dmul_rn((double)av, (double)av));
        }
        new_depth = -(float)(ab / aa);
    }
    // sigmaEstimate :73-104
    {
        const float l = length;
        const float lx = (sx - ex) / l, ly = (sy - ey) / l;
        const float alpha = (dmax - dmin) / l;
        const int rx = round_x86(mx), ry = round_x86(my);
        const long long idx = (long long)ry * w + rx; // H9: linear addressing when inside the buffer
        if (rx == INT_MIN || ry == INT_MIN || idx < 0 || idx >= (long long)w * h) {
            new_sigma = -1;
            return;
        }
        const float4 t = __ldg(born_tex + idx);
        const float gx = t.y, gy = t.z;
        if (is_invalid(gx) || is_invalid(gy)) {
            new_sigma = -1;
            return;
        }
        const float g_dot_l = fabsf(gx * lx + gy * ly);
        const float g_dot_l2 = g_dot_l * g_dot_l;
        const float gp2 = g_dot_l / l;
        const float epipolar = (0.5f * 0.5f) / (g_dot_l2 < DVO_EPS ? DVO_EPS : g_dot_l2);
        const float luminance = 2 * (0.5f * 0.5f) / (gp2 < DVO_EPS ? DVO_EPS : gp2);
        new_sigma = alpha * sqrtf(epipolar + luminance);
    }
}

// pose tables for Mapper::update (mapper.cpp:107): r_xi = concatenate(obj.m_xi, -born.m_xi) depends
// on the age only, so it is hoisted out of the pixel loop.  grid = (ceil(max_kf/64), S)
__global__ void k_pose_tables(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, MapState* __restrict__ mst,
                              const int* __restrict__ hist, PoseTab* __restrict__ tab)
{
    const int s = blockIdx.y;
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    MapState* m = mst + s;
    if (!m->do_update || a >= m->n_hist) return;
    const int born_slot = hist[(size_t)s * c.max_kf + (m->n_hist - 1 - a)]; // FrameHistory::operator[] frame.hpp:176
    const float* oxi = slots[m->obj_slot].xi + 6 * s;
    const float* bxi = slots[born_slot].xi + 6 * s;
    float a6[6], nb[6], r[6];
    for (int q = 0; q < 6; q++) { a6[q] = oxi[q]; nb[q] = -bxi[q]; }
    se3_concatenate(a6, nb, r);
    PoseTab* t = tab + (size_t)s * c.max_kf + a;
    se3_exp_neg12(r, t->T);
    t->t[0] = -r[0]; t->t[1] = -r[1]; t->t[2] = -r[2];
}

// exp(rel_xi) for Mapper::update's projection (mapper.cpp:94) and Implement::propagate (:239)
__global__ void k_map_prepare(const __grid_constant__ Consts c, MapState* __restrict__ mst, const float* __restrict__ rel_xi)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    float T[16];
    se3_exp(rel_xi + 6 * s, T);
    for (int q = 0; q < 12; q++) mst[s].Trel[q] = T[q];
    mst[s].rel_z = rel_xi[6 * s + 2];
    mst[s].valid_update = 0;
    mst[s].n_reject = 0;
    mst[s].ticket = 0;
}

// Mapper::update phase A (mapper.cpp:76-137): per ref pixel, epipolar observation + gate + fusion.
// Rejected pixels (Gaussian::update reset, gaussian.cpp:21-25) are only flagged here: their value
// depends on the row-major rank among all rejects (mt19937 order) -> phase B.
// grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_map_update(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                    MapState* __restrict__ mst, const int* __restrict__ hist,
                                                    const PoseTab* __restrict__ tab, uint8_t* __restrict__ flags,
                                                    unsigned int* __restrict__ blockcnt, int nblk)
{
    const int s = blockIdx.y;
    MapState* m = mst + s;
    if (!m->do_update) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    uint8_t fl = 0;
    bool reject = false;
    if (i < g.n) {
        const int y = i / g.w, x = i - y * g.w;
        if (!(x < c.map_x0 || x > c.map_x1 || y < c.map_y0 || y > c.map_y1)) { // mapper.cpp:90
            const SlotDev ref = slots[m->ref_slot];
            const SlotDev obj = slots[m->obj_slot];
            float2* ds = ref.ds[top] + (size_t)s * g.n;
            const float2 dsv = ds[i];
            const float d = dsv.x;
            float wx, wy;
            warp_point(m->Trel, g, (float)x, (float)y, d, wx, wy);
            const int px = round_x86(wx), py = round_x86(wy); // Point2f -> Point2i :94
            if (in_range(px, py, g.w, g.h)) {
                const int age = trunc_x86(ref.age[(size_t)s * g.n + i]); // :99
                if (age < 0 || age >= m->n_hist) {
                    atomicCAS(&m->err, 0, (int)DVO_ERR_HISTORY); // m_history.at() would throw
                } else {
                    const int born_slot = hist[(size_t)s * c.max_kf + (m->n_hist - 1 - age)];
                    const SlotDev born = slots[born_slot];
                    const PoseTab* pt = tab + (size_t)s * c.max_kf + age;
                    const float prior_d = d - m->rel_z; // :104
                    const float prior_s = dsv.y;
                    const float Kc[9] = {g.fx, 0.f, g.cx, 0.f, g.fy, g.cy, 0.f, 0.f, 1.f};
                    const float og = obj.gray[top][(size_t)s * g.n + (size_t)py * g.w + px];
                    float nd, ns;
                    implement_update(born.gray[top] + (size_t)s * g.n, born.tex[top] + (size_t)s * g.n, g, pt->T, pt->t, Kc, px, py,
                                     og, prior_d, prior_s, nd, ns);
                    fl |= 1;
                    if (nd > c.acc_d_lo && nd < c.acc_d_hi && ns > 0.0f && ns < c.acc_s_hi) { // :122
                        fl |= 2;
                        // Gaussian::update (gaussian.cpp:12-31)
                        float gd = prior_d, gs = prior_s;
                        if (gaussian_fuse(gd, gs, nd, ns)) {
                            fl |= 4;
                            ds[i] = make_float2(gd, gs);
                        } else {
                            fl |= 8;
                            reject = true;
                        }
                    }
                }
            }
        }
        flags[(size_t)s * g.n + i] = fl;
    }
    const int nrej = __syncthreads_count(reject);
    const int nval = __syncthreads_count((fl & 4) != 0);
    if (threadIdx.x == 0) {
        blockcnt[(size_t)s * nblk + blockIdx.x] = (unsigned)nrej;
        if (nval) atomicAdd(&m->valid_update, nval);
    }
}

// Mapper::update phase B: replay of the mt19937 stream for the rejected pixels in row-major order.
// rank = (rejects in preceding blocks) + (rejects before this thread in the block).
__global__ void __launch_bounds__(256) k_map_update_apply(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                          MapState* __restrict__ mst, const uint8_t* __restrict__ flags,
                                                          const unsigned int* __restrict__ blockcnt, int nblk,
                                                          const uint32_t* __restrict__ rng_raw)
{
    const int s = blockIdx.y;
    MapState* m = mst + s;
    if (!m->do_update) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned int s_base;
    __shared__ unsigned int s_part[8];
    // offset of this block = sum of preceding block counts
    unsigned int part = 0;
    for (int b = threadIdx.x; b < (int)blockIdx.x; b += blockDim.x) part += blockcnt[(size_t)s * nblk + b];
    for (int off = 16; off > 0; off >>= 1) part += __shfl_down_sync(0xffffffffu, part, off);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_part[warp] = part;
    const bool rej = (i < g.n) && (flags[(size_t)s * g.n + i] & 8);
    const unsigned int bal = __ballot_sync(0xffffffffu, rej);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = 0;
        for (int k = 0; k < 8; k++) t += s_part[k];
        s_base = t;
    }
    __syncthreads();
    if (rej) {
        unsigned int rank = s_base + __popc(bal & ((1u << lane) - 1u));
        for (int k = 0; k < warp; k++) rank += s_warp[k];
        const unsigned long long pos = m->rng_pos + rank;
        if (pos >= c.rng_table) {
            atomicCAS(&m->err, 0, (int)DVO_ERR_RNG);
        } else {
            const float u = dist_from_u32(__ldg(rng_raw + pos));
            const SlotDev ref = slots[m->ref_slot];
            ref.ds[top][(size_t)s * g.n + i] = make_float2(4.0f < u ? 4.0f : u, 0.5f); // gaussian.cpp:22-23
            ref.age[(size_t)s * g.n + i] = 0.f;                                        // mapper.cpp:126
        }
    }
    // the last block of the stream advances the replay position
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned int t = atomicAdd(&m->ticket, 1u);
        if (t == (unsigned)(nblk - 1)) {
            unsigned int total = 0;
            for (int b = 0; b < nblk; b++) total += blockcnt[(size_t)s * nblk + b];
            m->n_reject = total;
            m->rng_pos += total;
            m->ticket = 0;
        }
    }
}

// Implement::propagate (implement.cpp:217-256).  Pass 1: the serial reference lets the LAST source
// pixel in row-major order win a destination; atomicMax of the source index reproduces that
// without a race.  winner must be pre-filled with -1.
__global__ void __launch_bounds__(256) k_propagate_scatter(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                           const MapState* __restrict__ mst, int* __restrict__ winner)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_propagate) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const float rd = slots[m->ref_slot].ds[top][(size_t)s * g.n + i].x;
    if (fabsf(rd) < DVO_EPS) return; // :236
    float wx, wy;
    warp_point(m->Trel, g, (float)x, (float)y, rd, wx, wy); // :239
    const int px = round_x86(wx), py = round_x86(wy);
    if (!in_range(px, py, g.w, g.h)) return; // :240
    atomicMax(winner + (size_t)s * g.n + (size_t)py * g.w + px, i);
}

// Pass 2: gather from the winning source (or the 1/1/0 defaults, :229-231) into new_slot.
__global__ void __launch_bounds__(256) k_propagate_gather(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                          const MapState* __restrict__ mst, const int* __restrict__ winner)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_propagate) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const SlotDev ref = slots[m->ref_slot];
    const SlotDev dst = slots[m->new_slot];
    const int src = winner[(size_t)s * g.n + i];
    float2 out = make_float2(1.0f, 1.0f);
    float age = 0.f;
    if (src >= 0) {
        const float2 v = ref.ds[top][(size_t)s * g.n + src];
        const float d0 = v.x < 0.01f ? 0.01f : v.x; // :244
        const float d1 = d0 + m->rel_z;
        const float q = d1 / d0;
        const float q4 = q * (q * (q * q)); // math::pow(., 4)
        const float sg = sqrtf(q4 * (v.y * v.y) + (0.06f * 0.06f)); // :247
        out = make_float2(d1 < 0.0f ? 0.0f : d1, sg);
        age = ref.age[(size_t)s * g.n + src] + 1; // :252
    }
    dst.ds[top][(size_t)s * g.n + i] = out;
    dst.age[(size_t)s * g.n + i] = age;
}

// Implement::regularize (implement.cpp:156-180): 4-neighbour Gaussian fusion read from the input
// maps; result -> tmp (depth only).  Slot = mst[s].reg_slot.
__global__ void __launch_bounds__(256) k_regularize(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                    const MapState* __restrict__ mst, float* __restrict__ tmp)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_regularize) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const float2* ds = slots[m->reg_slot].ds[top] + (size_t)s * g.n;
    float2 v = ds[i];
    float gd = v.x, gs = v.y;
    if (x - 1 >= 0) { const float2 n = __ldg(ds + i - 1); gaussian_fuse(gd, gs, n.x, n.y); }      // {0,-1}
    if (x + 1 < g.w) { const float2 n = __ldg(ds + i + 1); gaussian_fuse(gd, gs, n.x, n.y); }     // {0, 1}
    if (y + 1 < g.h) { const float2 n = __ldg(ds + i + g.w); gaussian_fuse(gd, gs, n.x, n.y); }   // {1, 0}
    if (y - 1 >= 0) { const float2 n = __ldg(ds + i - g.w); gaussian_fuse(gd, gs, n.x, n.y); }    // {-1,0}
    tmp[(size_t)s * g.n + i] = gd < 6.0f ? gd : 6.0f; // cv::min(new_depth, 6.0) :178
}

// Frame::updateDepth (frame.cpp:56-61) after regularize: finest depth <- tmp, coarser depth AND
// sigma re-culled (this also covers Frame::updateDepthSigma of Mapper::update, mapper.cpp:135).
// grid = (ceil(n_top/256), S, L)
__global__ void __launch_bounds__(256) k_ds_rebuild(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                    const MapState* __restrict__ mst, const float* __restrict__ tmp)
{
    const int l = blockIdx.z, s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_regularize) return;
    const int top = c.L - 1;
    const LevelGeom g = c.lv[l];
    const LevelGeom gt = c.lv[top];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const SlotDev sl = slots[m->reg_slot];
    if (l == top) {
        reinterpret_cast<float*>(sl.ds[top] + (size_t)s * g.n + i)[0] = tmp[(size_t)s * g.n + i];
        return;
    }
    const int t = top - l;
    const int y = i / g.w, x = i - y * g.w;
    const size_t src = (size_t)(y << t) * gt.w + (x << t);
    const float d = tmp[(size_t)s * gt.n + src];
    const float sg = reinterpret_cast<const float*>(sl.ds[top] + (size_t)s * gt.n + src)[1];
    sl.ds[l][(size_t)s * g.n + i] = make_float2(get_pixel_value(d), get_pixel_value(sg));
}

// =============================================================================================
// VisualOdometry control (include/system/system.hpp:44-74, src/map/mapper.cpp:16-60) on device
// =============================================================================================
// After tracking: Frame::updateXi (frame.cpp:7-14), Mapper::needNewFrame (mapper.cpp:45-60).
__global__ void k_vo_decide(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, const TrackState* __restrict__ ts,
                            VoState* __restrict__ vo, MapState* __restrict__ mst, int frame_id, const float* __restrict__ forced_rel)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    VoState* v = vo + s;
    MapState* m = mst + s;
    const int cur_slot = v->cur_slot;
    const SlotDev cur = slots[cur_slot];
    m->valid_update = 0; m->n_reject = 0; m->ticket = 0;
    if (v->kf_slot < 0) { // first frame: m_history.setRefFrame(frame); return identity (system.hpp:49-54)
        for (int q = 0; q < 6; q++) { cur.xi[6 * s + q] = 0.f; cur.rel_xi[6 * s + q] = 0.f; }
        m->do_propagate = 0; m->do_update = 0; m->do_regularize = 0;
        v->need_new = 1; v->was_keyframe = 1; v->n_iters = 0;
        return;
    }
    float rel[6], kxi[6], xi[6];
    const SlotDev kf = slots[v->kf_slot];
    // forced_rel: the caller supplies the relative pose (teacher-forced parity runs, external pose
    // priors); the tracker's own result stays available in the iteration log.
    for (int q = 0; q < 6; q++) { rel[q] = forced_rel ? forced_rel[6 * s + q] : ts[s].xi[q]; kxi[q] = kf.xi[6 * s + q]; }
    se3_concatenate(kxi, rel, xi);
    for (int q = 0; q < 6; q++) { cur.xi[6 * s + q] = xi[q]; cur.rel_xi[6 * s + q] = rel[q]; }
    v->n_iters = ts[s].n_total;
    v->gn_iters += (unsigned long long)ts[s].n_total;
    v->gn_pixels += (unsigned long long)ts[s].px_total;
    bool need = false;
    if (norm_n(rel, 3) > (double)c.min_movement) need = true;       // mapper.cpp:48-52
    else if (frame_id - v->kf_frame_id >= c.max_forward) need = true; // mapper.cpp:53-56
    v->need_new = need ? 1 : 0;
    if (need) v->n_propagate++; else v->n_update++;
    v->was_keyframe = need ? 1 : 0;
    m->ref_slot = v->kf_slot; m->obj_slot = cur_slot; m->new_slot = cur_slot;
    m->n_hist = v->n_kf;
    m->do_propagate = need ? 1 : 0;
    m->do_update = need ? 0 : 1;
    m->do_regularize = 1;
    float T[16];
    se3_exp(rel, T);
    for (int q = 0; q < 12; q++) m->Trel[q] = T[q];
    m->rel_z = rel[2];
}

// FrameHistory::setRefFrame (frame.hpp:151-157) after propagate; regularize target = newest keyframe.
__global__ void k_vo_commit(const __grid_constant__ Consts c, VoState* __restrict__ vo, MapState* __restrict__ mst, int* __restrict__ hist,
                            int frame_id)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    VoState* v = vo + s;
    MapState* m = mst + s;
    const int cur_slot = v->cur_slot;
    if (v->need_new) {
        if (v->n_kf >= c.max_kf) {
            atomicCAS(&m->err, 0, (int)DVO_ERR_HISTORY);
        } else {
            hist[(size_t)s * c.max_kf + v->n_kf] = cur_slot;
            v->n_kf += 1;
            v->kf_slot = cur_slot;
            v->kf_frame_id = frame_id;
        }
    }
    m->reg_slot = v->kf_slot;
}

// se3::exp(frame->m_xi) for odometrize's return value (system.hpp:73)
__global__ void k_pose_out(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, const VoState* __restrict__ vo,
                           float* __restrict__ T_out)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    float T[16];
    se3_exp(slots[vo[s].cur_slot].xi + 6 * s, T);
    for (int q = 0; q < 16; q++) T_out[16 * s + q] = T[q];
}

// Start of a VO frame: the stream's current frame goes into the pool entry after its newest
// keyframe (entry k of the pool holds keyframe k of every stream: FrameHistory::m_history[k]).
__global__ void k_vo_begin(const __grid_constant__ Consts c, VoState* __restrict__ vo, MapState* __restrict__ mst)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    int k = vo[s].n_kf;
    if (k > c.max_kf) k = c.max_kf;
    vo[s].cur_slot = c.vo_base + k;
}

// xi / rel_xi of every stream's current frame, gathered for dvo_vo_state.  out: [S][12]
__global__ void k_vo_gather_xi(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, const VoState* __restrict__ vo,
                               float* __restrict__ out)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= c.S) return;
    const SlotDev& sd = slots[vo[s].cur_slot];
    for (int q = 0; q < 6; q++) { out[12 * s + q] = sd.xi[6 * s + q]; out[12 * s + 6 + q] = sd.rel_xi[6 * s + q]; }
}

// =============================================================================================
// Row-major mask compaction (decoupled 2-pass: block counts, then offsets + scatter).
// =============================================================================================
__global__ void __launch_bounds__(256) k_compact_count(const uint8_t* __restrict__ mask, int n, unsigned int* __restrict__ blockcnt)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = (i < n) && mask[i];
    const int cnt = __syncthreads_count(on);
    if (threadIdx.x == 0) blockcnt[blockIdx.x] = cnt;
}
__global__ void __launch_bounds__(256) k_compact_scatter(const uint8_t* __restrict__ mask, int n, const unsigned int* __restrict__ blockcnt,
                                                         int32_t* __restrict__ idx, unsigned int* __restrict__ total)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ unsigned int s_warp[8], s_part[8], s_base;
    unsigned int part = 0;
    for (int b = threadIdx.x; b < (int)blockIdx.x; b += blockDim.x) part += blockcnt[b];
    for (int off = 16; off > 0; off >>= 1) part += __shfl_down_sync(0xffffffffu, part, off);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_part[warp] = part;
    const bool on = (i < n) && mask[i];
    const unsigned int bal = __ballot_sync(0xffffffffu, on);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int t = 0;
        for (int k = 0; k < 8; k++) t += s_part[k];
        s_base = t;
        if (blockIdx.x == gridDim.x - 1) {
            unsigned int bt = 0;
            for (int k = 0; k < 8; k++) bt += s_warp[k];
            *total = t + bt;
        }
    }
    __syncthreads();
    if (on) {
        unsigned int rank = s_base + __popc(bal & ((1u << lane) - 1u));
        for (int k = 0; k < warp; k++) rank += s_warp[k];
        idx[rank] = i;
    }
}

} // namespace dvo
