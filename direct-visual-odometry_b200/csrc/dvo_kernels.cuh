// dvo_kernels.cuh — hand-written sm_100a kernels for the tracking + mapping hot path.
// Included once by dvo_cuda.cu.  Compiled with -fmad=false (see dvo_device.cuh).
#pragma once
#include "dvo_internal.cuh"

namespace dvo {

// =============================================================================================
// Level views: a coarser level is read through the finest image (dvo_internal.cuh).
// =============================================================================================
__device__ __forceinline__ LevelView level_view(const Consts& c, int l)
{
    LevelView v;
    v.w = c.lv[l].w; v.h = c.lv[l].h; v.t = c.L - 1 - l; v.W = c.lv[c.L - 1].w;
    return v;
}
__device__ __forceinline__ size_t top_index(const LevelView& v, int x, int y) { return (size_t)(y << v.t) * v.W + (x << v.t); }
// gray / depth / sigma of level pixel (x, y): the stored value at the finest level, Convert::getPixel
// (values <= -2 become exactly -2) on every coarser one.
__device__ __forceinline__ float level_value(const LevelView& v, float raw) { return v.t ? get_pixel_value(raw) : raw; }

// =============================================================================================
// Programmatic dependent launch (sm_90+): a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// while its predecessor in the stream still runs; pdl_wait() blocks until the predecessor has completed and its memory is
// visible, pdl_launch() lets the successor's CTAs be scheduled.  Both are no-ops for a kernel launched the ordinary way.
// =============================================================================================
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ int ld_volatile_i32(const int* p) { return *(const volatile int*)p; }

// the frame's first kernel clears the tracking -> mapping hand-over queue (dvo_internal.cuh)
__device__ __forceinline__ void ready_reset(ReadyQ* rq, int S)
{
    if (rq && blockIdx.x == 0 && blockIdx.y == 0) {
        for (int k = threadIdx.x; k < S; k += blockDim.x) rq->list[k] = -1;
        if (threadIdx.x == 0) { rq->n_reserved = 0; rq->n_finished = 0; rq->ticket = 0; }
    }
}

// =============================================================================================
// Frame construction: Frame ctor = Convert::cullImage(gray, culls) (frame.hpp:108-117,
// convert.cpp:7-20).  One thread per 4 finest-level pixels; the input row is read with a stride of
// 2^culls elements.  row_shift = culls when `in` holds every input row, 0 when the host staged only
// the rows that are sampled.   grid = (ceil(n_top / 1024), S)
// =============================================================================================
__global__ void __launch_bounds__(256) k_cull_gray(const __grid_constant__ Consts c, const float* __restrict__ in, int in_rows,
                                                   int row_shift, const SlotDev* __restrict__ slots, int fixed,
                                                   const VoState* __restrict__ vo, ReadyQ* rq)
{
    ready_reset(rq, c.S);
    const int s = blockIdx.y;
    const SlotDev& slot = slots[vo ? vo_cur_slot(c, vo[s].n_kf) : fixed];
    const LevelGeom g = c.lv[c.L - 1];
    const float* __restrict__ src = in + (size_t)s * c.in_w * in_rows;
    float* __restrict__ dst = slot.gray + (size_t)s * g.n;
    const int sh = c.culls;
    const int base = blockIdx.x * 1024 + threadIdx.x;
    float v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int i = base + k * 256;
        v[k] = 0.f;
        if (i < g.n) {
            const int y = i / g.w, x = i - y * g.w;
            v[k] = __ldcs(src + (size_t)(y << row_shift) * c.in_w + (x << sh)); // streamed once: evict-first
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int i = base + k * 256;
        if (i < g.n) dst[i] = (sh == 0) ? v[k] : get_pixel_value(v[k]);
    }
}

// The same from 8-bit gray: Loader::getNormalizedImages' convertTo(CV_32FC1, 1.0/255.0) (loader.cpp:61; OpenCV's cvt_32f for
// 8-bit input: one fp32 multiply by (float)(1.0/255.0)) fused with Convert::cullImage.  A warp reads 32 bytes spread over
// 32 << culls contiguous bytes of one input row.
__global__ void __launch_bounds__(256) k_cull_gray8(const __grid_constant__ Consts c, const uint8_t* __restrict__ in, int in_rows,
                                                    int row_shift, int invalid_code, const SlotDev* __restrict__ slots, int fixed,
                                                    const VoState* __restrict__ vo, ReadyQ* rq)
{
    ready_reset(rq, c.S);
    const int s = blockIdx.y;
    const SlotDev& slot = slots[vo ? vo_cur_slot(c, vo[s].n_kf) : fixed];
    const LevelGeom g = c.lv[c.L - 1];
    const uint8_t* __restrict__ src = in + (size_t)s * c.in_w * in_rows;
    float* __restrict__ dst = slot.gray + (size_t)s * g.n;
    const int sh = c.culls;
    const int base = blockIdx.x * 1024 + threadIdx.x;
    int v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int i = base + k * 256;
        v[k] = 0;
        if (i < g.n) {
            const int y = i / g.w, x = i - y * g.w;
            v[k] = __ldcs(src + (size_t)(y << row_shift) * c.in_w + (x << sh));
        }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int i = base + k * 256;
        if (i < g.n) dst[i] = (v[k] == invalid_code) ? DVO_INV : __fmul_rn((float)v[k], (float)(1.0 / 255.0));
    }
}

// Undistortion map -> source index of every finest-level pixel: cv::remap(INTER_NEAREST) reads source pixel
// (saturate_cast<short>(map_x), saturate_cast<short>(map_y)) = cvRound (round-half-even) clamped to short, BORDER_CONSTANT outside
// the image; Convert::cullImage then keeps destination pixel (x << culls, y << culls).  idx = sy * in_w + sx, or -1.  grid = ceil(n_top/256)
__global__ void __launch_bounds__(256) k_build_remap(const __grid_constant__ Consts c, const float* __restrict__ map_x,
                                                     const float* __restrict__ map_y, int* __restrict__ idx)
{
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t src = (size_t)(y << c.culls) * c.in_w + (x << c.culls);
    int sx = x << c.culls, sy = y << c.culls;
    if (map_x) {
        sx = max(-32768, min(32767, round_x86(map_x[src])));
        sy = max(-32768, min(32767, round_x86(map_y[src]))); // NaN: cvRound gives INT_MIN -> -32768 -> outside
    }
    idx[i] = (sx >= 0 && sx < c.in_w && sy >= 0 && sy < c.in_h) ? sy * c.in_w + sx : -1;
}

// Loader::getNormalizedUndistortedImages + Frame ctor from 8-bit BGR(A): gather through the index map, cvtColor(BGR2GRAY) in
// OpenCV's 15-bit fixed point, convertTo(1/255) as one fp32 multiply, INVALID outside the source image.  grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_cull_bgr8(const __grid_constant__ Consts c, const uint8_t* __restrict__ in, int channels,
                                                   const int* __restrict__ idx, const SlotDev* __restrict__ slots, int fixed,
                                                   const VoState* __restrict__ vo, ReadyQ* rq)
{
    ready_reset(rq, c.S);
    const int s = blockIdx.y;
    const SlotDev& slot = slots[vo ? vo_cur_slot(c, vo[s].n_kf) : fixed];
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int si = __ldg(idx + i);
    float v = DVO_INV;
    if (si >= 0) {
        const uint8_t* __restrict__ p = in + ((size_t)s * c.in_w * c.in_h + (size_t)si) * channels;
        const int b = __ldcs(p), gg = __ldcs(p + 1), r = __ldcs(p + 2);
        const int code = (b * 3735 + gg * 19235 + r * 9798 + (1 << 14)) >> 15;
        v = __fmul_rn((float)code, (float)(1.0 / 255.0));
    }
    slot.gray[(size_t)s * g.n + i] = v;
}

// ---------------------------------------------------------------------------------------------
// Kinect ingest: KinectLoader::getMappedImages + Transform::mapDepthtoGray + Frame(gray, depth, sigma, K, ...).
// ---------------------------------------------------------------------------------------------
struct KinectDev {
    int cw, ch, channels;   // colour image
    float fx, fy, cx, cy;   // colour intrinsics
    float dfx, dfy, dcx, dcy; // depth intrinsics at full resolution (dvo_params.K)
    float T[12];            // invT rows [R | t]
    const int* cidx;        // [ch][cw] source pixel of every UNDISTORTED colour pixel (cv::remap NEAREST), -1 = border (INVALID)
    const int* didx;        // [n_top] source pixel of every sampled undistorted depth pixel, -1 = border (depth 0)
};
// full-resolution variant of k_build_remap (the registration's taps land anywhere in the colour image)   grid = ceil(w h / 256)
__global__ void __launch_bounds__(256) k_build_remap_full(const float* __restrict__ map_x, const float* __restrict__ map_y, int w, int h,
                                                          int* __restrict__ idx)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= w * h) return;
    const int sx = max(-32768, min(32767, round_x86(map_x[i])));
    const int sy = max(-32768, min(32767, round_x86(map_y[i])));
    idx[i] = (sx >= 0 && sx < w && sy >= 0 && sy < h) ? sy * w + sx : -1;
}
// pixel (x, y) of the undistorted, normalised colour image (loader.cpp:110-113 applied to loader.cpp:143-144)
__device__ __forceinline__ float kinect_gray_at(const KinectDev& k, const uint8_t* __restrict__ img, int x, int y)
{
    const int si = __ldg(k.cidx + y * k.cw + x);
    if (si < 0) return DVO_INV;
    const uint8_t* __restrict__ p = img + (size_t)si * k.channels;
    const int code = (__ldg(p) * 3735 + __ldg(p + 1) * 19235 + __ldg(p + 2) * 9798 + (1 << 14)) >> 15;
    return __fmul_rn((float)code, (float)(1.0 / 255.0));
}
// grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_kinect_frame(const __grid_constant__ Consts c, const KinectDev k, const uint8_t* __restrict__ bgr,
                                                      const uint16_t* __restrict__ depth16, SlotDev slot)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const int X = x << c.culls, Y = y << c.culls; // the pixel of the full depth image that Convert::cullImage keeps
    const int di = __ldg(k.didx + i);
    const float depth = di >= 0 ? __fmul_rn((float)depth16[(size_t)s * c.in_w * c.in_h + di], (float)(1.0 / 5000.0)) : 0.f; // loader.cpp:145,113
    float gray = DVO_INV, sigma = 1.0f; // transform.cpp:60-61
    if (!(fabsf(depth) < 1e-6f)) {      // math::isEpsilon, transform.cpp:67
        // backProject with the FULL-resolution depth intrinsics (transform.cpp:25-28): registration happens before Frame culls
        const float Xc = depth * ((float)X - k.dcx) / k.dfx;
        const float Yc = depth * ((float)Y - k.dcy) / k.dfy;
        const float tx = (k.T[0] * Xc + k.T[1] * Yc + k.T[2] * depth) + k.T[3];   // transform.cpp:7-12
        const float ty = (k.T[4] * Xc + k.T[5] * Yc + k.T[6] * depth) + k.T[7];
        const float tz = (k.T[8] * Xc + k.T[9] * Yc + k.T[10] * depth) + k.T[11];
        const float px = tx * k.fx / tz + k.cx, py = ty * k.fy / tz + k.cy;        // transform.cpp:30-33
        // Convert::getSubpixel (convert.cpp:127-177) on the undistorted colour image
        const uint8_t* __restrict__ img = bgr + (size_t)s * k.cw * k.ch * k.channels;
        const int x0 = trunc_x86(px), y0 = trunc_x86(py);
        if (in_range(x0, y0, k.cw, k.ch)) {
            const float hh = px - (float)x0, v = py - (float)y0;
            float t[4];
            t[0] = t[1] = t[2] = t[3] = kinect_gray_at(k, img, x0, y0);
            const bool xin = x0 + 1 < k.cw, yin = y0 + 1 < k.ch;
            if (xin) t[1] = kinect_gray_at(k, img, x0 + 1, y0);
            if (yin) t[2] = kinect_gray_at(k, img, x0, y0 + 1);
            if (xin && yin) t[3] = kinect_gray_at(k, img, x0 + 1, y0 + 1);
            if (repair_taps(t)) gray = bilerp(t[0], t[1], t[2], t[3], hh, v);
        }
        sigma = 0.1f; // transform.cpp:73
    }
    // Frame(gray, depth, sigma, K, levels, culls): every plane point-sampled through Convert::getPixel (frame.hpp:91-106)
    const bool culled = c.culls > 0;
    slot.gray[(size_t)s * g.n + i] = culled ? get_pixel_value(gray) : gray;
    slot.ds[(size_t)s * g.n + i] = make_float2(culled ? get_pixel_value(depth) : depth, sigma);
    slot.age[(size_t)s * g.n + i] = 0.f;
}

// Transform::mapDepthtoGray (src/core/transform.cpp:53-78) on raw float images, the class-level form the reference's callers use
// (KinectLoader::getMappedImages, loader.cpp:116-119): per depth pixel backProject(depth_K) -> transform(invT) -> project(rgb_K) ->
// Convert::getSubpixel(gray); sigma 0.1 where mapped, 1.0 elsewhere.  grid = ceil(w h / 256)
struct MapD2G {
    int w, h, gw, gh;
    float dfx, dfy, dcx, dcy, fx, fy, cx, cy;
    float T[12];
};
__global__ void __launch_bounds__(256) k_map_depth_to_gray(const MapD2G P, const float* __restrict__ depth, const float* __restrict__ gray,
                                                           float* __restrict__ mapped, float* __restrict__ sigma)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.w * P.h) return;
    const int y = i / P.w, x = i - y * P.w;
    const float d = depth[i];
    float g = DVO_INV, sg = 1.0f; // transform.cpp:60-61
    if (!(fabsf(d) < DVO_EPS)) {  // math::isEpsilon :67
        const float Xc = d * ((float)x - P.dcx) / P.dfx, Yc = d * ((float)y - P.dcy) / P.dfy;
        const float tx = (P.T[0] * Xc + P.T[1] * Yc + P.T[2] * d) + P.T[3];
        const float ty = (P.T[4] * Xc + P.T[5] * Yc + P.T[6] * d) + P.T[7];
        const float tz = (P.T[8] * Xc + P.T[9] * Yc + P.T[10] * d) + P.T[11];
        const float px = tx * P.fx / tz + P.cx, py = ty * P.fy / tz + P.cy;
        const int x0 = trunc_x86(px), y0 = trunc_x86(py);
        if (in_range(x0, y0, P.gw, P.gh)) { // Convert::getSubpixel, convert.cpp:127-177
            const float hh = px - (float)x0, v = py - (float)y0;
            const float* p = gray + (size_t)y0 * P.gw + x0;
            float t[4];
            t[0] = t[1] = t[2] = t[3] = __ldg(p);
            const bool xin = x0 + 1 < P.gw, yin = y0 + 1 < P.gh;
            if (xin) t[1] = __ldg(p + 1);
            if (yin) t[2] = __ldg(p + P.gw);
            if (xin && yin) t[3] = __ldg(p + P.gw + 1);
            if (repair_taps(t)) g = bilerp(t[0], t[1], t[2], t[3], hh, v);
        }
        sg = 0.1f; // :73
    }
    mapped[i] = g;
    sigma[i] = sg;
}

// depth / sigma / age upload (Frame 6-arg ctor frame.hpp:91-106, updateDepthSigmaAge frame.cpp:47-54).
// src arrays are [S][sh][sw]; shift = culls when full-res, 0 otherwise.  grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_set_top(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, int fixed,
                                                 const VoState* __restrict__ vo, const float* __restrict__ depth,
                                                 const float* __restrict__ sigma, const float* __restrict__ age, int sw, int sh_,
                                                 int shift)
{
    const int s = blockIdx.y;
    const SlotDev& slot = slots[vo ? vo_cur_slot(c, vo[s].n_kf) : fixed];
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t src = (size_t)s * sw * sh_ + (size_t)(y << shift) * sw + (x << shift);
    float2* ds = slot.ds + (size_t)s * g.n;
    float2 v = ds[i];
    if (depth) { float d = depth[src]; v.x = shift ? get_pixel_value(d) : d; }
    if (sigma) { float d = sigma[src]; v.y = shift ? get_pixel_value(d) : d; }
    ds[i] = v;
    if (age) { float d = age[src]; slot.age[(size_t)s * g.n + i] = shift ? get_pixel_value(d) : d; }
}

// Convert::gradiate (convert.cpp:41-75) of level pixel (x, y), recomputed from the gray image.
__device__ __forceinline__ float grad_x_at(const float* __restrict__ gray, const LevelView& v, int x, int y)
{
    if (x - 1 <= -1 || x + 1 >= v.w) return DVO_INV;
    const float a = __ldg(gray + top_index(v, x - 1, y)), b = __ldg(gray + top_index(v, x + 1, y));
    if (is_invalid(a) || is_invalid(b)) return DVO_INV;
    return b - a;
}
__device__ __forceinline__ float grad_y_at(const float* __restrict__ gray, const LevelView& v, int x, int y)
{
    if (y - 1 <= -1 || y + 1 >= v.h) return DVO_INV;
    const float a = __ldg(gray + top_index(v, x, y - 1)), b = __ldg(gray + top_index(v, x, y + 1));
    if (is_invalid(a) || is_invalid(b)) return DVO_INV;
    return b - a;
}

// one image component of one level for download.  grid = (ceil(n_l/256), S)
__global__ void __launch_bounds__(256) k_extract(const __grid_constant__ Consts c, SlotDev slot, int l, int kind, float* __restrict__ out)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[l];
    const LevelView v = level_view(c, l);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t nt = (size_t)c.lv[c.L - 1].n;
    const size_t ti = (size_t)s * nt + top_index(v, x, y);
    float r;
    switch (kind) {
    case DVO_IMG_GRAY: r = level_value(v, slot.gray[ti]); break;
    case DVO_IMG_DEPTH: r = level_value(v, slot.ds[ti].x); break;
    case DVO_IMG_SIGMA: r = level_value(v, slot.ds[ti].y); break;
    case DVO_IMG_GRADX: r = grad_x_at(slot.gray + (size_t)s * nt, v, x, y); break;
    case DVO_IMG_GRADY: r = grad_y_at(slot.gray + (size_t)s * nt, v, x, y); break;
    default: r = slot.age[ti]; break;
    }
    out[(size_t)s * g.n + i] = r;
}

// ---------------------------------------------------------------------------------------------
// Draw::visualize* (src/core/draw.cpp:7-100), headless: the BGR bytes cv::imshow would be handed.
// OpenCV pieces restated (pinned against cv2 4.13, tests/golden/make_draw.py):
//   convertTo(CV_8UC1, a, b) = saturate_cast<uchar>(cvRound(fma(v, a, b)));
//   8-bit COLOR_HSV2BGR       = float HSV2RGB on (h * 6/180, s/255, v/255), fused 1 - s h, results * 255 truncated.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint8_t cv_convert_8u(float v, float a, float b)
{
    const int q = round_x86(__fmaf_rn(v, a, b));
    return (uint8_t)(q < 0 ? 0 : q > 255 ? 255 : q);
}
__device__ __forceinline__ uint8_t uchar_cast(float v) { return (uint8_t)(trunc_x86(v) & 0xff); } // static_cast<unsigned char>(float) on x86
__device__ __forceinline__ void cv_hsv2bgr_8u(uint8_t H, uint8_t S, uint8_t V, uint8_t* bgr)
{
    const float s = (float)S * (1.f / 255.f), v = (float)V * (1.f / 255.f);
    float h = (float)H * (6.f / 180.f);
    if (h >= 6.f) h -= 6.f;
    int sector = (int)floorf(h);
    h -= (float)sector;
    if ((unsigned)sector >= 6u) { sector = 0; h = 0.f; }
    const float t0 = v, t1 = v * (1.f - s), t2 = v * __fmaf_rn(-s, h, 1.f), t3 = v * __fmaf_rn(-s, 1.f - h, 1.f);
    float b, g, r; // sector_data = {1,3,0},{1,0,2},{3,0,1},{0,2,1},{0,1,3},{2,1,0}
    switch (sector) {
    case 0: b = t1; g = t3; r = t0; break;
    case 1: b = t1; g = t0; r = t2; break;
    case 2: b = t3; g = t0; r = t1; break;
    case 3: b = t0; g = t2; r = t1; break;
    case 4: b = t0; g = t1; r = t3; break;
    default: b = t2; g = t1; r = t0; break;
    }
    const float o[3] = {b, g, r};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int q = (int)(o[k] * 255.f);
        bgr[k] = (uint8_t)(q < 0 ? 0 : q > 255 ? 255 : q);
    }
}

// one pixel of Draw::visualize*: t = the image value the reference's function reads (gray, depth, sigma, a gradient, age),
// sg = sigma for visualizeDepth(depth, sigma)
__device__ __forceinline__ void vis_pixel(int kind, float t, float sg, uint8_t p[3])
{
    p[0] = p[1] = p[2] = 0;
    switch (kind) {
    case DVO_VIS_GRAY: // draw.cpp:7-19
        p[0] = p[1] = p[2] = cv_convert_8u(t, 255.f, 0.f);
        if (!is_valid(t)) p[0] = 255;
        break;
    case DVO_VIS_DEPTH: case DVO_VIS_DEPTH_SIGMA: { // draw.cpp:32-67
        if (t < 1e-6f) break;
        float hx = (t - 0.70f) * 70.0f;
        hx = hx < 0.0f ? 0.0f : (180.0f < hx ? 180.0f : hx); // std::clamp
        uint8_t vv = 255;
        if (kind == DVO_VIS_DEPTH_SIGMA) vv = uchar_cast(-500 * (sg < 0.5f ? sg : 0.5f) + 255);
        cv_hsv2bgr_8u(uchar_cast(hx), 255, vv, p);
        break;
    }
    case DVO_VIS_DEPTH_RAW: { // draw.cpp:21-29
        float q = t < 2.5f ? t : 2.5f;
        q = q > 0.5f ? q : 0.5f;
        p[0] = p[1] = p[2] = cv_convert_8u(q, 100.f, 0.f);
        break;
    }
    case DVO_VIS_SIGMA: p[0] = p[1] = p[2] = cv_convert_8u(t, -500.f, 255.f); break; // draw.cpp:69-75
    case DVO_VIS_GRADX: case DVO_VIS_GRADY: // draw.cpp:77-91
        if (is_invalid(t)) { p[0] = 255; break; }
        p[1] = uchar_cast(255 * (t > 0.0f ? t : 0.0f));
        p[2] = uchar_cast(-255 * (t < 0.0f ? t : 0.0f));
        break;
    default: p[0] = p[1] = p[2] = cv_convert_8u(t, 10.f, 0.f); break; // draw.cpp:94-100
    }
}

// one level of a frame slot.  kind: dvo_vis_kind.  grid = (ceil(n_l/256), S); out: [S][n_l][3]
__global__ void __launch_bounds__(256) k_visualize(const __grid_constant__ Consts c, SlotDev slot, int l, int kind, uint8_t* __restrict__ out)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[l];
    const LevelView v = level_view(c, l);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t nt = (size_t)c.lv[c.L - 1].n;
    const size_t ti = (size_t)s * nt + top_index(v, x, y);
    float t, sg = 0.f;
    switch (kind) {
    case DVO_VIS_GRAY: t = level_value(v, slot.gray[ti]); break;
    case DVO_VIS_DEPTH: case DVO_VIS_DEPTH_RAW: t = level_value(v, slot.ds[ti].x); break;
    case DVO_VIS_DEPTH_SIGMA: { const float2 d = slot.ds[ti]; t = level_value(v, d.x); sg = level_value(v, d.y); break; }
    case DVO_VIS_SIGMA: t = level_value(v, slot.ds[ti].y); break;
    case DVO_VIS_GRADX: t = grad_x_at(slot.gray + (size_t)s * nt, v, x, y); break;
    case DVO_VIS_GRADY: t = grad_y_at(slot.gray + (size_t)s * nt, v, x, y); break;
    default: t = slot.age[ti]; break;
    }
    uint8_t p[3];
    vis_pixel(kind, t, sg, p);
    uint8_t* o = out + ((size_t)s * g.n + i) * 3;
    o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
}

// a raw image, as the reference's callers pass it (Draw::visualizeGray(obj_gray), test/update.cpp:19).  aux = sigma or nullptr
__global__ void __launch_bounds__(256) k_visualize_image(int kind, const float* __restrict__ src, const float* __restrict__ aux, size_t n,
                                                         uint8_t* __restrict__ out)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint8_t p[3];
    vis_pixel(kind, src[i], aux ? aux[i] : 0.f, p);
    out[3 * i] = p[0]; out[3 * i + 1] = p[1]; out[3 * i + 2] = p[2];
}

// =============================================================================================
// Sampling of {gray, gradX, gradY} of one level at a warped point.  Convert::getSubpixel for the
// gray channel (with the invalid-tap repair automaton) and Convert::getSubpixelFromDense for the two
// gradients share x0, y0, h, v and the 2x2 footprint (src/core/convert.cpp:77-105,128-177); the
// gradients of the four footprint pixels are central differences of the 12 gray values around it.
//            x0-1  x0   x1   x0+2
//   y0-1            a0   a1
//   y0        bm    b0   b1   b2
//   y1        cm    c0   c1   c2
//   y0+2            d0   d1
// =============================================================================================
struct Sample3 {
    float I, gx, gy;
    bool in; // (x0, y0) in range
};
template <bool SMEM>
__device__ __forceinline__ float tap_ld(const float* p) { return SMEM ? *p : __ldg(p); }
template <bool WANT_GRAD, bool SMEM = false>
__device__ __forceinline__ Sample3 sample_level(const float* __restrict__ gray, const LevelView& lv, float px, float py)
{
    Sample3 r;
    r.I = r.gx = r.gy = DVO_INV;
    const int x0 = trunc_x86(px), y0 = trunc_x86(py);
    const int w = lv.w, h = lv.h;
    r.in = in_range(x0, y0, w, h);
    if (!r.in) return r;
    const float hh = px - (float)x0, v = py - (float)y0;
    const bool xin = (x0 + 1) < w, yin = (y0 + 1) < h;
    const float* p = gray + top_index(lv, x0, y0);
    const int dx = 1 << lv.t;
    const ptrdiff_t dy = (ptrdiff_t)lv.W << lv.t;
    const float b0 = level_value(lv, tap_ld<SMEM>(p));
    const float b1 = xin ? level_value(lv, tap_ld<SMEM>(p + dx)) : b0;
    const float c0 = yin ? level_value(lv, tap_ld<SMEM>(p + dy)) : b0;
    const float c1 = (xin && yin) ? level_value(lv, tap_ld<SMEM>(p + dy + dx)) : b0;
    if (WANT_GRAD) {
        // gradient of a footprint pixel: INVALID on the image border or next to an invalid tap
        const bool xl = x0 - 1 >= 0, xr = x0 + 2 < w, yu = y0 - 1 >= 0, yd = y0 + 2 < h;
        const float bm = xl ? tap_ld<SMEM>(p - dx) : DVO_INV;
        const float b2 = xr ? tap_ld<SMEM>(p + 2 * dx) : DVO_INV;
        const float a0 = yu ? tap_ld<SMEM>(p - dy) : DVO_INV;
        const float d0 = yd ? tap_ld<SMEM>(p + 2 * dy) : DVO_INV;
        float gx00 = DVO_INV, gy00 = DVO_INV; // texel (x0, y0)
        if (xl && xin && !(is_invalid(bm) || is_invalid(b1))) gx00 = b1 - bm;
        if (yu && yin && !(is_invalid(a0) || is_invalid(c0))) gy00 = c0 - a0;
        float gx10 = gx00, gy10 = gy00, gx01 = gx00, gy01 = gy00, gx11 = gx00, gy11 = gy00;
        if (xin) { // texel (x1, y0)
            const float a1 = yu ? tap_ld<SMEM>(p - dy + dx) : DVO_INV;
            gx10 = (xr && !(is_invalid(b0) || is_invalid(b2))) ? b2 - b0 : DVO_INV;
            gy10 = (yu && yin && !(is_invalid(a1) || is_invalid(c1))) ? c1 - a1 : DVO_INV;
        }
        if (yin) { // texel (x0, y1)
            const float cm = xl ? tap_ld<SMEM>(p + dy - dx) : DVO_INV;
            gx01 = (xl && xin && !(is_invalid(cm) || is_invalid(c1))) ? c1 - cm : DVO_INV;
            gy01 = (yd && !(is_invalid(b0) || is_invalid(d0))) ? d0 - b0 : DVO_INV;
        }
        if (xin && yin) { // texel (x1, y1)
            const float c2 = xr ? tap_ld<SMEM>(p + dy + 2 * dx) : DVO_INV;
            const float d1 = yd ? tap_ld<SMEM>(p + 2 * dy + dx) : DVO_INV;
            gx11 = (xr && !(is_invalid(c0) || is_invalid(c2))) ? c2 - c0 : DVO_INV;
            gy11 = (yd && !(is_invalid(b1) || is_invalid(d1))) ? d1 - b1 : DVO_INV;
        }
        r.gx = bilerp(gx00, gx10, gx01, gx11, hh, v);
        r.gy = bilerp(gy00, gy10, gy01, gy11, hh, v);
    }
    float g[4] = {b0, b1, c0, c1};
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
    const LevelView lv = level_view(c, l);
    __shared__ float T[12];
    if (threadIdx.x == 0) se3_exp_neg12(xi + 6 * s, T);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t nt = (size_t)c.lv[c.L - 1].n;
    const float d = level_value(lv, ref.ds[(size_t)s * nt + top_index(lv, x, y)].x);
    float p = DVO_INV;
    if (!(fabsf(d) < DVO_EPS)) {
        float wx, wy;
        warp_point(T, g, (float)x, (float)y, d, wx, wy);
        p = sample_level<false>(ref.gray + (size_t)s * nt, lv, wx, wy).I;
    }
    out[(size_t)s * g.n + i] = p;
}

// =============================================================================================
// 6x6 solve of the normal equations (replaces cv::solve(A,-B,DECOMP_SVD), optimize.cpp:97):
// Cholesky in fp64; rank-deficient systems fall back to a Jacobi eigen-decomposition with the
// SVD back-substitution threshold (w_i <= 2*FLT_EPSILON*sum(w) dropped) so the minimum-norm
// solution is returned like OpenCV's SVBkSb.
// =============================================================================================
__device__ __noinline__ void solve6_eigen(const double* H, const double* b, double* x)
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

// H: upper triangle, row-major packed (21 values: 00 01 .. 05 11 12 .. 55), as the reduction delivers it.
// Fully unrolled (everything in registers, no local-memory indexing) with one rsqrt per pivot: the diagonal of L is
// only ever used as a divisor, so 1/L_jj is what is kept.  This runs on ONE thread per stream while the stream's other
// threads wait, so its latency is on the critical path of every Gauss-Newton iteration.
__device__ __forceinline__ int tri6(int p, int q) { return p * 6 - (p * (p - 1)) / 2 + (q - p); } // p <= q
__device__ __noinline__ void solve6(const double* Hp, const double* b, double* x)
{
    double Lm[6][6], inv[6];
    double dmax = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) dmax = fmax(dmax, Hp[tri6(i, i)]);
    bool ok = dmax > 0;
    const double thr = dmax * 1e-11;
#pragma unroll
    for (int j = 0; j < 6; j++) {
        double sum = Hp[tri6(j, j)];
#pragma unroll
        for (int k = 0; k < j; k++) sum -= Lm[j][k] * Lm[j][k];
        ok = ok && (sum > thr);
        inv[j] = rsqrt(sum);
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            double s2 = Hp[tri6(j, i)];
#pragma unroll
            for (int k = 0; k < j; k++) s2 -= Lm[i][k] * Lm[j][k];
            Lm[i][j] = s2 * inv[j];
        }
    }
    if (!ok) {
        double H[36];
        for (int p = 0; p < 6; p++)
            for (int q = p; q < 6; q++) H[p * 6 + q] = H[q * 6 + p] = Hp[tri6(p, q)];
        solve6_eigen(H, b, x);
        return;
    }
    double yv[6];
#pragma unroll
    for (int i = 0; i < 6; i++) {
        double s2 = b[i];
#pragma unroll
        for (int k = 0; k < i; k++) s2 -= Lm[i][k] * yv[k];
        yv[i] = s2 * inv[i];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        double s2 = yv[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s2 -= Lm[k][i] * x[k];
        x[i] = s2 * inv[i];
    }
}

// =============================================================================================
// Tracking: Tracker::track (src/track/tracker.cpp:22-85) as ONE persistent launch per frame.
//   grid = (B, S): stream s owns CTAs (0..B-1, s); B = 1 when there are enough streams to fill the
//   GPU, otherwise the stream's pixels are split over B co-resident CTAs (cooperative launch) that
//   meet at a global-memory barrier once per Gauss-Newton iteration.
// Per iteration (Transform::warpImage + Track::optimize fused, transform.cpp:35-51,
// optimize.cpp:10-99): every thread walks its share of the level's (ROI) pixels — gates -> warp ->
// 12 gray taps -> residual, weight, 1x6 Jacobian -> 29 fp64 sums — then the sums are reduced in a
// fixed order (warp shuffle, shared memory, block partials) and the stream's leader thread solves the
// 6x6 system, composes the pose, applies the stop rules and publishes the next iteration's level and
// exp(-xi).  No host round trip, no empty launches.  In VO mode the leader finishes with
// Frame::updateXi + Mapper::needNewFrame + FrameHistory::setRefFrame and the pose table of
// Mapper::update (system.hpp:57-58, mapper.cpp:16-60,107).
// =============================================================================================
struct TrackArgs {
    TrackState* st;       // [S]
    double* partials;     // [S][B][32]
    int B;
    dvo_iter_log* log;    // [S][log_cap]
    int log_cap;
    uint8_t* mask;        // optional [S][n_level] (single-step mode; pre-zeroed)
    // granular mode (vo == nullptr)
    int obj_slot, ref_slot, mode, level0;
    const float* xi0;     // [S][6] or nullptr
    double stop_update;
    float stop_residual;
    // VO mode
    const SlotDev* slots;
    VoState* vo;
    MapState* ms;
    int* hist;
    PoseTab* tab;
    int frame_id;
    const float* forced_rel;
    float* T_out;         // optional [S][16]: se3::exp(frame->m_xi), odometrize's return value (system.hpp:73)
    int* status;          // first device-side error of any stream (dvo_status), read back with the poses
    ReadyQ* rq;           // tracking -> Mapper::update hand-over (nullptr: k_map_update runs after this kernel has completed)
    int* winner;          // [S][n_top] scatter scratch of Implement::propagate, all -1 between uses (nullptr: separate kernels)
};

struct TrackCtl { // per-CTA copy of the stream's iteration state
    int level, iter, done;
    float T[12];  // exp(-xi): what the pixels are warped with
    float Tp[16]; // exp(xi): left operand of the next se3::concatenate (leader CTA only)
    // the leader's copy of what the tail reads every iteration (TrackState in global memory is written, never read back)
    float xi[6];
    int n_total, px_total, mode;
    double stop_update;
    float stop_residual;
};

#ifdef DVO_TRK_PROF
__device__ unsigned int g_trk_stream_ns[1024]; // ns stream s spent in the last k_track launch
__device__ unsigned long long g_trk_prof[8]; // ns of stream 0: [0] pixel loop, [1] reduce, [2] gn_tail, [3] iterations, [4] whole kernel, [5] tail: solve, [6] tail: concatenate, [7] tail: exp + rest
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define TRK_PROF(i, t0) do { if (s == 0 && b == 0 && tid == 0) { const unsigned long long t1_ = gtimer(); g_trk_prof[i] += t1_ - (t0); (t0) = t1_; } } while (0)
#else
#define TRK_PROF(i, t0) do { } while (0)
#endif
#ifdef DVO_TRK_PROF
#define TAIL_PROF(i, t0) do { if (s == 0) { const unsigned long long t1_ = gtimer(); g_trk_prof[i] += t1_ - (t0); (t0) = t1_; } } while (0)
#else
#define TAIL_PROF(i, t0) do { } while (0)
#endif
// One Gauss-Newton tail: solve, compose, log, stop rules (optimize.cpp:92-98, tracker.cpp:46-73).
__device__ __noinline__ void gn_tail(const Consts& c, const TrackArgs& a, TrackState* st, int s, const double* stot, TrackCtl* ctl)
{
#ifdef DVO_TRK_PROF
    unsigned long long tq = gtimer();
#endif
    const int level = ctl->level;
    const LevelGeom& g = c.lv[level];
    const int valid = (int)stot[28];
    float upd[6] = {0, 0, 0, 0, 0, 0};
    float residual = -1.0f; // optimize.cpp:92-93
    if (valid > 0) {
        double xs[6];
        solve6(stot, stot + 21, xs);
        for (int q = 0; q < 6; q++) upd[q] = (float)xs[q];
        residual = (float)stot[27] / (float)valid; // optimize.cpp:98
    }
    TAIL_PROF(5, tq);
    float xi[6], xin[6];
    for (int q = 0; q < 6; q++) xi[q] = ctl->xi[q];
    const int n_total = ctl->n_total;
    const bool advance_pose = (ctl->mode == 0);
    bool moved = false;
    if (advance_pose) {
        se3_concatenate_T(ctl->Tp, upd, xin); // tracker.cpp:46; ctl->Tp = exp(xi) from the previous tail
        bool nan = false;
        for (int q = 0; q < 6; q++) nan |= isnan(xin[q]);
        if (!nan) { // tracker.cpp:47-51
            for (int q = 0; q < 6; q++) xi[q] = xin[q];
            moved = true;
        }
    }
    TAIL_PROF(6, tq);
    if (a.log && n_total < a.log_cap) {
        dvo_iter_log* L = a.log + (size_t)s * a.log_cap + n_total;
        L->level = level; L->iteration = ctl->iter; L->valid_pixels = valid; L->residual = residual;
        for (int q = 0; q < 6; q++) { L->update[q] = upd[q]; L->xi_after[q] = xi[q]; }
    }
    ctl->n_total = n_total + 1;
    ctl->px_total += g.n;
    st->n_total = ctl->n_total;
    st->px_total = ctl->px_total;
    if (!advance_pose) {
        ctl->done = 1;
    } else {
        for (int q = 0; q < 6; q++) { st->xi[q] = xi[q]; ctl->xi[q] = xi[q]; }
        // stop rules, tracker.cpp:68-70 (time budget never reached)
        const bool stop = (norm_n(upd, 6) < ctl->stop_update) || (residual < ctl->stop_residual);
        int iter = ctl->iter + 1;
        if (stop || iter >= c.max_iter) {
            iter = 0;
            ctl->level = level + 1;
            if (ctl->level >= c.L) ctl->done = 1;
        }
        ctl->iter = iter;
        // stuff.update(xi): the next iteration warps with exp(-xi) - not needed after the last iteration of the last level
        if (moved && !ctl->done) se3_exp(xi, ctl->Tp, ctl->T);
    }
    // publish for the other CTAs of the stream and for the host
    st->level = ctl->level; st->iter = ctl->iter; st->done = ctl->done;
    for (int q = 0; q < 12; q++) st->Tn[q] = ctl->T[q];
    TAIL_PROF(7, tq);
}

// pose table of Mapper::update (mapper.cpp:107): r_xi = concatenate(obj.m_xi, -born.m_xi) depends on the
// age only, so it is hoisted out of the pixel loop.  One entry per call.
__device__ inline void pose_table_entry(const Consts& c, const SlotDev* __restrict__ slots, const MapState* m, const int* __restrict__ hist,
                                        PoseTab* __restrict__ tab, int s, int age)
{
    const int born_slot = hist[(size_t)s * c.kf_stride + (m->n_hist - 1 - age)]; // FrameHistory::operator[] frame.hpp:176
    const float* oxi = slots[m->obj_slot].xi + 6 * s;
    const float* bxi = slots[born_slot].xi + 6 * s;
    float a6[6], nb[6], r[6];
    for (int q = 0; q < 6; q++) { a6[q] = oxi[q]; nb[q] = -bxi[q]; }
    se3_concatenate(a6, nb, r);
    PoseTab* t = tab + (size_t)s * c.kf_stride + age;
    se3_exp_neg12(r, t->T);
    t->t[0] = -r[0]; t->t[1] = -r[1]; t->t[2] = -r[2];
}

// After tracking (one thread per stream): Frame::updateXi (frame.cpp:7-14), Mapper::needNewFrame
// (mapper.cpp:45-60) and FrameHistory::setRefFrame (frame.hpp:151-157).  The keyframe list is appended
// here already; propagate / update work from the snapshot (ref_slot, n_hist) taken before.
__device__ __noinline__ void vo_decide(const Consts& c, const TrackArgs& a, const TrackState* st, int s)
{
    VoState* v = a.vo + s;
    MapState* m = a.ms + s;
    const int cur_slot = vo_cur_slot(c, v->n_kf);
    v->cur_slot = cur_slot;
    const SlotDev cur = a.slots[cur_slot];
    v->map_matched += (unsigned long long)(unsigned int)m->valid_update; // the previous frame's update, before its count is cleared
    m->valid_update = 0; m->n_reject = 0; m->ticket = 0;
    if (v->kf_slot < 0) { // first frame: m_history.setRefFrame(frame); return identity (system.hpp:49-54)
        for (int q = 0; q < 6; q++) { cur.xi[6 * s + q] = 0.f; cur.rel_xi[6 * s + q] = 0.f; }
        m->do_propagate = 0; m->do_update = 0; m->do_regularize = 0;
        v->need_new = 1; v->was_keyframe = 1; v->n_iters = 0;
        a.hist[(size_t)s * c.kf_stride] = cur_slot;
        v->n_kf = 1; v->kf_slot = cur_slot; v->kf_frame_id = a.frame_id;
        m->reg_slot = cur_slot;
        if (a.status) atomicMax(a.status + 1, 1);
        return;
    }
    float rel[6], kxi[6], xi[6];
    const SlotDev kf = a.slots[v->kf_slot];
    // forced_rel: the caller supplies the relative pose (teacher-forced parity runs, external pose
    // priors); the tracker's own result stays available in the iteration log.
    for (int q = 0; q < 6; q++) { rel[q] = a.forced_rel ? a.forced_rel[6 * s + q] : st->xi[q]; kxi[q] = kf.xi[6 * s + q]; }
    se3_concatenate(kxi, rel, xi);
    for (int q = 0; q < 6; q++) { cur.xi[6 * s + q] = xi[q]; cur.rel_xi[6 * s + q] = rel[q]; }
    v->n_iters = st->n_total;
    v->gn_iters += (unsigned long long)st->n_total;
    v->gn_pixels += (unsigned long long)st->px_total;
    bool need = false;
    if (norm_n(rel, 3) > (double)c.min_movement) need = true;          // mapper.cpp:48-52
    else if (a.frame_id - v->kf_frame_id >= c.max_forward) need = true; // mapper.cpp:53-56
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
    if (need) {
        if (v->n_kf >= c.max_kf) {
            atomicCAS(&m->err, 0, (int)DVO_ERR_HISTORY);
        } else {
            a.hist[(size_t)s * c.kf_stride + v->n_kf] = cur_slot;
            v->n_kf += 1;
            v->kf_slot = cur_slot;
            v->kf_frame_id = a.frame_id;
        }
    }
    m->reg_slot = v->kf_slot;
    if (a.status) atomicMax(a.status + 1, v->n_kf); // the host grows the keyframe pool by the largest count of any stream
    // a stream that has failed (history overflow) keeps tracking against its last good keyframe but is no longer mapped: its
    // maps would silently stop matching the reference (the reference itself would have thrown, frame.hpp:176)
    if (m->err) {
        m->do_propagate = 0; m->do_update = 0; m->do_regularize = 0;
        if (a.status) atomicCAS(a.status, 0, m->err);
    }
}

// Implement::propagate (implement.cpp:217-256) + Frame::updateDepthSigmaAge for ONE stream by ONE CTA, in the tracking kernel's
// epilogue: the streams that take a new keyframe are a third of all, their scatter + gather is 2 x 19 200 pixels, and doing it
// here overlaps it with the streams that still iterate instead of two more launches that every stream's CTAs attend.  Same
// arithmetic as k_propagate_scatter / k_propagate_gather (atomicMax of the source index: the LAST source in row-major order wins).
__device__ __forceinline__ void propagate_stream(const Consts& c, const SlotDev* __restrict__ slots, const MapState* m, int* __restrict__ winner,
                                                 int s, int tid, int nthreads);

__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) { return *(const volatile unsigned int*)p; }

// STAGE: coarse reference levels (strided views of the finest level) are gathered once per level into a dense shared-memory tile.
// Pays when a stream's Gauss-Newton chain is the critical path (few streams, CTAs of a stream split over the GPU: +11 % GN
// iterations/s on C1); costs co-residency with Mapper::update when 256 streams fill the GPU (early frames -12 %): host picks.
template <bool STAGE>
__device__ __forceinline__ void track_body(const Consts& c, const TrackArgs& a)
{
    const int s = blockIdx.y, b = blockIdx.x, B = a.B;
    const int tid = threadIdx.x;
    const bool leader = (b == 0);
    TrackState* __restrict__ st = a.st + s;

    __shared__ TrackCtl ctl;
    __shared__ __align__(16) double sC[TRK_THREADS / 32][64]; // per-warp X^T X
    __shared__ int scnt[TRK_THREADS / 32];
    __shared__ double stot[32];
    __shared__ __align__(16) float srow[TRK_THREADS / 32][32 * 8];  // per-warp X: 32 pixels x {J[0..5], w r, r}
    __shared__ const float* s_obj;
    __shared__ const float* s_ref;
    __shared__ const float2* s_ds;
    __shared__ unsigned int s_gen0;
    __shared__ float s_img[STAGE ? TRK_STAGE_PIXELS : 1]; // the reference level as a dense tile (coarse levels only)
    int staged_level = -1;

    // the ingest kernel before this one has completed (a no-op unless launched as a programmatic dependent); from here on
    // k_map_update's CTAs may be scheduled onto whatever this kernel leaves free (they wait per stream, see ReadyQ)
    pdl_wait();
    pdl_launch();

    // ---- prologue: Tracker::track's xi = 0, level 0 (tracker.cpp:28-40); every CTA of the stream
    // derives the same initial state, the leader also publishes it for the host.
    if (tid == 0) {
        int obj_slot = a.obj_slot, ref_slot = a.ref_slot;
        int done = 0;
        if (a.vo) {
            // All streams of a context are in lock-step, so "first frame" is a launch-wide fact passed by
            // the host (frame_id == 0) — it must not be read from VoState, which the stream's leader CTA
            // rewrites in its epilogue while a late sibling CTA may still be in this prologue.
            if (a.frame_id == 0) { obj_slot = ref_slot = c.vo_base; done = 1; } // nothing to track against
            else { obj_slot = vo_cur_slot(c, a.vo[s].n_kf); ref_slot = a.vo[s].kf_slot; }
        }
        const size_t off = (size_t)s * c.lv[c.L - 1].n;
        s_obj = a.slots[obj_slot].gray + off;
        s_ref = a.slots[ref_slot].gray + off;
        s_ds = a.slots[ref_slot].ds + off;
        float xi[6];
        for (int q = 0; q < 6; q++) xi[q] = a.xi0 ? a.xi0[6 * s + q] : 0.f;
        se3_exp(xi, ctl.Tp, ctl.T);
        ctl.level = a.level0; ctl.iter = 0; ctl.done = done;
        for (int q = 0; q < 6; q++) ctl.xi[q] = xi[q];
        ctl.n_total = 0; ctl.px_total = 0; ctl.mode = a.mode; ctl.stop_update = a.stop_update; ctl.stop_residual = a.stop_residual;
        s_gen0 = (B > 1) ? ld_volatile_u32(&st->gen) : 0u;
        if (leader) {
            for (int q = 0; q < 6; q++) st->xi[q] = xi[q];
            for (int q = 0; q < 12; q++) st->Tn[q] = ctl.T[q];
            st->level = ctl.level; st->iter = 0; st->done = done; st->n_total = 0; st->px_total = 0; st->mode = a.mode;
            st->stop_update = a.stop_update; st->stop_residual = a.stop_residual;
            st->obj_gray = s_obj; st->ref_gray = s_ref; st->ref_ds = s_ds;
        }
    }
    __syncthreads();
    const float* __restrict__ obj_gray = s_obj;
    const float* __restrict__ ref_gray = s_ref;
    const float2* __restrict__ ref_ds = s_ds;
    const unsigned int gen0 = s_gen0;
    unsigned int epoch = 0;
#ifdef DVO_TRK_PROF
    unsigned long long tp = gtimer();
    const unsigned long long tp_start = tp;
#endif
    while (!ctl.done) {
        const int level = ctl.level;
        const LevelGeom g = c.lv[level];
        const LevelView lv = level_view(c, level);
        const float* sT = ctl.T; // exp(-xi), read from shared memory (broadcast) where it is used
        float step = c.step_default; // optimize.cpp:22-26
        if (level == 1) step = c.step_level1;
        if (level == 2) step = c.step_level2;
        // pixels of this iteration: the whole level, or the ROI of optimize.cpp:33-36
        int rx0 = 0, rx1 = g.w - 1, ry0 = 0, ry1 = g.h - 1;
        if (level == c.roi_level) {
            rx0 = max(rx0, c.roi_x0); rx1 = min(rx1, c.roi_x1); ry0 = max(ry0, c.roi_y0); ry1 = min(ry1, c.roi_y1);
        }
        const int rw = rx1 - rx0 + 1, rh = ry1 - ry0 + 1;
        const int count = (rw > 0 && rh > 0) ? rw * rh : 0;
        const int per = (count + B - 1) / B;
        const int j0 = b * per, j1 = min(count, j0 + per);

        // Normal equations on the fp64 TENSOR CORES.  With X the 32 x 8 matrix of a warp-batch (row = pixel, columns J[0..5], w r, r)
        // the iteration needs X^T X: 21 J^T J + 6 J^T (w r) + sum r^2 are all entries of that 8 x 8 product.  Each lane leaves its
        // row as floats in shared memory and the warp issues 8 mma.sync.m8n8k4.f64 (4 pixels each): A (8 x 4, row-major) and B
        // (4 x 8, column-major) want X[k][r] on lane 4 r + k BOTH, so one conflict-free LDS.32 + F2D feeds both operands and the
        // 8 x 8 fp64 accumulator is 2 registers pairs per lane.  Exact fp64 FMAs like the scalar code, ~30 instructions per batch
        // instead of 110 (one accumulator per sum through shared memory) or 58 registers (29 accumulators per thread).
        // coarse levels are strided views of the finest level: with STAGE the reference level is gathered ONCE per level into a dense
        // shared-memory tile (cullImage's getPixel rule applied), so the 12 taps per pixel are LDS instead of L1 lookups
        const bool staged = STAGE && lv.t > 0 && g.n <= TRK_STAGE_PIXELS;
        if (staged && staged_level != level) {
            for (int i = tid; i < g.n; i += TRK_THREADS) {
                const int yy = i / g.w, xx = i - yy * g.w;
                s_img[i] = level_value(lv, __ldg(ref_gray + top_index(lv, xx, yy)));
            }
            staged_level = level;
            __syncthreads();
        }
        LevelView sv; sv.w = g.w; sv.h = g.h; sv.t = 0; sv.W = g.w;
        const int lane = tid & 31, warp = tid >> 5;
        double c0 = 0.0, c1 = 0.0; // C[lane >> 2][2 (lane & 3) + {0, 1}] of X^T X
        int nvalid = 0;
        float* rows = srow[warp];  // [32 pixels][8]
        const int frag_ofs = (lane & 3) * 8 + (lane >> 2); // X[k = lane & 3][r = lane >> 2] within a group of 4 pixels

        // A warp-batch is two dependent gathers deep (depth -> warp -> 12 gray taps) and a frame is only ~50 batches per warp, so
        // the first gather of batch k + 1 is requested before batch k is processed; pixel coordinates advance incrementally.
        int jn = j0 + warp * 32 + lane, xn = 0, yn = 0;
        float2 dsr_n = make_float2(0.f, 0.f);
        float i1_n = 0.f;
        if (jn < j1) {
            const int jy = jn / rw;
            yn = ry0 + jy; xn = rx0 + (jn - jy * rw);
            const size_t ti = top_index(lv, xn, yn);
            dsr_n = __ldg(ref_ds + ti);
            i1_n = __ldg(obj_gray + ti);
        }
        const int step_x = TRK_THREADS % max(rw, 1), step_y = TRK_THREADS / max(rw, 1);
        for (int jb = j0 + warp * 32; jb < j1; jb += TRK_THREADS) {
            const int j = jn;
            const int x = xn, y = yn;
            const float2 dsr = dsr_n;
            const float i1_raw = i1_n;
            jn += TRK_THREADS;
            xn += step_x; yn += step_y;
            if (xn > rx1) { xn -= rw; yn++; }
            if (jn < j1) {
                const size_t ti = top_index(lv, xn, yn);
                dsr_n = __ldg(ref_ds + ti);
                i1_n = __ldg(obj_gray + ti);
            }
            bool ok = false;
            float J[6], bw = 0.f, rr = 0.f;
            if (j < j1) {
                const float depth = level_value(lv, dsr.x);
                const float I1 = level_value(lv, i1_raw);
                ok = !(depth < c.min_depth_f) && !is_invalid(I1); // optimize.cpp:39,46
                if (ok) {
                    float X, Y, Z;
                    back_project(g, (float)x, (float)y, depth, X, Y, Z);
                    const float tx = (sT[0] * X + sT[1] * Y + sT[2] * Z) + sT[3];
                    const float ty = (sT[4] * X + sT[5] * Y + sT[6] * Z) + sT[7];
                    const float tz = (sT[8] * X + sT[9] * Y + sT[10] * Z) + sT[11];
                    const float wx = tx * g.fx / tz + g.cx;
                    const float wy = ty * g.fy / tz + g.cy;
                    const Sample3 sm = (STAGE && staged) ? sample_level<true, STAGE>(s_img, sv, wx, wy) : sample_level<true>(ref_gray, lv, wx, wy);
                    // I_2 valid (optimize.cpp:46); warped point inside the image (:52-56); gradients valid (:61)
                    ok = sm.in && !is_invalid(sm.I) && !(wx < 0 || wy < 0 || (float)g.w <= wx || (float)g.h <= wy)
                         && !(is_invalid(sm.gx) || is_invalid(sm.gy));
                    if (ok && c.min_grad2 > 0.0f) ok = !(sm.gx * sm.gx + sm.gy * sm.gy < c.min_grad2); // semi-dense selection (extra, default off)
                    if (ok) {
                        const float fgx = g.fx * sm.gx, fgy = g.fy * sm.gy; // optimize.cpp:67-77
                        const float xz = X / Z, yz = Y / Z;
                        J[0] = fgx / Z;
                        J[1] = fgy / Z;
                        J[2] = -(fgx * X + fgy * Y) / Z / Z;
                        J[3] = -fgx * xz * yz - fgy * (1.0f + yz * yz);
                        J[4] = fgx * (1.0f + xz * xz) + fgy * xz * yz;
                        J[5] = (-fgx * yz + fgy * xz);
                        const float r = sm.I - I1;
                        const float sg0 = level_value(lv, dsr.y);
                        const float sg = (sg0 < c.sigma_lo) ? c.sigma_lo : (c.sigma_hi < sg0) ? c.sigma_hi : sg0; // std::clamp :83
                        bw = r * (step / sg);
                        rr = r;
                        if (c.huber_delta > 0.0f) { // variance-normalised Huber weight (extra, default off): row and entry scaled by sqrt(w)
                            const float t = fabsf(r) / sg;
                            const float wh = (t <= c.huber_delta) ? 1.0f : c.huber_delta / t;
                            const float sw = sqrtf(wh);
#pragma unroll
                            for (int q = 0; q < 6; q++) J[q] = J[q] * sw;
                            bw = bw * sw;
                        }
                        nvalid++;
                    }
                }
                if (a.mask && ok) a.mask[(size_t)s * g.n + (size_t)y * g.w + x] = 1;
            }
            const unsigned vm = __ballot_sync(0xffffffffu, ok);
            if (vm) { // warp-uniform
                float4* rw4 = reinterpret_cast<float4*>(rows + lane * 8);
                if (ok) {
                    rw4[0] = make_float4(J[0], J[1], J[2], J[3]);
                    rw4[1] = make_float4(J[4], J[5], bw, rr);
                } else {
                    rw4[0] = rw4[1] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                __syncwarp();
#pragma unroll
                for (int m = 0; m < 8; m++) {
                    if ((vm >> (4 * m)) & 0xfu) { // warp-uniform: skip groups of 4 pixels without a valid one
                        const double xv = (double)rows[m * 32 + frag_ofs];
                        asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                            : "+d"(c0), "+d"(c1) : "d"(xv), "d"(xv));
                    }
                }
                __syncwarp();
            }
        }

        // ---- reduction in a fixed order (deterministic for a given B) ----
        {
            const int vn = __reduce_add_sync(0xffffffffu, nvalid);
            double2* cw = reinterpret_cast<double2*>(sC[warp] + (lane >> 2) * 8 + 2 * (lane & 3));
            *cw = make_double2(c0, c1);
            if (lane == 0) scnt[warp] = vn;
        }
        __syncthreads();
        TRK_PROF(0, tp);
        if (tid < NACC) {
            // sum k: 0..20 = (J^T J)[p][q], p <= q, row-major upper triangle; 21..26 = (J^T (w r))[q]; 27 = sum r^2; 28 = valid pixels
            int e = 63;
            if (tid < 21) {
                int p = 0, rem = tid;
                while (rem >= 6 - p) { rem -= 6 - p; p++; }
                e = p * 8 + p + rem;
            } else if (tid < 27) e = (tid - 21) * 8 + 6;
            double v = 0;
            if (tid == 28) {
                int n = 0;
#pragma unroll
                for (int wv = 0; wv < TRK_THREADS / 32; wv++) n += scnt[wv];
                v = (double)n;
            } else {
#pragma unroll
                for (int wv = 0; wv < TRK_THREADS / 32; wv++) v += sC[wv][e];
            }
            if (B == 1) stot[tid] = v;
            else a.partials[((size_t)s * B + b) * 32 + tid] = v;
        }
        epoch++;
        if (B > 1) {
            __threadfence();
            __syncthreads();
            if (tid == 0) {
                atomicAdd(&st->arrive, 1u);
                if (leader) { // wait until every CTA of the stream has delivered this iteration's partials
                    const unsigned int target = (gen0 + epoch) * (unsigned int)B;
                    while ((int)(ld_volatile_u32(&st->arrive) - target) < 0) __nanosleep(32);
                    __threadfence();
                }
            }
            __syncthreads();
            if (leader && tid < NACC) {
                const double* base = a.partials + (size_t)s * B * 32 + tid;
                double v = 0;
                for (int bb = 0; bb < B; bb++) v += __ldcg(base + (size_t)bb * 32);
                stot[tid] = v;
            }
        }
        __syncthreads();
        TRK_PROF(1, tp);
        if (leader) {
            if (tid == 0) {
                gn_tail(c, a, st, s, stot, &ctl);
                if (B > 1) {
                    __threadfence();
                    atomicExch(&st->gen, gen0 + epoch);
                }
            }
        } else if (tid == 0) {
            const unsigned int target = gen0 + epoch;
            while ((int)(ld_volatile_u32(&st->gen) - target) < 0) __nanosleep(32);
            __threadfence();
            ctl.level = __ldcg(&st->level); ctl.iter = __ldcg(&st->iter); ctl.done = __ldcg(&st->done);
            for (int q = 0; q < 12; q++) ctl.T[q] = __ldcg(&st->Tn[q]);
        }
        __syncthreads();
        TRK_PROF(2, tp);
#ifdef DVO_TRK_PROF
        if (s == 0 && b == 0 && tid == 0) g_trk_prof[3]++;
#endif
    }
#ifdef DVO_TRK_PROF
    if (s == 0 && b == 0 && tid == 0) g_trk_prof[4] += gtimer() - tp_start;
    if (b == 0 && tid == 0 && s < 1024) g_trk_stream_ns[s] = (unsigned int)(gtimer() - tp_start);
#endif

    // ---- VO epilogue (leader CTA of each stream) ----
    if (a.vo && leader) {
        if (tid == 0) {
            vo_decide(c, a, st, s);
            if (a.T_out) { // se3::exp(frame->m_xi), system.hpp:73 (identity on the first frame)
                float xi[6], T[16];
                const float* cx = a.slots[a.vo[s].cur_slot].xi + 6 * s;
                for (int q = 0; q < 6; q++) xi[q] = cx[q];
                se3_exp(xi, T);
                for (int q = 0; q < 16; q++) a.T_out[16 * s + q] = T[q];
            }
        }
        __syncthreads();
        const MapState* m = a.ms + s;
        const int do_update = m->do_update, do_propagate = m->do_propagate;
        if (do_update)
            for (int age = tid; age < m->n_hist; age += TRK_THREADS) pose_table_entry(c, a.slots, m, a.hist, a.tab, s, age);
        if (a.rq) {
            // publish: everything Mapper::update reads of this stream (MapState, history, pose table, frame poses) is written
            __threadfence();
            __syncthreads();
            if (tid == 0) {
                if (do_update) {
                    const int pos = atomicAdd(&a.rq->n_reserved, 1);
                    *(volatile int*)&a.rq->list[pos] = s;
                    __threadfence();
                }
                atomicAdd(&a.rq->n_finished, 1);
            }
        }
        if (a.winner && do_propagate) propagate_stream(c, a.slots, m, a.winner + (size_t)s * c.lv[c.L - 1].n, s, tid, TRK_THREADS);
    }
}
__global__ void __launch_bounds__(TRK_THREADS, TRK_MINB) k_track(const __grid_constant__ Consts c, const TrackArgs a) { track_body<false>(c, a); }
__global__ void __launch_bounds__(TRK_THREADS, TRK_MINB) k_track_staged(const __grid_constant__ Consts c, const TrackArgs a) { track_body<true>(c, a); }

// pose tables for the granular Mapper::update entry point.  grid = (ceil(max_kf/64), S)
__global__ void k_pose_tables(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, MapState* __restrict__ mst,
                              const int* __restrict__ hist, PoseTab* __restrict__ tab)
{
    const int s = blockIdx.y;
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    MapState* m = mst + s;
    if (!m->do_update || a >= m->n_hist) return;
    pose_table_entry(c, slots, m, hist, tab, s, a);
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
// `cv::norm(Point2f(x, y)) < L` (implement.cpp:120) for a float-valued L, without the fp64 square root
// on the common path.  L*L is exact in double (24-bit L); sqrt is monotonic and correctly rounded, so
// d2 >= L*L gives sqrt(d2) >= L, and d2 <= L*L*(1 - 2^-49) gives sqrt(d2) < L - ulp(L); only the sliver
// in between (and NaNs) needs the real square root.  Bit-identical to the plain comparison.
struct NormLess {
    double L, hi, lo;
    __device__ __forceinline__ explicit NormLess(float length) : L((double)length)
    {
        hi = __dmul_rn(L, L);
        lo = __dmul_rn(hi, 1.0 - 1.7763568394002505e-15);
    }
    __device__ __forceinline__ bool operator()(float x, float y) const
    {
        const double d2 = __dadd_rn(__dmul_rn((double)x, (double)x), __dmul_rn((double)y, (double)y));
        if (d2 >= hi) return false;
        if (d2 <= lo) return true;
        return sqrt(d2) < L;
    }
};

// ---------------------------------------------------------------------------------------------
// doMatching (implement.cpp:106-152) as a WARP-COOPERATIVE search.  The reference walks each pixel's
// epipolar segment serially; segment lengths differ wildly between neighbouring pixels (1 .. 102 steps,
// measured lane efficiency of a one-pixel-per-lane loop: 0.26-0.37), so the (pixel, step) pairs of the 32
// pixels of a warp are flattened and dealt out 32 at a time:
//   1. every lane walks its own segment WITHOUT sampling (pt += dir, the reference's float accumulation,
//      and the `norm(pt - start) < length` / 102-step stop rules) to learn its step count, leaving a
//      checkpoint of pt every 8 steps in shared memory;
//   2. the work unit is a CHUNK of 8 consecutive steps of one pixel, starting exactly on a checkpoint; an
//      inclusive warp scan of the chunk counts orders all chunks of the warp;
//   3. lane i of round r takes chunk (32 r + i): a head-lane table gives (pixel p, chunk c); the lane runs the
//      reference's loop body for steps 8c+1 .. 8c+8 from the checkpoint (same additions in the same order as
//      the serial walk; the tap pt + dir is bit-for-bit the next step's centre tap and is carried over);
//   4. the weighted SSD (float accumulator, double terms, implement.cpp:134) is packed with the step index
//      into one 64-bit key (ssd >= 0: its bit pattern is monotonic); the lane keeps the smallest key of its
//      chunk, a segmented warp min (redux.sync) keeps, per pixel, the smallest ssd and among equal ones the
//      earliest step — the reference's strict `<` rule.
// Results are bit-identical to the serial walk.
// ---------------------------------------------------------------------------------------------
#ifdef DVO_MAP_STATS
__device__ unsigned long long g_map_stats[8]; // [0] steps, [1] lane slots spent on them
__device__ unsigned int g_map_hist[128];
#endif

#ifndef DVO_CP_SHIFT
#define DVO_CP_SHIFT 3
#endif
constexpr int SEARCH_CP_SHIFT = DVO_CP_SHIFT;     // checkpoint every 2^shift steps
constexpr int SEARCH_NCP = (102 >> SEARCH_CP_SHIFT) + 1; // steps 0, 8, ..., 96
#ifndef DVO_SEARCH_UNROLL
#define DVO_SEARCH_UNROLL 1
#endif
constexpr int SEARCH_UNROLL = DVO_SEARCH_UNROLL; // steps of a chunk whose loads are in flight together
struct SearchSmem { // per warp
    float2 cp[SEARCH_NCP][32];
    float dirx[32], diry[32], og[32];
    const float4* born[32];
    int nsteps[32];
    int2 owner[32]; // per round: run head lane -> {pixel, first chunk index of the pixel}
    unsigned long long best[32];
};

#define SEARCH_TAP(img, w, h, wf, hf, x, y) subpixel_quad(img, w, wf, hf, x, y)
__device__ __forceinline__ float search_ssd(float s0, float s1, float s2, float og)
{
    // implement.cpp:122-139: a tap that is invalid ends the accumulation with ssd = 2.0 * 3
    if (is_invalid(s0) || is_invalid(s1) || is_invalid(s2)) return 2.0f * 3;
    const float d0 = s0 - og, d1 = s1 - og, d2 = s2 - og;
    // first term: (float)(0.0 + (double)(1/3) * (double)a), a = d0 * d0 >= 0, equals the correctly rounded float quotient a / 3 for
    // EVERY non-negative finite float (a / 3 is never within 2^-27 relative of a float rounding boundary, the double product is within
    // 2^-52 of it; checked exhaustively over all 2^31 bit patterns, tests/test_oracle_known_answers.py samples it).  a / 3 itself =
    // Markstein's two-FMA correction of a * RN(1/3): three fp32 instructions instead of two conversions and two fp64 operations.
    const float a0 = d0 * d0;
    const float q0 = __fmul_rn(a0, 0.3333333432674407958984375f);
    float ssd = __fmaf_rn(__fmaf_rn(-3.0f, q0, a0), 0.3333333432674407958984375f, q0); // :134
    ssd = (float)__dadd_rn((double)ssd, __dmul_rn(1.0 * 2 / 3, (double)(d1 * d1)));
    // third term: weight 3/3 = 1.0, i.e. (float)((double)ssd + (double)(d2 * d2)).  For two non-negative floats that equals the
    // float sum: the double sum is exact unless the exponents are > 29 bits apart, and then both roundings return ssd itself.
    ssd = __fadd_rn(ssd, d2 * d2);
    return ssd;
}

// All 32 lanes of the warp must call this (active = false for lanes without a segment).  w, h are warp-uniform.
__device__ __forceinline__ void warp_search(SearchSmem& sm, bool active, const float4* __restrict__ born_quad, int w, int h, float sx,
                                            float sy, float dirx, float diry, float length, float og, float& bestx, float& besty,
                                            float& min_ssd)
{
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    // 1. dry walk: step count + checkpoints.  The stop rule `norm(pt - start) < length` (fp64) cannot fire
    // during the first (int)length - 2 steps: |dir| <= 1 + 3 * 2^-24 and, for |start| < 2^14, the float
    // accumulation drifts by < 0.1 px per 102 steps — so those steps only accumulate pt (bit-identical walk).
    int n = 0;
    sm.cp[0][lane] = make_float2(sx, sy);
    if (active) {
        float ptx = sx, pty = sy;
        int nskip = 0;
        if (fabsf(sx) < 16384.f && fabsf(sy) < 16384.f && length < 16384.f) nskip = min((int)length - 2, 100);
        // whole checkpoint intervals first: 2^shift additions per coordinate back to back (n is a multiple of the interval here)
        for (; n + (1 << SEARCH_CP_SHIFT) <= nskip;) {
#pragma unroll
            for (int u = 0; u < (1 << SEARCH_CP_SHIFT); u++) {
                ptx += dirx;
                pty += diry;
            }
            n += 1 << SEARCH_CP_SHIFT;
            sm.cp[n >> SEARCH_CP_SHIFT][lane] = make_float2(ptx, pty);
        }
        for (; n < nskip;) {
            ptx += dirx;
            pty += diry;
            n++;
            if ((n & ((1 << SEARCH_CP_SHIFT) - 1)) == 0) sm.cp[n >> SEARCH_CP_SHIFT][lane] = make_float2(ptx, pty);
        }
        const NormLess shorter_than(length);
        while (shorter_than(ptx - sx, pty - sy)) { // :120
            ptx += dirx;
            pty += diry;
            n++;
            if ((n & ((1 << SEARCH_CP_SHIFT) - 1)) == 0) sm.cp[n >> SEARCH_CP_SHIFT][lane] = make_float2(ptx, pty);
            if (n > 101) break; // `if (count++ > 100) break` :141 -> at most 102 steps
        }
    }
    // 2. work unit = one CHUNK of 2^SEARCH_CP_SHIFT consecutive steps of one pixel (it starts exactly on a
    // checkpoint); inclusive scan of the chunk counts
    const int nch = (n + (1 << SEARCH_CP_SHIFT) - 1) >> SEARCH_CP_SHIFT;
    int incl = nch;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(full, incl, off);
        if (lane >= off) incl += v;
    }
    const int excl = incl - nch;
    const int total = __shfl_sync(full, incl, 31);
    sm.dirx[lane] = dirx; sm.diry[lane] = diry; sm.og[lane] = og; sm.born[lane] = born_quad; sm.nsteps[lane] = n;
    const unsigned long long key0 = (unsigned long long)__float_as_uint(2.0f * 3) << 32; // min_ssd = 6, best = start (step 0)
    sm.best[lane] = key0;
    __syncwarp();
#ifdef DVO_MAP_STATS
    if (lane == 0) { atomicAdd(&g_map_stats[0], (unsigned long long)__reduce_add_sync(full, n)); atomicAdd(&g_map_stats[1], (unsigned long long)((total + 31) / 32) * 32 << SEARCH_CP_SHIFT); }
#endif
    // 3./4. flattened chunks, 32 per round
    for (int base = 0; base < total; base += 32) {
        const int wi = base + lane;
        const bool on = wi < total;
        // which pixel owns chunk wi: pixel q's run starts at lane (excl_q - base); runs are contiguous and ordered
        const int p0 = __popc(__ballot_sync(full, incl <= base)); // the pixel that owns chunk `base`
        const int st = excl - base;
        const unsigned hb = (nch > 0 && st >= 0 && st < 32) ? (1u << st) : 0u;
        const unsigned heads = __reduce_or_sync(full, hb) | 1u;
        if (hb) sm.owner[st] = make_int2(lane, excl);
        else if (lane == p0) sm.owner[0] = make_int2(lane, excl);
        __syncwarp();
        const int hp = 31 - __clz(heads & (0xffffffffu >> (31 - lane))); // head lane of this lane's run
        if (on) {
            const int2 o = sm.owner[hp];
            const int p = o.x;
            const int ch = wi - o.y;                       // chunk of pixel p: steps ch*2^k + 1 ... (ch+1)*2^k
            const int b0 = ch << SEARCH_CP_SHIFT;
            const int cnt = min(1 << SEARCH_CP_SHIFT, sm.nsteps[p] - b0);
            const float2 cpt = sm.cp[ch][p];
            float ptx = cpt.x, pty = cpt.y;
            const float dx = sm.dirx[p], dy = sm.diry[p], ogp = sm.og[p];
            const float4* __restrict__ img = sm.born[p];
            // the reference's loop body (implement.cpp:121-140) for steps b0+1 .. b0+cnt.  It samples pt - dir, pt, pt + dir at
            // every step ("TODO: 1/N" in the reference).  pt + dir is, bit for bit, the next step's pt (pt += dir), so the centre
            // tap is the previous step's third tap; and pt - dir rounds back to the previous pt unless the forward addition hit a
            // rounding tie or crossed a binade (fl(fl(p + d) - d) == p when the rounding error of p + d is below half an ulp of p),
            // which is CHECKED per step: equal coordinates -> the tap of two steps ago is reused, otherwise it is sampled.  One new
            // tap per step instead of three, results bit-identical.
            float prevx = ptx, prevy = pty;                    // pt of the previous step (the checkpoint itself for k = 1)
            const float wf = (float)w, hf = (float)h;
            float c_prev = SEARCH_TAP(img, w, h, wf, hf, ptx, pty); // tap at prev
            ptx += dx;
            pty += dy;
            float c_cur = SEARCH_TAP(img, w, h, wf, hf, ptx + dx * 0.0f, pty + dy * 0.0f); // tap at pt (k = 1)
            unsigned long long kbest = ~0ull;
#pragma unroll(SEARCH_UNROLL)
            for (int k = 1; k <= cnt; k++) {
                const float s2 = SEARCH_TAP(img, w, h, wf, hf, ptx + dx * 1.0f, pty + dy * 1.0f);
                const float bx = ptx + dx * -1.0f, by = pty + dy * -1.0f;
                const float s0 = (bx == prevx && by == prevy) ? c_prev : SEARCH_TAP(img, w, h, wf, hf, bx, by);
                const unsigned long long key = ((unsigned long long)__float_as_uint(search_ssd(s0, c_cur, s2, ogp)) << 32) | (unsigned)(b0 + k);
                if (key < kbest) kbest = key;
                c_prev = c_cur; c_cur = s2;
                prevx = ptx; prevy = pty;
                ptx += dx;
                pty += dy;
            }
            // lanes [hp, rend) form this lane's run: smallest ssd of the run, at its earliest step
            const unsigned above = (hp == 31) ? 0u : (heads >> (hp + 1)) << (hp + 1);
            const int end = above ? __ffs(above) - 1 : 32;
            const int rend = min(end, min(total - base, 32));
            const unsigned seg = ((rend == 32) ? 0xffffffffu : ((1u << rend) - 1u)) & ~((1u << hp) - 1u);
            const unsigned hi = (unsigned)(kbest >> 32), lo = (unsigned)kbest;
            const unsigned m = __reduce_min_sync(seg, hi);
            const unsigned bm = __reduce_min_sync(seg, hi == m ? lo : 0xffffffffu);
            if (lane == hp) {
                const unsigned long long key = ((unsigned long long)m << 32) | bm;
                if (key < sm.best[p]) sm.best[p] = key;
            }
        }
        __syncwarp();
    }
    // 5. this lane's result
    const unsigned long long kb = sm.best[lane];
    min_ssd = __uint_as_float((unsigned)(kb >> 32));
    const int bb = (int)(unsigned)kb;
    const float2 c = sm.cp[bb >> SEARCH_CP_SHIFT][lane];
    bestx = c.x; besty = c.y;
    for (int r = bb & ((1 << SEARCH_CP_SHIFT) - 1); r > 0; r--) { bestx += dirx; besty += diry; }
    __syncwarp(); // the shared arrays are reused by the warp's next call
}

// Epipolar segment of Implement::update (EpipolarSegment, implement.cpp:23-47).
struct Segment {
    float dmin, dmax, sx, sy, ex, ey, length;
};
__device__ __forceinline__ Segment make_segment(const LevelGeom& g, const float* __restrict__ T, int px, int py, float depth, float sigma)
{
    Segment e;
    e.dmin = (depth - sigma) < 0.10f ? 0.10f : (depth - sigma); // :30
    e.dmax = depth + sigma;
    warp_point(T, g, (float)px, (float)py, e.dmax, e.sx, e.sy); // start :31
    warp_point(T, g, (float)px, (float)py, e.dmin, e.ex, e.ey); // end   :32
    e.length = (float)point_norm(e.sx - e.ex, e.sy - e.ey);     // :33
    return e;
}

// Implement::update after doMatching: bounds check (:196-200), depthEstimate (:49-71), sigmaEstimate (:73-104).
__device__ inline void implement_finish(const float* __restrict__ born_gray, const LevelGeom& g, const float* __restrict__ T,
                                        const float* __restrict__ tneg, int px, int py, const Segment& e, float bestx, float besty,
                                        float min_ssd, float& new_depth, float& new_sigma)
{
    const int w = g.w, h = g.h;
    const float dmin = e.dmin, dmax = e.dmax, sx = e.sx, sy = e.sy, ex = e.ex, ey = e.ey, length = e.length;
    float mx = bestx, my = besty;
    if ((double)min_ssd > 3 * 0.1) { mx = -1; my = -1; } // :145
    if (mx < 0 || my < 0 || mx > (float)w || my > (float)h) { // :196-200
        new_depth = -1; new_sigma = -1;
        return;
    }
    // depthEstimate :49-71
    {
        const float Kc[9] = {g.fx, 0.f, g.cx, 0.f, g.fy, g.cy, 0.f, 0.f, 1.f};
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
        // ref_grad_x(ref_x_i): the gradient image's element at that linear index
        LevelView tv; tv.w = w; tv.h = h; tv.t = 0; tv.W = w;
        const int yy = (int)(idx / w), xx = (int)(idx - (long long)yy * w);
        const float gx = grad_x_at(born_gray, tv, xx, yy), gy = grad_y_at(born_gray, tv, xx, yy);
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

// exp(rel_xi) for Mapper::update's projection (mapper.cpp:94) and Implement::propagate (:239) — granular API
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

// The ROI of Mapper::update (mapper.cpp:90) clipped to the image; threads are laid out over it in
// row-major order, which is also the reference's visiting order restricted to the ROI.
#ifndef DVO_MAPU_THREADS
#define DVO_MAPU_THREADS 64
#endif
constexpr int MAPU_THREADS = DVO_MAPU_THREADS; // small blocks: a block waits for its slowest epipolar search
struct MapRoi {
    int x0, y0, w, h, n; // origin, size, pixel count
    int nblk;            // blocks of MAPU_THREADS covering it
};

// Mapper::update phase A (mapper.cpp:76-137): per ref pixel, epipolar observation + gate + fusion.
// Rejected pixels (Gaussian::update reset, gaussian.cpp:21-25) are only flagged here: their value
// depends on the row-major rank among all rejects (mt19937 order) -> phase B.
// grid = (roi.nblk, S)
#ifndef DVO_MAPU_MINB
#define DVO_MAPU_MINB 16
#endif
__global__ void __launch_bounds__(MAPU_THREADS, DVO_MAPU_MINB) k_map_update(const __grid_constant__ Consts c, const MapRoi roi, const SlotDev* __restrict__ slots,
                                                       MapState* __restrict__ mst, const int* __restrict__ hist,
                                                       const PoseTab* __restrict__ tab, uint8_t* __restrict__ flags,
                                                       unsigned int* __restrict__ blockcnt, ReadyQ* rq, int* status, int fold_scan)
{
    __shared__ SearchSmem s_search[MAPU_THREADS / 32];
    __shared__ int s_sel[2];
    // Which (stream, tile) this CTA serves.  Ordinary launch: (blockIdx.y, blockIdx.x) of streams that update.  Overlapped with
    // the tracking kernel (rq != nullptr, programmatic dependent launch): the CTA takes a ticket and serves tile (ticket % nblk)
    // of the (ticket / nblk)-th stream to FINISH tracking with an update to do, waiting for that list entry if tracking has not
    // got that far - tickets are handed out in start order and streams are listed in completion order, so a CTA only ever waits
    // when every tile of every finished stream is already taken, i.e. when its SM slot would have idled anyway.
    int s = blockIdx.y, tile = blockIdx.x;
    if (rq) {
        if (threadIdx.x == 0) {
            const unsigned int t = atomicAdd(&rq->ticket, 1u);
            const int q = (int)(t / (unsigned int)roi.nblk);
            int v;
            unsigned long long t0 = 0;
            for (;;) {
                v = ld_volatile_i32(&rq->list[q]);
                if (v >= 0) break;
                // The producers (k_track's CTAs) are all resident before this kernel may start, so the wait is bounded by one
                // tracking run (< 1 ms).  Should that ever not hold (a tool serialising the two launches the wrong way round),
                // give up after 2 s with an error instead of hanging the GPU.
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > 2000000000ull) { if (status) atomicCAS(status, 0, (int)DVO_ERR_STATE); v = -1; break; }
                if (ld_volatile_i32(&rq->n_finished) >= c.S) { // every stream has decided: the list is final
                    __threadfence();
                    v = ld_volatile_i32(&rq->list[q]);
                    break;
                }
                __nanosleep(200);
            }
            __threadfence(); // acquire: the stream's MapState / history / pose table were written before its list entry
            s_sel[0] = v; s_sel[1] = (int)(t - (unsigned int)q * (unsigned int)roi.nblk);
        }
        __syncthreads();
        s = s_sel[0]; tile = s_sel[1];
        if (s < 0) return;
    }
    MapState* m = mst + s;
    if (!rq && !m->do_update) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int j = tile * blockDim.x + threadIdx.x;
    uint8_t fl = 0;
    bool reject = false;
    // --- per pixel: candidate gate, keyframe lookup, epipolar segment (mapper.cpp:90-107, implement.cpp:23-47)
    bool cand = false;
    int i = 0, px = 0, py = 0;
    float prior_d = 0.f, prior_s = 0.f, og = 0.f;
    const float4* born_quad = nullptr;
    int born_slot = 0;
    const PoseTab* pt = nullptr;
    Segment seg{};
    float2* ds = nullptr;
    if (j < roi.n) {
        const int ry = j / roi.w;
        const int x = roi.x0 + (j - ry * roi.w), y = roi.y0 + ry;
        i = y * g.w + x;
        const int n_hist = __ldcg(&m->n_hist);
        const SlotDev ref = slots[__ldcg(&m->ref_slot)];
        ds = ref.ds + (size_t)s * g.n;
        const float2 dsv = ds[i];
        const float d = dsv.x;
        float wx, wy;
        {
            float Trel[12];
#pragma unroll
            for (int q = 0; q < 12; q++) Trel[q] = __ldcg(&m->Trel[q]);
            warp_point(Trel, g, (float)x, (float)y, d, wx, wy);
        }
        px = round_x86(wx); py = round_x86(wy); // Point2f -> Point2i :94
        if (in_range(px, py, g.w, g.h)) {
            const int age = trunc_x86(ref.age[(size_t)s * g.n + i]); // :99
            if (age < 0 || age >= n_hist) {
                atomicCAS(&m->err, 0, (int)DVO_ERR_HISTORY); // m_history.at() would throw
            } else {
                born_slot = __ldcg(hist + (size_t)s * c.kf_stride + (n_hist - 1 - age));
                pt = tab + (size_t)s * c.kf_stride + age;
                prior_d = d - __ldcg(&m->rel_z); // :104
                prior_s = dsv.y;
                og = __ldg(slots[__ldcg(&m->obj_slot)].gray + (size_t)s * g.n + (size_t)py * g.w + px);
                born_quad = slots[born_slot].quad + (size_t)s * g.n;
                seg = make_segment(g, pt->T, px, py, prior_d, prior_s);
                cand = true;
            }
        }
    }
    // --- epipolar search, shared by the warp (implement.cpp:106-152)
    float bestx, besty, min_ssd;
    warp_search(s_search[threadIdx.x >> 5], cand, born_quad, g.w, g.h, seg.sx, seg.sy, (seg.ex - seg.sx) / seg.length,
                (seg.ey - seg.sy) / seg.length, seg.length, og, bestx, besty, min_ssd);
    // --- per pixel: depth / sigma estimate, gate, fusion (implement.cpp:196-212, mapper.cpp:122-131)
    if (cand) {
        float nd, ns;
        implement_finish(slots[born_slot].gray + (size_t)s * g.n, g, pt->T, pt->t, px, py, seg, bestx, besty, min_ssd, nd, ns);
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
    if (j < roi.n) flags[(size_t)s * g.n + i] = fl;
    const int nrej = __syncthreads_count(reject);
    const int nval = __syncthreads_count((fl & 4) != 0);
    // The stream's LAST phase-A CTA turns the per-CTA reject counts into exclusive offsets (the row-major rank base of phase B)
    // and their total - no separate scan launch, and nothing quadratic in the CTA count (129 600 CTAs at 3840x2160).
    unsigned int* bc = blockcnt + (size_t)s * roi.nblk;
    __shared__ unsigned int s_last;
    if (!fold_scan) { // large grids (1080p, 4K single streams): k_map_scan does it with a thousand threads
        if (threadIdx.x == 0) {
            bc[tile] = (unsigned)nrej;
            if (nval) atomicAdd(&m->valid_update, nval);
        }
        return;
    }
    if (threadIdx.x == 0) {
        bc[tile] = (unsigned)nrej;
        if (nval) atomicAdd(&m->valid_update, nval);
        __threadfence();
        s_last = (atomicAdd(&m->ticket_a, 1u) == (unsigned int)(roi.nblk - 1)) ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    unsigned int* s_sum = reinterpret_cast<unsigned int*>(s_search); // the search scratch is free now
    const int per = (roi.nblk + MAPU_THREADS - 1) / MAPU_THREADS;
    const int b0 = threadIdx.x * per, b1 = min(roi.nblk, b0 + per);
    unsigned int sum = 0;
    for (int b = b0; b < b1; b++) sum += __ldcg(bc + b);
    s_sum[threadIdx.x] = sum;
    __syncthreads();
    unsigned int off = 0;
    for (int k = 0; k < (int)threadIdx.x; k++) off += s_sum[k];
    for (int b = b0; b < b1; b++) {
        const unsigned int v = __ldcg(bc + b);
        bc[b] = off;
        off += v;
    }
    if (threadIdx.x == MAPU_THREADS - 1) { m->n_reject = off; m->ticket_a = 0; }
}

// The scan of the per-CTA reject counts as a kernel of its own, for grids too large for one 64-thread CTA to scan in passing
// (27 000 CTAs at 1920x1080, 129 600 at 3840x2160).  grid = S, 1024 threads.
__global__ void __launch_bounds__(1024) k_map_scan(const __grid_constant__ Consts c, const MapRoi roi, MapState* __restrict__ mst,
                                                   unsigned int* __restrict__ blockcnt)
{
    const int s = blockIdx.x;
    MapState* m = mst + s;
    if (!m->do_update) return;
    unsigned int* bc = blockcnt + (size_t)s * roi.nblk;
    const int per = (roi.nblk + 1023) / 1024;
    const int b0 = threadIdx.x * per, b1 = min(roi.nblk, b0 + per);
    unsigned int sum = 0;
#pragma unroll 8
    for (int b = b0; b < b1; b++) sum += bc[b];
    // exclusive scan of the 1024 partial sums: warp scans + a scan of the 32 warp totals
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    __shared__ unsigned int s_warp[32];
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        unsigned int w = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned int v = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += v; }
        s_warp[lane] = w; // inclusive
    }
    __syncthreads();
    unsigned int off = (incl - sum) + (warp ? s_warp[warp - 1] : 0u);
    for (int b = b0; b < b1; b++) {
        const unsigned int v = bc[b];
        bc[b] = off;
        off += v;
    }
    if (threadIdx.x == 1023) m->n_reject = s_warp[31];
}

// Mapper::update phase B: replay of the mt19937 stream for the rejected pixels in row-major order.
// rank = (rejects in the preceding phase-A CTAs, scanned by the stream's last phase-A CTA) + (rejects before this thread in its phase-A CTA).
// One CTA covers APPLY_SUB phase-A CTAs.   grid = (ceil(roi.nblk / APPLY_SUB), S)
constexpr int APPLY_SUB = 256 / MAPU_THREADS;
__global__ void __launch_bounds__(MAPU_THREADS * APPLY_SUB) k_map_update_apply(const __grid_constant__ Consts c, const MapRoi roi,
                                                          const SlotDev* __restrict__ slots, MapState* __restrict__ mst,
                                                          const uint8_t* __restrict__ flags, const unsigned int* __restrict__ blockcnt,
                                                          const uint32_t* __restrict__ rng_raw, int* __restrict__ status)
{
    const int s = blockIdx.y;
    MapState* m = mst + s;
    if (!m->do_update) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    int i = -1;
    if (j < roi.n) {
        const int ry = j / roi.w;
        i = (roi.y0 + ry) * g.w + roi.x0 + (j - ry * roi.w);
    }
    constexpr int WPS = MAPU_THREADS / 32; // warps per phase-A CTA
    __shared__ unsigned int s_warp[WPS * APPLY_SUB];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool rej = (i >= 0) && (flags[(size_t)s * g.n + i] & 8);
    const unsigned int bal = __ballot_sync(0xffffffffu, rej);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (rej) {
        const int sub = blockIdx.x * APPLY_SUB + warp / WPS; // the phase-A CTA this pixel belonged to
        unsigned int rank = blockcnt[(size_t)s * roi.nblk + sub] + __popc(bal & ((1u << lane) - 1u));
        for (int k = (warp / WPS) * WPS; k < warp; k++) rank += s_warp[k];
        const unsigned long long pos = m->rng_pos + rank;
        if (pos >= c.rng_table) {
            atomicCAS(&m->err, 0, (int)DVO_ERR_RNG);
        } else {
            const float u = dist_from_u32(__ldg(rng_raw + pos));
            const SlotDev ref = slots[m->ref_slot];
            ref.ds[(size_t)s * g.n + i] = make_float2(4.0f < u ? 4.0f : u, 0.5f); // gaussian.cpp:22-23
            ref.age[(size_t)s * g.n + i] = 0.f;                                   // mapper.cpp:126
        }
    }
    // the last block of the stream advances the replay position
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        unsigned int t = atomicAdd(&m->ticket, 1u);
        if (t == gridDim.x - 1) {
            m->rng_pos += m->n_reject;
            m->ticket = 0;
            // the host sizes the replay table by the furthest position of any stream (read back asynchronously)
            if (status) atomicMax(status + 2, (int)(m->rng_pos < 0x7fffffffull ? m->rng_pos : 0x7fffffffull));
        }
    }
}

// Implement::propagate for one stream by one CTA (declared above; called from the tracking kernel's epilogue).  winner: the
// stream's [n_top] scratch, all -1 on entry and on exit.
__device__ __forceinline__ void propagate_stream(const Consts& c, const SlotDev* __restrict__ slots, const MapState* m, int* __restrict__ winner,
                                                 int s, int tid, int nthreads)
{
    const LevelGeom g = c.lv[c.L - 1];
    const SlotDev ref = slots[m->ref_slot];
    const SlotDev dst = slots[m->new_slot];
    const float2* __restrict__ rds = ref.ds + (size_t)s * g.n;
    const float* __restrict__ rage = ref.age + (size_t)s * g.n;
    float T[12];
#pragma unroll
    for (int q = 0; q < 12; q++) T[q] = m->Trel[q];
    const float rel_z = m->rel_z;
    // One CTA walks the whole level, so the loop is latency-bound: four independent pixels per thread and round, loads first.
    constexpr int U = 4;
    for (int i0 = tid; i0 < g.n; i0 += U * nthreads) {
        float rd[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + u * nthreads;
            rd[u] = i < g.n ? rds[i].x : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + u * nthreads;
            if (i >= g.n || fabsf(rd[u]) < DVO_EPS) continue; // :236
            const int y = i / g.w, x = i - y * g.w;
            float wx, wy;
            warp_point(T, g, (float)x, (float)y, rd[u], wx, wy); // :239
            const int px = round_x86(wx), py = round_x86(wy);
            if (!in_range(px, py, g.w, g.h)) continue; // :240
            atomicMax(winner + (size_t)py * g.w + px, i);
        }
    }
    __syncthreads(); // every scatter of this stream is at L2 before the first gather reads it (ld.cg below)
    for (int i0 = tid; i0 < g.n; i0 += U * nthreads) {
        int src[U];
        float2 v[U];
        float ag[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + u * nthreads;
            src[u] = i < g.n ? __ldcg(winner + i) : -1;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            v[u] = make_float2(0.f, 0.f); ag[u] = 0.f;
            if (src[u] >= 0) { v[u] = rds[src[u]]; ag[u] = rage[src[u]]; }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + u * nthreads;
            if (i >= g.n) continue;
            float2 out = make_float2(1.0f, 1.0f); // :229-231
            float age = 0.f;
            if (src[u] >= 0) {
                winner[i] = -1;
                const float d0 = v[u].x < 0.01f ? 0.01f : v[u].x; // :244
                const float d1 = d0 + rel_z;
                const float q = d1 / d0;
                const float q4 = q * (q * (q * q)); // math::pow(., 4)
                const float sg = sqrtf(q4 * (v[u].y * v[u].y) + (0.06f * 0.06f)); // :247
                out = make_float2(d1 < 0.0f ? 0.0f : d1, sg);
                age = ag[u] + 1; // :252
            }
            dst.ds[(size_t)s * g.n + i] = out;
            dst.age[(size_t)s * g.n + i] = age;
        }
    }
}

// Implement::propagate (implement.cpp:217-256).  Pass 1: the serial reference lets the LAST source
// pixel in row-major order win a destination; atomicMax of the source index reproduces that
// without a race.  `winner` is all -1 between calls (pass 2 restores it).
__global__ void __launch_bounds__(256) k_propagate_scatter(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                           const MapState* __restrict__ mst, int* __restrict__ winner)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_propagate) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const float rd = slots[m->ref_slot].ds[(size_t)s * g.n + i].x;
    if (fabsf(rd) < DVO_EPS) return; // :236
    float wx, wy;
    warp_point(m->Trel, g, (float)x, (float)y, rd, wx, wy); // :239
    const int px = round_x86(wx), py = round_x86(wy);
    if (!in_range(px, py, g.w, g.h)) return; // :240
    atomicMax(winner + (size_t)s * g.n + (size_t)py * g.w + px, i);
}

// Pass 2: gather from the winning source (or the 1/1/0 defaults, :229-231) into new_slot
// (= Frame::updateDepthSigmaAge, frame.cpp:47-54: the coarser levels are views of this one).
__global__ void __launch_bounds__(256) k_propagate_gather(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                          const MapState* __restrict__ mst, int* __restrict__ winner)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_propagate) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const SlotDev ref = slots[m->ref_slot];
    const SlotDev dst = slots[m->new_slot];
    const int src = winner[(size_t)s * g.n + i];
    float2 out = make_float2(1.0f, 1.0f);
    float age = 0.f;
    if (src >= 0) {
        winner[(size_t)s * g.n + i] = -1;
        const float2 v = ref.ds[(size_t)s * g.n + src];
        const float d0 = v.x < 0.01f ? 0.01f : v.x; // :244
        const float d1 = d0 + m->rel_z;
        const float q = d1 / d0;
        const float q4 = q * (q * (q * q)); // math::pow(., 4)
        const float sg = sqrtf(q4 * (v.y * v.y) + (0.06f * 0.06f)); // :247
        out = make_float2(d1 < 0.0f ? 0.0f : d1, sg);
        age = ref.age[(size_t)s * g.n + src] + 1; // :252
    }
    dst.ds[(size_t)s * g.n + i] = out;
    dst.age[(size_t)s * g.n + i] = age;
}

// Implement::regularize (implement.cpp:156-180): 4-neighbour Gaussian fusion read from the input
// maps; result -> tmp (depth only).  Slot = mst[s].reg_slot.
// build_quad != 0 (VO loop): a stream that has just taken a NEW keyframe (do_propagate) also gets that keyframe's 2x2-footprint
// image written here (SlotDev::quad) - the kernel visits every pixel of the new keyframe anyway, and the epipolar searches of
// later frames sample born keyframes through it.
__global__ void __launch_bounds__(256) k_regularize(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                    const MapState* __restrict__ mst, float* __restrict__ tmp, int build_quad)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_regularize) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const SlotDev slot = slots[m->reg_slot];
    const float2* ds = slot.ds + (size_t)s * g.n;
    float2 v = ds[i];
    float gd = v.x, gs = v.y;
    if (x - 1 >= 0) { const float2 n = __ldg(ds + i - 1); gaussian_fuse(gd, gs, n.x, n.y); }      // {0,-1}
    if (x + 1 < g.w) { const float2 n = __ldg(ds + i + 1); gaussian_fuse(gd, gs, n.x, n.y); }     // {0, 1}
    if (y + 1 < g.h) { const float2 n = __ldg(ds + i + g.w); gaussian_fuse(gd, gs, n.x, n.y); }   // {1, 0}
    if (y - 1 >= 0) { const float2 n = __ldg(ds + i - g.w); gaussian_fuse(gd, gs, n.x, n.y); }    // {-1,0}
    tmp[(size_t)s * g.n + i] = gd < 6.0f ? gd : 6.0f; // cv::min(new_depth, 6.0) :178
    if (build_quad && m->do_propagate) slot.quad[(size_t)s * g.n + i] = make_quad(slot.gray + (size_t)s * g.n, g.w, g.h, x, y);
}

// SlotDev::quad of one slot (fixed >= 0) or of every stream's current KEYFRAME (fixed < 0: after frame 0 of a VO run, whose
// epilogue has already made that frame the first keyframe).
// grid = (ceil(n_top/256), S)
__global__ void __launch_bounds__(256) k_build_quad(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots, int fixed,
                                                    const VoState* __restrict__ vo)
{
    const int s = blockIdx.y;
    const SlotDev slot = slots[fixed >= 0 ? fixed : vo[s].kf_slot];
    const LevelGeom g = c.lv[c.L - 1];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    slot.quad[(size_t)s * g.n + i] = make_quad(slot.gray + (size_t)s * g.n, g.w, g.h, x, y);
}

// Frame::updateDepth (frame.cpp:56-61): depth <- tmp; sigma stays.  Two pixels per thread so that
// whole float4 {d, s, d, s} are rewritten.  grid = (ceil(n_top/512), S)
__global__ void __launch_bounds__(256) k_regularize_commit(const __grid_constant__ Consts c, const SlotDev* __restrict__ slots,
                                                           const MapState* __restrict__ mst, const float* __restrict__ tmp)
{
    const int s = blockIdx.y;
    const MapState* m = mst + s;
    if (!m->do_regularize) return;
    const LevelGeom g = c.lv[c.L - 1];
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i >= g.n) return;
    float2* ds = slots[m->reg_slot].ds + (size_t)s * g.n;
    const float* t = tmp + (size_t)s * g.n;
    const bool vec = (i + 1 < g.n) && ((((size_t)s * g.n + i) & 1) == 0);
    if (vec) {
        float4 v = *reinterpret_cast<float4*>(ds + i);
        const float2 d = *reinterpret_cast<const float2*>(t + i);
        v.x = d.x; v.z = d.y;
        *reinterpret_cast<float4*>(ds + i) = v;
    } else {
        ds[i].x = t[i];
        if (i + 1 < g.n) ds[i + 1].x = t[i + 1];
    }
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

// math::se3::{exp, log, concatenate} (src/math/se3.cpp:76-131) for n operands.  op 0: exp(a) -> out[16];
// 1: log(a[16]) -> out[6]; 2: concatenate(a, b) -> out[6].
__global__ void k_se3_ops(int op, const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (op == 0) {
        float xi[6], T[16];
        for (int q = 0; q < 6; q++) xi[q] = a[6 * i + q];
        se3_exp(xi, T);
        for (int q = 0; q < 16; q++) out[16 * i + q] = T[q];
    } else if (op == 1) {
        float T[16], xi[6];
        for (int q = 0; q < 16; q++) T[q] = a[16 * i + q];
        se3_log(T, xi);
        for (int q = 0; q < 6; q++) out[6 * i + q] = xi[q];
    } else {
        float x[6], y[6], r[6];
        for (int q = 0; q < 6; q++) { x[q] = a[6 * i + q]; y[q] = b[6 * i + q]; }
        se3_concatenate(x, y, r);
        for (int q = 0; q < 6; q++) out[6 * i + q] = r[q];
    }
}

// Map::Implement::update (implement.cpp:182-214) for a list of points.  grid = (ceil(n/64), S)
__global__ void __launch_bounds__(64) k_update_points(const __grid_constant__ Consts c, SlotDev obj, SlotDev born, const float* __restrict__ rxi,
                                                      int n, const int* __restrict__ px, const int* __restrict__ py,
                                                      const float* __restrict__ depth, const float* __restrict__ sigma,
                                                      float* __restrict__ out_d, float* __restrict__ out_s)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[c.L - 1];
    __shared__ float T[12];
    __shared__ float tneg[3];
    if (threadIdx.x == 0) {
        se3_exp_neg12(rxi + 6 * s, T); // exp(-xi): implement.cpp:56,193
        for (int q = 0; q < 3; q++) tneg[q] = -rxi[6 * s + q];
    }
    __syncthreads();
    __shared__ SearchSmem s_search[2];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const size_t k = (size_t)s * n + (i < n ? i : 0);
    const int x = px[k], y = py[k];
    const bool cand = i < n && in_range(x, y, g.w, g.h);
    const float* born_gray = born.gray + (size_t)s * g.n;
    float og = 0.f;
    Segment seg{};
    if (cand) {
        og = __ldg(obj.gray + (size_t)s * g.n + (size_t)y * g.w + x);
        seg = make_segment(g, T, x, y, depth[k], sigma[k]);
    }
    float bestx, besty, min_ssd;
    warp_search(s_search[threadIdx.x >> 5], cand, born.quad + (size_t)s * g.n, g.w, g.h, seg.sx, seg.sy, (seg.ex - seg.sx) / seg.length,
                (seg.ey - seg.sy) / seg.length, seg.length, og, bestx, besty, min_ssd);
    float nd = -1.f, ns = -1.f;
    if (cand) implement_finish(born_gray, g, T, tneg, x, y, seg, bestx, besty, min_ssd, nd, ns);
    if (i >= n) return;
    out_d[k] = nd; out_s[k] = ns;
}

// =============================================================================================
// Row-major mask compaction, batched over streams: per-CTA counts -> one scan CTA per stream (exclusive offsets + total) ->
// scatter.  Linear in the CTA count (the first version summed all preceding counts in every CTA).  mask: [S][n] bytes,
// idx: [S][n], blockcnt: [S][nblk], total: [S].
// =============================================================================================
__global__ void __launch_bounds__(256) k_compact_count(const uint8_t* __restrict__ mask, int n, unsigned int* __restrict__ blockcnt)
{
    const int s = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool on = (i < n) && mask[(size_t)s * n + i];
    const int cnt = __syncthreads_count(on);
    if (threadIdx.x == 0) blockcnt[(size_t)s * gridDim.x + blockIdx.x] = cnt;
}
__global__ void __launch_bounds__(256) k_compact_scan(int nblk, unsigned int* __restrict__ blockcnt, int* __restrict__ total)
{
    const int s = blockIdx.x;
    __shared__ unsigned int s_sum[256];
    unsigned int* bc = blockcnt + (size_t)s * nblk;
    const int per = (nblk + 255) / 256;
    const int b0 = threadIdx.x * per, b1 = min(nblk, b0 + per);
    unsigned int sum = 0;
    for (int b = b0; b < b1; b++) sum += bc[b];
    s_sum[threadIdx.x] = sum;
    __syncthreads();
    unsigned int off = 0;
    for (int k = 0; k < (int)threadIdx.x; k++) off += s_sum[k];
    for (int b = b0; b < b1; b++) {
        const unsigned int v = bc[b];
        bc[b] = off;
        off += v;
    }
    if (threadIdx.x == 255) total[s] = (int)off;
}
__global__ void __launch_bounds__(256) k_compact_scatter(const uint8_t* __restrict__ mask, int n, const unsigned int* __restrict__ blockcnt,
                                                         int32_t* __restrict__ idx)
{
    const int s = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    __shared__ unsigned int s_warp[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool on = (i < n) && mask[(size_t)s * n + i];
    const unsigned int bal = __ballot_sync(0xffffffffu, on);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (on) {
        unsigned int rank = blockcnt[(size_t)s * gridDim.x + blockIdx.x] + __popc(bal & ((1u << lane) - 1u));
        for (int k = 0; k < warp; k++) rank += s_warp[k];
        idx[(size_t)s * n + rank] = i;
    }
}

// Semi-dense pixel selection by gradient threshold (north_star; no reference counterpart): mask of the level pixels whose gray and
// both Convert::gradiate values are valid and gx^2 + gy^2 >= thr2.  grid = (ceil(n_l/256), S)
__global__ void __launch_bounds__(256) k_select_mask(const __grid_constant__ Consts c, SlotDev slot, int l, float thr2, uint8_t* __restrict__ mask)
{
    const int s = blockIdx.y;
    const LevelGeom g = c.lv[l];
    const LevelView v = level_view(c, l);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.n) return;
    const int y = i / g.w, x = i - y * g.w;
    const size_t nt = (size_t)c.lv[c.L - 1].n;
    const float* gray = slot.gray + (size_t)s * nt;
    const float I = level_value(v, gray[top_index(v, x, y)]);
    const float gx = grad_x_at(gray, v, x, y), gy = grad_y_at(gray, v, x, y);
    mask[(size_t)s * g.n + i] = (!is_invalid(I) && !is_invalid(gx) && !is_invalid(gy) && !(gx * gx + gy * gy < thr2)) ? 1 : 0;
}

} // namespace dvo
