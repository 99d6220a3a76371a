// dvo_synth.cuh — synthetic sequence renderer: an analytic
// textured box room ray-cast per pixel, with a smooth 6-DoF camera trajectory per stream.  Pure
// arithmetic, no graphics API.  It only MAKES INPUTS for benchmarks and stress tests (so that 256
// streams x N frames need not cross PCIe); it is not on the tracking/mapping path and has no
// reference counterpart (the reference reads PNG files).
#pragma once
#include "dvo_internal.cuh"

namespace dvo {

struct SynthParams {
    int w, h;
    float fx, fy, cx, cy;
    unsigned long long seed0; // stream s renders the scene / trajectory of seed0 + s * seed_stride
    unsigned long long seed_stride;
    float trans_step, rot_step; // per-frame motion scale [m], [rad]
    int frame;
};

constexpr int SY_PLANES = 5, SY_WAVES = 8;

__device__ __forceinline__ unsigned long long sy_mix(unsigned long long z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
// uniform in [0,1) from (seed, counter)
__device__ __forceinline__ float sy_u01(unsigned long long seed, unsigned int k)
{
    return (float)(sy_mix(seed * 0x100000001B3ull + k) >> 40) * (1.0f / 16777216.0f);
}

struct SynthScene {
    float fu[SY_PLANES][SY_WAVES], fv[SY_PLANES][SY_WAVES], ph[SY_PLANES][SY_WAVES], amp[SY_PLANES][SY_WAVES];
    float chk[SY_PLANES];
    float R[9], t[3]; // camera-to-world
};

// camera-to-world pose of frame i: low-frequency sinusoids in all six degrees of freedom
__device__ inline void sy_pose(unsigned long long seed, int i, float trans_step, float rot_step, float* R, float* t)
{
    float v[6];
    for (int k = 0; k < 6; k++) {
        const float f = 0.02f + 0.06f * sy_u01(seed, 1000 + k);
        const float p = 6.2831853f * sy_u01(seed, 1010 + k);
        const float a = (k < 3 ? trans_step : rot_step) / (6.2831853f * f);
        v[k] = a * (sinf(6.2831853f * f * (float)i + p) - sinf(p));
    }
    t[0] = v[0]; t[1] = v[1]; t[2] = v[2];
    const float th = sqrtf(v[3] * v[3] + v[4] * v[4] + v[5] * v[5]);
    if (th < 1e-9f) {
        for (int k = 0; k < 9; k++) R[k] = (k % 4 == 0) ? 1.f : 0.f;
        return;
    }
    const float kx = v[3] / th, ky = v[4] / th, kz = v[5] / th;
    const float c = cosf(th), s = sinf(th), c1 = 1.f - c;
    R[0] = c + c1 * kx * kx;      R[1] = c1 * kx * ky - s * kz; R[2] = c1 * kx * kz + s * ky;
    R[3] = c1 * kx * ky + s * kz; R[4] = c + c1 * ky * ky;      R[5] = c1 * ky * kz - s * kx;
    R[6] = c1 * kx * kz - s * ky; R[7] = c1 * ky * kz + s * kx; R[8] = c + c1 * kz * kz;
}

// grid = (ceil(w*h/256), S).  gray: [S][h][w]; depth (optional): z in the camera frame.
// gray8 (optional): the same image as 8-bit codes, round-half-even of 255 g - what a camera / cv::imread + cvtColor delivers
__global__ void __launch_bounds__(256) k_synth_render(const SynthParams P, float* __restrict__ gray, float* __restrict__ depth,
                                                      uint8_t* __restrict__ gray8)
{
    __shared__ SynthScene sc;
    const int s = blockIdx.y;
    const unsigned long long seed = P.seed0 + (unsigned long long)s * P.seed_stride;
    for (int k = threadIdx.x; k < SY_PLANES * SY_WAVES; k += blockDim.x) {
        const int p = k / SY_WAVES, wv = k % SY_WAVES;
        const float sgu = sy_u01(seed, 4 * k + 0) < 0.5f ? -1.f : 1.f, sgv = sy_u01(seed, 4 * k + 1) < 0.5f ? -1.f : 1.f;
        sc.fu[p][wv] = sgu * (1.5f + 7.5f * sy_u01(seed, 4 * k + 2));
        sc.fv[p][wv] = sgv * (1.5f + 7.5f * sy_u01(seed, 4 * k + 3));
        sc.ph[p][wv] = 6.2831853f * sy_u01(seed, 500 + k);
        sc.amp[p][wv] = 0.3f + 0.7f * sy_u01(seed, 600 + k);
    }
    if (threadIdx.x < SY_PLANES) sc.chk[threadIdx.x] = 0.8f + 1.2f * sy_u01(seed, 700 + threadIdx.x);
    if (threadIdx.x == 32) sy_pose(seed, P.frame, P.trans_step, P.rot_step, sc.R, sc.t);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.w * P.h) return;
    const int y = i / P.w, x = i - y * P.w;
    const float zb = 2.6f, xw = 1.9f, yw = 1.4f;
    const float dcx = ((float)x - P.cx) / P.fx, dcy = ((float)y - P.cy) / P.fy;
    const float dw[3] = {sc.R[0] * dcx + sc.R[1] * dcy + sc.R[2], sc.R[3] * dcx + sc.R[4] * dcy + sc.R[5],
                         sc.R[6] * dcx + sc.R[7] * dcy + sc.R[8]};
    float best = 1e30f;
    int bp = -1;
    float bu = 0.f, bv = 0.f;
    const int axis[SY_PLANES] = {2, 0, 0, 1, 1};
    const float val[SY_PLANES] = {zb, xw, -xw, yw, -yw};
#pragma unroll
    for (int p = 0; p < SY_PLANES; p++) {
        const int ax = axis[p];
        const float lam = (val[p] - sc.t[ax]) / dw[ax];
        if (!(lam > 1e-6f && lam < best)) continue;
        const float X = sc.t[0] + lam * dw[0], Y = sc.t[1] + lam * dw[1], Z = sc.t[2] + lam * dw[2];
        if (fabsf(X) > xw + 1e-4f || fabsf(Y) > yw + 1e-4f || Z > zb + 1e-4f) continue;
        best = lam; bp = p;
        if (ax == 2) { bu = X; bv = Y; } else if (ax == 0) { bu = Z; bv = Y; } else { bu = X; bv = Z; }
    }
    float g = 0.f, d = 0.f;
    if (bp >= 0) {
        float tsum = 0.f, asum = 0.f;
#pragma unroll
        for (int wv = 0; wv < SY_WAVES; wv++) {
            tsum += sc.amp[bp][wv] * sinf(sc.fu[bp][wv] * bu + sc.fv[bp][wv] * bv + sc.ph[bp][wv]);
            asum += sc.amp[bp][wv];
        }
        const float ck = sinf(sc.chk[bp] * 3.14159265f * bu) * sinf(sc.chk[bp] * 3.14159265f * bv);
        g = 0.5f + 0.33f * (tsum / asum) + (ck > 0.f ? 0.10f : -0.10f);
        g = fminf(fmaxf(g, 0.02f), 0.98f);
        d = best; // the camera-frame ray has unit z, so lambda == z
    }
    if (gray) gray[(size_t)s * P.w * P.h + i] = g;
    if (gray8) gray8[(size_t)s * P.w * P.h + i] = (uint8_t)__float2int_rn(fminf(fmaxf(g * 255.0f, 0.0f), 255.0f));
    if (depth) depth[(size_t)s * P.w * P.h + i] = d;
}

// ground truth of the renderer for tests and error reporting: out[0..11] = camera-to-world [R | t] rows of frame `frame`,
// out[12 ..] = the scene of `seed`: fu, fv, ph, amp as [SY_PLANES][SY_WAVES] each, then chk[SY_PLANES]
constexpr int SY_INFO_FLOATS = 12 + 4 * SY_PLANES * SY_WAVES + SY_PLANES;
__global__ void k_synth_info(unsigned long long seed, int frame, float trans_step, float rot_step, float* __restrict__ out)
{
    if (threadIdx.x == 0) {
        float R[9], t[3];
        sy_pose(seed, frame, trans_step, rot_step, R, t);
        for (int r = 0; r < 3; r++) {
            for (int q = 0; q < 3; q++) out[r * 4 + q] = R[r * 3 + q];
            out[r * 4 + 3] = t[r];
        }
    }
    const int NW = SY_PLANES * SY_WAVES;
    for (int k = threadIdx.x; k < NW; k += blockDim.x) {
        const float sgu = sy_u01(seed, 4 * k + 0) < 0.5f ? -1.f : 1.f, sgv = sy_u01(seed, 4 * k + 1) < 0.5f ? -1.f : 1.f;
        out[12 + k] = sgu * (1.5f + 7.5f * sy_u01(seed, 4 * k + 2));
        out[12 + NW + k] = sgv * (1.5f + 7.5f * sy_u01(seed, 4 * k + 3));
        out[12 + 2 * NW + k] = 6.2831853f * sy_u01(seed, 500 + k);
        out[12 + 3 * NW + k] = 0.3f + 0.7f * sy_u01(seed, 600 + k);
    }
    if (threadIdx.x < SY_PLANES) out[12 + 4 * NW + threadIdx.x] = 0.8f + 1.2f * sy_u01(seed, 700 + threadIdx.x);
}

} // namespace dvo
