// dvo_device.cuh — device-side restatement of the reference's scalar geometry / sampling / se3
// algebra.  Compiled with -fmad=false: every fp32 expression is evaluated in the reference's
// order with IEEE round-to-nearest and NO fused multiply-add, so validity masks and integer
// coordinates match the CPU oracle bit for bit.  FMA is used only where explicitly
// requested (fma()/__fmaf_rn) in reductions that have no reference counterpart.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define DVO_INV (-2.0f)
#define DVO_EPS (1e-6f)

namespace dvo {

struct LevelGeom {
    int w, h, n;
    float fx, fy, cx, cy;
};

__device__ __forceinline__ bool is_valid(float v) { return DVO_INV < v; }    // include/math/util.hpp:9
__device__ __forceinline__ bool is_invalid(float v) { return v <= DVO_INV; } // include/math/util.hpp:10

// (int)float as x86 cvttss2si does it: NaN / out-of-range -> INT_MIN (src/core/convert.cpp:82-83).
__device__ __forceinline__ int trunc_x86(float v)
{
    if (!(v > -2147483648.0f && v < 2147483648.0f)) return INT_MIN;
    return __float2int_rz(v);
}
// cv::saturate_cast<int>(float) = cvRound = cvtss2si: round-half-even, NaN / out-of-range -> INT_MIN
// (cv::Point2f -> cv::Point2i, src/map/mapper.cpp:94, implement.cpp:84,240,250).
__device__ __forceinline__ int round_x86(float v)
{
    if (!(v > -2147483648.0f && v < 2147483648.0f)) return INT_MIN;
    return __float2int_rn(v);
}
__device__ __forceinline__ bool in_range(int x, int y, int w, int h) // include/math/util.hpp:46-55
{
    return !(x < 0 || w <= x || y < 0 || h <= y);
}

// Transform::backProject, src/core/transform.cpp:25-28
__device__ __forceinline__ void back_project(const LevelGeom& g, float x, float y, float depth, float& X, float& Y, float& Z)
{
    X = depth * (x - g.cx) / g.fx;
    Y = depth * (y - g.cy) / g.fy;
    Z = depth;
}
// Transform::warp = project(transform(backProject)), src/core/transform.cpp:7-33.  T = 12 floats
// (rows of [R|t]); the reference re-exponentiates the twist per pixel, the value is the same.
__device__ __forceinline__ void warp_point(const float* __restrict__ T, const LevelGeom& g, float x, float y, float depth,
                                           float& wx, float& wy)
{
    float X, Y, Z;
    back_project(g, x, y, depth, X, Y, Z);
    float tx = (T[0] * X + T[1] * Y + T[2] * Z) + T[3];
    float ty = (T[4] * X + T[5] * Y + T[6] * Z) + T[7];
    float tz = (T[8] * X + T[9] * Y + T[10] * Z) + T[11];
    wx = tx * g.fx / tz + g.cx;
    wy = ty * g.fy / tz + g.cy;
}

__device__ __forceinline__ float bilerp(float g0, float g1, float g2, float g3, float h, float v)
{
    return (g0 * (1.f - h) + g1 * h) * (1.f - v) + (g2 * (1.f - h) + g3 * h) * v; // convert.cpp:103-104
}

// The invalid-tap repair automaton of Convert::getSubpixel (src/core/convert.cpp:155-173),
// reproduced literally.  Returns false when the function returns INVALID.
__device__ __forceinline__ bool repair_taps(float g[4])
{
    int valid = 0, id = 0;
    float last = -1.f;
    for (int guard = 0; guard < 32; guard++) {
        if (is_valid(g[id])) {
            valid++;
            last = g[id];
        } else if (last > 0) {
            g[id] = last;
            valid++;
        }
        if (valid == 4) return true;
        if (id == 3 && valid == 0) return false;
        id = (id + 1) & 3;
    }
    return true;
}

// Convert::getPixel (src/core/convert.cpp:107-125) applied by cullImage: values <= -2 become -2.
__device__ __forceinline__ float get_pixel_value(float g) { return is_valid(g) ? g : DVO_INV; }

// Convert::getSubpixelFromDense on a plain float image (src/core/convert.cpp:77-105)
__device__ __forceinline__ float subpixel_dense(const float* __restrict__ img, int w, int h, float px, float py)
{
    int x0 = trunc_x86(px), y0 = trunc_x86(py);
    if (!in_range(x0, y0, w, h)) return DVO_INV;
    int x1 = x0 + 1, y1 = y0 + 1;
    float hh = px - (float)x0, v = py - (float)y0;
    float g0 = __ldg(img + (size_t)y0 * w + x0);
    float g1 = g0, g2 = g0, g3 = g0;
    bool xin = x1 < w, yin = y1 < h;
    if (xin) g1 = __ldg(img + (size_t)y0 * w + x1);
    if (yin) g2 = __ldg(img + (size_t)y1 * w + x0);
    if (xin && yin) g3 = __ldg(img + (size_t)y1 * w + x1);
    return bilerp(g0, g1, g2, g3, hh, v);
}

// Convert::getSubpixelFromDense through the 2x2-footprint image (SlotDev::quad): the border rule is baked into the footprints, so
// a tap is one range test, one 16-byte load and the bilinear arithmetic.  Bit-identical to subpixel_dense.
__device__ __forceinline__ float subpixel_quad(const float4* __restrict__ quad, int w, float wf, float hf, float px, float py)
{
    if (!(px > -1.0f && px < wf && py > -1.0f && py < hf)) return DVO_INV;
    const int x0 = __float2int_rz(px), y0 = __float2int_rz(py);
    const float hh = px - (float)x0, v = py - (float)y0;
    const float4 q = __ldg(quad + (y0 * w + x0));
    return bilerp(q.x, q.y, q.z, q.w, hh, v);
}
__device__ __forceinline__ float4 make_quad(const float* __restrict__ img, int w, int h, int x, int y)
{
    const float* p = img + (size_t)y * w + x;
    const bool xin = x + 1 < w, yin = y + 1 < h;
    const float g0 = p[0];
    return make_float4(g0, xin ? p[1] : g0, yin ? p[w] : g0, (xin && yin) ? p[w + 1] : g0);
}

// The same, split into position, loads and arithmetic, so that a caller can put several samples' loads in
// flight before consuming any of them, or take the four texels from elsewhere when (x0, y0) coincide.
struct TapPos {
    int x0, y0;
    float hh, v;
    bool in; // (x0, y0) inside the image
};
struct Texels {
    float g0, g1, g2, g3;
};
__device__ __forceinline__ TapPos tap_pos(int w, int h, float px, float py)
{
    TapPos t;
    t.x0 = trunc_x86(px); t.y0 = trunc_x86(py);
    t.in = in_range(t.x0, t.y0, w, h);
    t.hh = px - (float)t.x0; t.v = py - (float)t.y0;
    return t;
}
__device__ __forceinline__ Texels tap_fetch(const float* __restrict__ img, int w, int h, const TapPos& t)
{
    Texels g;
    const float* p = img + (size_t)t.y0 * w + t.x0;
    const bool xin = t.x0 + 1 < w, yin = t.y0 + 1 < h;
    g.g0 = __ldg(p);
    g.g1 = xin ? __ldg(p + 1) : g.g0;
    g.g2 = yin ? __ldg(p + w) : g.g0;
    g.g3 = (xin && yin) ? __ldg(p + w + 1) : g.g0;
    return g;
}
__device__ __forceinline__ float tap_value(const TapPos& t, const Texels& g) { return t.in ? bilerp(g.g0, g.g1, g.g2, g.g3, t.hh, t.v) : DVO_INV; }

// ---------------------------------------------------------------------------------------------
// se3 / so3 — src/math/se3.cpp.  Transcendentals: the reference calls glibc (double for
// Rodrigues/acos, float overloads elsewhere); here every one is evaluated in double and rounded,
// which agrees with glibc's (almost always correctly rounded) results except in ~1e-4 of cases
// by one ulp — within the 1e-5 pose tolerance, masks are tested with identical T.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void mat3_mul(const float* a, const float* b, float* d)
{
    float t[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) t[i * 3 + j] = a[i * 3 + 0] * b[0 * 3 + j] + a[i * 3 + 1] * b[1 * 3 + j] + a[i * 3 + 2] * b[2 * 3 + j];
#pragma unroll
    for (int i = 0; i < 9; i++) d[i] = t[i];
}
__device__ __forceinline__ void mat3_vec(const float* a, const float* v, float* d)
{
    float t0 = a[0] * v[0] + a[1] * v[1] + a[2] * v[2];
    float t1 = a[3] * v[0] + a[4] * v[1] + a[5] * v[2];
    float t2 = a[6] * v[0] + a[7] * v[1] + a[8] * v[2];
    d[0] = t0; d[1] = t1; d[2] = t2;
}
__device__ __forceinline__ void hat(const float* w, float* M) // se3.cpp:8-15
{
    M[0] = 0;     M[1] = -w[2]; M[2] = w[1];
    M[3] = w[2];  M[4] = 0;     M[5] = -w[0];
    M[6] = -w[1]; M[7] = w[0];  M[8] = 0;
}
__device__ __forceinline__ float norm3(const float* v) // (float)cv::norm: double sum of squares
{
    double s = 0;
    s = __dadd_rn(s, __dmul_rn((double)v[0], (double)v[0]));
    s = __dadd_rn(s, __dmul_rn((double)v[1], (double)v[1]));
    s = __dadd_rn(s, __dmul_rn((double)v[2], (double)v[2]));
    return (float)sqrt(s);
}
__device__ __forceinline__ double norm_n(const float* v, int n)
{
    double s = 0;
    for (int i = 0; i < n; i++) s = __dadd_rn(s, __dmul_rn((double)v[i], (double)v[i]));
    return sqrt(s);
}
// cv::Rodrigues vector -> matrix (so3::exp, se3.cpp:21-28)
__device__ inline void rodrigues(const float* w, float* R)
{
    double rx = w[0], ry = w[1], rz = w[2];
    double theta = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(rx, rx), __dmul_rn(ry, ry)), __dmul_rn(rz, rz)));
    if (theta < 2.220446049250313e-16) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    double c, s;
    sincos(theta, &s, &c); // same reduction and kernels as sin() and cos(); one range reduction instead of two
    const double c1 = 1. - c, it = 1. / theta;
    rx = __dmul_rn(rx, it); ry = __dmul_rn(ry, it); rz = __dmul_rn(rz, it);
    double rrt[9] = {__dmul_rn(rx, rx), __dmul_rn(rx, ry), __dmul_rn(rx, rz), __dmul_rn(rx, ry), __dmul_rn(ry, ry),
                     __dmul_rn(ry, rz), __dmul_rn(rx, rz), __dmul_rn(ry, rz), __dmul_rn(rz, rz)};
    double r_x[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
#pragma unroll
    for (int k = 0; k < 9; k++) {
        double eye = (k == 0 || k == 4 || k == 8) ? 1.0 : 0.0;
        R[k] = (float)__dadd_rn(__dadd_rn(__dmul_rn(c, eye), __dmul_rn(c1, rrt[k])), __dmul_rn(s, r_x[k]));
    }
}
__device__ __forceinline__ float cosf_ref(float x) { return (float)cos((double)x); }
__device__ __forceinline__ float sinf_ref(float x) { return (float)sin((double)x); }

// se3::exp (se3.cpp:70-98); T: 16 floats row-major.  Tn12 (optional) receives exp(-xi) as 12 floats from the same
// transcendentals: cv::Rodrigues(-w) is bit for bit the transpose of cv::Rodrigues(w) (theta, c, s, r r^T are even in w,
// [r]x is odd and rounding is symmetric), hat(-w)^2 == hat(w)^2 exactly, and the a/b, c/d factors depend on |w| only; the
// V matrix and the translation of exp(-xi) are then evaluated with the negated operands in the reference's order.
__device__ inline void se3_exp(const float* xi, float* T, float* Tn12 = nullptr)
{
    const float* v = xi;
    const float* w = xi + 3;
    float w_hat[9];
    hat(w, w_hat);
    float wl = norm3(w);
    float R[9];
    rodrigues(w, R);
    float t[3], tn[3];
    if (wl > 1e-6f) {
        double sn, cs;
        sincos((double)wl, &sn, &cs);
        float a = 1.0f - (float)cs;
        float b = wl * wl;
        float c = wl - (float)sn;
        float d = wl * wl * wl;
        float w_hat2[9], V[9];
        mat3_mul(w_hat, w_hat, w_hat2);
        // `eye + w_hat*a/b + (w_hat*w_hat)*c/d` is a lazy cv::MatExpr in the reference: `/ s` folds
        // 1.0/s into a double alpha, the first sum is cv::scaleAdd, the scaled product is cv::gemm's
        // (float)(t*(float)alpha), the last sum cv::add (DESIGN.md "third-party arithmetic").
        const double alpha1 = __dmul_rn((double)a, __ddiv_rn(1.0, (double)b));
        const double alpha2 = __dmul_rn((double)c, __ddiv_rn(1.0, (double)d));
        const float alpha1f = (float)alpha1;
#pragma unroll
        for (int k = 0; k < 9; k++) {
            float eye = (k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f;
            float m1 = w_hat[k] * alpha1f + eye;
            float m2 = w_hat2[k] * (float)alpha2;
            V[k] = m1 + m2;
        }
        mat3_vec(V, v, t);
        if (Tn12) {
            float wn[3] = {-w[0], -w[1], -w[2]}, vn[3] = {-v[0], -v[1], -v[2]}, wn_hat[9];
            hat(wn, wn_hat);
#pragma unroll
            for (int k = 0; k < 9; k++) {
                float eye = (k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f;
                float m1 = wn_hat[k] * alpha1f + eye;
                float m2 = w_hat2[k] * (float)alpha2;
                V[k] = m1 + m2;
            }
            mat3_vec(V, vn, tn);
        }
    } else {
        t[0] = v[0]; t[1] = v[1]; t[2] = v[2];
        tn[0] = -v[0]; tn[1] = -v[1]; tn[2] = -v[2];
    }
    T[0] = R[0]; T[1] = R[1]; T[2] = R[2];  T[3] = t[0];
    T[4] = R[3]; T[5] = R[4]; T[6] = R[5];  T[7] = t[1];
    T[8] = R[6]; T[9] = R[7]; T[10] = R[8]; T[11] = t[2];
    T[12] = 0;   T[13] = 0;   T[14] = 0;    T[15] = 1;
    if (Tn12) {
        Tn12[0] = R[0]; Tn12[1] = R[3]; Tn12[2] = R[6];  Tn12[3] = tn[0];
        Tn12[4] = R[1]; Tn12[5] = R[4]; Tn12[6] = R[7];  Tn12[7] = tn[1];
        Tn12[8] = R[2]; Tn12[9] = R[5]; Tn12[10] = R[8]; Tn12[11] = tn[2];
    }
}
// so3::log (se3.cpp:31-43)
__device__ inline void so3_log(const float* R, float* w)
{
    w[0] = w[1] = w[2] = 0.0f;
    double tr = 0;
    tr = __dadd_rn(tr, (double)R[0]); tr = __dadd_rn(tr, (double)R[4]); tr = __dadd_rn(tr, (double)R[8]);
    float wl = (float)acos(__dmul_rn(__dadd_rn(tr, -1.0), 0.5));
    if (wl > 1e-6f) {
        float tmp[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
        float k = 1.0f / (2.0f * sinf_ref(wl));
        const float alpha = (float)__dmul_rn((double)k, (double)wl); // lazy MatExpr: k * tmp * wl -> tmp * (float)(k*wl)
#pragma unroll
        for (int i = 0; i < 3; i++) w[i] = tmp[i] * alpha;
    }
}
// se3::log (se3.cpp:101-124)
__device__ inline void se3_log(const float* T, float* xi)
{
    float R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    float t[3] = {T[3], T[7], T[11]};
    float w[3];
    so3_log(R, w);
    float w_hat[9];
    hat(w, w_hat);
    float wl = norm3(w);
    float V_inv[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (wl > 1e-6f) {
        double snh, csh;
        sincos((double)(wl * 0.5f), &snh, &csh);
        float k = 1.0f - (wl * (float)csh) / (2.0f * (float)snh);
        float b = wl * wl;
        float w_hat2[9];
        mat3_mul(w_hat, w_hat, w_hat2);
        const double alpha2 = __dmul_rn((double)k, __ddiv_rn(1.0, (double)b)); // lazy MatExpr, see se3_exp
#pragma unroll
        for (int i = 0; i < 9; i++) {
            float eye = (i == 0 || i == 4 || i == 8) ? 1.0f : 0.0f;
            float m1 = w_hat[i] * -0.5f + eye;
            float m2 = w_hat2[i] * (float)alpha2;
            V_inv[i] = m1 + m2;
        }
    }
    float v[3];
    mat3_vec(V_inv, t, v);
    xi[0] = v[0]; xi[1] = v[1]; xi[2] = v[2];
    xi[3] = w[0]; xi[4] = w[1]; xi[5] = w[2];
}
// se3::concatenate (se3.cpp:127-131)
__device__ inline void se3_concatenate(const float* a, const float* b, float* out)
{
    float Ta[16], Tb[16], Tc[16];
    se3_exp(a, Ta);
    se3_exp(b, Tb);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++)
            Tc[i * 4 + j] = Ta[i * 4 + 0] * Tb[0 * 4 + j] + Ta[i * 4 + 1] * Tb[1 * 4 + j] + Ta[i * 4 + 2] * Tb[2 * 4 + j]
                            + Ta[i * 4 + 3] * Tb[3 * 4 + j];
    se3_log(Tc, out);
}
// se3::concatenate when exp(a) is already at hand (Ta: 16 floats, as se3_exp(a) returns it)
__device__ inline void se3_concatenate_T(const float* Ta, const float* b, float* out)
{
    float Tb[16], Tc[16];
    se3_exp(b, Tb);
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++)
            Tc[i * 4 + j] = Ta[i * 4 + 0] * Tb[0 * 4 + j] + Ta[i * 4 + 1] * Tb[1 * 4 + j] + Ta[i * 4 + 2] * Tb[2 * 4 + j]
                            + Ta[i * 4 + 3] * Tb[3 * 4 + j];
    se3_log(Tc, out);
}
// exp(-xi) as 12 floats (what Transform::warp(-xi, ...) uses)
__device__ inline void se3_exp_neg12(const float* xi, float* T12)
{
    float n[6], T[16];
#pragma unroll
    for (int i = 0; i < 6; i++) n[i] = -xi[i];
    se3_exp(n, T);
#pragma unroll
    for (int i = 0; i < 12; i++) T12[i] = T[i];
}

// Gaussian::operator() (src/math/gaussian.cpp:33-50); returns accepted
__device__ __forceinline__ bool gaussian_fuse(float& depth, float& sigma, float d, float s)
{
    float v1 = sigma * sigma;
    float v2 = s * s;
    float v = v1 + v2;
    float diff = fabsf(d - depth);
    float m = d < diff ? d : diff;
    float gain = ((double)m < 0.8) ? 0.5f + m / 0.8f * 0.5f : 1.0f;
    float mx = (sigma < s) ? s : sigma;
    if (diff > gain * mx) return false;
    depth = (v2 * depth + v1 * d) / v;
    sigma = sqrtf((v1 * v2) / v);
    return true;
}

// libstdc++ uniform_real_distribution<float>(2.0, 0.5) from one mt19937 draw (gaussian.cpp:9,22)
__device__ __forceinline__ float dist_from_u32(uint32_t raw)
{
    float sum = __uint2float_rn(raw);
    float ret = sum / 4294967296.0f;
    if (ret >= 1.0f) ret = 0.99999994f; // nextafter(1, 0)
    return ret * (0.5f - 2.0f) + 2.0f;
}

} // namespace dvo
