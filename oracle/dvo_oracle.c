/*
 * dvo_oracle.c — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).  See dvo_oracle.h.
 *
 * Plain-C restatement of the reference's tracking + mapping hot path, serial row-major,
 * IEEE-strict fp32.  Build with:  gcc -O2 -ffp-contract=off -std=c11 -fPIC -shared
 * (no -ffast-math, no FMA contraction — the arithmetic order below is the specification the
 * CUDA kernels are tested against bit-for-bit where masks/indices are concerned).
 *
 * Third-party arithmetic restated here because OpenCV C++ is not vendored in the reference:
 *   cv::Rodrigues, cv::norm, cv::trace, cv::solve(DECOMP_SVD) (JacobiSVD + SVBkSb, float),
 *   small-matrix cv::gemm (fp32 left-to-right dot products), Point_<float> -> Point_<int>
 *   conversion (= cvRound, round-half-even), libstdc++ std::mt19937 +
 *   uniform_real_distribution<float>.   MatExpr scalar ops are evaluated eagerly, element-wise,
 *   in fp32, left to right (documented choice; see DESIGN.md "third-party arithmetic").
 */
#include "dvo_oracle.h"

#include <float.h>
#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------ */
/* helpers                                                                                    */
/* ------------------------------------------------------------------------------------------ */

static inline int orc_is_valid(float v) { return ORC_INVALID < v; }    /* include/math/util.hpp:9 */
static inline int orc_is_invalid(float v) { return v <= ORC_INVALID; } /* include/math/util.hpp:10 */

/* saturate_cast<int>(float) == cvRound: round-half-even; x86 cvtss2si yields INT_MIN for NaN and
 * out-of-range inputs.  Used wherever a cv::Point2f is implicitly converted to cv::Point2i
 * (src/map/mapper.cpp:94, src/map/implement.cpp:84,240,250-252). */
static inline int orc_round(float v)
{
    if (!(v > -2147483648.0f && v < 2147483648.0f)) return INT_MIN;
    return (int)lrintf(v);
}
/* C cast (int)float: truncation; x86 cvttss2si yields INT_MIN for NaN / out-of-range
 * (src/core/convert.cpp:82-83,133-134). */
static inline int orc_trunc(float v)
{
    if (!(v > -2147483648.0f && v < 2147483648.0f)) return INT_MIN;
    return (int)v;
}
static inline int orc_in_range(int x, int y, int w, int h) /* include/math/util.hpp:46-55 */
{
    if (x < 0 || w <= x) return 0;
    if (y < 0 || h <= y) return 0;
    return 1;
}

void orc_params_default(orc_params* p)
{
    p->min_residual = 5e-3f;
    p->min_update = 5e-4f;
    p->max_iterations = 15;
    p->roi_level = 2;
    p->roi_x0 = 20; p->roi_x1 = 140; p->roi_y0 = 20; p->roi_y1 = 100;
    p->min_depth = 0.20;
    p->sigma_lo = 0.01f; p->sigma_hi = 0.5f;
    p->step_default = 2.0f; p->step_level1 = 1.5f; p->step_level2 = 1.0f;
    p->map_roi_x0 = 16; p->map_roi_x1 = 144; p->map_roi_y0 = 12; p->map_roi_y1 = 108;
    p->min_movement = 0.02f;
    p->max_forward = 6;
    p->accept_depth_lo = 0.2f; p->accept_depth_hi = 6.0f; p->accept_sigma_hi = 0.5f;
    p->huber_delta = 0.0f; p->min_gradient = 0.0f;
}

/* small cv::gemm: fp32 dot products evaluated left to right, no FMA. */
static void mat3_mul(const float* a, const float* b, float* d) /* 3x3 * 3x3 */
{
    float t[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            t[i * 3 + j] = a[i * 3 + 0] * b[0 * 3 + j] + a[i * 3 + 1] * b[1 * 3 + j] + a[i * 3 + 2] * b[2 * 3 + j];
    memcpy(d, t, sizeof t);
}
static void mat3_vec(const float* a, const float* v, float* d) /* 3x3 * 3x1 */
{
    float t[3];
    for (int i = 0; i < 3; i++) t[i] = a[i * 3 + 0] * v[0] + a[i * 3 + 1] * v[1] + a[i * 3 + 2] * v[2];
    d[0] = t[0]; d[1] = t[1]; d[2] = t[2];
}
static void mat4_mul(const float* a, const float* b, float* d) /* 4x4 * 4x4 */
{
    float t[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++)
            t[i * 4 + j] = a[i * 4 + 0] * b[0 * 4 + j] + a[i * 4 + 1] * b[1 * 4 + j] + a[i * 4 + 2] * b[2 * 4 + j]
                           + a[i * 4 + 3] * b[3 * 4 + j];
    memcpy(d, t, sizeof t);
}

/* ------------------------------------------------------------------------------------------ */
/* math — src/math/se3.cpp                                                                    */
/* ------------------------------------------------------------------------------------------ */

/* cv::norm(Mat1f) (NORM_L2): sum of squares in double, sequential, then sqrt (validated against
 * cv2.norm bit-for-bit in tests/test_oracle_known_answers.py). */
double orc_norm(const float* v, int n)
{
    double s = 0;
    for (int i = 0; i < n; i++) s += (double)v[i] * (double)v[i];
    return sqrt(s);
}

static void orc_hat(const float w[3], float M[9]) /* src/math/se3.cpp:8-15 */
{
    M[0] = 0;     M[1] = -w[2]; M[2] = w[1];
    M[3] = w[2];  M[4] = 0;     M[5] = -w[0];
    M[6] = -w[1]; M[7] = w[0];  M[8] = 0;
}

/* so3::exp = cv::Rodrigues(vector -> matrix), src/math/se3.cpp:21-28.  OpenCV's algorithm in
 * double, result rounded to float (bit-exact against cv2.Rodrigues, see tests). */
void orc_rodrigues(const float w[3], float R[9])
{
    double rx = w[0], ry = w[1], rz = w[2];
    double theta = sqrt(rx * rx + ry * ry + rz * rz);
    if (theta < DBL_EPSILON) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    double c = cos(theta), s = sin(theta), c1 = 1. - c, itheta = 1. / theta;
    rx *= itheta; ry *= itheta; rz *= itheta;
    double rrt[9] = {rx * rx, rx * ry, rx * rz, rx * ry, ry * ry, ry * rz, rx * rz, ry * rz, rz * rz};
    double r_x[9] = {0, -rz, ry, rz, 0, -rx, -ry, rx, 0};
    double eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    for (int k = 0; k < 9; k++) R[k] = (float)((c * eye[k] + c1 * rrt[k]) + s * r_x[k]);
}

/* so3::log, src/math/se3.cpp:31-43.  cv::trace sums the diagonal in double; the acos is the
 * double overload; std::sin(float) is the float overload. */
void orc_so3_log(const float R[9], float w[3])
{
    w[0] = w[1] = w[2] = 0.0f;
    double tr = 0;
    tr += R[0]; tr += R[4]; tr += R[8];
    float w_length = (float)acos((tr - 1.0f) * 0.5f);
    if (w_length > 1e-6f) {
        float tmp[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
        /* `k * tmp * w_length` is a lazy cv::MatExpr: the two scalars fold into one double alpha
         * (matop.cpp MatOp_AddEx::multiply) applied by convertTo as src * (float)alpha. */
        float k = 1.0f / (2.0f * sinf(w_length));
        float alpha = (float)((double)k * (double)w_length);
        for (int i = 0; i < 3; i++) w[i] = tmp[i] * alpha;
    }
}

/* se3::exp, src/math/se3.cpp:70-98 */
void orc_se3_exp(const float xi[6], float T[16])
{
    const float* v = xi;
    const float* w = xi + 3;
    float w_hat[9];
    orc_hat(w, w_hat);
    float w_length = (float)orc_norm(w, 3);
    float R[9];
    orc_rodrigues(w, R);
    float t[3];
    if (w_length > 1e-6f) {
        float a = 1.0f - cosf(w_length);
        float b = w_length * w_length;
        float c = w_length - sinf(w_length);
        float d = w_length * w_length * w_length;
        float w_hat2[9], V[9];
        mat3_mul(w_hat, w_hat, w_hat2);
        /* eye + w_hat*a/b + (w_hat*w_hat)*c/d as OpenCV evaluates the lazy MatExpr (matop.cpp):
         * `/ s` multiplies alpha by 1.0/s in double; eye + alpha1*w_hat is cv::scaleAdd
         * (w_hat*(float)alpha1 + eye); the scaled product is cv::gemm's (float)(t*(float)alpha2) with
         * the fp32 dot product t; the final sum is cv::add. */
        const double alpha1 = (double)a * (1.0 / (double)b);
        const double alpha2 = (double)c * (1.0 / (double)d);
        for (int k = 0; k < 9; k++) {
            float eye = (k == 0 || k == 4 || k == 8) ? 1.0f : 0.0f;
            float m1 = w_hat[k] * (float)alpha1 + eye;
            float m2 = w_hat2[k] * (float)alpha2; /* gemm: (float)(t * (float)alpha) */
            V[k] = m1 + m2;
        }
        mat3_vec(V, v, t);
    } else {
        t[0] = v[0]; t[1] = v[1]; t[2] = v[2];
    }
    T[0] = R[0]; T[1] = R[1]; T[2] = R[2];   T[3] = t[0];
    T[4] = R[3]; T[5] = R[4]; T[6] = R[5];   T[7] = t[1];
    T[8] = R[6]; T[9] = R[7]; T[10] = R[8];  T[11] = t[2];
    T[12] = 0;   T[13] = 0;   T[14] = 0;     T[15] = 1;
}

/* se3::log, src/math/se3.cpp:101-124 */
void orc_se3_log(const float T[16], float xi[6])
{
    float R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    float t[3] = {T[3], T[7], T[11]};
    float w[3];
    orc_so3_log(R, w);
    float w_hat[9];
    orc_hat(w, w_hat);
    float w_length = (float)orc_norm(w, 3);
    float V_inv[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (w_length > 1e-6f) {
        float k = 1.0f - (w_length * cosf(w_length * 0.5f)) / (2.0f * sinf(w_length * 0.5f));
        float b = w_length * w_length;
        float w_hat2[9];
        mat3_mul(w_hat, w_hat, w_hat2);
        /* eye - 0.5f*w_hat + k*(w_hat*w_hat)/b: scaleAdd(w_hat, -0.5, eye) + gemm(alpha = k*(1/b)) */
        const double alpha2 = (double)k * (1.0 / (double)b);
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

/* se3::concatenate, src/math/se3.cpp:127-131 */
void orc_se3_concatenate(const float a[6], const float b[6], float out[6])
{
    float Ta[16], Tb[16], Tc[16];
    orc_se3_exp(a, Ta);
    orc_se3_exp(b, Tb);
    mat4_mul(Ta, Tb, Tc);
    orc_se3_log(Tc, out);
}

static int orc_test_xi(const float xi[6]) /* include/math/util.hpp:34-44 */
{
    for (int i = 0; i < 6; i++)
        if (isnan(xi[i])) return 0;
    return 1;
}

/* Convert::inversePose (sic: [R^T | -t ; 0 0 0 0]), src/core/convert.cpp:31-39 */
void orc_inverse_pose(const float T[16], float out[16])
{
    memset(out, 0, 16 * sizeof(float));
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) out[i * 4 + j] = T[j * 4 + i];
    for (int i = 0; i < 3; i++) out[i * 4 + 3] = -T[i * 4 + 3];
}

/* ------------------------------------------------------------------------------------------ */
/* gaussian — src/math/gaussian.cpp                                                           */
/* ------------------------------------------------------------------------------------------ */

void orc_rng_seed(orc_rng* r, uint32_t seed) /* std::mt19937 (default_seed 5489) */
{
    r->mt[0] = seed;
    for (int i = 1; i < 624; i++) r->mt[i] = 1812433253u * (r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) + (uint32_t)i;
    r->idx = 624;
    r->draws = 0;
}
uint32_t orc_rng_next_u32(orc_rng* r)
{
    if (r->idx >= 624) {
        for (int i = 0; i < 624; i++) {
            uint32_t y = (r->mt[i] & 0x80000000u) | (r->mt[(i + 1) % 624] & 0x7fffffffu);
            r->mt[i] = r->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        r->idx = 0;
    }
    uint32_t y = r->mt[r->idx++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    r->draws++;
    return y;
}
/* libstdc++ uniform_real_distribution<float>(2.0, 0.5) fed by one 32-bit draw
 * (generate_canonical<float,24>: float(raw)/2^32, clamped below 1), src/math/gaussian.cpp:9,22. */
float orc_dist_from_u32(uint32_t raw)
{
    float sum = (float)raw;          /* rounds to 24 bits */
    float ret = sum / 4294967296.0f; /* exact */
    if (ret >= 1.0f) ret = nextafterf(1.0f, 0.0f);
    return ret * (0.5f - 2.0f) + 2.0f;
}
float orc_rng_dist(orc_rng* r) { return orc_dist_from_u32(orc_rng_next_u32(r)); }

/* Gaussian::update, src/math/gaussian.cpp:12-31 */
int orc_gaussian_update(float* depth, float* sigma, float d, float s, orc_rng* r)
{
    float v1 = (*sigma) * (*sigma);
    float v2 = s * s;
    float v = v1 + v2;
    float diff = fabsf(d - *depth);
    float m = d < diff ? d : diff; /* std::min(d, diff) */
    float gain = ((double)m < 0.8) ? 0.5f + m / 0.8f * 0.5f : 1.0f;
    float mx = (*sigma < s) ? s : *sigma; /* std::max(sigma, s) */
    if (diff > gain * mx) {
        float u = orc_rng_dist(r);
        *depth = (4.0f < u) ? 4.0f : u; /* std::min(dist(engine), 4.0f) */
        *sigma = 0.5f;
        return 0;
    }
    *depth = (v2 * (*depth) + v1 * d) / v;
    *sigma = sqrtf((v1 * v2) / v);
    return 1;
}

/* Gaussian::operator(), src/math/gaussian.cpp:33-50 */
int orc_gaussian_fuse(float* depth, float* sigma, float d, float s)
{
    float v1 = (*sigma) * (*sigma);
    float v2 = s * s;
    float v = v1 + v2;
    float diff = fabsf(d - *depth);
    float m = d < diff ? d : diff;
    float gain = ((double)m < 0.8) ? 0.5f + m / 0.8f * 0.5f : 1.0f;
    float mx = (*sigma < s) ? s : *sigma;
    if (diff > gain * mx) return 0;
    *depth = (v2 * (*depth) + v1 * d) / v;
    *sigma = sqrtf((v1 * v2) / v);
    return 1;
}

/* ------------------------------------------------------------------------------------------ */
/* core — src/core/transform.cpp, src/core/convert.cpp                                        */
/* ------------------------------------------------------------------------------------------ */

/* Transform::warp = project(transform(backProject)), src/core/transform.cpp:7-33, T is 4x4. */
void orc_warp_point(const float T[16], const float K[9], float x, float y, float depth, float out[2])
{
    const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    /* backProject :27 */
    float X = depth * (x - cx) / fx;
    float Y = depth * (y - cy) / fy;
    float Z = depth;
    /* transform :10 (gemm R*x then + t) */
    float tx = (T[0] * X + T[1] * Y + T[2] * Z) + T[3];
    float ty = (T[4] * X + T[5] * Y + T[6] * Z) + T[7];
    float tz = (T[8] * X + T[9] * Y + T[10] * Z) + T[11];
    /* project :22 */
    out[0] = tx * fx / tz + cx;
    out[1] = ty * fy / tz + cy;
}

float orc_get_pixel(const float* img, int w, int h, int x, int y) /* src/core/convert.cpp:107-125 */
{
    if (orc_in_range(x, y, w, h)) {
        float g = img[(size_t)y * w + x];
        if (orc_is_valid(g)) return g;
    }
    return ORC_INVALID;
}

float orc_get_subpixel_dense(const float* img, int w, int h, float px, float py) /* convert.cpp:77-105 */
{
    int x0 = orc_trunc(px), y0 = orc_trunc(py);
    if (!orc_in_range(x0, y0, w, h)) return ORC_INVALID;
    int x1 = x0 + 1, y1 = y0 + 1;
    float hh = px - (float)x0, v = py - (float)y0;
    float g[4];
    g[0] = g[1] = g[2] = g[3] = img[(size_t)y0 * w + x0];
    if (orc_in_range(x1, y0, w, h)) g[1] = img[(size_t)y0 * w + x1];
    if (orc_in_range(x0, y1, w, h)) g[2] = img[(size_t)y1 * w + x0];
    if (orc_in_range(x1, y1, w, h)) g[3] = img[(size_t)y1 * w + x1];
    return (g[0] * (1.f - hh) + g[1] * hh) * (1.f - v) + (g[2] * (1.f - hh) + g[3] * hh) * v;
}

float orc_get_subpixel(const float* img, int w, int h, float px, float py) /* convert.cpp:128-177 */
{
    int x0 = orc_trunc(px), y0 = orc_trunc(py);
    if (!orc_in_range(x0, y0, w, h)) return ORC_INVALID;
    int x1 = x0 + 1, y1 = y0 + 1;
    float hh = px - (float)x0, v = py - (float)y0;
    float g[4];
    g[0] = g[1] = g[2] = g[3] = img[(size_t)y0 * w + x0];
    if (orc_in_range(x1, y0, w, h)) g[1] = img[(size_t)y0 * w + x1];
    if (orc_in_range(x0, y1, w, h)) g[2] = img[(size_t)y1 * w + x0];
    if (orc_in_range(x1, y1, w, h)) g[3] = img[(size_t)y1 * w + x1];
    int valid = 0, id = 0;
    float last = -1;
    for (;;) { /* the reference's repair automaton, reproduced literally */
        if (orc_is_valid(g[id])) {
            valid++;
            last = g[id];
        } else if (last > 0) {
            g[id] = last;
            valid++;
        }
        if (valid == 4) break;
        if (id == 3 && valid == 0) return ORC_INVALID;
        id = (id + 1) % 4;
    }
    return (g[0] * (1.f - hh) + g[1] * hh) * (1.f - v) + (g[2] * (1.f - hh) + g[3] * hh) * v;
}

void orc_cull_image(const float* src, int w, int h, int times, float* dst) /* convert.cpp:7-20 */
{
    int red = 1 << times;
    int dw = w / red, dh = h / red;
    if (times == 0) {
        memcpy(dst, src, (size_t)w * h * sizeof(float));
        return;
    }
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) dst[(size_t)y * dw + x] = orc_get_pixel(src, w, h, x * red, y * red);
}

void orc_cull_intrinsic(const float K[9], int times, float out[9]) /* convert.cpp:22-29 */
{
    if (times == 0) {
        memcpy(out, K, 9 * sizeof(float));
        return;
    }
    float s = (float)(1.0 / (double)(1 << times)); /* Mat / int -> scale by 1/s */
    for (int i = 0; i < 9; i++) out[i] = K[i] * s;
    out[8] = 1;
}

void orc_gradiate(const float* gray, int w, int h, int is_x, float* out) /* convert.cpp:41-75 */
{
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            float p = ORC_INVALID;
            if (is_x) {
                if (!(x - 1 <= -1 || x + 1 >= w)) {
                    float a = orc_get_pixel(gray, w, h, x - 1, y);
                    float b = orc_get_pixel(gray, w, h, x + 1, y);
                    if (!(orc_is_invalid(a) || orc_is_invalid(b))) p = b - a;
                }
            } else {
                if (!(y - 1 <= -1 || y + 1 >= h)) {
                    float a = orc_get_pixel(gray, w, h, x, y - 1);
                    float b = orc_get_pixel(gray, w, h, x, y + 1);
                    if (!(orc_is_invalid(a) || orc_is_invalid(b))) p = b - a;
                }
            }
            out[(size_t)y * w + x] = p;
        }
}

/* Transform::warpImage, src/core/transform.cpp:35-51 */
void orc_warp_image(const float xi[6], const float* gray, const float* depth, const float K[9],
                    int w, int h, float* out)
{
    float nxi[6], T[16];
    for (int i = 0; i < 6; i++) nxi[i] = -xi[i];
    orc_se3_exp(nxi, T);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            float d = depth[(size_t)y * w + x];
            float p = ORC_INVALID;
            if (!(fabsf(d) < ORC_EPSILON)) {
                float wp[2];
                orc_warp_point(T, K, (float)x, (float)y, d, wp);
                p = orc_get_subpixel(gray, w, h, wp[0], wp[1]);
            }
            out[(size_t)y * w + x] = p;
        }
}

/* ------------------------------------------------------------------------------------------ */
/* cv::solve(A, b, x, DECOMP_SVD) for float — OpenCV modules/core/src/lapack.cpp              */
/* (JacobiSVDImpl_<float> one-sided Jacobi on A^T + SVBkSbImpl_), restated from the published */
/* algorithm.  OpenCV is an un-vendored dependency of the reference ("OpenCV >=3",            */
/* README.md:8); call site src/track/optimize.cpp:97.                                         */
/* ------------------------------------------------------------------------------------------ */
int orc_solve_svd(const float* A, const float* b, int m, int n, float* x)
{
    /* At: n x m (rows = columns of A) */
    float* At = (float*)malloc((size_t)n * m * sizeof(float));
    float* Vt = (float*)malloc((size_t)n * n * sizeof(float));
    double* W = (double*)malloc((size_t)n * sizeof(double));
    float* Wf = (float*)malloc((size_t)n * sizeof(float));
    if (!At || !Vt || !W || !Wf) return -1;
    for (int i = 0; i < m; i++)
        for (int j = 0; j < n; j++) At[(size_t)j * m + i] = A[(size_t)i * n + j];

    const float eps = FLT_EPSILON * 2;
    const double minval = FLT_MIN;
    int max_iter = m > 30 ? m : 30;
    if (max_iter > 1000) max_iter = 1000; /* converges in < 20 sweeps; bound kept finite */

    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) {
            float t = At[(size_t)i * m + k];
            sd += (double)t * t;
        }
        W[i] = sd;
        for (int k = 0; k < n; k++) Vt[i * n + k] = 0;
        Vt[i * n + i] = 1;
    }
    for (int iter = 0; iter < max_iter; iter++) {
        int changed = 0;
        for (int i = 0; i < n - 1; i++)
            for (int j = i + 1; j < n; j++) {
                float *Ai = At + (size_t)i * m, *Aj = At + (size_t)j * m;
                double a = W[i], p = 0, bb = W[j];
                for (int k = 0; k < m; k++) p += (double)Ai[k] * Aj[k];
                if (fabs(p) <= eps * sqrt(a * bb)) continue;
                p *= 2;
                double beta = a - bb, gamma = hypot(p, beta);
                float c, s;
                if (beta < 0) {
                    double delta = (gamma - beta) * 0.5;
                    s = (float)sqrt(delta / gamma);
                    c = (float)(p / (gamma * s * 2));
                } else {
                    c = (float)sqrt((gamma + beta) / (gamma * 2));
                    s = (float)(p / (gamma * c * 2));
                }
                a = bb = 0;
                for (int k = 0; k < m; k++) {
                    float t0 = c * Ai[k] + s * Aj[k];
                    float t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0; Aj[k] = t1;
                    a += (double)t0 * t0; bb += (double)t1 * t1;
                }
                W[i] = a; W[j] = bb;
                changed = 1;
                float *Vi = Vt + i * n, *Vj = Vt + j * n;
                for (int k = 0; k < n; k++) {
                    float t0 = c * Vi[k] + s * Vj[k];
                    float t1 = -s * Vi[k] + c * Vj[k];
                    Vi[k] = t0; Vj[k] = t1;
                }
            }
        if (!changed) break;
    }
    for (int i = 0; i < n; i++) {
        double sd = 0;
        for (int k = 0; k < m; k++) {
            float t = At[(size_t)i * m + k];
            sd += (double)t * t;
        }
        W[i] = sqrt(sd);
    }
    for (int i = 0; i < n - 1; i++) { /* sort descending */
        int j = i;
        for (int k = i + 1; k < n; k++)
            if (W[j] < W[k]) j = k;
        if (i != j) {
            double tw = W[i]; W[i] = W[j]; W[j] = tw;
            for (int k = 0; k < m; k++) { float t = At[(size_t)i * m + k]; At[(size_t)i * m + k] = At[(size_t)j * m + k]; At[(size_t)j * m + k] = t; }
            for (int k = 0; k < n; k++) { float t = Vt[i * n + k]; Vt[i * n + k] = Vt[j * n + k]; Vt[j * n + k] = t; }
        }
    }
    for (int i = 0; i < n; i++) Wf[i] = (float)W[i];
    /* left singular vectors: u_i = At_i / w_i (zero singular values -> row left as zero: it is
     * skipped by the threshold in the back-substitution below, so OpenCV's random completion of
     * U for null singular values cannot influence x). */
    for (int i = 0; i < n; i++) {
        double sd = W[i];
        float s = (float)(sd > minval ? 1 / sd : 0.);
        for (int k = 0; k < m; k++) At[(size_t)i * m + k] *= s;
    }
    /* SVBkSb: x = V * diag(1/w) * U^T b, threshold eps * sum(w) (double) */
    double threshold = 0;
    for (int i = 0; i < n; i++) threshold += Wf[i];
    threshold *= (double)(float)(FLT_EPSILON * 2);
    for (int j = 0; j < n; j++) x[j] = 0;
    int nm = n; /* the reference's A always has w*h >= n rows (zero rows for invalid pixels) */
    for (int i = 0; i < nm; i++) {
        double wi = Wf[i];
        if (fabs(wi) <= threshold) continue;
        wi = 1 / wi;
        double s = 0;
        for (int j = 0; j < m; j++) s += At[(size_t)i * m + j] * b[j]; /* float product, double sum */
        s *= wi;
        for (int j = 0; j < n; j++) x[j] = (float)(x[j] + s * Vt[i * n + j]);
    }
    free(At); free(Vt); free(W); free(Wf);
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* track — src/track/optimize.cpp:10-99                                                       */
/* ------------------------------------------------------------------------------------------ */
/* sys_A / sys_B (optional, caller-allocated w*h*6 / w*h floats): the stacked system cv::solve is handed, rows of valid pixels
 * only, B already negated (optimize.cpp:87-97) - lets a test solve the very same system in fp64. */
void orc_optimize_system(const float* obj_gray, const float* warped_gray, const float* ref_depth,
                         const float* ref_sigma, const float* grad_x, const float* grad_y,
                         const float K[9], const float xi[6], int w, int h, int level_index,
                         const orc_params* P, float upd[6], float* residual_out, int* valid_out,
                         uint8_t* mask, float* sys_A, float* sys_B);

void orc_optimize(const float* obj_gray, const float* warped_gray, const float* ref_depth,
                  const float* ref_sigma, const float* grad_x, const float* grad_y,
                  const float K[9], const float xi[6], int w, int h, int level_index,
                  const orc_params* P, float upd[6], float* residual_out, int* valid_out,
                  uint8_t* mask)
{
    orc_optimize_system(obj_gray, warped_gray, ref_depth, ref_sigma, grad_x, grad_y, K, xi, w, h, level_index, P, upd, residual_out,
                        valid_out, mask, NULL, NULL);
}

void orc_optimize_system(const float* obj_gray, const float* warped_gray, const float* ref_depth,
                         const float* ref_sigma, const float* grad_x, const float* grad_y,
                         const float K[9], const float xi[6], int w, int h, int level_index,
                         const orc_params* P, float upd[6], float* residual_out, int* valid_out,
                         uint8_t* mask, float* sys_A, float* sys_B)
{
    float residual = 0;
    const float fx = K[0], fy = K[4];
    int valid_pixels = 0;
    size_t max_size = (size_t)w * h;
    float* A = (float*)malloc(max_size * 6 * sizeof(float)); /* compacted rows; zero rows of the
        reference's N x 6 matrix add exact zeros to every sum in the Jacobi SVD, so dropping them
        is bit-identical. */
    float* B = (float*)malloc(max_size * sizeof(float));

    float step = P->step_default; /* :22-26 */
    if (level_index == 1) step = P->step_level1;
    if (level_index == 2) step = P->step_level2;

    float nxi[6], T[16];
    for (int i = 0; i < 6; i++) nxi[i] = -xi[i];
    orc_se3_exp(nxi, T); /* Transform::warp(-xi, ...) re-exponentiates per pixel; same value */

    if (mask) memset(mask, 0, max_size);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            size_t id = (size_t)y * w + x;
            float depth = ref_depth[id];
            if (level_index == P->roi_level) { /* :33-36 */
                if (x < P->roi_x0 || x > P->roi_x1 || y < P->roi_y0 || y > P->roi_y1) continue;
            }
            if ((double)depth < P->min_depth) continue; /* :39 */
            float I_1 = obj_gray[id];
            float I_2 = warped_gray[id];
            if (orc_is_invalid(I_1) || orc_is_invalid(I_2)) continue; /* :46 */
            float wp[2];
            orc_warp_point(T, K, (float)x, (float)y, depth, wp); /* :51 */
            if (wp[0] < 0 || wp[1] < 0 || (float)w <= wp[0] || (float)h <= wp[1]) continue;
            float gx = orc_get_subpixel_dense(grad_x, w, h, wp[0], wp[1]);
            float gy = orc_get_subpixel_dense(grad_y, w, h, wp[0], wp[1]);
            if (orc_is_invalid(gx) || orc_is_invalid(gy)) continue; /* :61 */
            if (P->min_gradient > 0.0f && gx * gx + gy * gy < P->min_gradient * P->min_gradient) continue; /* extra, default off */

            /* jacobian :67-77 */
            float X = depth * ((float)x - K[2]) / K[0];
            float Y = depth * ((float)y - K[5]) / K[4];
            float Z = depth;
            float fgx = fx * gx, fgy = fy * gy;
            float xz = X / Z, yz = Y / Z;
            float* J = A + (size_t)valid_pixels * 6;
            J[0] = fgx / Z;
            J[1] = fgy / Z;
            J[2] = -(fgx * X + fgy * Y) / Z / Z;
            J[3] = -fgx * xz * yz - fgy * (1.0f + yz * yz);
            J[4] = fgx * (1.0f + xz * xz) + fgy * xz * yz;
            J[5] = (-fgx * yz + fgy * xz);

            float r = I_2 - I_1;
            residual += r * r;
            float s = ref_sigma[id]; /* std::clamp(sigma, 0.01f, 0.5f) :83 */
            float sigma = (s < P->sigma_lo) ? P->sigma_lo : (P->sigma_hi < s) ? P->sigma_hi : s;
            float weight = step / sigma;
            B[valid_pixels] = -(r * weight); /* cv::solve(A, -B, ...) :97 */
            if (P->huber_delta > 0.0f) { /* extra, default off: variance-normalised Huber weight, rows scaled by sqrt(w) */
                float t = fabsf(r) / sigma;
                float wh = (t <= P->huber_delta) ? 1.0f : P->huber_delta / t;
                float sw = sqrtf(wh);
                for (int q = 0; q < 6; q++) J[q] = J[q] * sw;
                B[valid_pixels] = B[valid_pixels] * sw;
            }
            valid_pixels++;
            if (mask) mask[id] = 1;
        }

    *valid_out = valid_pixels;
    if (sys_A) memcpy(sys_A, A, (size_t)valid_pixels * 6 * sizeof(float));
    if (sys_B) memcpy(sys_B, B, (size_t)valid_pixels * sizeof(float));
    if (valid_pixels == 0) { /* :92-93 */
        for (int i = 0; i < 6; i++) upd[i] = 0;
        *residual_out = -1;
    } else {
        float xs[6];
        orc_solve_svd(A, B, valid_pixels, 6, xs);
        for (int i = 0; i < 6; i++) upd[i] = -xs[i]; /* :98 */
        *residual_out = residual / (float)valid_pixels;
    }
    free(A);
    free(B);
}

/* ------------------------------------------------------------------------------------------ */
/* frames — include/system/frame.hpp, src/system/frame.cpp                                    */
/* ------------------------------------------------------------------------------------------ */
typedef struct orc_scene {
    int w, h;
    float *gray, *depth, *sigma, *gx, *gy; /* gx/gy lazily computed */
    float K[9];
} orc_scene;

struct orc_frame {
    int id, levels, culls;
    int w0, h0;     /* input size */
    orc_scene* sc;  /* index 0 = coarsest ... levels-1 = finest (frame.cpp:30-37) */
    float* age;     /* finest size */
    float xi[6], rel_xi[6];
    orc_frame* ref;
    int refcount;
};

static int g_latest_id = -1; /* Frame::latest_id, src/system/frame.cpp:5 */

static void scene_alloc(orc_scene* s, int w, int h)
{
    s->w = w; s->h = h;
    size_t n = (size_t)w * h;
    s->gray = (float*)malloc(n * sizeof(float));
    s->depth = (float*)malloc(n * sizeof(float));
    s->sigma = (float*)malloc(n * sizeof(float));
    s->gx = NULL; s->gy = NULL;
}

static void frame_recull_depth_sigma(orc_frame* f, int do_depth, int do_sigma) /* frame.cpp:39-61 */
{
    orc_scene* top = &f->sc[f->levels - 1];
    for (int i = 0; i < f->levels - 1; i++) {
        int times = f->levels - 1 - i;
        if (do_depth) orc_cull_image(top->depth, top->w, top->h, times, f->sc[i].depth);
        if (do_sigma) orc_cull_image(top->sigma, top->w, top->h, times, f->sc[i].sigma);
    }
}

/* Frame ctors, include/system/frame.hpp:91-117.  depth/sigma may be NULL (4-arg ctor: the
 * reference fills a cv::randn prior that is never read except on the very first keyframe,
 * which callers inject through orc_frame_set_depth_sigma_age — SURVEY H6). */
orc_frame* orc_frame_create(const float* gray, const float* depth, const float* sigma, int w, int h,
                            const float K[9], int levels, int culls)
{
    orc_frame* f = (orc_frame*)calloc(1, sizeof(orc_frame));
    f->id = ++g_latest_id;
    f->levels = levels; f->culls = culls; f->w0 = w; f->h0 = h;
    f->sc = (orc_scene*)calloc((size_t)levels, sizeof(orc_scene));
    f->refcount = 1;
    int bw = w >> culls, bh = h >> culls;
    if (culls == 0) { bw = w; bh = h; }
    orc_scene* top = &f->sc[levels - 1];
    scene_alloc(top, bw, bh);
    orc_cull_image(gray, w, h, culls, top->gray);
    if (depth) orc_cull_image(depth, w, h, culls, top->depth);
    else for (size_t i = 0; i < (size_t)bw * bh; i++) top->depth[i] = 1.5f;
    if (sigma) orc_cull_image(sigma, w, h, culls, top->sigma);
    else for (size_t i = 0; i < (size_t)bw * bh; i++) top->sigma[i] = 0.5f;
    orc_cull_intrinsic(K, culls, top->K);
    for (int i = 0; i < levels - 1; i++) {
        int times = levels - 1 - i;
        int red = 1 << times;
        scene_alloc(&f->sc[i], bw / red, bh / red);
        orc_cull_image(top->gray, bw, bh, times, f->sc[i].gray);
        orc_cull_intrinsic(top->K, times, f->sc[i].K);
    }
    frame_recull_depth_sigma(f, 1, 1);
    f->age = (float*)calloc((size_t)bw * bh, sizeof(float));
    return f;
}

void orc_frame_destroy(orc_frame* f)
{
    if (!f) return;
    if (--f->refcount > 0) return;
    for (int i = 0; i < f->levels; i++) {
        free(f->sc[i].gray); free(f->sc[i].depth); free(f->sc[i].sigma); free(f->sc[i].gx); free(f->sc[i].gy);
    }
    free(f->sc);
    free(f->age);
    free(f);
}

int orc_frame_level_size(const orc_frame* f, int level, int* w, int* h)
{
    if (level < 0 || level >= f->levels) return -1;
    *w = f->sc[level].w; *h = f->sc[level].h;
    return 0;
}

static const float* scene_gx(orc_scene* s) /* frame.hpp:52-57 */
{
    if (!s->gx) {
        s->gx = (float*)malloc((size_t)s->w * s->h * sizeof(float));
        orc_gradiate(s->gray, s->w, s->h, 1, s->gx);
    }
    return s->gx;
}
static const float* scene_gy(orc_scene* s) /* frame.hpp:58-63 */
{
    if (!s->gy) {
        s->gy = (float*)malloc((size_t)s->w * s->h * sizeof(float));
        orc_gradiate(s->gray, s->w, s->h, 0, s->gy);
    }
    return s->gy;
}

const float* orc_frame_image(orc_frame* f, int level, int which)
{
    if (level < 0 || level >= f->levels) return NULL;
    orc_scene* s = &f->sc[level];
    switch (which) {
    case 0: return s->gray;
    case 1: return s->depth;
    case 2: return s->sigma;
    case 3: return scene_gx(s);
    case 4: return scene_gy(s);
    case 5: return level == f->levels - 1 ? f->age : NULL;
    }
    return NULL;
}
void orc_frame_K(const orc_frame* f, int level, float K[9]) { memcpy(K, f->sc[level].K, 9 * sizeof(float)); }
void orc_frame_get_xi(const orc_frame* f, float xi[6], float rel_xi[6])
{
    memcpy(xi, f->xi, sizeof f->xi);
    memcpy(rel_xi, f->rel_xi, sizeof f->rel_xi);
}

/* Frame::updateDepthSigmaAge / updateDepthSigma / updateDepth, src/system/frame.cpp:39-61 */
void orc_frame_set_depth_sigma_age(orc_frame* f, const float* depth, const float* sigma, const float* age)
{
    orc_scene* top = &f->sc[f->levels - 1];
    size_t n = (size_t)top->w * top->h;
    if (depth && depth != top->depth) memcpy(top->depth, depth, n * sizeof(float));
    if (sigma && sigma != top->sigma) memcpy(top->sigma, sigma, n * sizeof(float));
    if (age && age != f->age) memcpy(f->age, age, n * sizeof(float));
    frame_recull_depth_sigma(f, depth != NULL, sigma != NULL);
}

/* ------------------------------------------------------------------------------------------ */
/* Tracker::track — src/track/tracker.cpp:22-85 (200 ms budget treated as never reached, H1)   */
/* ------------------------------------------------------------------------------------------ */
int orc_track(orc_frame* obj, orc_frame* ref, const orc_params* P, float xi_out[6],
              orc_iter_log* log, int log_cap, double stop_update, float stop_residual)
{
    float xi[6] = {0, 0, 0, 0, 0, 0};
    int n_log = 0;
    for (int level = 0; level < ref->levels; level++) {
        orc_scene* os = &obj->sc[level];
        orc_scene* rs = &ref->sc[level];
        const int w = rs->w, h = rs->h;
        const float* gx = scene_gx(rs);
        const float* gy = scene_gy(rs);
        float* warped = (float*)malloc((size_t)w * h * sizeof(float));
        orc_warp_image(xi, rs->gray, rs->depth, rs->K, w, h, warped); /* Stuff ctor, optimize.hpp:19 */
        for (int iteration = 0; iteration < P->max_iterations; iteration++) {
            float upd[6], residual;
            int valid;
            orc_optimize(os->gray, warped, rs->depth, rs->sigma, gx, gy, rs->K, xi, w, h, level, P,
                         upd, &residual, &valid, NULL);
            float updated[6];
            orc_se3_concatenate(xi, upd, updated);
            if (log && n_log < log_cap) {
                orc_iter_log* L = &log[n_log];
                L->level = level; L->iteration = iteration; L->valid_pixels = valid; L->residual = residual;
                memcpy(L->update, upd, sizeof upd);
                memcpy(L->xi_before, xi, sizeof xi);
            }
            if (orc_test_xi(updated)) memcpy(xi, updated, sizeof xi); /* :47-51 */
            if (log && n_log < log_cap) memcpy(log[n_log].xi_after, xi, sizeof xi);
            n_log++;
            orc_warp_image(xi, rs->gray, rs->depth, rs->K, w, h, warped); /* stuff.update :52 */
            if (orc_norm(upd, 6) < stop_update || residual < stop_residual) break; /* :68-70 */
        }
        free(warped);
    }
    memcpy(xi_out, xi, sizeof xi);
    return n_log;
}

/* ------------------------------------------------------------------------------------------ */
/* map — src/map/implement.cpp                                                                */
/* ------------------------------------------------------------------------------------------ */
static const float luminance_variance = 0.5f * 0.5f; /* implement.cpp:12-13 */
static const float epipolar_variance = 0.5f * 0.5f;  /* implement.cpp:14-15 */
static const float predict_variance = 0.06f * 0.06f; /* implement.cpp:17-18 */
static const double MATCHING_THRESHOLD_RATIO = 0.1;  /* implement.cpp:20 */

typedef struct {
    float min, max;
    float start[2], end[2];
    float length;
} orc_segment;

/* cv::norm(Point2f) = sqrt((double)x*x + (double)y*y) */
static double point_norm(float x, float y) { return sqrt((double)x * x + (double)y * y); }

/* Implement::update, src/map/implement.cpp:182-214 (+ EpipolarSegment :23-47, doMatching
 * :106-152, depthEstimate :49-71, sigmaEstimate :73-104). */
void orc_map_update_pixel(const float* obj_gray, const float* ref_gray, const float* ref_gradx,
                          const float* ref_grady, int w, int h, const float relative_xi[6],
                          const float K[9], int px, int py, float depth, float sigma,
                          float* new_depth, float* new_sigma)
{
    float es_xi[6], Tes[16];
    for (int i = 0; i < 6; i++) es_xi[i] = -relative_xi[i]; /* :193 */
    orc_se3_exp(es_xi, Tes);

    orc_segment es;
    es.min = (depth - sigma) < 0.10f ? 0.10f : (depth - sigma); /* std::max(depth - sigma, 0.10f) :30 */
    es.max = depth + sigma;
    orc_warp_point(Tes, K, (float)px, (float)py, es.max, es.start); /* :31 */
    orc_warp_point(Tes, K, (float)px, (float)py, es.min, es.end);   /* :32 */
    es.length = (float)point_norm(es.start[0] - es.end[0], es.start[1] - es.end[1]); /* :33 */

    /* doMatching :106-152 */
    const float og = obj_gray[(size_t)py * w + px]; /* obj_gray(x_i) :195 */
    float dir[2] = {(es.end[0] - es.start[0]) / es.length, (es.end[1] - es.start[1]) / es.length};
    float pt[2] = {es.start[0], es.start[1]};
    float best[2] = {pt[0], pt[1]};
    const int N = 3;
    const int center = (N + 1) / 2;
    float min_ssd = 2.0f * N;
    int count = 0;
    while (point_norm(pt[0] - es.start[0], pt[1] - es.start[1]) < es.length) {
        float ssd = 0;
        pt[0] += dir[0];
        pt[1] += dir[1];
        for (int i = 0; i < N; i++) {
            float k = (float)(i - N / 2);
            float tx = pt[0] + dir[0] * k; /* int * Point2f -> saturate_cast<float>(b.x * a) */
            float ty = pt[1] + dir[1] * k;
            float sub = orc_get_subpixel_dense(ref_gray, w, h, tx, ty);
            if (orc_is_invalid(sub)) {
                ssd = 2.0f * N;
                break;
            }
            float diff = sub - og;
            ssd = (float)((double)ssd + 1.0 * (N - abs(i - center)) / N * (double)(diff * diff)); /* :134 */
        }
        if (ssd < min_ssd) {
            best[0] = pt[0]; best[1] = pt[1];
            min_ssd = ssd;
        }
        if (count++ > 100) break;
    }
    float m[2];
    if ((double)min_ssd > N * MATCHING_THRESHOLD_RATIO) { /* :145 */
        m[0] = -1; m[1] = -1;
    } else {
        m[0] = best[0]; m[1] = best[1];
    }
    if (m[0] < 0 || m[1] < 0 || m[0] > (float)w || m[1] > (float)h) { /* :196-200 */
        *new_depth = -1; *new_sigma = -1;
        return;
    }

    /* depthEstimate(ref_x_i = m, obj_x_i = x_i, K, xi = relative_xi) :49-71 */
    {
        float x_q[3] = {1.0f * ((float)px - K[2]) / K[0], 1.0f * ((float)py - K[5]) / K[4], 1.0f};
        float t[3] = {-relative_xi[0], -relative_xi[1], -relative_xi[2]};
        const float* T = Tes; /* se3::exp(-xi) */
        float R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
        float x_i[3] = {m[0], m[1], 1.0f};
        double r3q = 0; /* Mat::dot -> double */
        r3q += (double)R[6] * x_q[0]; r3q += (double)R[7] * x_q[1]; r3q += (double)R[8] * x_q[2];
        float KR[9], KRq[3], Kt[3];
        mat3_mul(K, R, KR);
        mat3_vec(KR, x_q, KRq);
        mat3_vec(K, t, Kt);
        /* `s*x_i - (K*R)*x_q` and `t(2)*x_i - K*t` fold into cv::gemm(A, B, -1, x_i, s)
         * (MatOp_GEMM::subtract): d = (float)(t*(float)alpha + c*(float)beta) evaluated in double. */
        float a[3], b[3];
        for (int i = 0; i < 3; i++) {
            a[i] = (float)((double)KRq[i] * -1.0 + (double)x_i[i] * (double)(float)r3q);
            b[i] = (float)((double)Kt[i] * -1.0 + (double)x_i[i] * (double)t[2]);
        }
        double ab = 0, aa = 0;
        for (int i = 0; i < 3; i++) { ab += (double)a[i] * b[i]; aa += (double)a[i] * a[i]; }
        *new_depth = -(float)(ab / aa);
    }

    /* sigmaEstimate :73-104 */
    {
        float l = es.length;
        float lx = (es.start[0] - es.end[0]) / l, ly = (es.start[1] - es.end[1]) / l;
        const float alpha = (es.max - es.min) / l;
        int rx = orc_round(m[0]), ry = orc_round(m[1]); /* Point2f -> Point2i :84 */
        long idx = (long)ry * w + rx;                    /* H9: linear addressing when in-buffer */
        if (rx == INT_MIN || ry == INT_MIN || idx < 0 || idx >= (long)w * h) {
            *new_sigma = -1;
            return;
        }
        float gx = ref_gradx[idx], gy = ref_grady[idx];
        if (orc_is_invalid(gx) || orc_is_invalid(gy)) {
            *new_sigma = -1;
            return;
        }
        float g_dot_l = fabsf(gx * lx + gy * ly);
        float g_dot_l2 = g_dot_l * g_dot_l;
        float gp2 = g_dot_l / l;
        float epipolar = epipolar_variance / (g_dot_l2 < ORC_EPSILON ? ORC_EPSILON : g_dot_l2);
        float luminance = 2 * luminance_variance / (gp2 < ORC_EPSILON ? ORC_EPSILON : gp2);
        *new_sigma = alpha * sqrtf(epipolar + luminance);
    }
}

/* Implement::propagate, src/map/implement.cpp:217-256 (serial: later sources overwrite) */
void orc_propagate(const float* ref_depth, const float* ref_sigma, const float* ref_age,
                   const float xi[6], const float K[9], int w, int h,
                   float* depth, float* sigma, float* age)
{
    const float tz = xi[2];
    float T[16];
    orc_se3_exp(xi, T);
    size_t n = (size_t)w * h;
    for (size_t i = 0; i < n; i++) { depth[i] = 1.0f; sigma[i] = 1.0f; age[i] = 0.0f; }
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            size_t id = (size_t)y * w + x;
            float rd = ref_depth[id];
            if (fabsf(rd) < ORC_EPSILON) continue;
            float wp[2];
            orc_warp_point(T, K, (float)x, (float)y, rd, wp);
            int wx = orc_round(wp[0]), wy = orc_round(wp[1]);
            if (!orc_in_range(wx, wy, w, h)) continue;
            float s = ref_sigma[id];
            float d0 = rd < 0.01f ? 0.01f : rd; /* std::max(rd, 0.01f) */
            float d1 = d0 + tz;
            float q = d1 / d0;
            float q4 = q * (q * (q * q)); /* math::pow(base, 4): base * pow(base, 3) ... */
            s = sqrtf(q4 * (s * s) + predict_variance);
            size_t o = (size_t)wy * w + wx;
            depth[o] = d1 < 0.0f ? 0.0f : d1; /* std::max(d1, 0.0f) */
            sigma[o] = s;
            age[o] = ref_age[id] + 1;
        }
}

/* Implement::regularize, src/map/implement.cpp:156-180 */
void orc_regularize(const float* depth, const float* sigma, int w, int h, float* out)
{
    static const int off[4][2] = {{0, -1}, {0, 1}, {1, 0}, {-1, 0}}; /* {dy, dx} :160 */
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            float gd = depth[(size_t)y * w + x], gs = sigma[(size_t)y * w + x];
            for (int k = 0; k < 4; k++) {
                int nx = x + off[k][1], ny = y + off[k][0];
                if (!orc_in_range(nx, ny, w, h)) continue;
                orc_gaussian_fuse(&gd, &gs, depth[(size_t)ny * w + nx], sigma[(size_t)ny * w + nx]);
            }
            out[(size_t)y * w + x] = gd < 6.0f ? gd : 6.0f; /* cv::min(new_depth, 6.0) :178 */
        }
}

/* Mapper::update, src/map/mapper.cpp:76-137, on explicit arrays. */
int orc_mapper_update(float* depth, float* sigma, float* age, int w, int h, const float* obj_gray,
                      const float* const* born_gray, const float* const* born_gx,
                      const float* const* born_gy, const float* hist_xi, int n_hist,
                      const float obj_xi[6], const float rel_xi[6], const float K[9],
                      const orc_params* P, orc_rng* rng, uint8_t* mask, int* valid_update_out)
{
    int valid_update = 0;
    float T[16];
    orc_se3_exp(rel_xi, T); /* Transform::warp(xi, ...) :94 */
    if (mask) memset(mask, 0, (size_t)w * h);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            if (x < P->map_roi_x0 || x > P->map_roi_x1 || y < P->map_roi_y0 || y > P->map_roi_y1) continue; /* :90 */
            size_t id = (size_t)y * w + x;
            float d = depth[id];
            float wp[2];
            orc_warp_point(T, K, (float)x, (float)y, d, wp);
            int wx = orc_round(wp[0]), wy = orc_round(wp[1]); /* Point2f -> Point2i :94 */
            if (!orc_in_range(wx, wy, w, h)) continue;
            int a = (int)age[id]; /* :99 */
            if (a < 0 || a >= n_hist) return -1; /* m_history.at() would throw :176 */
            float pd = d - rel_xi[2]; /* :104 */
            float ps = sigma[id];
            float nb[6], r_xi[6];
            for (int i = 0; i < 6; i++) nb[i] = -hist_xi[a * 6 + i];
            orc_se3_concatenate(obj_xi, nb, r_xi); /* :107 */
            float nd, ns;
            orc_map_update_pixel(obj_gray, born_gray[a], born_gx[a], born_gy[a], w, h, r_xi, K, wx, wy, pd, ps, &nd, &ns);
            if (mask) mask[id] |= 1;
            if (nd > P->accept_depth_lo && nd < P->accept_depth_hi && ns > 0.0f && ns < P->accept_sigma_hi) { /* :122 */
                float gd = pd, gs = ps;
                if (mask) mask[id] |= 2;
                if (!orc_gaussian_update(&gd, &gs, nd, ns, rng)) {
                    age[id] = 0;
                    if (mask) mask[id] |= 8;
                } else {
                    valid_update++;
                    if (mask) mask[id] |= 4;
                }
                depth[id] = gd;
                sigma[id] = gs;
            }
        }
    if (valid_update_out) *valid_update_out = valid_update;
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* VisualOdometry — include/system/system.hpp:44-74 ; Mapper::estimate src/map/mapper.cpp:16-60 */
/* ------------------------------------------------------------------------------------------ */
struct orc_vo {
    float K[9];
    int levels, culls;
    orc_params P;
    orc_frame** hist; /* m_history (push_back order) */
    int n_hist, cap_hist;
    orc_frame* last;
    orc_rng rng; /* file-scope engine of gaussian.cpp:8 */
    int last_was_keyframe, last_valid_update;
    uint8_t* upd_mask;
    size_t upd_mask_n;
    orc_iter_log log[256]; /* Gauss-Newton iterations of the last frame's tracking */
    int n_log;
};

orc_vo* orc_vo_create(const float K[9], int levels, int culls, const orc_params* P)
{
    orc_vo* vo = (orc_vo*)calloc(1, sizeof(orc_vo));
    memcpy(vo->K, K, sizeof vo->K);
    vo->levels = levels; vo->culls = culls;
    if (P) vo->P = *P; else orc_params_default(&vo->P);
    orc_rng_seed(&vo->rng, 5489u);
    g_latest_id = -1;
    return vo;
}
void orc_vo_destroy(orc_vo* vo)
{
    if (!vo) return;
    if (vo->last) orc_frame_destroy(vo->last);
    for (int i = 0; i < vo->n_hist; i++) orc_frame_destroy(vo->hist[i]);
    free(vo->hist);
    free(vo->upd_mask);
    free(vo);
}
static void vo_push_history(orc_vo* vo, orc_frame* f) /* FrameHistory::setRefFrame frame.hpp:151-157 */
{
    if (vo->n_hist == vo->cap_hist) {
        vo->cap_hist = vo->cap_hist ? vo->cap_hist * 2 : 16;
        vo->hist = (orc_frame**)realloc(vo->hist, (size_t)vo->cap_hist * sizeof(orc_frame*));
    }
    f->refcount++;
    vo->hist[vo->n_hist++] = f;
}
orc_frame* orc_vo_ref_frame(orc_vo* vo) { return vo->n_hist ? vo->hist[vo->n_hist - 1] : NULL; }
orc_frame* orc_vo_last_frame(orc_vo* vo) { return vo->last; }
int orc_vo_history_size(const orc_vo* vo) { return vo->n_hist; }
int orc_vo_last_was_keyframe(const orc_vo* vo) { return vo->last_was_keyframe; }
int orc_vo_last_valid_update(const orc_vo* vo) { return vo->last_valid_update; }
uint64_t orc_vo_rng_draws(const orc_vo* vo) { return vo->rng.draws; }
const uint8_t* orc_vo_last_update_mask(const orc_vo* vo) { return vo->upd_mask; }
int orc_vo_last_log(const orc_vo* vo, orc_iter_log* out, int cap)
{
    int n = vo->n_log < cap ? vo->n_log : cap;
    memcpy(out, vo->log, (size_t)n * sizeof(orc_iter_log));
    return vo->n_log;
}

int orc_vo_odometrize(orc_vo* vo, const float* gray, int w, int h, const float* init_depth,
                      const float* init_sigma, float T_out[16])
{
    static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    orc_frame* frame = orc_frame_create(gray, NULL, NULL, w, h, vo->K, vo->levels, vo->culls); /* :47 */
    if (init_depth || init_sigma) orc_frame_set_depth_sigma_age(frame, init_depth, init_sigma, NULL);
    if (vo->last) orc_frame_destroy(vo->last);
    vo->last = frame;
    vo->last_was_keyframe = 0;
    vo->last_valid_update = 0;
    vo->n_log = 0;
    orc_frame* ref = orc_vo_ref_frame(vo);
    if (!ref) { /* :49-54 */
        vo_push_history(vo, frame);
        vo->last_was_keyframe = 1;
        memcpy(T_out, I16, sizeof I16);
        return 0;
    }
    /* Tracking :57-58 */
    float rel[6];
    vo->n_log = orc_track(frame, ref, &vo->P, rel, vo->log, 256, (double)vo->P.min_update, vo->P.min_residual);
    if (vo->n_log > 256) vo->n_log = 256;
    memcpy(frame->rel_xi, rel, sizeof rel); /* Frame::updateXi frame.cpp:7-14 */
    frame->ref = ref;
    orc_se3_concatenate(ref->xi, rel, frame->xi);

    /* Mapping: Mapper::estimate mapper.cpp:16-33 */
    orc_scene* rtop = &ref->sc[ref->levels - 1];
    orc_scene* ftop = &frame->sc[frame->levels - 1];
    const int fw = rtop->w, fh = rtop->h;
    int need_new = 0; /* needNewFrame :45-60 */
    if (orc_norm(frame->rel_xi, 3) > vo->P.min_movement) need_new = 1;
    else if (frame->id - ref->id >= vo->P.max_forward) need_new = 1;
    if (need_new) {
        size_t n = (size_t)fw * fh;
        float* d = (float*)malloc(n * sizeof(float));
        float* s = (float*)malloc(n * sizeof(float));
        float* a = (float*)malloc(n * sizeof(float));
        orc_propagate(rtop->depth, rtop->sigma, ref->age, frame->rel_xi, ftop->K, fw, fh, d, s, a); /* :62-74 */
        orc_frame_set_depth_sigma_age(frame, d, s, a);
        free(d); free(s); free(a);
        vo_push_history(vo, frame);
        vo->last_was_keyframe = 1;
    } else {
        int nh = vo->n_hist;
        const float** bg = (const float**)malloc(sizeof(float*) * nh);
        const float** bgx = (const float**)malloc(sizeof(float*) * nh);
        const float** bgy = (const float**)malloc(sizeof(float*) * nh);
        float* hxi = (float*)malloc(sizeof(float) * 6 * nh);
        for (int a = 0; a < nh; a++) { /* frame_history[age] = m_history.at(size-age-1) :176 */
            orc_frame* born = vo->hist[nh - a - 1];
            orc_scene* bt = &born->sc[born->levels - 1];
            bg[a] = bt->gray; bgx[a] = scene_gx(bt); bgy[a] = scene_gy(bt);
            memcpy(hxi + 6 * a, born->xi, 6 * sizeof(float));
        }
        if (vo->upd_mask_n != (size_t)fw * fh) {
            free(vo->upd_mask);
            vo->upd_mask = (uint8_t*)malloc((size_t)fw * fh);
            vo->upd_mask_n = (size_t)fw * fh;
        }
        int rc = orc_mapper_update(rtop->depth, rtop->sigma, ref->age, fw, fh, ftop->gray, bg, bgx, bgy, hxi, nh,
                                   frame->xi, frame->rel_xi, ftop->K, &vo->P, &vo->rng, vo->upd_mask,
                                   &vo->last_valid_update);
        free(bg); free(bgx); free(bgy); free(hxi);
        if (rc) return rc;
        frame_recull_depth_sigma(ref, 1, 1); /* ref->updateDepthSigma :135 */
    }
    /* regularize(history.getRefFrame()) :30,139-144 */
    {
        orc_frame* kf = orc_vo_ref_frame(vo);
        orc_scene* kt = &kf->sc[kf->levels - 1];
        float* nd = (float*)malloc((size_t)kt->w * kt->h * sizeof(float));
        orc_regularize(kt->depth, kt->sigma, kt->w, kt->h, nd);
        orc_frame_set_depth_sigma_age(kf, nd, NULL, NULL); /* updateDepth frame.cpp:56-61 */
        free(nd);
    }
    orc_se3_exp(frame->xi, T_out); /* :73 */
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* Draw — src/core/draw.cpp (headless: the BGR bytes cv::imshow would get)                    */
/* ------------------------------------------------------------------------------------------ */

/* cv::Mat::convertTo(CV_8UC1, a, b) on CV_32F: saturate_cast<uchar>(src * a + b) in float, rounded half to even (cvRound).
 * OpenCV's vector body uses v_fma where the CPU has FMA3 (cv2 4.13 here does), its scalar tail a separate multiply and add;
 * the two differ only for b != 0 (visualizeSigma).  The fused form is the one restated. */
uint8_t orc_convert_8u(float v, float a, float b)
{
    const float r = fmaf(v, a, b);
    if (!(r > -2147483648.0f && r < 2147483648.0f)) return 0; /* cvRound -> INT_MIN -> saturates to 0 */
    const long q = lrintf(r);
    return (uint8_t)(q < 0 ? 0 : q > 255 ? 255 : q);
}

/* cv::cvtColor(CV_8UC3, COLOR_HSV2BGR), H in [0, 180]: OpenCV 4.x HSV2RGB_b = float HSV2RGB with hscale 6/180 on
 * (h, s / 255, v / 255), tab[] with fused 1 - s h, results * 255 TRUNCATED (v_trunc in the vector body). */
void orc_hsv2bgr_8u(uint8_t H, uint8_t S, uint8_t V, uint8_t bgr[3])
{
    static const int sector_data[6][3] = {{1, 3, 0}, {1, 0, 2}, {3, 0, 1}, {0, 2, 1}, {0, 1, 3}, {2, 1, 0}};
    const float s = (float)S * (1.f / 255.f), v = (float)V * (1.f / 255.f);
    float h = (float)H * (6.f / 180.f);
    if (h >= 6.f) h -= 6.f;
    int sector = (int)floorf(h);
    h -= (float)sector;
    if ((unsigned)sector >= 6u) { sector = 0; h = 0.f; }
    float tab[4];
    tab[0] = v;
    tab[1] = v * (1.f - s);
    tab[2] = v * fmaf(-s, h, 1.f);
    tab[3] = v * fmaf(-s, 1.f - h, 1.f);
    for (int k = 0; k < 3; k++) {
        const float x = tab[sector_data[sector][k]] * 255.f;
        const int q = (int)x;
        bgr[k] = (uint8_t)(q < 0 ? 0 : q > 255 ? 255 : q);
    }
}

/* static_cast<unsigned char>(float) as x86 compiles it: cvttss2si, low byte */
static uint8_t uchar_cast(float v) { return (uint8_t)(orc_trunc(v) & 0xff); }

void orc_visualize(int kind, const float* src, const float* aux, int n, uint8_t* bgr)
{
    for (int i = 0; i < n; i++) {
        const float t = src[i];
        uint8_t* p = bgr + 3 * (size_t)i;
        switch (kind) {
        case 0: { /* draw.cpp:7-19 */
            p[0] = p[1] = p[2] = orc_convert_8u(t, 255.f, 0.f);
            if (!(t > ORC_INVALID)) p[0] = 255;
            break;
        }
        case 1: case 2: { /* draw.cpp:32-67 */
            if (t < 1e-6f) { p[0] = p[1] = p[2] = 0; break; }
            float x = (t - 0.70f) * 70.0f;
            x = x < 0.0f ? 0.0f : (180.0f < x ? 180.0f : x); /* std::clamp */
            const uint8_t hh = uchar_cast(x);
            uint8_t vv = 255;
            if (kind == 2) { const float sg = aux[i]; vv = uchar_cast(-500 * (sg < 0.5f ? sg : 0.5f) + 255); }
            orc_hsv2bgr_8u(hh, 255, vv, p);
            break;
        }
        case 3: { /* draw.cpp:21-29: cv::min(., 2.5f), cv::max(., 0.5f), convertTo(CV_8UC1, 100) */
            float x = t < 2.5f ? t : 2.5f; /* cv::min / cv::max with a scalar: std::min(a, b) = b < a ? b : a */
            x = x > 0.5f ? x : 0.5f;
            if (t != t) x = 0.5f;
            p[0] = p[1] = p[2] = orc_convert_8u(x, 100.f, 0.f);
            break;
        }
        case 4: p[0] = p[1] = p[2] = orc_convert_8u(t, -500.f, 255.f); break; /* draw.cpp:69-75 */
        case 5: { /* draw.cpp:77-91 */
            p[0] = p[1] = p[2] = 0;
            if (t <= ORC_INVALID) { p[0] = 255; break; }
            p[1] = uchar_cast(255 * (t > 0.0f ? t : 0.0f));
            p[2] = uchar_cast(-255 * (t < 0.0f ? t : 0.0f));
            break;
        }
        default: p[0] = p[1] = p[2] = orc_convert_8u(t, 10.f, 0.f); break; /* draw.cpp:94-100 */
        }
    }
}
