/*
 * dvo_oracle.h — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the per-pixel hot path of KYabuuchi/direct-visual-odometry
 * (tracking: warp + Gauss-Newton step; mapping: epipolar update, propagate, regularize),
 * executed serially in row-major order with IEEE-strict fp32 (build: -O2 -ffp-contract=off).
 * Every function cites the reference file:line it follows (paths relative to the reference root).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libdvo_cuda.so) never links or calls it.
 *
 * Parity pinning: the reference's own tests assert nothing, so this restatement is
 * pinned (a) against the reference's own translation units compiled over an OpenCV-API shim
 * (oracle/_ref, see oracle/Makefile) and (b) against opencv-python 4.13 known answers for the
 * third-party pieces (Rodrigues, norm, solve(DECOMP_SVD)); see tests/test_oracle_*.py.
 */
#ifndef DVO_ORACLE_H
#define DVO_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_INVALID (-2.0f) /* include/math/util.hpp:7 */
#define ORC_EPSILON (1e-6f) /* include/math/util.hpp:6 */

/* Compile-time constants of the reference gathered in one POD (defaults = the literals). */
typedef struct orc_params {
    /* Tracker::track — src/track/tracker.cpp:16-19 */
    float min_residual;   /* 5e-3 */
    float min_update;     /* 5e-4 */
    int max_iterations;   /* 15   */
    /* Track::optimize — src/track/optimize.cpp:22-39,83 */
    int roi_level;                       /* 2: pyramid index the GN ROI applies to */
    int roi_x0, roi_x1, roi_y0, roi_y1;  /* 20,140,20,100 (inclusive) */
    double min_depth;                    /* 0.20 (double literal, float promoted) */
    float sigma_lo, sigma_hi;            /* 0.01, 0.5 */
    float step_default, step_level1, step_level2; /* 2.0, 1.5, 1.0 */
    /* Mapper — src/map/mapper.cpp:12-13,90,122 */
    int map_roi_x0, map_roi_x1, map_roi_y0, map_roi_y1; /* 16,144,12,108 (inclusive) */
    float min_movement; /* 0.02 */
    int max_forward;    /* 6 */
    float accept_depth_lo, accept_depth_hi, accept_sigma_hi; /* 0.2, 6.0, 0.5 */
    /* Extras north_star names that the REFERENCE DOES NOT HAVE; 0 = off = the reference's behaviour.  Same definition
     * on the GPU (dvo_params.huber_delta / min_gradient): a pixel enters Track::optimize only if gx^2 + gy^2 >= min_gradient^2
     * (sampled gradients, fp32), and its row of A and entry of B are scaled by sqrt(w), w = 1 if |r| / sigma <= huber_delta else
     * huber_delta / (|r| / sigma), sigma the clamped sigma of optimize.cpp:83. */
    float huber_delta, min_gradient;
} orc_params;

void orc_params_default(orc_params* p);

/* ---- math (src/math/se3.cpp) ---- */
void orc_rodrigues(const float w[3], float R[9]);
void orc_so3_log(const float R[9], float w[3]);
void orc_se3_exp(const float xi[6], float T[16]);
void orc_se3_log(const float T[16], float xi[6]);
void orc_se3_concatenate(const float a[6], const float b[6], float out[6]);
double orc_norm(const float* v, int n); /* cv::norm(Mat1f) NORM_L2 */

/* ---- gaussian (src/math/gaussian.cpp) ---- */
typedef struct orc_rng { uint32_t mt[624]; int idx; uint64_t draws; } orc_rng;
void orc_rng_seed(orc_rng* r, uint32_t seed); /* std::mt19937 default seed = 5489 */
uint32_t orc_rng_next_u32(orc_rng* r);
float orc_rng_dist(orc_rng* r);               /* uniform_real_distribution<float>(2.0,0.5) */
float orc_dist_from_u32(uint32_t raw);
int orc_gaussian_update(float* depth, float* sigma, float d, float s, orc_rng* r);
int orc_gaussian_fuse(float* depth, float* sigma, float d, float s);

/* ---- core (src/core/transform.cpp, src/core/convert.cpp) ---- */
void orc_warp_point(const float T[16], const float K[9], float x, float y, float depth, float out[2]);
float orc_get_pixel(const float* img, int w, int h, int x, int y);
float orc_get_subpixel(const float* img, int w, int h, float px, float py);
float orc_get_subpixel_dense(const float* img, int w, int h, float px, float py);
void orc_cull_image(const float* src, int w, int h, int times, float* dst); /* dst (w>>t)x(h>>t) */
void orc_cull_intrinsic(const float K[9], int times, float out[9]);
void orc_gradiate(const float* gray, int w, int h, int is_x, float* out);
void orc_warp_image(const float xi[6], const float* gray, const float* depth, const float K[9],
                    int w, int h, float* out);
void orc_inverse_pose(const float T[16], float out[16]);

/* ---- track (src/track/optimize.cpp, tracker.cpp) ---- */
int orc_solve_svd(const float* A, const float* b, int m, int n, float* x); /* cv::solve DECOMP_SVD */
/* One Gauss-Newton step.  mask (optional, w*h bytes) receives the contributing-pixel mask. */
void orc_optimize(const float* obj_gray, const float* warped_gray, const float* ref_depth,
                  const float* ref_sigma, const float* grad_x, const float* grad_y,
                  const float K[9], const float xi[6], int w, int h, int level_index,
                  const orc_params* P, float upd[6], float* residual, int* valid_pixels,
                  uint8_t* mask);
void orc_optimize_system(const float* obj_gray, const float* warped_gray, const float* ref_depth,
                  const float* ref_sigma, const float* grad_x, const float* grad_y,
                  const float K[9], const float xi[6], int w, int h, int level_index,
                  const orc_params* P, float upd[6], float* residual, int* valid_pixels,
                  uint8_t* mask, float* sys_A, float* sys_B);

typedef struct orc_iter_log {
    int level, iteration, valid_pixels;
    float residual;
    float update[6];
    float xi_before[6];
    float xi_after[6];
} orc_iter_log;

/* ---- frames / system (include/system/frame.hpp, src/system/frame.cpp, include/system/system.hpp) ---- */
typedef struct orc_frame orc_frame;
typedef struct orc_vo orc_vo;

orc_frame* orc_frame_create(const float* gray, const float* depth, const float* sigma, int w, int h,
                            const float K[9], int levels, int culls);
void orc_frame_destroy(orc_frame* f);
int orc_frame_level_size(const orc_frame* f, int level, int* w, int* h);
/* which: 0 gray 1 depth 2 sigma 3 gradX 4 gradY 5 age(finest only) ; returns pointer to internal data */
const float* orc_frame_image(orc_frame* f, int level, int which);
void orc_frame_K(const orc_frame* f, int level, float K[9]);
void orc_frame_set_depth_sigma_age(orc_frame* f, const float* depth, const float* sigma, const float* age);
void orc_frame_get_xi(const orc_frame* f, float xi[6], float rel_xi[6]);

/* Tracker::track (time budget treated as never reached). log may be NULL; returns #iterations. */
int orc_track(orc_frame* obj, orc_frame* ref, const orc_params* P, float xi_out[6],
              orc_iter_log* log, int log_cap, double stop_update, float stop_residual);

/* ---- map (src/map/implement.cpp, mapper.cpp) ---- */
void orc_map_update_pixel(const float* obj_gray, const float* ref_gray, const float* ref_gradx,
                          const float* ref_grady, int w, int h, const float relative_xi[6],
                          const float K[9], int px, int py, float depth, float sigma,
                          float* new_depth, float* new_sigma);
void orc_propagate(const float* ref_depth, const float* ref_sigma, const float* ref_age,
                   const float xi[6], const float K[9], int w, int h,
                   float* depth, float* sigma, float* age);
void orc_regularize(const float* depth, const float* sigma, int w, int h, float* out);

/* ---- full VO (include/system/system.hpp:44-74) ---- */
orc_vo* orc_vo_create(const float K[9], int levels, int culls, const orc_params* P);
void orc_vo_destroy(orc_vo* vo);
/* init_depth/init_sigma (base-level size, may be NULL except for frame 0) stand in for the
 * cv::randn prior of Scene's ctor (include/system/frame.hpp:18-21), see SURVEY H6. */
int orc_vo_odometrize(orc_vo* vo, const float* gray, int w, int h, const float* init_depth,
                      const float* init_sigma, float T_out[16]);
orc_frame* orc_vo_ref_frame(orc_vo* vo);
orc_frame* orc_vo_last_frame(orc_vo* vo);
int orc_vo_history_size(const orc_vo* vo);
int orc_vo_last_was_keyframe(const orc_vo* vo);
int orc_vo_last_valid_update(const orc_vo* vo);
uint64_t orc_vo_rng_draws(const orc_vo* vo);
/* masks of the last Mapper::update: 1 candidate (entered Implement::update), 2 accepted gate,
 * 4 fused ok, 8 reset.  Size = finest w*h. */
const uint8_t* orc_vo_last_update_mask(const orc_vo* vo);
/* Gauss-Newton iteration records of the last frame's tracking; returns their number */
int orc_vo_last_log(const orc_vo* vo, orc_iter_log* out, int cap);
/* Mapper::update as a standalone op (src/map/mapper.cpp:76-137) on explicit arrays.
 * born_* are arrays of n_hist pointers indexed by age; hist_xi is n_hist x 6 (index = age). */
int orc_mapper_update(float* depth, float* sigma, float* age, int w, int h, const float* obj_gray,
                      const float* const* born_gray, const float* const* born_gx,
                      const float* const* born_gy, const float* hist_xi, int n_hist,
                      const float obj_xi[6], const float rel_xi[6], const float K[9],
                      const orc_params* P, orc_rng* rng, uint8_t* mask, int* valid_update);

/* Draw::visualize* (src/core/draw.cpp:7-100): float image -> 8-bit BGR, n pixels.  kind: 0 visualizeGray, 1 visualizeDepth(depth),
 * 2 visualizeDepth(depth, sigma), 3 visualizeDepthRaw, 4 visualizeSigma, 5 visualizeGradient, 6 visualizeAge.  `aux` = sigma for
 * kind 2, else unused.  OpenCV's convertTo(CV_8UC1, a, b) and 8-bit COLOR_HSV2BGR are restated from OpenCV 4.x (third party, not
 * in /root/reference) and pinned against cv2 4.13 by tests/golden/make_draw.py. */
void orc_hsv2bgr_8u(uint8_t h, uint8_t s, uint8_t v, uint8_t bgr[3]);
uint8_t orc_convert_8u(float v, float a, float b);
void orc_visualize(int kind, const float* src, const float* aux, int n, uint8_t* bgr);

#ifdef __cplusplus
}
#endif
#endif
