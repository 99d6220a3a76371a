/*
 * dvo_cuda.h — C ABI of libdvo_cuda.so, the B200 (sm_100a) implementation of the per-pixel hot path
 * of KYabuuchi/direct-visual-odometry (tracking + mapping).  Plain pointers and sizes only.
 *
 * The reference has no FFI/plugin registry: its boundary is the C++ class API in include/
 * (SURVEY.md §8-b).  Each entry point below names the reference interface it replaces
 * (file:line relative to the reference root).  The C++ classes in include/dvo/ keep the
 * reference's names/signatures and forward to these functions.
 *
 * Conventions
 *  - A context processes `n_streams` independent sequences in lock-step: every image argument is
 *    an array of n_streams images, stream-major ([stream][row][col], float32, row-major).  With
 *    n_streams == 1 the calls are exactly the reference's single-sequence operations.
 *  - Host pointers unless the name ends in _dev.  The library owns device memory and one CUDA
 *    stream per context; calls are asynchronous until an output pointer must be filled.
 *  - A context is bound to the CUDA device current at dvo_create and is not thread-safe.
 *  - Returns 0 on success, a negative dvo_status otherwise; never throws across the ABI.
 *  - There is NO CPU fallback: without a CUDA device dvo_create fails with DVO_ERR_CUDA.
 */
#ifndef DVO_CUDA_H
#define DVO_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DVO_MAX_LEVELS 8
#define DVO_INVALID (-2.0f) /* include/math/util.hpp:7 */

typedef enum dvo_status {
    DVO_OK = 0,
    DVO_ERR_CUDA = -1,     /* CUDA runtime error (see dvo_last_error) */
    DVO_ERR_ARG = -2,      /* bad argument */
    DVO_ERR_SLOT = -3,     /* bad / exhausted frame slot */
    DVO_ERR_HISTORY = -4,  /* age >= history size: std::vector::at would throw (frame.hpp:176) */
    DVO_ERR_RNG = -5,      /* mt19937 replay table cannot grow further (2^32 draws) */
    DVO_ERR_STATE = -6     /* call sequence error */
} dvo_status;

/* Geometry + the reference's compile-time constants (defaults = the reference's literals). */
typedef struct dvo_params {
    /* geometry: System::Frame ctor (include/system/frame.hpp:91-117) */
    int n_streams;      /* independent sequences per call (>=1) */
    int width, height;  /* input image size */
    int levels, culls;  /* Frame(gray, K, levels, culls): 3,2 in odometrize (system.hpp:47); 4,1 in test/track.cpp:24 */
    float K[9];         /* input intrinsics, row-major 3x3 */
    int max_frames;     /* frame-slot capacity */
    int max_keyframes;  /* per-stream history capacity (FrameHistory is never pruned, frame.hpp:179-185) */
    /* Tracker::track constants (src/track/tracker.cpp:16-19); the 200 ms budget is never reached */
    float min_residual; /* 5e-3 */
    float min_update;   /* 5e-4 */
    int max_iterations; /* 15 */
    /* Track::optimize (src/track/optimize.cpp:22-39,83) */
    int roi_level;                      /* 2 */
    int roi_x0, roi_x1, roi_y0, roi_y1; /* 20,140,20,100 inclusive */
    double min_depth;                   /* 0.20 (double literal) */
    float sigma_lo, sigma_hi;           /* 0.01, 0.5 */
    float step_default, step_level1, step_level2; /* 2.0, 1.5, 1.0 */
    /* Mapper (src/map/mapper.cpp:12-13,90,122) */
    int map_roi_x0, map_roi_x1, map_roi_y0, map_roi_y1; /* 16,144,12,108 inclusive */
    float min_movement; /* 0.02 */
    int max_forward;    /* 6 */
    float accept_depth_lo, accept_depth_hi, accept_sigma_hi; /* 0.2, 6.0, 0.5 */
    /* INITIAL length (draws) of the mt19937 replay table for Gaussian::update's reset (src/math/gaussian.cpp:8-9,22); the table
     * grows on demand (the reference's engine is unbounded) */
    int rng_table;      /* default 1<<22 */
    /* Extras that north_star names and the REFERENCE DOES NOT HAVE: 0 = off = the reference's behaviour (default).
     * min_gradient: semi-dense selection - a pixel enters Track::optimize only if gx^2 + gy^2 >= min_gradient^2 (the gradients
     * sampled at its warped position, fp32).  huber_delta: variance-normalised Huber weight - its row of A and entry of B
     * (optimize.cpp:87-89) are scaled by sqrt(w), w = 1 if |r| / sigma <= huber_delta else huber_delta / (|r| / sigma), sigma the
     * clamped sigma of optimize.cpp:83.  The oracle implements the same definition (orc_params), so parity holds with them on. */
    float huber_delta, min_gradient;
} dvo_params;

typedef struct dvo_ctx dvo_ctx;

/* One Gauss-Newton iteration record (what Tracker::track prints, src/track/tracker.cpp:56-61). */
typedef struct dvo_iter_log {
    int level, iteration, valid_pixels;
    float residual;
    float update[6];
    float xi_after[6];
} dvo_iter_log;

typedef enum dvo_image_kind {
    DVO_IMG_GRAY = 0, DVO_IMG_DEPTH = 1, DVO_IMG_SIGMA = 2, DVO_IMG_GRADX = 3, DVO_IMG_GRADY = 4,
    DVO_IMG_AGE = 5 /* finest level only */
} dvo_image_kind;

/* ---- context ------------------------------------------------------------------------------ */
int dvo_device_count(void);
int dvo_set_device(int device); /* cudaSetDevice for the calling thread; call before dvo_create */
void dvo_params_default(dvo_params* p); /* geometry 640x480, levels 3, culls 2, n_streams 1 */
int dvo_create(dvo_ctx** out, const dvo_params* p);
void dvo_destroy(dvo_ctx* ctx);
const char* dvo_last_error(const dvo_ctx* ctx);
int dvo_sync(dvo_ctx* ctx);
int dvo_level_size(const dvo_ctx* ctx, int level, int* w, int* h); /* Scene::cols/rows, frame.hpp:41-42 */
int dvo_level_K(const dvo_ctx* ctx, int level, float K[9]);        /* Convert::cullIntrinsic, convert.cpp:22-29 */
/* pinned host memory for zero-staging uploads / downloads */
void* dvo_host_alloc(size_t bytes);
void dvo_host_free(void* p);

/* ---- frames: System::Frame (include/system/frame.hpp:72-144, src/system/frame.cpp) ---------- */
/* Frame(gray, K, levels, culls): decimation pyramid (Convert::cullImage convert.cpp:7-20) +
 * gradients (Convert::gradiate convert.cpp:41-75) built on device.  gray: [n_streams][height][width]. */
int dvo_frame_create(dvo_ctx* ctx, int* slot, const float* gray);
int dvo_frame_create_dev(dvo_ctx* ctx, int* slot, const float* gray_dev);
/* The same from 8-BIT gray, the form the reference's loader holds right after cv::cvtColor(BGR2GRAY)
 * (src/core/loader.cpp:59-61,143-144): the device performs convertTo(CV_32FC1, 1.0/255.0) — for 8-bit input
 * OpenCV multiplies in fp32, (float)code * (float)(1.0/255.0) — and then Convert::cullImage, so a quarter of the
 * bytes cross PCIe.  gray8: [n_streams][height][width] uint8.  invalid_code (0..255, or -1 for none) marks pixels
 * that are INVALID (-2.0, include/math/util.hpp:7): the border cv::remap(BORDER_CONSTANT, INVALID) leaves
 * (loader.cpp:39-40) cannot be expressed in 8 bits otherwise. */
int dvo_frame_create_gray8(dvo_ctx* ctx, int* slot, const uint8_t* gray8, int invalid_code);
/* The whole of Loader::getNormalizedUndistortedImages (src/core/loader.cpp:32-42,55-63) on the device, from the 8-bit BGR (or
 * BGRA: the Kinect PNGs are 4-channel) frame cv::imread returns:
 *   cv::cvtColor(BGR2GRAY)            8-bit fixed point, (3735 B + 19235 G + 9798 R + 2^14) >> 15 (OpenCV 4's RGB2Gray<uchar>)
 *   convertTo(CV_32FC1, 1.0/255.0)    one fp32 multiply
 *   cv::remap(map_x, map_y, INTER_NEAREST, BORDER_CONSTANT, INVALID)   source pixel (cvRound(map_x), cvRound(map_y)), -2.0 outside
 *   Convert::cullImage                only the sampled pixels are ever computed
 * dvo_set_undistort_map takes the two CV_32FC1 maps ([height][width] each) exactly as cv::initUndistortRectifyMap produced them
 * (Loader::createUndistortMap, loader.cpp:15-30); NULL, NULL selects the identity (getNormalizedImages only).  The bgr8 entry
 * points need channels = 3 or 4; bgr8: [n_streams][height][width][channels] uint8. */
int dvo_set_undistort_map(dvo_ctx* ctx, const float* map_x, const float* map_y);
int dvo_frame_create_bgr8(dvo_ctx* ctx, int* slot, const uint8_t* bgr8, int channels);
int dvo_frame_release(dvo_ctx* ctx, int slot);
/* full_res != 0: arrays are input-sized and culled like the 6-arg Frame ctor (frame.hpp:91-106);
 * full_res == 0: arrays are finest-level sized (Frame::updateDepthSigmaAge, frame.cpp:47-54).
 * Any of depth/sigma/age may be NULL (left unchanged).  The coarser levels are re-culled. */
int dvo_frame_set_depth(dvo_ctx* ctx, int slot, const float* depth, const float* sigma, const float* age, int full_res);
int dvo_frame_download(dvo_ctx* ctx, int slot, int level, int kind, float* out);
int dvo_frame_set_xi(dvo_ctx* ctx, int slot, const float* xi /*[n_streams][6]*/, const float* rel_xi /*nullable*/);
/* ---- headless outputs -------------------------------------------------------------
 * Draw::visualize* (include/core/draw.hpp:9-23, src/core/draw.cpp:7-100) of one level of a frame slot, computed on the device:
 * the 8-bit BGR image cv::imshow would be handed (invalid pixels blue, depth as hue, ...).  bgr_out: [n_streams][h_l][w_l][3]. */
typedef enum dvo_vis_kind {
    DVO_VIS_GRAY = 0,        /* visualizeGray(gray)            draw.cpp:7-19  */
    DVO_VIS_DEPTH = 1,       /* visualizeDepth(depth)          draw.cpp:32-48 */
    DVO_VIS_DEPTH_SIGMA = 2, /* visualizeDepth(depth, sigma)   draw.cpp:50-67 */
    DVO_VIS_DEPTH_RAW = 3,   /* visualizeDepthRaw(depth)       draw.cpp:21-29 */
    DVO_VIS_SIGMA = 4,       /* visualizeSigma(sigma)          draw.cpp:69-75 */
    DVO_VIS_GRADX = 5,       /* visualizeGradient(gradX)       draw.cpp:77-91 */
    DVO_VIS_GRADY = 6,       /* visualizeGradient(gradY)                      */
    DVO_VIS_AGE = 7          /* visualizeAge(age), finest level draw.cpp:94-100 */
} dvo_vis_kind;
int dvo_visualize(dvo_ctx* ctx, int slot, int level, int kind, uint8_t* bgr_out);
/* The same functions with the reference's own argument: a raw float image on the host (test/update.cpp:19-27,
 * src/map/mapper.cpp:37-42 pass frame->gray(), frame->depth(), ...).  src / aux: [height][width]; aux = sigma, only for
 * DVO_VIS_DEPTH_SIGMA; for DVO_VIS_GRADX / GRADY src is the gradient image itself.  Needs no context; DVO_ERR_CUDA without a device. */
int dvo_visualize_image(int kind, const float* src, const float* aux, int width, int height, uint8_t* bgr_out);
/* Convert::inversePose (src/core/convert.cpp:31-39): R^T and -t (not -R^T t) in a 4x4 whose last row stays ZERO, as the reference
 * builds it; main.cpp:50-51 pushes that matrix on the trajectory it draws.  Host arithmetic (16 floats). */
int dvo_inverse_pose(const float T[16], float out[16]);
/* The trajectory main.cpp:50-53,76-86 hands to the GLFW drawer, as a file: one line per pose, `frame,t00,...,t33` (%.9g).
 * append != 0 adds to an existing file; first_frame numbers the first row. */
int dvo_write_trajectory_csv(const char* path, const float* T /*[n][16]*/, int n, int first_frame, int append);
/* 8-bit BGR image (dvo_visualize's layout, one stream) -> PNG file (RGB, stored deflate blocks: no zlib dependency). */
int dvo_write_png(const char* path, const uint8_t* bgr, int width, int height);

int dvo_frame_get_xi(dvo_ctx* ctx, int slot, float* xi, float* rel_xi /*nullable*/);

/* ---- Kinect ingest: Core::KinectLoader::getMappedImages (src/core/loader.cpp:90-147) + Transform::mapDepthtoGray
 * (src/core/transform.cpp:53-78) + the 6-argument Frame ctor (include/system/frame.hpp:91-106), on the device -------------------
 * The context's width x height is the DEPTH camera's image (512 x 424 for Kinect v2) and dvo_params.K its intrinsics.  Per
 * finest-level pixel:  depth = remap_nearest(depth16 * (float)(1.0/5000.0), depth maps, border 0);  |depth| < 1e-6 -> gray INVALID,
 * sigma 1.0;  otherwise  x_c = invT * backProject(depth_K, x, depth),  gray = Convert::getSubpixel(undistorted colour image,
 * project(rgb_K, x_c)) with sigma 0.1, where the undistorted colour image is never materialised: each of the four taps goes
 * through the colour remap index to the 8-bit BGR(A) pixel (cvtColor + convertTo as in dvo_frame_create_bgr8).
 * maps: CV_32FC1 as cv::initUndistortRectifyMap returns them (KinectLoader::createUndistortMap, loader.cpp:77-88); invT: row-major
 * 4x4 (Calibration::ExtrinsicParams::invT, params_struct.hpp:38-46). */
typedef struct dvo_kinect_params {
    int color_width, color_height, channels; /* 1920, 1080, 4 for the bundled data */
    float rgb_K[9];                          /* colour intrinsics */
    float invT[16];                          /* depth camera -> colour camera */
} dvo_kinect_params;
int dvo_kinect_configure(dvo_ctx* ctx, const dvo_kinect_params* kp, const float* rgb_map_x, const float* rgb_map_y,
                         const float* depth_map_x, const float* depth_map_y);
/* bgr8: [n_streams][color_height][color_width][channels] uint8; depth16: [n_streams][height][width] uint16 (1/5000 m). */
int dvo_frame_create_kinect(dvo_ctx* ctx, int* slot, const uint8_t* bgr8, const uint16_t* depth16);

/* ---- tracking ------------------------------------------------------------------------------- */
/* Transform::warpImage(xi, gray, depth, K) (src/core/transform.cpp:35-51) of ref_slot at `level`.
 * xi: [n_streams][6]; out: [n_streams][h][w]. */
int dvo_warp_image(dvo_ctx* ctx, int ref_slot, int level, const float* xi, float* out);
/* Track::optimize(Stuff{obj.at(level), ref.at(level), xi, level}) (src/track/optimize.cpp:10-99;
 * the warped image of Stuff is computed in-register).  Outputs per stream: upd[6], residual, valid
 * pixel count, and (optional) the contributing-pixel mask [n_streams][h][w]. */
int dvo_gn_step(dvo_ctx* ctx, int obj_slot, int ref_slot, int level, const float* xi, float* upd,
                float* residual, int* valid, uint8_t* mask);
/* Tracker::track(obj, ref) (src/track/tracker.cpp:22-85), device resident: ONE persistent kernel
 * launch runs every level and iteration (solve + se3 composition + stop rules on device), no host round
 * trip.  stop_update/stop_residual < 0 select the reference constants.  log: [n_streams][log_cap]. */
int dvo_track(dvo_ctx* ctx, int obj_slot, int ref_slot, float* xi_out, dvo_iter_log* log, int log_cap,
              int* n_iters, double stop_update, float stop_residual);

/* ---- mapping -------------------------------------------------------------------------------- */
/* Mapper::propagate (src/map/mapper.cpp:62-74) = Implement::propagate (implement.cpp:217-256) +
 * Frame::updateDepthSigmaAge on new_slot.  rel_xi: [n_streams][6]. */
int dvo_map_propagate(dvo_ctx* ctx, int ref_slot, int new_slot, const float* rel_xi);
/* Mapper::update (src/map/mapper.cpp:76-137): history_slots[age] = keyframe born `age` keyframes
 * ago (FrameHistory::operator[], frame.hpp:176); poses are the slots' xi (dvo_frame_set_xi).
 * Updates ref_slot's depth/sigma/age in place and re-culls.  mask (optional) per pixel:
 * 1 candidate | 2 accepted by the gate | 4 fused | 8 reset.  valid_updates: [n_streams]. */
int dvo_map_update(dvo_ctx* ctx, int obj_slot, int ref_slot, const int* history_slots, int n_hist,
                   int* valid_updates, uint8_t* mask);
/* Mapper::regularize (src/map/mapper.cpp:139-144) = Implement::regularize (implement.cpp:156-180)
 * + Frame::updateDepth. */
int dvo_map_regularize(dvo_ctx* ctx, int slot);
/* Map::Implement::update (include/map/implement.hpp:10-19, src/map/implement.cpp:182-214) for a batch
 * of n points per stream: epipolar search of obj_slot's pixel (px, py) in born_slot's finest-level
 * gray along the segment of prior (depth, sigma) under relative_xi, then depthEstimate + sigmaEstimate.
 * relative_xi: [n_streams][6]; px, py: [n_streams][n]; depth, sigma: [n_streams][n];
 * out_depth, out_sigma: [n_streams][n], (-1, -1) where the reference returns failure. */
int dvo_map_update_points(dvo_ctx* ctx, int obj_slot, int born_slot, const float* relative_xi, int n, const int* px,
                          const int* py, const float* depth, const float* sigma, float* out_depth, float* out_sigma);
/* mt19937 replay position (draws consumed so far by Gaussian::update resets), per stream. */
int dvo_rng_position(dvo_ctx* ctx, uint64_t* pos /*[n_streams]*/);

/* ---- se3 algebra: math::se3 (include/math/se3.hpp, src/math/se3.cpp:76-131) ---------------------
 * Evaluated by the device code the tracker itself uses (cv::Rodrigues restated, lazy MatExpr order),
 * so host-side pose bookkeeping (Frame::updateXi, frame.cpp:7-14) agrees with the kernels bit for bit.
 * n twists / matrices per call; xi: [n][6] (v, omega), T: [n][16] row-major 4x4. */
int dvo_se3_exp(dvo_ctx* ctx, const float* xi, float* T, int n);                        /* se3::exp   se3.cpp:76-99 */
int dvo_se3_log(dvo_ctx* ctx, const float* T, float* xi, int n);                        /* se3::log   se3.cpp:102-124 */
int dvo_se3_concatenate(dvo_ctx* ctx, const float* a, const float* b, float* out, int n); /* log(exp(a)*exp(b)) se3.cpp:127-131 */

/* ---- scalar boundary helpers: Transform::{transform, project, backProject} (include/core/transform.hpp:8-16, src/core/transform.cpp:7-28)
 * and math::Gaussian (include/math/gaussian.hpp:8-17, src/math/gaussian.cpp:12-50).  Host arithmetic in the reference's fp32 order,
 * compiled inside the library so that a caller's -ffast-math / FMA contraction cannot change a rounding.  Transform::warp is their
 * composition with se3::exp (include/dvo/core/transform.hpp). */
int dvo_transform_point(const float T[16], const float x[3], float out[3]);   /* R x + t of a row-major 4x4 */
int dvo_project_point(const float K[9], const float p[3], float out[2]);      /* (X fx / Z + cx, Y fy / Z + cy) */
int dvo_back_project_point(const float K[9], const float xy[2], float depth, float out[3]);
/* Gaussian::operator()(d, s): 1 = fused into (*depth, *sigma), 0 = refused (unchanged). */
int dvo_gaussian_fuse(float* depth, float* sigma, float d, float s);
/* Gaussian::update(d, s): like operator() but a refused observation RESETS the hypothesis to min(dist(engine), 4) / 0.5 from a
 * process-wide default-seeded std::mt19937 (gaussian.cpp:8-9,21-25); dvo_gaussian_reseed() restarts that engine. */
int dvo_gaussian_update(float* depth, float* sigma, float d, float s);
void dvo_gaussian_reseed(void);
/* Transform::mapDepthtoGray (include/core/transform.hpp:17-22, src/core/transform.cpp:53-78) on raw host images, computed on the
 * device: depth [h][w] (the depth camera), gray [gh][gw] (the colour camera), invT row-major 4x4 (depth -> colour);
 * mapped / sigma: [h][w].  Needs no context; DVO_ERR_CUDA without a device. */
int dvo_map_depth_to_gray(const float* depth, int w, int h, const float* gray, int gw, int gh, const float rgb_K[9],
                          const float depth_K[9], const float invT[16], float* mapped, float* sigma);

/* ---- mask compaction ----------------------------------------------------------- */
/* Row-major compaction of a byte mask on device: idx receives the linear indices of non-zero
 * entries in increasing order; returns their number in *count. */
int dvo_mask_compact(dvo_ctx* ctx, const uint8_t* mask, int n, int32_t* idx, int* count);
/* dvo_mask_compact for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * mask_dev ([n] bytes), idx_dev ([n] int32) and count_dev ([1] int) are DEVICE buffers of the caller (e.g. from dvo_dev_alloc, made
 * once, outside the loop); the result is the same as dvo_mask_compact's: idx_dev[0 .. *count_dev) = the linear indices of the
 * non-zero mask bytes in increasing order (entries past the count are left untouched).  The call allocates and frees nothing: its
 * scratch (one count per 256 mask bytes) is a static device array of DVO_NOALLOC_MAX_N / 256 * 4 = 128 KiB that every GPU
 * loading the library holds until exit; n > DVO_NOALLOC_MAX_N is DVO_ERR_ARG.  The kernels are enqueued on the context's stream
 * and the call returns without waiting: dvo_sync before reading the outputs.  ALL calls on a device share the scratch array: calls
 * on one context are ordered by its stream; calls from different contexts of the same device that overlap on the GPU race. */
#define DVO_NOALLOC_MAX_N (1 << 23) /* 8 388 608 mask bytes: one 3840x2160 level */
int dvo_mask_compact_noalloc(dvo_ctx* ctx, const uint8_t* mask_dev, int n, int32_t* idx_dev, int* count_dev);
/* Semi-dense pixel selection of one pyramid level of a frame by gradient threshold (north_star; no reference counterpart), selected
 * and compacted on the device: pixel (x, y) is kept iff its gray, Convert::gradiate x and y values are all valid and
 * gx^2 + gy^2 >= min_gradient^2.  idx ([n_streams][w_l h_l], row-major linear indices in increasing order, the first count[s]
 * entries of stream s are meaningful), count: [n_streams]. */
int dvo_select_pixels(dvo_ctx* ctx, int slot, int level, float min_gradient, int32_t* idx, int* count);

/* ---- full VO: System::VisualOdometry (include/system/system.hpp:12-104) --------------------- */
int dvo_vo_reset(dvo_ctx* ctx);
/* Pre-allocate the keyframe pool for n_keyframes keyframes per stream (otherwise entries are
 * allocated on first need inside dvo_vo_push, which synchronises the device). */
int dvo_vo_reserve(dvo_ctx* ctx, int n_keyframes);
/* VisualOdometry::odometrize(gray) (system.hpp:44-74) for every stream; the whole frame (pyramid,
 * <= levels*15 GN iterations, keyframe decision, propagate | update, regularize) is enqueued
 * without host synchronisation.  init_depth/init_sigma ([n_streams][h_top][w_top], nullable
 * after the first frame) replace the cv::randn prior of Scene's ctor (frame.hpp:18-21); the FIRST frame of a sequence must
 * bring init_depth (DVO_ERR_STATE otherwise: without a prior every pixel fails optimize's depth gate).
 * T_out ([n_streams][16], nullable) = se3::exp(frame->m_xi); passing NULL keeps the call async.
 * Per frame the library launches: ingest, k_track (all levels x iterations + keyframe decision + Implement::propagate + the pose),
 * k_map_update (Mapper::update; as a programmatic dependent of k_track, so finished streams are mapped while others still
 * iterate), its reset pass, k_regularize (+ Frame::updateDepth).  A device-side error of any stream (DVO_ERR_HISTORY: keyframe
 * capacity exhausted) is returned by the call that reads the poses back; that stream is tracked but no longer mapped. */
int dvo_vo_push(dvo_ctx* ctx, const float* gray, const float* init_depth, const float* init_sigma, float* T_out);
int dvo_vo_push_dev(dvo_ctx* ctx, const float* gray_dev, const float* init_depth, const float* init_sigma, float* T_out);
/* Pipelined form of dvo_vo_push for throughput: dvo_vo_submit stages `gray` (pinned host memory from
 * dvo_host_alloc, kept untouched until the matching collect) on a copy stream into one of two device
 * staging buffers and enqueues the frame behind it; it returns without waiting, so the next frame's
 * host->device copy overlaps this frame's kernels.  T_out ([n_streams][16], pinned, nullable) is filled
 * by the time the matching dvo_vo_collect returns.  At most 2 frames may be in flight; collect returns
 * them in submission order.  Results are identical to dvo_vo_push. */
int dvo_vo_submit(dvo_ctx* ctx, const float* gray, float* T_out);
int dvo_vo_collect(dvo_ctx* ctx); /* also returns a device-side error flagged up to the collected frame (when it had a T_out) */
/* dvo_vo_push / dvo_vo_submit from 8-bit gray frames (see dvo_frame_create_gray8): Loader::getNormalizedImages'
 * convertTo (loader.cpp:61) runs on the device.  Results are bit-identical to feeding the float frames
 * (float)code * (float)(1.0/255.0) to dvo_vo_push. */
int dvo_vo_push_gray8(dvo_ctx* ctx, const uint8_t* gray8, int invalid_code, const float* init_depth, const float* init_sigma,
                      float* T_out);
/* the same with the 8-bit frames already in device memory ([n_streams][height][width] uint8, e.g. from dvo_dev_alloc) */
int dvo_vo_push_gray8_dev(dvo_ctx* ctx, const uint8_t* gray8_dev, int invalid_code, const float* init_depth, const float* init_sigma,
                          float* T_out);
int dvo_vo_submit_gray8(dvo_ctx* ctx, const uint8_t* gray8, int invalid_code, float* T_out);
/* ... and from 8-bit BGR / BGRA frames through the undistortion map (see dvo_frame_create_bgr8).  The whole frame crosses PCIe
 * (3 or 4 bytes per input pixel), so this form is for hosts that cannot spare the cores for cvtColor + remap. */
int dvo_vo_push_bgr8(dvo_ctx* ctx, const uint8_t* bgr8, int channels, const float* init_depth, const float* init_sigma, float* T_out);
int dvo_vo_submit_bgr8(dvo_ctx* ctx, const uint8_t* bgr8, int channels, float* T_out);
/* Supply the relative pose (frame-from-keyframe twist, [n_streams][6]) the NEXT dvo_vo_push uses for
 * Frame::updateXi (frame.cpp:7-14) and mapping instead of its own tracking result; the tracker still
 * runs and its iterations stay readable through dvo_vo_last_log.  Used for teacher-forced parity
 * runs (the reference's iteration is chaotic at keyframe switches) and external pose
 * priors.  NULL cancels. */
int dvo_vo_force_pose(dvo_ctx* ctx, const float* rel_xi);
/* per-stream state of the last pushed frame */
int dvo_vo_state(dvo_ctx* ctx, float* xi /*[S][6]*/, float* rel_xi /*[S][6]*/, int* was_keyframe /*[S]*/,
                 int* n_keyframes /*[S]*/, int* valid_updates /*[S]*/, int* n_iters /*[S]*/);
/* cumulative per-stream work counters since dvo_vo_reset (any pointer may be NULL), [n_streams] each:
 * Gauss-Newton iterations, pixels swept by them (sum of level sizes), frames mapped by
 * Mapper::propagate and by Mapper::update. */
int dvo_vo_counters(dvo_ctx* ctx, uint64_t* gn_iters, uint64_t* gn_pixels, uint64_t* n_propagate, uint64_t* n_update);
/* pixels matched and fused by Mapper::update since dvo_vo_reset, per stream: the sum of its `valid_update` counts
 * (src/map/mapper.cpp:122-128,135) - the epipolar search's own unit of work (matched pixels / s). */
int dvo_vo_matched(dvo_ctx* ctx, uint64_t* matched /*[n_streams]*/);
/* Gauss-Newton iteration records of the last pushed frame's tracking (what Tracker::track prints,
 * src/track/tracker.cpp:56-61).  log: [n_streams][log_cap]; n_iters: [n_streams]. */
int dvo_vo_last_log(dvo_ctx* ctx, dvo_iter_log* log, int log_cap, int* n_iters);
/* slot of stream s's current keyframe / last frame (for dvo_frame_download) */
int dvo_vo_ref_slot(dvo_ctx* ctx, int stream, int* slot);
int dvo_vo_last_slot(dvo_ctx* ctx, int stream, int* slot);
/* first error flagged on device by any stream (history overflow, rng exhaustion), 0 if none */
int dvo_vo_error(dvo_ctx* ctx);

/* ---- synthetic inputs (no reference counterpart: the reference reads PNGs) ---- */
/* Renders frame `frame` of n_streams synthetic sequences (analytic textured box room, smooth 6-DoF
 * trajectory; stream s uses seed0 + s) at the context's input size and intrinsics, on device.
 * gray_dev: [n_streams][height][width] device floats in [0,1]; depth_dev (nullable): ground-truth z. */
int dvo_synth_render_dev(dvo_ctx* ctx, float* gray_dev, float* depth_dev, int frame, uint64_t seed0, float trans_step,
                         float rot_step_deg);
/* The same image as 8-bit codes (round-half-even of 255 g: what a camera / cv::imread + cvtColor hand the loader), stream s using
 * seed0 + s * seed_stride (sharded sequence ids).  gray8_dev: [n_streams][height][width] device bytes. */
int dvo_synth_render_gray8_dev(dvo_ctx* ctx, uint8_t* gray8_dev, float* depth_dev, int frame, uint64_t seed0, uint64_t seed_stride,
                               float trans_step, float rot_step_deg);
/* Ground truth of that renderer: the camera-to-world pose of `frame` of the sequence with seed `seed` (row-major 4x4), and the
 * scene itself (out: fu, fv, phase, amplitude as [5 planes][8 waves] each, then the 5 checker frequencies = 165 floats), so that
 * tests can re-render a frame independently and benchmarks can report the trajectory error against ground truth (C3). */
int dvo_synth_pose(dvo_ctx* ctx, int frame, uint64_t seed, float trans_step, float rot_step_deg, float* T_wc /*[16]*/);
int dvo_synth_scene(dvo_ctx* ctx, uint64_t seed, float* out, int cap);
/* plain device memory for inputs that stay resident (benchmarks), and copies on the context's stream */
void* dvo_dev_alloc(size_t bytes);
void dvo_dev_free(void* p);
int dvo_copy_to_host(dvo_ctx* ctx, void* host, const void* dev, size_t bytes);   /* synchronises */
int dvo_copy_to_dev(dvo_ctx* ctx, void* dev, const void* host, size_t bytes);    /* synchronises */

/* ---- measurement helpers (CUDA events on the context's stream) -------------------------------- */
int dvo_timer_start(dvo_ctx* ctx);
int dvo_timer_stop(dvo_ctx* ctx, float* ms); /* synchronises */
/* any number of marks on the context's stream (id 0..65535) and the device time between two of them */
int dvo_event_record(dvo_ctx* ctx, int id);
int dvo_event_elapsed(dvo_ctx* ctx, int id0, int id1, float* ms); /* synchronises on id1 */
/* number of kernels this context has launched since creation */
uint64_t dvo_launch_count(const dvo_ctx* ctx);
/* Device time per kernel between kernel_timing(1, ...) and kernel_timing(0, out, cap, &n): CUDA events
 * are recorded around every launch of the hot-path kernels on the context's stream while enabled
 * (this serialises nothing but adds two event records per launch, so use it in a separate pass). */
typedef struct dvo_kernel_time {
    char name[32];
    float total_ms;
    int launches;
} dvo_kernel_time;
int dvo_kernel_timing(dvo_ctx* ctx, int enable, dvo_kernel_time* out, int cap, int* n_out);
/* the tracking kernel's share of the above (k_track: one launch = Tracker::track of every stream) */
int dvo_gn_timing(dvo_ctx* ctx, int enable, float* total_ms, int* launches);

#ifdef __cplusplus
}
#endif
#endif
