// oracle/ref_capi.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A plain-C entry layer over the REFERENCE'S OWN classes and functions (KYabuuchi/direct-visual-
// odometry), whose translation units are compiled unmodified from /root/reference against the
// OpenCV-API shim in oracle/shim (see oracle/Makefile, target _ref/libdvo_ref.so).  It exists so
// that tests/ can (1) pin the plain-C restatement oracle/dvo_oracle.c against the reference code
// itself and (2) generate the golden vectors under tests/golden/.  Nothing here re-implements the
// hot path: every function forwards to Transform:: / Convert:: / Track:: / Map:: / System:: / math::.
//
// Determinism: forEach is serial row-major (shim), build flags are IEEE-strict, and
// shim/core/timer.hpp shadows the reference's Timer so that Tracker::track's wall-clock stop rule
// (src/track/tracker.cpp:70) never fires.
#include <opencv2/opencv.hpp>

#include <cstdio>
#include <iomanip>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

// read-only access to VisualOdometry::m_history for the map read-back (layout is unchanged)
#define private public
#include "system/system.hpp"
#undef private
#include "core/convert.hpp"
#include "core/transform.hpp"
#include "map/implement.hpp"
#include "map/mapper.hpp"
#include "math/math.hpp"
#include "track/optimize.hpp"
#include "track/tracker.hpp"

// src/math/gaussian.cpp keeps its std::mt19937 in an anonymous namespace; including the source file
// here (instead of compiling it on its own) is the only way to re-seed it between test cases.
#include "src/math/gaussian.cpp"

// Draw:: is display only (src/core/draw.cpp is not compiled: no highgui on this box)
namespace Draw
{
cv::Mat visualizeGray(const cv::Mat1f& m) { return m; }
cv::Mat visualizeDepth(const cv::Mat1f& m) { return m; }
cv::Mat visualizeDepth(const cv::Mat1f& m, const cv::Mat1f&) { return m; }
cv::Mat visualizeDepthRaw(const cv::Mat1f& m) { return m; }
cv::Mat visualizeGradient(const cv::Mat1f& m) { return m; }
cv::Mat visualizeSigma(const cv::Mat1f& m) { return m; }
cv::Mat visualizeAge(const cv::Mat1f& m) { return m; }
cv::Mat merge(std::vector<cv::Mat>& v) { return v.empty() ? cv::Mat() : v[0]; }
void showImage(const std::string&, std::vector<cv::Mat>&) {}
} // namespace Draw

namespace
{
cv::Mat1f wrap(const float* p, int w, int h) // deep copy into a Mat1f
{
    cv::Mat1f m(h, w);
    std::memcpy(m.data, p, sizeof(float) * (size_t)w * h);
    return m;
}
cv::Mat1f vec(const float* p, int n)
{
    cv::Mat1f m(n, 1);
    for (int i = 0; i < n; i++) m(i) = p[i];
    return m;
}
void put(const cv::Mat& m, float* out)
{
    for (int y = 0; y < m.rows; y++)
        for (int x = 0; x < m.cols; x++) *out++ = m.data[(size_t)y * m.step + x];
}

// The reference reports through std::cout only.  Captured at 9 significant digits every float
// round-trips exactly, so the per-iteration records of Tracker::track (tracker.cpp:56-61) and
// "valid update" of Mapper::update (mapper.cpp:136) are read back from its own prints.
struct Capture {
    std::ostringstream ss;
    std::streambuf* old;
    std::streamsize prec;
    Capture() : old(std::cout.rdbuf(ss.rdbuf())), prec(std::cout.precision(9)) {}
    ~Capture() { std::cout.rdbuf(old); std::cout.precision(prec); }
};
} // namespace

extern "C" {

typedef struct ref_iter_log {
    int level, iteration, valid_pixels;
    float residual;
    float update[6];
} ref_iter_log;

static int parse_track_log(const std::string& text, ref_iter_log* log, int cap, int* valid_update)
{
    std::istringstream in(text);
    std::string line;
    int level = -1, n = 0;
    if (valid_update) *valid_update = -1;
    while (std::getline(in, line)) {
        int a, b, c;
        if (std::sscanf(line.c_str(), "LEVEL: %d ROW: %d COL: %d", &a, &b, &c) == 3) level = a;
        ref_iter_log L;
        double nrm;
        if (std::sscanf(line.c_str(), "itr: %d r: %f upd: [%f, %f, %f, %f, %f, %f] norm: %lf rows : %d", &L.iteration, &L.residual,
                        &L.update[0], &L.update[1], &L.update[2], &L.update[3], &L.update[4], &L.update[5], &nrm, &L.valid_pixels) == 10) {
            L.level = level;
            if (log && n < cap) log[n] = L;
            n++;
        }
        if (valid_update && std::sscanf(line.c_str(), "valid update: %d pixel", &a) == 1) *valid_update = a;
    }
    return n;
}

// ---- math (src/math/se3.cpp) -----------------------------------------------------------------
void ref_so3_exp(const float w[3], float R[9]) { put(math::so3::exp(vec(w, 3)), R); }
void ref_so3_log(const float R[9], float w[3]) { put(math::so3::log(wrap(R, 3, 3)), w); }
void ref_se3_exp(const float xi[6], float T[16]) { put(math::se3::exp(vec(xi, 6)), T); }
void ref_se3_log(const float T[16], float xi[6]) { put(math::se3::log(wrap(T, 4, 4)), xi); }
void ref_se3_concatenate(const float a[6], const float b[6], float out[6]) { put(math::se3::concatenate(vec(a, 6), vec(b, 6)), out); }
double ref_norm(const float* v, int n) { return cv::norm(vec(v, n)); }

// ---- gaussian (src/math/gaussian.cpp) --------------------------------------------------------
void ref_rng_reset(void) { math::engine.seed(std::mt19937::default_seed); math::dist.reset(); }
int ref_gaussian_update(float* depth, float* sigma, float d, float s)
{
    math::Gaussian g(*depth, *sigma);
    const bool ok = g.update(d, s);
    *depth = g.depth; *sigma = g.sigma;
    return ok ? 1 : 0;
}
int ref_gaussian_fuse(float* depth, float* sigma, float d, float s)
{
    math::Gaussian g(*depth, *sigma);
    const bool ok = g(d, s);
    *depth = g.depth; *sigma = g.sigma;
    return ok ? 1 : 0;
}

// ---- core (src/core/transform.cpp, src/core/convert.cpp) --------------------------------------
void ref_warp_point(const float xi[6], const float K[9], float x, float y, float depth, float out[2])
{
    const cv::Point2f p = Transform::warp(vec(xi, 6), cv::Point2f(x, y), depth, wrap(K, 3, 3));
    out[0] = p.x; out[1] = p.y;
}
float ref_get_pixel(const float* img, int w, int h, int x, int y) { return Convert::getPixel(wrap(img, w, h), cv::Point2i(x, y)); }
float ref_get_subpixel(const float* img, int w, int h, float px, float py) { return Convert::getSubpixel(wrap(img, w, h), cv::Point2f(px, py)); }
float ref_get_subpixel_dense(const float* img, int w, int h, float px, float py)
{
    return Convert::getSubpixelFromDense(wrap(img, w, h), cv::Point2f(px, py));
}
// batched variants: one image, many points
void ref_get_subpixel_many(const float* img, int w, int h, const float* pts, int n, int dense, float* out)
{
    const cv::Mat1f m = wrap(img, w, h);
    for (int i = 0; i < n; i++) {
        const cv::Point2f p(pts[2 * i], pts[2 * i + 1]);
        out[i] = dense ? Convert::getSubpixelFromDense(m, p) : Convert::getSubpixel(m, p);
    }
}
void ref_cull_image(const float* src, int w, int h, int times, float* dst) { put(Convert::cullImage(wrap(src, w, h), times), dst); }
void ref_cull_intrinsic(const float K[9], int times, float out[9]) { put(Convert::cullIntrinsic(wrap(K, 3, 3), times), out); }
void ref_gradiate(const float* gray, int w, int h, int is_x, float* out) { put(Convert::gradiate(wrap(gray, w, h), is_x != 0), out); }
void ref_warp_image(const float xi[6], const float* gray, const float* depth, const float K[9], int w, int h, float* out)
{
    put(Transform::warpImage(vec(xi, 6), wrap(gray, w, h), wrap(depth, w, h), wrap(K, 3, 3)), out);
}
void ref_inverse_pose(const float T[16], float out[16]) { put(Convert::inversePose(wrap(T, 4, 4)), out); }

// ---- track (src/track/optimize.cpp) -----------------------------------------------------------
// Track::optimize(Stuff{obj, ref, xi, level_index}); Stuff computes warped_gray and the gradients itself.
void ref_optimize(const float* obj_gray, const float* ref_gray, const float* ref_depth, const float* ref_sigma, const float K[9],
                  const float xi[6], int w, int h, int level_index, float upd[6], float* residual, int* valid_pixels)
{
    const cv::Mat1f Km = wrap(K, 3, 3);
    const cv::Mat1f rd = wrap(ref_depth, w, h), rs = wrap(ref_sigma, w, h);
    auto obj = std::make_shared<System::Scene>(wrap(obj_gray, w, h), rd, rs, Km);
    auto ref = std::make_shared<System::Scene>(wrap(ref_gray, w, h), rd, rs, Km);
    Track::Stuff stuff(obj, ref, vec(xi, 6), level_index);
    const Track::Outcome o = Track::optimize(stuff);
    put(o.xi_update, upd);
    *residual = o.residual;
    *valid_pixels = o.valid_pixels;
}

// ---- frames (include/system/frame.hpp, src/system/frame.cpp) ----------------------------------
typedef std::shared_ptr<System::Frame> pFrame;
struct ref_frame { pFrame f; };

ref_frame* ref_frame_create(const float* gray, const float* depth, const float* sigma, int w, int h, const float K[9], int levels, int culls)
{
    ref_frame* r = new ref_frame;
    if (depth && sigma)
        r->f = std::make_shared<System::Frame>(wrap(gray, w, h), wrap(depth, w, h), wrap(sigma, w, h), wrap(K, 3, 3), levels, culls);
    else
        r->f = std::make_shared<System::Frame>(wrap(gray, w, h), wrap(K, 3, 3), levels, culls);
    return r;
}
void ref_frame_destroy(ref_frame* r) { delete r; }
int ref_frame_level_size(const ref_frame* r, int level, int* w, int* h)
{
    if (level < 0 || level >= r->f->levels) return -1;
    *w = r->f->at(level)->cols; *h = r->f->at(level)->rows;
    return 0;
}
// which: 0 gray 1 depth 2 sigma 3 gradX 4 gradY 5 age (finest only)
int ref_frame_image(ref_frame* r, int level, int which, float* out)
{
    if (level < 0 || level >= r->f->levels) return -1;
    auto sc = r->f->at(level);
    switch (which) {
    case 0: put(sc->gray(), out); break;
    case 1: put(sc->depth(), out); break;
    case 2: put(sc->sigma(), out); break;
    case 3: put(sc->gradX(), out); break;
    case 4: put(sc->gradY(), out); break;
    case 5: put(r->f->age(), out); break;
    default: return -1;
    }
    return 0;
}
void ref_frame_K(const ref_frame* r, int level, float K[9]) { put(r->f->at(level)->K(), K); }
void ref_frame_set_depth_sigma_age(ref_frame* r, const float* depth, const float* sigma, const float* age)
{
    const int w = r->f->top()->cols, h = r->f->top()->rows;
    const cv::Mat1f d = depth ? wrap(depth, w, h) : cv::Mat1f(r->f->depth().clone());
    const cv::Mat1f s = sigma ? wrap(sigma, w, h) : cv::Mat1f(r->f->sigma().clone());
    if (age) r->f->updateDepthSigmaAge(d, s, wrap(age, w, h));
    else r->f->updateDepthSigma(d, s);
}
void ref_frame_get_xi(const ref_frame* r, float xi[6], float rel_xi[6])
{
    if (xi) put(r->f->m_xi, xi);
    if (rel_xi) put(r->f->m_relative_xi, rel_xi);
}
void ref_frame_set_xi(ref_frame* r, const float xi[6], const float rel_xi[6], ref_frame* ref)
{
    if (xi) r->f->m_xi = vec(xi, 6);
    if (rel_xi) r->f->m_relative_xi = vec(rel_xi, 6);
    if (ref) r->f->m_ref_frame = ref->f;
}

// Tracker::track (src/track/tracker.cpp:22-85); returns the number of GN iterations
int ref_track(ref_frame* obj, ref_frame* ref, float xi_out[6], ref_iter_log* log, int log_cap)
{
    Track::Tracker tracker;
    cv::Mat1f xi;
    std::string text;
    {
        Capture cap;
        xi = tracker.track(obj->f, ref->f);
        text = cap.ss.str();
    }
    put(xi, xi_out);
    return parse_track_log(text, log, log_cap, nullptr);
}

// ---- map (src/map/implement.cpp, src/map/mapper.cpp) ------------------------------------------
void ref_map_update_pixel(const float* obj_gray, const float* ref_gray, const float* ref_gradx, const float* ref_grady, int w, int h,
                          const float relative_xi[6], const float K[9], int px, int py, float depth, float sigma, float* new_depth,
                          float* new_sigma)
{
    auto [d, s] = Map::Implement::update(wrap(obj_gray, w, h), wrap(ref_gray, w, h), wrap(ref_gradx, w, h), wrap(ref_grady, w, h),
                                         vec(relative_xi, 6), wrap(K, 3, 3), cv::Point2i(px, py), depth, sigma);
    *new_depth = d; *new_sigma = s;
}
void ref_propagate(const float* ref_depth, const float* ref_sigma, const float* ref_age, const float xi[6], const float K[9], int w, int h,
                   float* depth, float* sigma, float* age)
{
    auto [d, s, a] = Map::Implement::propagate(wrap(ref_depth, w, h), wrap(ref_sigma, w, h), wrap(ref_age, w, h), vec(xi, 6), wrap(K, 3, 3));
    put(d, depth); put(s, sigma); put(a, age);
}
void ref_regularize(const float* depth, const float* sigma, int w, int h, float* out)
{
    put(Map::Implement::regularize(wrap(depth, w, h), wrap(sigma, w, h)), out);
}
// Mapper::update(history, obj) (src/map/mapper.cpp:76-137).  hist[age] = keyframe born `age`
// keyframes ago (hist[0] = the current keyframe, which receives the update); obj's m_xi,
// m_relative_xi and m_ref_frame must have been set (ref_frame_set_xi).  Returns "valid update".
int ref_mapper_update(ref_frame* const* hist, int n_hist, ref_frame* obj)
{
    System::FrameHistory history;
    std::string text;
    int valid = -1;
    {
        Capture cap;
        for (int a = n_hist - 1; a >= 0; a--) history.setRefFrame(hist[a]->f);
        Map::Mapper mapper;
        mapper.update(history, obj->f);
        text = cap.ss.str();
    }
    parse_track_log(text, nullptr, 0, &valid);
    return valid;
}

// ---- full VO (include/system/system.hpp) ------------------------------------------------------
struct ref_vo {
    std::unique_ptr<System::VisualOdometry> vo;
    pFrame last;
    int last_was_keyframe = 0, last_valid_update = -1;
    std::vector<ref_iter_log> log;
    ref_frame ref_view, last_view;
};
ref_vo* ref_vo_create(const float K[9])
{
    ref_vo* v = new ref_vo;
    Capture cap;
    v->vo.reset(new System::VisualOdometry(wrap(K, 3, 3)));
    return v;
}
void ref_vo_destroy(ref_vo* v) { delete v; }
// VisualOdometry::odometrize(gray) (system.hpp:44-74).  init_depth (finest-level sized, i.e.
// (w>>2) x (h>>2), values >= 0.5) replaces the cv::randn draw of Scene's constructor
// (frame.hpp:18-21) for THIS frame; sigma is the constructor's 0.5.
int ref_vo_odometrize(ref_vo* v, const float* gray, int w, int h, const float* init_depth, float T_out[16])
{
    if (init_depth) {
        cv::shim::randn_source().data = init_depth;
        cv::shim::randn_source().n = (size_t)(w >> 2) * (h >> 2);
    }
    const int n_before = v->vo->m_history.size();
    std::string text;
    cv::Mat1f T;
    {
        Capture cap;
        T = v->vo->odometrize(wrap(gray, w, h));
        text = cap.ss.str();
    }
    cv::shim::randn_source().data = nullptr;
    cv::shim::randn_source().n = 0;
    put(T, T_out);
    v->last_was_keyframe = v->vo->m_history.size() > n_before ? 1 : 0;
    v->log.assign(256, ref_iter_log());
    const int n = parse_track_log(text, v->log.data(), 256, &v->last_valid_update);
    v->log.resize(n < 256 ? n : 256);
    return 0;
}
int ref_vo_history_size(const ref_vo* v) { return v->vo->m_history.size(); }
int ref_vo_last_was_keyframe(const ref_vo* v) { return v->last_was_keyframe; }
int ref_vo_last_valid_update(const ref_vo* v) { return v->last_valid_update; }
int ref_vo_last_log(const ref_vo* v, ref_iter_log* out, int cap)
{
    const int n = (int)v->log.size();
    for (int i = 0; i < n && i < cap; i++) out[i] = v->log[i];
    return n;
}
// keyframe born `age` keyframes ago (0 = the current keyframe); the handle is owned by the caller
ref_frame* ref_vo_history_frame(ref_vo* v, int age)
{
    if (age < 0 || age >= v->vo->m_history.size()) return nullptr;
    ref_frame* r = new ref_frame;
    r->f = v->vo->m_history[(size_t)age];
    return r;
}

// ---- shim self-checks: the restated OpenCV pieces, exposed so that tests can compare them with
// opencv-python known answers (tests/test_oracle_known_answers.py) -------------------------------
void ref_shim_randn(unsigned long long seed, int n, double mean, double stddev, float* out)
{
    cv::theRNG().state = seed;
    cv::Mat m(1, n);
    cv::randn(m, mean, stddev);
    put(m, out);
}
void ref_shim_gemm(const float* A, int m, int k, const float* B, int n, double alpha, const float* Cm, double beta, float* out)
{
    const cv::Mat1f c = Cm ? wrap(Cm, n, m) : cv::Mat1f();
    put(cv::shim::gemm(wrap(A, k, m), wrap(B, n, k), alpha, Cm ? &c : nullptr, beta), out);
}
void ref_shim_solve_svd(const float* A, const float* b, int m, int n, float* x)
{
    cv::Mat X;
    cv::solve(wrap(A, n, m), wrap(b, 1, m), X, cv::DECOMP_SVD);
    put(X, x);
}
// MatExpr lowering of `A*alpha + B*beta` (cv::add / subtract / scaleAdd / addWeighted)
void ref_shim_addex(const float* A, const float* B, int n, double alpha, double beta, float* out)
{
    put(cv::MatExpr::addex(wrap(A, n, 1), B ? wrap(B, n, 1) : cv::Mat1f(), alpha, beta).eval(), out);
}

} // extern "C"
