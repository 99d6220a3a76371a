// tests/cpp/dropin_main.cpp — exercises the drop-in C++ classes (include/dvo/) the way the reference's
// own callers do (main.cpp:33,49; test/track.cpp:45-57; test/warp.cpp:42; test/update.cpp:86-95;
// test/propagate.cpp:57-62; test/regularize.cpp:38) and dumps every result for tests/test_gpu_cpp_dropin.py,
// which checks them against the oracle.
//   usage: dropin_main in.bin out.bin
//   in.bin : int32 n, w, h; float K[9]; n gray frames [h][w]; depth0 [h][w]; sigma0 [h][w];
//            prior depth [h/4][w/4]; prior sigma [h/4][w/4]
//   out.bin: sequence of (int32 count, float[count]) records, in the order written below
#include "dvo/core/convert.hpp"
#include "dvo/core/draw.hpp"
#include "dvo/core/transform.hpp"
#include "dvo/math/gaussian.hpp"
#include "dvo/multi_gpu.hpp"
#include "dvo/map/implement.hpp"
#include "dvo/map/mapper.hpp"
#include "dvo/system/system.hpp"
#include "dvo/track/tracker.hpp"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

using dvo::Mat1f;

static FILE* g_out = nullptr;
static void put(const float* p, int n)
{
    fwrite(&n, sizeof(int), 1, g_out);
    fwrite(p, sizeof(float), (size_t)n, g_out);
}
static void put(const Mat1f& m) { put(dvo::data(m), (int)m.total()); }
static void put(float v) { put(&v, 1); }

static Mat1f read_mat(FILE* f, int h, int w)
{
    Mat1f m(h, w);
    if (fread(dvo::data(m), sizeof(float), (size_t)h * w, f) != (size_t)h * w) {
        fprintf(stderr, "short read\n");
        exit(2);
    }
    return m;
}

int main(int argc, char** argv)
{
    if (argc < 3) return 2;
    FILE* f = fopen(argv[1], "rb");
    g_out = fopen(argv[2], "wb");
    if (!f || !g_out) return 2;
    int hdr[3];
    if (fread(hdr, sizeof(int), 3, f) != 3) return 2;
    const int n = hdr[0], w = hdr[1], h = hdr[2];
    Mat1f K = read_mat(f, 3, 3);
    std::vector<Mat1f> gray;
    for (int i = 0; i < n; i++) gray.push_back(read_mat(f, h, w));
    Mat1f depth0 = read_mat(f, h, w), sigma0 = read_mat(f, h, w);
    Mat1f prior_d = read_mat(f, h / 4, w / 4), prior_s = read_mat(f, h / 4, w / 4);
    fclose(f);

    try {
        // 1. test/track.cpp: Frame(..., 4, 1) pair, Tracker::track
        auto ref = std::make_shared<System::Frame>(gray[0], depth0, sigma0, K, 4, 1);
        auto obj = std::make_shared<System::Frame>(gray[2], K, 4, 1);
        Track::Tracker tracker;
        Mat1f xi = tracker.track(obj, ref);
        put(xi);
        put((float)tracker.iterations().size());
        // 2. one Track::optimize step at pyramid index 1 from xi = 0
        Track::Stuff stuff(obj->at(1), ref->at(1), math::se3::xi(), 1);
        Track::Outcome oc = Track::optimize(stuff);
        put(oc.xi_update); put(oc.residual); put((float)oc.valid_pixels);
        // 3. Transform::warpImage on the finest level's raw images
        Mat1f top_gray = ref->gray(), top_depth = ref->depth(), top_sigma = ref->sigma();
        put(Transform::warpImage(xi, top_gray, top_depth, ref->K()));
        put(stuff.warped_gray());
        // 4. Map::Implement::regularize / propagate on raw maps
        put(Map::Implement::regularize(top_depth, top_sigma));
        Mat1f age0 = Mat1f::zeros(top_depth.rows, top_depth.cols);
        Mat1f pxi = math::se3::xi({0.01f, -0.01f, 0.f, 0.f, 0.f, 0.f});  // test/propagate.cpp:51
        auto prop = Map::Implement::propagate(top_depth, top_sigma, age0, pxi, ref->K());
        put(std::get<0>(prop)); put(std::get<1>(prop)); put(std::get<2>(prop));
        // 5. Map::Implement::update for one pixel (test/update.cpp:86-95)
        Mat1f uxi = math::se3::xi({-0.03f, 0.004f, 0.002f, 0.002f, -0.003f, 0.001f});
        auto du = Map::Implement::update(obj->gray(), top_gray, ref->gradX(), ref->gradY(), uxi, ref->K(),
                                         dvo::Point2i(top_gray.cols / 2, top_gray.rows / 2), 1.6f, 0.5f);
        put(du.first); put(du.second);
        // 6. se3 round trip
        put(math::se3::log(math::se3::exp(xi)));
        // 7. main.cpp: VisualOdometry(K) + odometrize per frame
        System::VisualOdometry vo(K);
        vo.setInitialDepth(prior_d, prior_s);
        for (int i = 0; i < n; i++) put(vo.odometrize(gray[i]));
        put(vo.keyframeImage(DVO_IMG_DEPTH)); put(vo.keyframeImage(DVO_IMG_SIGMA)); put(vo.keyframeImage(DVO_IMG_AGE));
        // 8. the same loop composed from the classes, as system.hpp:44-74 is written
        System::Frame::latestId() = -1;
        System::FrameHistory history;
        Map::Mapper mapper;
        for (int i = 0; i < n; i++) {
            auto frame = std::make_shared<System::Frame>(gray[i], K, 3, 2);
            auto kf = history.getRefFrame();
            if (kf == nullptr) {
                frame->updateDepthSigma(prior_d, prior_s);
                history.setRefFrame(frame);
                put(math::se3::T());
                continue;
            }
            Mat1f rel = tracker.track(frame, kf);
            frame->updateXi(rel, kf);
            mapper.estimate(history, frame);
            put(math::se3::exp(frame->m_xi));
        }
        auto kf = history.getRefFrame();
        put(kf->depth()); put(kf->sigma()); put(kf->age());
        // 9. tracking-only mode
        System::VisualOdometry vo2(K);
        put(vo2.odometrizeUsingDepth(gray[0], depth0, sigma0));
        put(vo2.odometrizeUsingDepth(gray[1], depth0, sigma0));
        // 10. headless outputs: Draw::visualize* of the keyframe of step 8, Convert::inversePose, trajectory file, PNG
        auto put_img = [](const Draw::Image& im) {
            std::vector<float> f(im.data.begin(), im.data.end());
            put(f.data(), (int)f.size());
        };
        const System::Scene& top = *kf->top();
        put_img(Draw::visualizeGray(top)); put_img(Draw::visualizeDepth(top)); put_img(Draw::visualizeDepthSigma(top));
        put_img(Draw::visualizeSigma(top)); put_img(Draw::visualizeGradientX(top)); put_img(Draw::visualizeAge(top));
        Draw::Image both = Draw::merge(Draw::visualizeDepthRaw(top), Draw::visualizeGradientY(top));
        put((float)both.rows);
        const std::string base = std::string(argv[2]);
        both.writePng(base + ".png");
        dvo::TrajectoryWriter traj(base + ".csv");
        Mat1f Tlast = math::se3::exp(kf->m_xi);
        traj.push(Convert::inversePose(Tlast)); traj.push(Tlast);
        put(Convert::inversePose(Tlast));
        // the reference's own call forms (src/map/mapper.cpp:37-42): raw images
        put_img(Draw::visualizeGray(kf->gray())); put_img(Draw::visualizeDepth(kf->depth(), kf->sigma()));
        put_img(Draw::visualizeDepth(kf->depth())); put_img(Draw::visualizeAge(kf->age())); put_img(Draw::visualizeGradient(kf->gradX()));
        // 11. test/update.cpp:51-116, literally (on images culled x8 so that the per-pixel calls stay few): gradients, a noisy
        //     depth prior, Map::Implement::update per pixel + math::Gaussian::operator(), then Map::Implement::regularize
        {
            const Mat1f ref_gray = Convert::cullImage(gray[0], 3), obj_gray = Convert::cullImage(gray[2], 3);
            const Mat1f Kc = Convert::cullIntrinsic(K, 3);
            Mat1f ref_gradx = Convert::gradiate(ref_gray, true);
            Mat1f ref_grady = Convert::gradiate(ref_gray, false);
            const Mat1f uxi2 = math::se3::xi({-0.1f, 0, 0, 0, 0, 0});  // test/update.cpp:58
            Mat1f ref_depth = Convert::cullImage(prior_d, 1).clone();  // stands in for the cv::randn noise of :64-68
            Mat1f ref_sigma(ref_depth.rows, ref_depth.cols, 0.5f);     // :71
            put(ref_gray); put(Kc); put(ref_gradx); put(ref_grady);
            for (int y = 0; y < ref_depth.rows; y++)
                for (int x = 0; x < ref_depth.cols; x++) {             // ref_depth.forEach, row-major
                    const dvo::Point2i x_i(x, y);
                    float& depth = ref_depth(x_i);
                    float sigma = ref_sigma(x_i);
                    depth = std::max(depth, 0.1f);
                    auto [new_depth, new_sigma] = Map::Implement::update(obj_gray, ref_gray, ref_gradx, ref_grady, uxi2, Kc, x_i, depth, sigma);
                    if (new_depth > 0 and new_depth < 6.0) {
                        math::Gaussian g(depth, sigma);
                        g(new_depth, new_sigma);
                        ref_depth(x_i) = g.depth;
                        ref_sigma(x_i) = g.sigma;
                    }
                }
            put(ref_depth); put(ref_sigma);
            ref_depth = Map::Implement::regularize(ref_depth, ref_sigma);  // :115
            put(ref_depth);
            // math::Gaussian::update: the reset branch draws from the process-wide mt19937
            dvo_gaussian_reseed();
            math::Gaussian g1(1.0f, 0.1f), g2(1.0f, 0.1f);
            const bool f1 = g1.update(3.0f, 0.1f), f2 = g2.update(1.05f, 0.2f);
            put(f1 ? 1.f : 0.f); put(g1.depth); put(g1.sigma); put(f2 ? 1.f : 0.f); put(g2.depth); put(g2.sigma);
        }
        // 12. test/warp.cpp:27-42: cullImage / cullIntrinsic, then Transform::warpImage on the culled images
        {
            Mat1f gray_image = Convert::cullImage(gray[0], 1), depth_image = Convert::cullImage(depth0, 1);
            Mat1f Kw = Convert::cullIntrinsic(K, 1);
            std::array<float, 6> params_f = {0.02f, -0.01f, 0.03f, 0.01f, -0.02f, 0.005f};
            Mat1f wxi = math::se3::xi(params_f);
            put(math::se3::exp(wxi));
            put(Transform::warpImage(wxi, gray_image, depth_image, Kw));
        }
        // 13. the scalar boundary: math::testXi (test/track.cpp:55), Transform::{backProject, transform, project, warp}, mapDepthtoGray
        {
            Mat1f bad = math::se3::xi(); bad(3) = std::nanf("");
            put(math::testXi(xi) ? 1.f : 0.f); put(math::testXi(bad) ? 1.f : 0.f);
            const Transform::Point2f px(123.25f, 77.5f);
            const Transform::Point3f xc = Transform::backProject(K, px, 1.7f);
            const Mat1f Tm = math::se3::exp(xi);
            const Transform::Point3f xt = Transform::transform(Tm, xc), xt2 = Transform::transform(xi, xc);
            const Transform::Point2f pj = Transform::project(K, xt), wp = Transform::warp(xi, px, 1.7f, K);
            const float v[] = {xc.x, xc.y, xc.z, xt.x, xt.y, xt.z, xt2.x, xt2.y, xt2.z, pj.x, pj.y, wp.x, wp.y};
            put(v, 13);
            Mat1f invT = math::se3::exp(math::se3::xi({0.05f, -0.01f, 0.0f, 0.01f, 0.02f, -0.01f}));
            auto md = Transform::mapDepthtoGray(Convert::cullImage(depth0, 1), gray[1], K, Convert::cullIntrinsic(K, 1), invT);
            put(invT); put(md.first); put(md.second);
        }
        // 14. dvo::MultiGpu: 3 sequences over the node's GPUs (one on this box), id mod nGPU; == the single-context run of step 7
        {
            dvo_params p = dvo::make_params(w, h, 3, 2, K);
            p.max_frames = 2; p.max_keyframes = n + 2;
            dvo::MultiGpu mg(3, p);
            std::vector<float> g3((size_t)3 * w * h), d3((size_t)3 * prior_d.total()), s3(d3.size()), T3(3 * 16);
            put((float)mg.world());
            for (int i = 0; i < n; i++) {
                for (int q = 0; q < 3; q++) {
                    std::memcpy(&g3[(size_t)q * w * h], dvo::data(gray[i]), sizeof(float) * (size_t)w * h);
                    std::memcpy(&d3[q * prior_d.total()], dvo::data(prior_d), sizeof(float) * prior_d.total());
                    std::memcpy(&s3[q * prior_s.total()], dvo::data(prior_s), sizeof(float) * prior_s.total());
                }
                mg.push(g3.data(), i == 0 ? d3.data() : nullptr, i == 0 ? s3.data() : nullptr, T3.data());
                put(T3.data(), 48);
            }
        }
    } catch (const std::exception& e) {
        fprintf(stderr, "dropin_main: %s\n", e.what());
        return 1;
    }
    fclose(g_out);
    dvo::Device::reset();
    return 0;
}
