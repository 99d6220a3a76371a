// tests/cpp/dropin_main.cpp — exercises the drop-in C++ classes (include/dvo/) the way the reference's
// own callers do (main.cpp:33,49; test/track.cpp:45-57; test/warp.cpp:42; test/update.cpp:86-95;
// test/propagate.cpp:57-62; test/regularize.cpp:38) and dumps every result for tests/test_gpu_cpp_dropin.py,
// which checks them against the oracle.
//   usage: dropin_main in.bin out.bin
//   in.bin : int32 n, w, h; float K[9]; n gray frames [h][w]; depth0 [h][w]; sigma0 [h][w];
//            prior depth [h/4][w/4]; prior sigma [h/4][w/4]
//   out.bin: sequence of (int32 count, float[count]) records, in the order written below
#include "dvo/core/draw.hpp"
#include "dvo/core/transform.hpp"
#include "dvo/map/implement.hpp"
#include "dvo/map/mapper.hpp"
#include "dvo/system/system.hpp"
#include "dvo/track/tracker.hpp"

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
    } catch (const std::exception& e) {
        fprintf(stderr, "dropin_main: %s\n", e.what());
        return 1;
    }
    fclose(g_out);
    dvo::Device::reset();
    return 0;
}
