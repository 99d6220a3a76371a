// dvo/core/draw.hpp — Draw::visualize* (include/core/draw.hpp:9-23, src/core/draw.cpp:7-100) and Convert::inversePose
// (src/core/convert.cpp:31-39) over the C ABI, headless: the reference hands the 8-bit BGR image to cv::imshow and the inverted
// poses to its GLFW drawer (main.cpp:50-53,76-86); here the same bytes come back as a Draw::Image (optionally written as a PNG)
// and the poses go to a CSV file.  The colour mapping runs on the device: either from a raw image, with the reference's own
// signatures (visualizeGray(const Mat1f&), visualizeDepth(depth, sigma), ...), or straight from a frame's slot through the
// overloads that take the System::Scene (one pyramid level of a Frame) whose image the reference would have passed.
#pragma once
#include "../system/frame.hpp"

#include <cstdint>
#include <string>
#include <vector>

namespace Draw
{
struct Image {  // CV_8UC3, BGR, row-major
    int rows = 0, cols = 0;
    std::vector<uint8_t> data;
    const uint8_t* at(int r, int c) const { return data.data() + 3 * ((size_t)r * cols + c); }
    void writePng(const std::string& path) const
    {
        if (dvo_write_png(path.c_str(), data.data(), cols, rows)) throw dvo::Error(DVO_ERR_ARG, "dvo_write_png: " + path);
    }
};

inline Image visualize(const System::Scene& scene, dvo_vis_kind kind)
{
    Image out;
    out.rows = scene.rows; out.cols = scene.cols;
    out.data.resize((size_t)3 * scene.rows * scene.cols);
    const auto& slot = scene.slot();
    slot->ctx->check(dvo_visualize(slot->ctx->h, slot->id, scene.level(), kind, out.data.data()));
    return out;
}
inline Image visualizeGray(const System::Scene& s) { return visualize(s, DVO_VIS_GRAY); }                // draw.cpp:7-19
inline Image visualizeDepth(const System::Scene& s) { return visualize(s, DVO_VIS_DEPTH); }              // draw.cpp:32-48
inline Image visualizeDepthSigma(const System::Scene& s) { return visualize(s, DVO_VIS_DEPTH_SIGMA); }   // visualizeDepth(depth, sigma), draw.cpp:50-67
inline Image visualizeDepthRaw(const System::Scene& s) { return visualize(s, DVO_VIS_DEPTH_RAW); }       // draw.cpp:21-29
inline Image visualizeSigma(const System::Scene& s) { return visualize(s, DVO_VIS_SIGMA); }              // draw.cpp:69-75
inline Image visualizeGradientX(const System::Scene& s) { return visualize(s, DVO_VIS_GRADX); }          // visualizeGradient(gradX), draw.cpp:77-91
inline Image visualizeGradientY(const System::Scene& s) { return visualize(s, DVO_VIS_GRADY); }
inline Image visualizeAge(const System::Scene& s) { return visualize(s, DVO_VIS_AGE); }                  // finest level only, draw.cpp:94-100

// The reference's own signatures: a raw image (test/update.cpp:19-27, src/map/mapper.cpp:37-42)
inline Image visualize(const dvo::Mat1f& image, dvo_vis_kind kind, const dvo::Mat1f* sigma = nullptr)
{
    dvo::require(!image.empty(), "visualize: empty image");
    dvo::require(sigma == nullptr || (sigma->rows == image.rows && sigma->cols == image.cols), "visualize: depth/sigma size mismatch");
    Image out;
    out.rows = image.rows; out.cols = image.cols;
    out.data.resize((size_t)3 * image.rows * image.cols);
    dvo::Mat1f k0, k1;
    const int rc = dvo_visualize_image(kind, dvo::pixels(image, k0), sigma ? dvo::pixels(*sigma, k1) : nullptr, image.cols, image.rows, out.data.data());
    if (rc) throw dvo::Error(rc, "dvo_visualize_image failed");
    return out;
}
inline Image visualizeGray(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_GRAY); }
inline Image visualizeDepth(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_DEPTH); }
inline Image visualizeDepth(const dvo::Mat1f& depth, const dvo::Mat1f& sigma) { return visualize(depth, DVO_VIS_DEPTH_SIGMA, &sigma); }
inline Image visualizeDepthRaw(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_DEPTH_RAW); }
inline Image visualizeSigma(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_SIGMA); }
inline Image visualizeGradient(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_GRADX); }
inline Image visualizeAge(const dvo::Mat1f& m) { return visualize(m, DVO_VIS_AGE); }

// Draw::merge(img1, img2) (draw.hpp:33-39): vconcat
inline Image merge(const Image& a, const Image& b)
{
    dvo::require(a.cols == b.cols, "merge: widths differ");
    Image m;
    m.cols = a.cols; m.rows = a.rows + b.rows;
    m.data = a.data;
    m.data.insert(m.data.end(), b.data.begin(), b.data.end());
    return m;
}
}  // namespace Draw

namespace Convert
{
using dvo::Mat1f;
// R^T and -t in a 4x4 whose last row stays zero, exactly as the reference builds it (convert.cpp:31-39)
inline Mat1f inversePose(const Mat1f& T)
{
    dvo::require(T.rows == 4 && T.cols == 4, "inversePose: 4x4 expected");
    float in[16], out[16];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) in[i * 4 + j] = T(i, j);
    dvo_inverse_pose(in, out);
    Mat1f r(4, 4);
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) r(i, j) = out[i * 4 + j];
    return r;
}
}  // namespace Convert

namespace dvo
{
// the trajectory main.cpp:50-53 accumulates for Graphic::draw, as a CSV file (frame, 16 values per line)
class TrajectoryWriter
{
public:
    explicit TrajectoryWriter(std::string path) : m_path(std::move(path)) {}
    void push(const Mat1f& T)
    {
        float t[16];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 4; j++) t[i * 4 + j] = T(i, j);
        if (dvo_write_trajectory_csv(m_path.c_str(), t, 1, m_n, m_n > 0)) throw Error(DVO_ERR_ARG, "cannot write " + m_path);
        m_n++;
    }
    int size() const { return m_n; }

private:
    std::string m_path;
    int m_n = 0;
};
}  // namespace dvo
