// Stub written for this repo (NOT OpenCV): the only OpenCV names the reference's render path touches
// are in Material::upload_texture (Material.hpp:29-43).  Texture *decoding* is out of scope;
// harnesses feed raw texels, so imread returns an empty image and GpuMat copies nothing.
#pragma once
#include <cstddef>
#include <string>
#define CV_8UC3 16
namespace cv {
enum { IMREAD_COLOR = 1 };
struct Mat { int cols = 0, rows = 0; };
inline Mat imread(const std::string&, int) { return Mat(); }
namespace cuda {
struct GpuMat {
    GpuMat(int, int, int, void*, size_t) {}
    void upload(const Mat&) {}
    void download(Mat&) const {}
};
}  // namespace cuda
}  // namespace cv
