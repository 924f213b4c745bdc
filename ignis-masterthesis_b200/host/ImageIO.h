// ImageIO.h -- headless image output for igcli: OpenEXR (uncompressed scanlines, three FLOAT channels; the container
// the reference writes, src/runtime/ImageIO / frontend/common/IO.cpp) and PFM.
#pragma once
#include <cstddef>
#include <string>

namespace IG {
// rgb: width*height*3 floats, row 0 = top row (the film's order).  The extension picks the format (.pfm, else EXR).
bool saveImageRGB(const std::string& path, const float* rgb, size_t width, size_t height, float scale);
} // namespace IG
