// PngReader.h -- 8-bit PNG decoding for image textures (see PngReader.cpp).
#pragma once

#include <cstdint>
#include <string>
#include <vector>

namespace IG {
struct PngImage {
    uint32_t             width = 0, height = 0;
    bool                 opaque = true; // the file carries no alpha channel
    std::vector<uint8_t> rgba;          // top row first, 4 bytes per pixel
};
bool loadPng8(const std::string& path, PngImage& out, std::string& error);
} // namespace IG
