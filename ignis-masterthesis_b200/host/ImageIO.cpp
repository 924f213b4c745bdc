#include "ImageIO.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace IG {
namespace {
void put(std::vector<uint8_t>& b, const void* p, size_t n) { b.insert(b.end(), (const uint8_t*)p, (const uint8_t*)p + n); }
void put_str(std::vector<uint8_t>& b, const char* s) { put(b, s, std::strlen(s) + 1); }
void put_i32(std::vector<uint8_t>& b, int32_t v) { put(b, &v, 4); }
void attr(std::vector<uint8_t>& b, const char* name, const char* type, const void* data, int32_t size)
{
    put_str(b, name);
    put_str(b, type);
    put_i32(b, size);
    put(b, data, size);
}

bool save_exr(const std::string& path, const float* rgb, size_t w, size_t h, float scale)
{
    std::vector<uint8_t> head;
    const uint32_t magic = 20000630u, version = 2u; // single-part scanline file
    put(head, &magic, 4);
    put(head, &version, 4);
    std::vector<uint8_t> ch;
    for (const char* name : { "B", "G", "R" }) { // channels are stored in alphabetical order
        put_str(ch, name);
        put_i32(ch, 2); // FLOAT
        const uint8_t plinear[4] = { 0, 0, 0, 0 };
        put(ch, plinear, 4);
        put_i32(ch, 1);
        put_i32(ch, 1);
    }
    ch.push_back(0);
    attr(head, "channels", "chlist", ch.data(), (int32_t)ch.size());
    const uint8_t no_compression = 0;
    attr(head, "compression", "compression", &no_compression, 1);
    const int32_t window[4] = { 0, 0, (int32_t)w - 1, (int32_t)h - 1 };
    attr(head, "dataWindow", "box2i", window, 16);
    attr(head, "displayWindow", "box2i", window, 16);
    const uint8_t increasing_y = 0;
    attr(head, "lineOrder", "lineOrder", &increasing_y, 1);
    const float one = 1.0f, centre[2] = { 0.0f, 0.0f };
    attr(head, "pixelAspectRatio", "float", &one, 4);
    attr(head, "screenWindowCenter", "v2f", centre, 8);
    attr(head, "screenWindowWidth", "float", &one, 4);
    head.push_back(0);

    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const uint64_t line_bytes = 8 + 3 * 4 * (uint64_t)w;
    const uint64_t first      = head.size() + 8 * (uint64_t)h;
    bool ok = std::fwrite(head.data(), 1, head.size(), f) == head.size();
    for (size_t y = 0; ok && y < h; ++y) {
        const uint64_t off = first + y * line_bytes;
        ok = std::fwrite(&off, 8, 1, f) == 1;
    }
    std::vector<float> line(3 * w);
    for (size_t y = 0; ok && y < h; ++y) {
        const int32_t rec[2] = { (int32_t)y, (int32_t)(3 * 4 * w) };
        for (size_t x = 0; x < w; ++x) {
            const float* p  = rgb + 3 * (y * w + x);
            line[x]         = p[2] * scale;
            line[w + x]     = p[1] * scale;
            line[2 * w + x] = p[0] * scale;
        }
        ok = std::fwrite(rec, 4, 2, f) == 2 && std::fwrite(line.data(), 4, line.size(), f) == line.size();
    }
    return std::fclose(f) == 0 && ok;
}

bool save_pfm(const std::string& path, const float* rgb, size_t w, size_t h, float scale)
{
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "PF\n%zu %zu\n-1.0\n", w, h); // negative scale = little endian; rows bottom to top
    std::vector<float> line(3 * w);
    bool ok = true;
    for (size_t y = 0; ok && y < h; ++y) {
        const float* src = rgb + 3 * w * (h - 1 - y);
        for (size_t i = 0; i < 3 * w; ++i) line[i] = src[i] * scale;
        ok = std::fwrite(line.data(), 4, line.size(), f) == line.size();
    }
    return std::fclose(f) == 0 && ok;
}
} // namespace

bool saveImageRGB(const std::string& path, const float* rgb, size_t width, size_t height, float scale)
{
    const bool pfm = path.size() > 4 && path.compare(path.size() - 4, 4, ".pfm") == 0;
    return pfm ? save_pfm(path, rgb, width, height, scale) : save_exr(path, rgb, width, height, scale);
}
} // namespace IG
