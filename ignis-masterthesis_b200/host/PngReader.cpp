// PngReader.cpp -- 8-bit PNG decoding for image textures (the reference loads LDR images through stb_image,
// src/runtime/Image.cpp:565-655; stb is not available here, zlib is).  Non-interlaced grey, grey + alpha, RGB, RGBA and
// palette images at 8 bits per sample; anything else is refused.
#include "PngReader.h"

#include <cstring>
#include <fstream>
#include <sstream>
#include <zlib.h>

namespace IG {
namespace {
uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
int      paeth(int a, int b, int c)
{
    const int p = a + b - c, pa = p > a ? p - a : a - p, pb = p > b ? p - b : b - p, pc = p > c ? p - c : c - p;
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}
} // namespace

bool loadPng8(const std::string& path, PngImage& out, std::string& error)
{
    std::ifstream in(path, std::ios::binary);
    if (!in) {
        error = "cannot open " + path;
        return false;
    }
    std::ostringstream ss;
    ss << in.rdbuf();
    const std::string file = ss.str();
    const uint8_t*    d    = reinterpret_cast<const uint8_t*>(file.data());
    static const uint8_t magic[8] = { 0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n' };
    if (file.size() < 33 || std::memcmp(d, magic, 8) != 0) {
        error = path + " is not a PNG file";
        return false;
    }
    uint32_t             w = 0, h = 0;
    int                  depth = 0, ctype = 0, interlace = 0;
    std::vector<uint8_t> idat, palette, trns;
    size_t               pos = 8;
    while (pos + 12 <= file.size()) {
        const uint32_t len = be32(d + pos);
        const char*    tag = file.data() + pos + 4;
        if (pos + 12 + (size_t)len > file.size()) break;
        const uint8_t* body = d + pos + 8;
        if (!std::memcmp(tag, "IHDR", 4) && len >= 13) {
            w         = be32(body);
            h         = be32(body + 4);
            depth     = body[8];
            ctype     = body[9];
            interlace = body[12];
        } else if (!std::memcmp(tag, "PLTE", 4)) {
            palette.assign(body, body + len);
        } else if (!std::memcmp(tag, "tRNS", 4)) {
            trns.assign(body, body + len);
        } else if (!std::memcmp(tag, "IDAT", 4)) {
            idat.insert(idat.end(), body, body + len);
        } else if (!std::memcmp(tag, "IEND", 4)) {
            break;
        }
        pos += 12 + (size_t)len;
    }
    if (w == 0 || h == 0 || depth != 8 || interlace != 0) {
        error = path + ": only non-interlaced 8-bit PNG images are supported";
        return false;
    }
    int spp = 0; // samples per pixel in the file
    switch (ctype) {
    case 0: spp = 1; break;
    case 2: spp = 3; break;
    case 3: spp = 1; break;
    case 4: spp = 2; break;
    case 6: spp = 4; break;
    default: error = path + ": unknown PNG colour type"; return false;
    }
    const size_t         stride = (size_t)w * spp;
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf               raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size()) {
        error = path + ": corrupt PNG data";
        return false;
    }
    std::vector<uint8_t> px(stride * h);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t  ft   = raw[(stride + 1) * y];
        const uint8_t* src  = raw.data() + (stride + 1) * y + 1;
        uint8_t*       dst  = px.data() + stride * y;
        const uint8_t* prev = y ? dst - stride : nullptr;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= (size_t)spp ? dst[i - spp] : 0, b = prev ? prev[i] : 0, c = (prev && i >= (size_t)spp) ? prev[i - spp] : 0;
            int       v = src[i];
            switch (ft) {
            case 0: break;
            case 1: v += a; break;
            case 2: v += b; break;
            case 3: v += (a + b) / 2; break;
            case 4: v += paeth(a, b, c); break;
            default: error = path + ": bad PNG filter"; return false;
            }
            dst[i] = (uint8_t)v;
        }
    }
    // channel semantics of the reference's loader: grey -> rgb replicated, opaque when the file has no alpha
    out.width  = w;
    out.height = h;
    out.opaque = !(ctype == 4 || ctype == 6 || (ctype == 3 && !trns.empty()));
    out.rgba.resize((size_t)w * h * 4);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        uint8_t r, g, b, a = 255;
        const uint8_t* p = px.data() + i * spp;
        switch (ctype) {
        case 0: r = g = b = p[0]; break;
        case 2: r = p[0], g = p[1], b = p[2]; break;
        case 3: {
            const size_t k = p[0];
            if (3 * k + 2 >= palette.size()) {
                r = g = b = 0;
            } else {
                r = palette[3 * k], g = palette[3 * k + 1], b = palette[3 * k + 2];
            }
            if (k < trns.size()) a = trns[k];
            break;
        }
        case 4: r = g = b = p[0], a = p[1]; break;
        default: r = p[0], g = p[1], b = p[2], a = p[3]; break;
        }
        out.rgba[4 * i + 0] = r;
        out.rgba[4 * i + 1] = g;
        out.rgba[4 * i + 2] = b;
        out.rgba[4 * i + 3] = a;
    }
    return true;
}
} // namespace IG
