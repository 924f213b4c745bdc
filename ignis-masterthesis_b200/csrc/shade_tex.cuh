// shade_tex.cuh -- textures, the GGX distribution of visible normals and bump mapping (sm_100a).
// Continuation of shade.cuh: same citation (paths relative to src/artic) and arithmetic conventions.  The functions
// here are rare next to the diffuse / specular paths of the shade kernel, so they are __noinline__: their registers
// are paid only by the lanes that call them.  They take and return VALUES only: a reference parameter would pin the
// caller's object (the RNG, the BSDF, the kernel's scene argument) in local memory for the whole kernel.
#pragma once
#include "shade.cuh"

namespace igb {
namespace shade {

// ---- texture/checkerboard.art, texture/image.art, driver/image.art ---------------------------------------
__device__ __forceinline__ float wrapf(float v, float mn, float mx) // math::wrap, core/math.art:88-91
{
    const float range = mx - mn;
    return range <= flt_eps ? mn : v - (range * floorf((v - mn) / range));
}
__device__ __forceinline__ float fractf(float x) { return x - floorf(x); } // core/math.art:74
__device__ __forceinline__ int32_t tex_border(uint32_t mode, int32_t x, int32_t w) // texture/image.art:9-40
{
    // inside the image every border mode is the identity except mirror (which flips even periods): skip the integer
    // division (~25 instructions; a bicubic lookup wraps 32 coordinates)
    if (mode != IGB200_WRAP_MIRROR && (uint32_t)x < (uint32_t)w) return x;
    if (mode == IGB200_WRAP_CLAMP) return min(max(x, 0), w - 1);
    if (mode == IGB200_WRAP_MIRROR) {
        const int32_t t = x < 0 ? -1 - x : x;
        const int32_t i = t / w, k = t - i * w;
        return (i & 1) == 0 ? w - 1 - k : k;
    }
    const int32_t t = x % w;
    return t < 0 ? t + w : t;
}
#ifndef IGB_TEX_LUT
#define IGB_TEX_LUT 1 // measured on B200 (profiles/r02_ab_wave_stream_texlut.txt): C5 3342 -> 3441 Msamples/s, image tests unchanged
#endif
// image_rgba_unpack (driver/image.art:8-14) divides each byte by 255.  An IEEE division is ~10 instructions and a bicubic,
// bump-mapped hit unpacks 192 bytes; the 256 possible quotients are tabulated instead (hex float literals of
// float32(i) / float32(255), i.e. bit-identical to the division the oracle performs) and read through the L1.
static __device__ const float kUnorm8[256] = {
    0x0.0p+0f, 0x1.0101020000000p-8f, 0x1.0101020000000p-7f, 0x1.8181820000000p-7f, 0x1.0101020000000p-6f, 0x1.4141420000000p-6f, 0x1.8181820000000p-6f, 0x1.c1c1c20000000p-6f,
    0x1.0101020000000p-5f, 0x1.2121220000000p-5f, 0x1.4141420000000p-5f, 0x1.6161620000000p-5f, 0x1.8181820000000p-5f, 0x1.a1a1a20000000p-5f, 0x1.c1c1c20000000p-5f, 0x1.e1e1e20000000p-5f,
    0x1.0101020000000p-4f, 0x1.1111120000000p-4f, 0x1.2121220000000p-4f, 0x1.3131320000000p-4f, 0x1.4141420000000p-4f, 0x1.5151520000000p-4f, 0x1.6161620000000p-4f, 0x1.7171720000000p-4f,
    0x1.8181820000000p-4f, 0x1.9191920000000p-4f, 0x1.a1a1a20000000p-4f, 0x1.b1b1b20000000p-4f, 0x1.c1c1c20000000p-4f, 0x1.d1d1d20000000p-4f, 0x1.e1e1e20000000p-4f, 0x1.f1f1f20000000p-4f,
    0x1.0101020000000p-3f, 0x1.09090a0000000p-3f, 0x1.1111120000000p-3f, 0x1.19191a0000000p-3f, 0x1.2121220000000p-3f, 0x1.29292a0000000p-3f, 0x1.3131320000000p-3f, 0x1.39393a0000000p-3f,
    0x1.4141420000000p-3f, 0x1.49494a0000000p-3f, 0x1.5151520000000p-3f, 0x1.59595a0000000p-3f, 0x1.6161620000000p-3f, 0x1.69696a0000000p-3f, 0x1.7171720000000p-3f, 0x1.79797a0000000p-3f,
    0x1.8181820000000p-3f, 0x1.89898a0000000p-3f, 0x1.9191920000000p-3f, 0x1.99999a0000000p-3f, 0x1.a1a1a20000000p-3f, 0x1.a9a9aa0000000p-3f, 0x1.b1b1b20000000p-3f, 0x1.b9b9ba0000000p-3f,
    0x1.c1c1c20000000p-3f, 0x1.c9c9ca0000000p-3f, 0x1.d1d1d20000000p-3f, 0x1.d9d9da0000000p-3f, 0x1.e1e1e20000000p-3f, 0x1.e9e9ea0000000p-3f, 0x1.f1f1f20000000p-3f, 0x1.f9f9fa0000000p-3f,
    0x1.0101020000000p-2f, 0x1.0505060000000p-2f, 0x1.09090a0000000p-2f, 0x1.0d0d0e0000000p-2f, 0x1.1111120000000p-2f, 0x1.1515160000000p-2f, 0x1.19191a0000000p-2f, 0x1.1d1d1e0000000p-2f,
    0x1.2121220000000p-2f, 0x1.2525260000000p-2f, 0x1.29292a0000000p-2f, 0x1.2d2d2e0000000p-2f, 0x1.3131320000000p-2f, 0x1.3535360000000p-2f, 0x1.39393a0000000p-2f, 0x1.3d3d3e0000000p-2f,
    0x1.4141420000000p-2f, 0x1.4545460000000p-2f, 0x1.49494a0000000p-2f, 0x1.4d4d4e0000000p-2f, 0x1.5151520000000p-2f, 0x1.5555560000000p-2f, 0x1.59595a0000000p-2f, 0x1.5d5d5e0000000p-2f,
    0x1.6161620000000p-2f, 0x1.6565660000000p-2f, 0x1.69696a0000000p-2f, 0x1.6d6d6e0000000p-2f, 0x1.7171720000000p-2f, 0x1.7575760000000p-2f, 0x1.79797a0000000p-2f, 0x1.7d7d7e0000000p-2f,
    0x1.8181820000000p-2f, 0x1.8585860000000p-2f, 0x1.89898a0000000p-2f, 0x1.8d8d8e0000000p-2f, 0x1.9191920000000p-2f, 0x1.9595960000000p-2f, 0x1.99999a0000000p-2f, 0x1.9d9d9e0000000p-2f,
    0x1.a1a1a20000000p-2f, 0x1.a5a5a60000000p-2f, 0x1.a9a9aa0000000p-2f, 0x1.adadae0000000p-2f, 0x1.b1b1b20000000p-2f, 0x1.b5b5b60000000p-2f, 0x1.b9b9ba0000000p-2f, 0x1.bdbdbe0000000p-2f,
    0x1.c1c1c20000000p-2f, 0x1.c5c5c60000000p-2f, 0x1.c9c9ca0000000p-2f, 0x1.cdcdce0000000p-2f, 0x1.d1d1d20000000p-2f, 0x1.d5d5d60000000p-2f, 0x1.d9d9da0000000p-2f, 0x1.ddddde0000000p-2f,
    0x1.e1e1e20000000p-2f, 0x1.e5e5e60000000p-2f, 0x1.e9e9ea0000000p-2f, 0x1.ededee0000000p-2f, 0x1.f1f1f20000000p-2f, 0x1.f5f5f60000000p-2f, 0x1.f9f9fa0000000p-2f, 0x1.fdfdfe0000000p-2f,
    0x1.0101020000000p-1f, 0x1.0303040000000p-1f, 0x1.0505060000000p-1f, 0x1.0707080000000p-1f, 0x1.09090a0000000p-1f, 0x1.0b0b0c0000000p-1f, 0x1.0d0d0e0000000p-1f, 0x1.0f0f100000000p-1f,
    0x1.1111120000000p-1f, 0x1.1313140000000p-1f, 0x1.1515160000000p-1f, 0x1.1717180000000p-1f, 0x1.19191a0000000p-1f, 0x1.1b1b1c0000000p-1f, 0x1.1d1d1e0000000p-1f, 0x1.1f1f200000000p-1f,
    0x1.2121220000000p-1f, 0x1.2323240000000p-1f, 0x1.2525260000000p-1f, 0x1.2727280000000p-1f, 0x1.29292a0000000p-1f, 0x1.2b2b2c0000000p-1f, 0x1.2d2d2e0000000p-1f, 0x1.2f2f300000000p-1f,
    0x1.3131320000000p-1f, 0x1.3333340000000p-1f, 0x1.3535360000000p-1f, 0x1.3737380000000p-1f, 0x1.39393a0000000p-1f, 0x1.3b3b3c0000000p-1f, 0x1.3d3d3e0000000p-1f, 0x1.3f3f400000000p-1f,
    0x1.4141420000000p-1f, 0x1.4343440000000p-1f, 0x1.4545460000000p-1f, 0x1.4747480000000p-1f, 0x1.49494a0000000p-1f, 0x1.4b4b4c0000000p-1f, 0x1.4d4d4e0000000p-1f, 0x1.4f4f500000000p-1f,
    0x1.5151520000000p-1f, 0x1.5353540000000p-1f, 0x1.5555560000000p-1f, 0x1.5757580000000p-1f, 0x1.59595a0000000p-1f, 0x1.5b5b5c0000000p-1f, 0x1.5d5d5e0000000p-1f, 0x1.5f5f600000000p-1f,
    0x1.6161620000000p-1f, 0x1.6363640000000p-1f, 0x1.6565660000000p-1f, 0x1.6767680000000p-1f, 0x1.69696a0000000p-1f, 0x1.6b6b6c0000000p-1f, 0x1.6d6d6e0000000p-1f, 0x1.6f6f700000000p-1f,
    0x1.7171720000000p-1f, 0x1.7373740000000p-1f, 0x1.7575760000000p-1f, 0x1.7777780000000p-1f, 0x1.79797a0000000p-1f, 0x1.7b7b7c0000000p-1f, 0x1.7d7d7e0000000p-1f, 0x1.7f7f800000000p-1f,
    0x1.8181820000000p-1f, 0x1.8383840000000p-1f, 0x1.8585860000000p-1f, 0x1.8787880000000p-1f, 0x1.89898a0000000p-1f, 0x1.8b8b8c0000000p-1f, 0x1.8d8d8e0000000p-1f, 0x1.8f8f900000000p-1f,
    0x1.9191920000000p-1f, 0x1.9393940000000p-1f, 0x1.9595960000000p-1f, 0x1.9797980000000p-1f, 0x1.99999a0000000p-1f, 0x1.9b9b9c0000000p-1f, 0x1.9d9d9e0000000p-1f, 0x1.9f9fa00000000p-1f,
    0x1.a1a1a20000000p-1f, 0x1.a3a3a40000000p-1f, 0x1.a5a5a60000000p-1f, 0x1.a7a7a80000000p-1f, 0x1.a9a9aa0000000p-1f, 0x1.ababac0000000p-1f, 0x1.adadae0000000p-1f, 0x1.afafb00000000p-1f,
    0x1.b1b1b20000000p-1f, 0x1.b3b3b40000000p-1f, 0x1.b5b5b60000000p-1f, 0x1.b7b7b80000000p-1f, 0x1.b9b9ba0000000p-1f, 0x1.bbbbbc0000000p-1f, 0x1.bdbdbe0000000p-1f, 0x1.bfbfc00000000p-1f,
    0x1.c1c1c20000000p-1f, 0x1.c3c3c40000000p-1f, 0x1.c5c5c60000000p-1f, 0x1.c7c7c80000000p-1f, 0x1.c9c9ca0000000p-1f, 0x1.cbcbcc0000000p-1f, 0x1.cdcdce0000000p-1f, 0x1.cfcfd00000000p-1f,
    0x1.d1d1d20000000p-1f, 0x1.d3d3d40000000p-1f, 0x1.d5d5d60000000p-1f, 0x1.d7d7d80000000p-1f, 0x1.d9d9da0000000p-1f, 0x1.dbdbdc0000000p-1f, 0x1.ddddde0000000p-1f, 0x1.dfdfe00000000p-1f,
    0x1.e1e1e20000000p-1f, 0x1.e3e3e40000000p-1f, 0x1.e5e5e60000000p-1f, 0x1.e7e7e80000000p-1f, 0x1.e9e9ea0000000p-1f, 0x1.ebebec0000000p-1f, 0x1.ededee0000000p-1f, 0x1.efeff00000000p-1f,
    0x1.f1f1f20000000p-1f, 0x1.f3f3f40000000p-1f, 0x1.f5f5f60000000p-1f, 0x1.f7f7f80000000p-1f, 0x1.f9f9fa0000000p-1f, 0x1.fbfbfc0000000p-1f, 0x1.fdfdfe0000000p-1f, 0x1.0000000000000p+0f,
};
__device__ __forceinline__ float unorm8(uint32_t byte) { return __ldg(kUnorm8 + byte); }
__device__ __forceinline__ float4 tex_texel(const igb200_texture& t, const uint8_t* blob, int32_t x, int32_t y)
{
    const size_t i = (size_t)y * t.width + (size_t)x;
    if (t.format == IGB200_IMAGE_RGBA32F) return __ldg(reinterpret_cast<const float4*>(blob + t.offset) + i);
    const uint32_t q = __ldg(reinterpret_cast<const uint32_t*>(blob + t.offset) + i); // image_rgba_unpack
#if IGB_TEX_LUT
    return make_float4(unorm8(q & 0xFFu), unorm8((q >> 8) & 0xFFu), unorm8((q >> 16) & 0xFFu), t.opaque ? 1.0f : unorm8(q >> 24));
#else
    return make_float4((float)(q & 0xFFu) / 255.0f, (float)((q & 0xFF00u) >> 8) / 255.0f, (float)((q & 0xFF0000u) >> 16) / 255.0f,
                       t.opaque ? 1.0f : (float)((q & 0xFF000000u) >> 24) / 255.0f);
#endif
}
static __device__ __noinline__ Color tex_eval(const igb200_texture* textures, const uint8_t* blob, uint32_t id, float u, float v)
{
    const igb200_texture& t = textures[id];
    const float* m  = t.transform; // mat3x3_transform_point_affine, core/matrix.art:223-226
    const float  tu = __fmaf_rn(m[0], u, __fmaf_rn(m[3], v, m[6] * 1.0f));
    const float  tv = __fmaf_rn(m[1], u, __fmaf_rn(m[4], v, m[7] * 1.0f));
    if (t.type == IGB200_TEX_CHECKERBOARD) { // texture/checkerboard.art:4-13
        const float sx = tu * t.scale[0], sy = tv * t.scale[1];
        const bool  px = ((int32_t)wrapf(sx, 0.0f, 2.0f) % 2) == 0;
        const bool  py = ((int32_t)wrapf(sy, 0.0f, 2.0f) % 2) == 0;
        return (px != py) ? Color{ t.color0[0], t.color0[1], t.color0[2] } : Color{ t.color1[0], t.color1[1], t.color1[2] };
    }
    const int32_t w = (int32_t)t.width, h = (int32_t)t.height;
    if (t.filter == IGB200_FILTER_NEAREST) { // :59-68,81-89
        const float  fu = tu * (float)w, fv = tv * (float)h;
        const float4 p  = tex_texel(t, blob, tex_border(t.wrap_u, (int32_t)floorf(fu), w), tex_border(t.wrap_v, (int32_t)floorf(fv), h));
        return Color{ p.x, p.y, p.z };
    }
    const float   fu = tu * (float)w - 0.5f, fv = tv * (float)h - 0.5f; // map_uv_to_image_pixel :70-79
    const int32_t ix = (int32_t)floorf(fu), iy = (int32_t)floorf(fv);
    const float   fx = fractf(fu), fy = fractf(fv);
    if (t.filter == IGB200_FILTER_BILINEAR) { // :91-106
        const int32_t x0 = tex_border(t.wrap_u, ix, w), y0 = tex_border(t.wrap_v, iy, h);
        const int32_t x1 = tex_border(t.wrap_u, ix + 1, w), y1 = tex_border(t.wrap_v, iy + 1, h);
        const float4  p00 = tex_texel(t, blob, x0, y0), p10 = tex_texel(t, blob, x1, y0);
        const float4  p01 = tex_texel(t, blob, x0, y1), p11 = tex_texel(t, blob, x1, y1);
        auto lerp = [](float a, float b, float k) { return (1.0f - k) * a + k * b; };
        return Color{ lerp(lerp(p00.x, p10.x, fx), lerp(p01.x, p11.x, fx), fy), lerp(lerp(p00.y, p10.y, fx), lerp(p01.y, p11.y, fx), fy),
                      lerp(lerp(p00.z, p10.z, fx), lerp(p01.z, p11.z, fx), fy) };
    }
    // make_bicubic_filter :108-146
    auto w0 = [](float a) { return (a * (a * (-a + 3.0f) - 3.0f) + 1.0f) / 6.0f; };
    auto w1 = [](float a) { return (a * a * (3.0f * a - 6.0f) + 4.0f) / 6.0f; };
    auto w2 = [](float a) { return (a * (a * (-3.0f * a + 3.0f) + 3.0f) + 1.0f) / 6.0f; };
    auto w3 = [](float a) { return (a * a * a) / 6.0f; };
    auto g0 = [&](float a) { return w0(a) + w1(a); };
    auto g1 = [&](float a) { return w2(a) + w3(a); };
    auto h0 = [&](float a) { return (w1(a) / g0(a)) - 1.0f; };
    auto h1 = [&](float a) { return (w3(a) / g1(a)) + 1.0f; };
    const float   g0x = g0(fx), g0y = g0(fy), g1x = g1(fx), g1y = g1(fy);
    const int32_t ix0 = (int32_t)floorf((float)ix + h0(fx) + 0.5f), iy0 = (int32_t)floorf((float)iy + h0(fy) + 0.5f);
    const int32_t ix1 = (int32_t)floorf((float)ix + h1(fx) + 0.5f), iy1 = (int32_t)floorf((float)iy + h1(fy) + 0.5f);
    const int32_t x0 = tex_border(t.wrap_u, ix0, w), y0 = tex_border(t.wrap_v, iy0, h);
    const int32_t x1 = tex_border(t.wrap_u, ix1, w), y1 = tex_border(t.wrap_v, iy1, h);
    const float4  p00 = tex_texel(t, blob, x0, y0), p10 = tex_texel(t, blob, x1, y0);
    const float4  p01 = tex_texel(t, blob, x0, y1), p11 = tex_texel(t, blob, x1, y1);
    const float   a = g0x * g0y, b = g1x * g0y, c = g0x * g1y, d = g1x * g1y;
    return Color{ (p00.x * a + p10.x * b) + (p01.x * c + p11.x * d), (p00.y * a + p10.y * b) + (p01.y * c + p11.y * d),
                  (p00.z * a + p10.z * b) + (p01.z * c + p11.z * d) };
}

// ---- core/microfacet.art: make_vndf_ggx_distribution :383-407 ------------------------------------------------
__device__ __forceinline__ Vec3 to_local(const Mat3& l, Vec3 v) { return Vec3{ dot3(l.c0, v), dot3(l.c1, v), dot3(l.c2, v) }; } // shading.art:9
__device__ __forceinline__ Vec3 to_world(const Mat3& l, Vec3 v) { return vadd(vadd(vmulf(l.c0, v.x), vmulf(l.c1, v.y)), vmulf(l.c2, v.z)); }
__device__ __forceinline__ float diff_of_prod(float a, float b, float c, float d) // core/common.art:132-137
{
    const float cd = c * d, diff = __fmaf_rn(a, b, -cd), err = __fmaf_rn(-c, d, cd);
    return diff + err;
}
__device__ __forceinline__ float g_1_smith(const Mat3& l, Vec3 w, float au, float av) // :159-177
{
    const float cosZ = dot3(l.c2, w);
    if (fabsf(cosZ) <= flt_eps) return 0.0f;
    const float kx = au * dot3(l.c0, w), ky = av * dot3(l.c1, w);
    const float a2 = kx * kx + ky * ky;
    if (a2 <= flt_eps) return 1.0f;
    const float k2 = a2 / (cosZ * cosZ);
    return 2.0f / (1.0f + sqrtf(1.0f + k2));
}
__device__ __forceinline__ float ndf_ggx(const Mat3& l, Vec3 m, float au, float av) // :191-201
{
    const float cosZ = dot3(l.c2, m), kx = dot3(l.c0, m) / au, ky = dot3(l.c1, m) / av;
    const float k = kx * kx + ky * ky + cosZ * cosZ;
    return safe_div(1.0f, flt_pi * au * av * k * k);
}
__device__ __forceinline__ float pdf_vndf_ggx(const Mat3& l, Vec3 w, Vec3 h, float au, float av) // :378-381
{
    const float cosZ = fabsf(dot3(l.c2, w));
    return safe_div(g_1_smith(l, w, au, av) * fabsf(dot3(w, h)) * ndf_ggx(l, h, au, av), cosZ);
}
static __device__ __noinline__ Vec3 sample_vndf_ggx(float u0, float u1, Mat3 l, Vec3 vN, float au, float av) // :324-376
{
    const Vec3  vL = to_local(l, vN);
    const Vec3  sL = vnormalize(Vec3{ au * vL.x, av * vL.y, vL.z });
    const float sin_t = safe_sqrt(1.0f - sL.z * sL.z); // shading::sin_cos_phi
    float       sin_phi = 0.0f, cos_phi = 1.0f;
    if (fabsf(sin_t) > flt_eps) { sin_phi = sL.y / sin_t; cos_phi = sL.x / sin_t; }
    const float cos_theta = fabsf(sL.z);
    float       pcx, pcy; // square_to_concentric_disk, core/warp.art:2-22
    {
        const float a = 2.0f * u0 - 1.0f, b = 2.0f * u1 - 1.0f;
        if (a == 0.0f && b == 0.0f) { pcx = 0.0f; pcy = 0.0f; }
        else if (a * a > b * b) { const float phi = (flt_pi / 4.0f) * safe_div(b, a); pcx = cosf(phi) * a; pcy = sinf(phi) * a; }
        else { const float phi = (flt_pi / 2.0f) - (flt_pi / 4.0f) * safe_div(a, b); pcx = cosf(phi) * b; pcy = sinf(phi) * b; }
    }
    const float s  = 0.5f * (1.0f + cos_theta);
    const float y  = (1.0f - s) * safe_sqrt(1.0f - pcx * pcx) + s * pcy;
    const float z  = safe_sqrt(1.0f - y * y - pcx * pcx);
    const float st = safe_sqrt(1.0f - cos_theta * cos_theta);
    const float nm = safe_div(1.0f, sum_of_prod(st, y, cos_theta, z));
    const float slx = diff_of_prod(cos_theta, y, st, z) * nm, sly = pcx * nm;
    const float s2x = (cos_phi * slx - sin_phi * sly) * au, s2y = (sin_phi * slx + cos_phi * sly) * av;
    const Vec3  Nh  = !isfinite(s2x) ? Vec3{ 0.0f, 0.0f, 0.0f } : vnormalize(Vec3{ -s2x, -s2y, 1.0f });
    return to_world(l, Nh);
}

// ---- core/sampling.art:120-165, core/matrix.art:247-270,110-113 ------------------------------------------------
__device__ __forceinline__ Vec3 vreflect(Vec3 v, Vec3 n) { return vsub(vmulf(n, 2.0f * dot3(n, v)), v); }
static __device__ __noinline__ Vec3 ensure_valid_reflection(Vec3 Ng, Vec3 I, Vec3 N)
{
    const Vec3  R         = vreflect(I, N);
    const float threshold = fminf(0.9f * dot3(Ng, I), 0.01f);
    if (dot3(Ng, R) >= threshold) return N;
    const float NdotNg = dot3(N, Ng);
    const Vec3  X      = vnormalize(vsub(N, vmulf(Ng, NdotNg)));
    const float Ix = dot3(I, X), Iz = dot3(I, Ng), Ix2 = Ix * Ix, Iz2 = Iz * Iz, a = Ix2 + Iz2;
    const float b = safe_sqrt(Ix2 * (a - threshold * threshold)), c = Iz * threshold + a;
    const float fac = 0.5f / a, N1_z2 = fac * (b + c), N2_z2 = fac * (-b + c);
    const bool  valid1 = (N1_z2 > 1e-5f) && (N1_z2 <= (1.0f + 1e-5f)), valid2 = (N2_z2 > 1e-5f) && (N2_z2 <= (1.0f + 1e-5f));
    float nx, ny;
    if (valid1 && valid2) {
        const float n1x = safe_sqrt(1.0f - N1_z2), n1y = safe_sqrt(N1_z2), n2x = safe_sqrt(1.0f - N2_z2), n2y = safe_sqrt(N2_z2);
        const float R1 = 2.0f * (n1x * Ix + n1y * Iz) * n1y - Iz, R2 = 2.0f * (n2x * Ix + n2y * Iz) * n2y - Iz;
        const bool  valid3 = R1 >= 1e-5f, valid4 = R2 >= 1e-5f;
        const bool  first  = (valid3 && valid4) ? (R1 < R2) : (R1 > R2);
        nx = first ? n1x : n2x; ny = first ? n1y : n2y;
    } else if (valid1 || valid2) {
        const float Nz2 = valid1 ? N1_z2 : N2_z2;
        nx = safe_sqrt(1.0f - Nz2); ny = safe_sqrt(Nz2);
    } else { nx = 0.0f; ny = 1.0f; }
    return vadd(vmulf(X, nx), vmulf(Ng, ny));
}
__device__ __forceinline__ Mat3 align_vectors(Vec3 a, Vec3 b)
{
    const Vec3  ax   = cross3(b, a);
    const float cosA = dot3(a, b);
    if (cosA <= -1.0f) return Mat3{ Vec3{ -1.0f, 0.0f, 0.0f }, Vec3{ 0.0f, -1.0f, 0.0f }, Vec3{ 0.0f, 0.0f, -1.0f } };
    const float k = 1.0f / (1.0f + cosA);
    return Mat3{ Vec3{ (ax.x * ax.x * k) + cosA, (ax.y * ax.x * k) - ax.z, (ax.z * ax.x * k) + ax.y },
                 Vec3{ (ax.x * ax.y * k) + ax.z, (ax.y * ax.y * k) + cosA, (ax.z * ax.y * k) - ax.x },
                 Vec3{ (ax.x * ax.z * k) - ax.y, (ax.y * ax.z * k) + ax.x, (ax.z * ax.z * k) + cosA } };
}
__device__ __forceinline__ Mat3 mat3_matmul(const Mat3& a, const Mat3& b) { return Mat3{ mat3_mul(a, b.c0), mat3_mul(a, b.c1), mat3_mul(a, b.c2) }; }

// make_bumpmap -> make_normal_set (bsdf/map.art:36-44,64-68) with texture_dx / texture_dy (texture/common.art:28-38):
// only the shading frame changes (adjoint = false).
static __device__ __noinline__ Mat3 bump_frame(const igb200_texture* textures, const uint8_t* blob, uint32_t map, float strength, float tu,
                                               float tv, Vec3 face_normal, Vec3 ray_dir, Mat3 local)
{
    const float delta = 0.001f;
    const float t0 = tex_eval(textures, blob, map, tu, tv).r;
    const float dx = (tex_eval(textures, blob, map, tu + delta, tv).r - t0) * (1.0f / delta);
    const float dy = (tex_eval(textures, blob, map, tu, tv + delta).r - t0) * (1.0f / delta);
    const Vec3  N  = vnormalize(vsub(local.c2, vmulf(vadd(vmulf(local.c0, dx), vmulf(local.c1, dy)), strength)));
    const Vec3  nn = ensure_valid_reflection(face_normal, vneg(ray_dir), vnormalize(N));
    return mat3_matmul(align_vectors(local.c2, nn), local);
}

// ---- make_rough_base_conductor_bsdf with kd = black (bsdf/conductor.art:34-110) ----------------------------------
// core/fresnel.art:29-36
__device__ __forceinline__ float conductor_factor(float n, float k, float cos_i)
{
    const float f   = n * n + k * k;
    const float d1  = f * cos_i * cos_i;
    const float d2  = 2.0f * n * cos_i;
    const float R_s = safe_div(d1 - d2, d1 + d2);
    const float R_p = safe_div(f - d2 + cos_i * cos_i, f + d2 + cos_i * cos_i);
    return clampf((R_s * R_s + R_p * R_p) * 0.5f, 0.0f, 1.0f);
}
struct RoughParams { Mat3 local; float au, av; Color ks, eta, k; };
__device__ __forceinline__ Color rough_eval_inl(const RoughParams& r, Vec3 in_dir, Vec3 out_dir) // :53-66
{
    const Vec3  N     = r.local.c2;
    const float cos_o = fabsf(dot3(out_dir, N)), cos_i = fabsf(dot3(in_dir, N));
    if (cos_o <= flt_eps || cos_i <= flt_eps) return Color{ 0.0f, 0.0f, 0.0f };
    const Vec3  H = vnormalize(vadd(in_dir, out_dir));
    const float D = ndf_ggx(r.local, H, r.au, r.av);
    const float G = g_1_smith(r.local, in_dir, r.au, r.av) * g_1_smith(r.local, out_dir, r.au, r.av);
    const float c = fabsf(dot3(out_dir, H));
    const Color F{ conductor_factor(r.eta.r, r.k.r, c), conductor_factor(r.eta.g, r.k.g, c), conductor_factor(r.eta.b, r.k.b, c) };
    const Color IF{ 1.0f - F.r, 1.0f - F.g, 1.0f - F.b };
    const Color black{ 0.0f, 0.0f, 0.0f };
    return cmulf(cadd(cmul(black, IF), cmul(r.ks, F)), D * G / (4.0f * cos_o));
}
static __device__ __noinline__ Color rough_eval(RoughParams r, Vec3 in_dir, Vec3 out_dir) { return rough_eval_inl(r, in_dir, out_dir); }
static __device__ __noinline__ float rough_pdf(Mat3 local, float au, float av, Vec3 in_dir, Vec3 out_dir) // :70-75
{
    const Vec3  H       = vnormalize(vadd(in_dir, out_dir));
    const float cos_h_o = fabsf(dot3(out_dir, H));
    return pdf_vndf_ggx(local, out_dir, H, au, av) * safe_div(1.0f, 4.0f * cos_h_o);
}
struct RoughSample { Vec3 in_dir; float pdf; Color col; bool ok; };
static __device__ __noinline__ RoughSample rough_sample(RoughParams r, float u0, float u1, Vec3 out_dir) // :76-97
{
    RoughSample out{ Vec3{ 0.0f, 0.0f, 0.0f }, 0.0f, Color{ 0.0f, 0.0f, 0.0f }, false };
    const Vec3  n     = r.local.c2;
    const float cos_o = fabsf(dot3(out_dir, n));
    if (cos_o <= flt_eps) return out;
    const Vec3  m    = sample_vndf_ggx(u0, u1, r.local, out_dir, r.au, r.av);
    const float spdf = pdf_vndf_ggx(r.local, out_dir, m, r.au, r.av);
    if (dot3(m, m) <= flt_eps) return out;
    const Vec3  oH     = vnormalize(m);
    const Vec3  H      = signbit(dot3(oH, out_dir)) ? vneg(oH) : oH;
    const Vec3  in_dir = vreflect(out_dir, H);
    const float cos_i  = fabsf(dot3(in_dir, n));
    if (cos_i <= flt_eps) return out;
    const float cos_h_o = fabsf(dot3(out_dir, H));
    const float pdfv    = spdf * (1.0f / (4.0f * cos_h_o));
    out = RoughSample{ in_dir, pdfv, cmulf(rough_eval_inl(r, in_dir, out_dir), safe_div(1.0f, pdfv)), true };
    return out;
}

} // namespace shade
} // namespace igb
