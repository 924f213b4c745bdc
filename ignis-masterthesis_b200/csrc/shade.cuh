// shade.cuh -- device-side shading library of the wavefront path tracer (sm_100a).
//
// Replaces the slice of the reference's Artic shading library the supported scenes need;
// each function names the Artic function it implements (paths relative to src/artic).  The arithmetic contract
// is the one of trace_core.cuh / oracle.h: vec3_dot is an explicit fma chain (core/vector.art:99), everything
// else is un-contracted IEEE f32 (this file is compiled with -fmad=false).
#pragma once
#include "trace_core.cuh"

namespace igb {
namespace shade {

using core::Vec3;
using core::dot3;
using core::cross3;
using core::dot4;

constexpr float flt_eps    = 1.1920928955e-07f; // core/common.art
constexpr float flt_max    = 3.4028234664e+38f;
constexpr float flt_pi     = 3.14159265359f;
constexpr float flt_inv_pi = 0.31830988618379067154f;

struct Color { float r, g, b; };
__device__ __forceinline__ Color cmul(Color a, Color b) { return Color{ a.r * b.r, a.g * b.g, a.b * b.b }; }
__device__ __forceinline__ Color cmulf(Color a, float f) { return Color{ a.r * f, a.g * f, a.b * f }; }
__device__ __forceinline__ Color cadd(Color a, Color b) { return Color{ a.r + b.r, a.g + b.g, a.b + b.b }; }

__device__ __forceinline__ Vec3  vadd(Vec3 a, Vec3 b) { return Vec3{ a.x + b.x, a.y + b.y, a.z + b.z }; }
__device__ __forceinline__ Vec3  vsub(Vec3 a, Vec3 b) { return Vec3{ a.x - b.x, a.y - b.y, a.z - b.z }; }
__device__ __forceinline__ Vec3  vneg(Vec3 a) { return Vec3{ -a.x, -a.y, -a.z }; }
__device__ __forceinline__ Vec3  vmulf(Vec3 a, float f) { return Vec3{ a.x * f, a.y * f, a.z * f }; }
__device__ __forceinline__ float vlen(Vec3 a) { return __fsqrt_rn(dot3(a, a)); }
__device__ __forceinline__ Vec3  vnormalize(Vec3 a) { return vmulf(a, __fdiv_rn(1.0f, vlen(a))); }
__device__ __forceinline__ float clampf(float v, float l, float u) { return fminf(u, fmaxf(l, v)); }
__device__ __forceinline__ float safe_div(float a, float b) { return fabsf(b) <= flt_eps ? 0.0f : __fdiv_rn(a, b); }
__device__ __forceinline__ float safe_sqrt(float a) { return __fsqrt_rn(fmaxf(0.0f, a)); }
__device__ __forceinline__ float lerp2(float a, float b, float c, float k1, float k2) { return (1.0f - k1 - k2) * a + k1 * b + k2 * c; }
__device__ __forceinline__ Vec3  lerp2v(Vec3 a, Vec3 b, Vec3 c, float u, float v)
{
    return Vec3{ lerp2(a.x, b.x, c.x, u, v), lerp2(a.y, b.y, c.y, u, v), lerp2(a.z, b.z, c.z, u, v) };
}
// sum_of_prod, core/common.art:140-145
__device__ __forceinline__ float sum_of_prod(float a, float b, float c, float d)
{
    const float cd  = c * d;
    const float sum = __fmaf_rn(a, b, cd);
    const float err = __fmaf_rn(c, d, -cd);
    return sum + err;
}

// ---- core/matrix.art ------------------------------------------------------------------------------
struct Mat3 { Vec3 c0, c1, c2; };
__device__ __forceinline__ Vec3 mat3_mul(const Mat3& m, Vec3 v) // mat3x3_mul
{
    return Vec3{ dot3(Vec3{ m.c0.x, m.c1.x, m.c2.x }, v), dot3(Vec3{ m.c0.y, m.c1.y, m.c2.y }, v),
                 dot3(Vec3{ m.c0.z, m.c1.z, m.c2.z }, v) };
}
// make_orthonormal_mat3x3, core/matrix.art:20-28
__device__ __forceinline__ Mat3 make_orthonormal(Vec3 n)
{
    const float sign = copysignf(1.0f, n.z);
    const float a    = __fdiv_rn(-1.0f, sign + n.z);
    const float b    = n.x * n.y * a;
    return Mat3{ Vec3{ 1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x }, Vec3{ b, sign + n.y * n.y * a, -n.y }, n };
}
// mat3x4_transform_point over a column-major 3x4 stored as 12 consecutive floats
__device__ __forceinline__ Vec3 xform_point(const float* m, Vec3 p)
{
    return Vec3{ dot4(m[0], m[3], m[6], m[9], p.x, p.y, p.z, 1.0f), dot4(m[1], m[4], m[7], m[10], p.x, p.y, p.z, 1.0f),
                 dot4(m[2], m[5], m[8], m[11], p.x, p.y, p.z, 1.0f) };
}

// ---- core/random.art ------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t hash_combine(uint32_t h, uint32_t d)
{
    h = (h * 16777619u) ^ (d & 0xFFu);
    h = (h * 16777619u) ^ ((d >> 8) & 0xFFu);
    h = (h * 16777619u) ^ ((d >> 16) & 0xFFu);
    h = (h * 16777619u) ^ ((d >> 24) & 0xFFu);
    return h;
}
__device__ __forceinline__ uint32_t sample_tea_u32(uint32_t v0, uint32_t v1)
{
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v1;
}
__device__ __forceinline__ uint32_t create_random_seed(uint32_t sample, uint32_t iter, uint32_t frame, uint32_t x, uint32_t y,
                                                       uint32_t user)
{
    uint32_t h = 0x811C9DC5u;
    h = hash_combine(h, sample);
    h = hash_combine(h, iter);
    h = hash_combine(h, frame);
    h = hash_combine(h, x);
    h = hash_combine(h, y);
    h = hash_combine(h, user);
    return h;
}
struct Rnd { // create_random_generator_live, core/random.art:82-87; `counter` is the stream's rnd column
    uint32_t seed, counter;
    __device__ __forceinline__ uint32_t next_u32() { return sample_tea_u32(seed, counter++); }
    __device__ __forceinline__ float    next_f32() { return __uint_as_float((next_u32() & 0x7FFFFFu) | 0x3F800000u) - 1.0f; }
    __device__ __forceinline__ int32_t  next_i32(int32_t s, int32_t e) // next_u32_range, core/random.art:45-62
    {
        const uint32_t range = (uint32_t)(e - s);
        if (range == 0xFFFFFFFFu) return (int32_t)next_u32() + s;
        const uint32_t erange = range + 1u, scaling = 0xFFFFFFFFu / erange, past = erange * scaling;
        uint32_t       ret = next_u32();
        while (ret >= past) ret = next_u32();
        return (int32_t)(ret / scaling) + s;
    }
};

// ---- core/sampling.art, core/warp.art ---------------------------------------------------------------
__device__ __forceinline__ Vec3 sample_cosine_hemisphere(float u, float v, float& pdf) // core/sampling.art:62-74
{
    const float c = safe_sqrt(v), s = safe_sqrt(1.0f - v), phi = 2.0f * flt_pi * u;
    pdf           = __fdiv_rn(c, flt_pi);
    return Vec3{ s * cosf(phi), s * sinf(phi), c };
}
__device__ __forceinline__ Vec3 equal_area_square_to_sphere(float px, float py) // core/warp.art:63-91
{
    const float u = 2.0f * px - 1.0f, v = 2.0f * py - 1.0f;
    const float au = fabsf(u), av = fabsf(v);
    const float signedDistance = 1.0f - (au + av);
    const float d = fabsf(signedDistance), r = 1.0f - d;
    const float phi      = __fdiv_rn((r == 0.0f ? 1.0f : __fdiv_rn(av - au, r) + 1.0f) * flt_pi, 4.0f);
    const float cosTheta = copysignf(1.0f - r * r, signedDistance);
    const float sinTheta = safe_sqrt(2.0f - r * r) * r;
    const float cosPhi = copysignf(cosf(phi), u), sinPhi = copysignf(sinf(phi), v);
    return Vec3{ cosPhi * sinTheta, sinPhi * sinTheta, cosTheta };
}

// ---- driver/pdf.art -------------------------------------------------------------------------------------
struct Pdf {
    float value;
    int   area; // 1: area measure, 0: solid angle or delta (as_solid returns the value)
    __device__ __forceinline__ float as_solid(float cos, float dist2) const { return area ? __fdiv_rn(value * dist2, cos) : value; }
};

// ---- driver/surface_element.art ---------------------------------------------------------------------------
struct Surf {
    bool  is_entering;
    Vec3  point, face_normal;
    float inv_area;
    float pu, pv;        // prim_coords
    const float2* tex;   // tex_coords are fetched on demand (surf_tex_coords): only textured BSDFs read them.
    int32_t tix, tiy, tiz; // tex == nullptr: tex_coords = prim_coords (spheres)
    Mat3  local;
};
// SurfaceElement::tex_coords (shapes/trimesh.art:31-33, shapes/sphere.art:64)
__device__ __forceinline__ void surf_tex_coords(const Surf& s, float& tu, float& tv)
{
    if (!s.tex) { tu = s.pu; tv = s.pv; return; }
    const float2 t0 = __ldg(s.tex + s.tix), t1 = __ldg(s.tex + s.tiy), t2 = __ldg(s.tex + s.tiz);
    tu = lerp2(t0.x, t1.x, t2.x, s.pu, s.pv);
    tv = lerp2(t0.y, t1.y, t2.y, s.pu, s.pv);
}

} // namespace shade
} // namespace igb
