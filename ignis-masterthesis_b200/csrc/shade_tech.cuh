// shade_tech.cuh -- lights, the uniform light selector and the path / debug technique callbacks (sm_100a).
// Continuation of shade_lib.cuh; same citation and arithmetic conventions.
#pragma once
#include "shade_lib.cuh"

namespace igb {
namespace shade {

struct DirectSample { Vec3 pos, dir; Color intensity; Pdf pdf; float cos, dist; };

__device__ __forceinline__ bool light_delta(uint32_t type)
{
    return type == IGB200_LIGHT_POINT || type == IGB200_LIGHT_SPOT || type == IGB200_LIGHT_DIRECTIONAL || type == IGB200_LIGHT_SUN;
}
__device__ __forceinline__ bool light_infinite(uint32_t type)
{
    return type == IGB200_LIGHT_ENV || type == IGB200_LIGHT_DIRECTIONAL || type == IGB200_LIGHT_ENV_TEXTURED || type == IGB200_LIGHT_ENV_FUNCTION ||
           type == IGB200_LIGHT_SUN;
}
__device__ __forceinline__ Color light_color(const igb200_light& l) { return Color{ l.color[0], l.color[1], l.color[2] }; }

// make_plane_area_emitter, light/area.art:107-232
struct SQ { Vec3 o, n; float x0, y0, z0, x1, y1, b0, b1, k, s; };
struct PlaneEmitter {
    Vec3  origin, normal, ex, ey;
    float inv_area, width, height;
    __device__ __forceinline__ explicit PlaneEmitter(const igb200_light& l)
    {
        origin = Vec3{ l.v[0], l.v[1], l.v[2] };
        const Vec3 x_axis{ l.v[3], l.v[4], l.v[5] }, y_axis{ l.v[6], l.v[7], l.v[8] };
        normal   = Vec3{ l.v[9], l.v[10], l.v[11] };
        inv_area = __fdiv_rn(1.0f, l.v[12]);
        width    = vlen(x_axis);
        height   = vlen(y_axis);
        ex       = vmulf(x_axis, __fdiv_rn(1.0f, width));
        ey       = vmulf(y_axis, __fdiv_rn(1.0f, height));
    }
    __device__ __forceinline__ SQ compute_sq(Vec3 from) const // light/area.art:117-160
    {
        const Vec3  dir = vsub(origin, from);
        const float x0 = dot3(dir, ex), y0 = dot3(dir, ey), z0_ = dot3(dir, normal);
        const float x1 = x0 + width, y1 = y0 + height;
        const bool  pos = !signbit(z0_);
        const float z0  = pos ? -z0_ : z0_;
        const Vec3  n   = pos ? vneg(normal) : normal;
        const float diff[4] = { x0 - x1, y1 - y0, x1 - x0, y0 - y1 };
        const float nz_[4]  = { y0 * diff[0], x1 * diff[1], y1 * diff[2], x0 * diff[3] };
        float       nz[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) nz[i] = __fdiv_rn(nz_[i], __fsqrt_rn(diff[i] * diff[i] * (z0 * z0) + nz_[i] * nz_[i]));
        const float g0 = acosf(clampf(-nz[0] * nz[1], -1.0f, 1.0f)), g1 = acosf(clampf(-nz[1] * nz[2], -1.0f, 1.0f));
        const float g2 = acosf(clampf(-nz[2] * nz[3], -1.0f, 1.0f)), g3 = acosf(clampf(-nz[3] * nz[0], -1.0f, 1.0f));
        const float k = 2.0f * flt_pi - g2 - g3, s = g0 + g1 - k;
        return SQ{ from, n, x0, y0, z0, x1, y1, nz[0], nz[2], k, s };
    }
    __device__ __forceinline__ void sample_direct(float ux, float uy, Vec3 from, Vec3& p, Pdf& pdf, float& weight) const // :162-204
    {
        const SQ    sq = compute_sq(from);
        const float au = __fmaf_rn(ux, sq.s, sq.k);
        const float fu = __fdiv_rn(__fmaf_rn(cosf(au), sq.b0, -sq.b1), sinf(au));
        const float cu = clampf(__fdiv_rn(copysignf(1.0f, fu), __fsqrt_rn(sum_of_prod(fu, fu, sq.b0, sq.b0))), -1.0f, 1.0f);
        const float xu = clampf(__fdiv_rn(-(cu * sq.z0), __fsqrt_rn(__fmaf_rn(-cu, cu, 1.0f))), sq.x0, sq.x1);
        const float d   = __fsqrt_rn(sum_of_prod(xu, xu, sq.z0, sq.z0));
        const float h0  = __fdiv_rn(sq.y0, __fsqrt_rn(sum_of_prod(d, d, sq.y0, sq.y0)));
        const float h1  = __fdiv_rn(sq.y1, __fsqrt_rn(sum_of_prod(d, d, sq.y1, sq.y1)));
        const float hv  = __fmaf_rn(uy, h1 - h0, h0);
        const float hv2 = hv * hv;
        const float yv  = hv2 < 1.0f - 1e-6f ? __fdiv_rn(hv * d, __fsqrt_rn(1.0f - hv2)) : sq.y1;
        p      = vadd(sq.o, vadd(vmulf(ex, xu), vadd(vmulf(ey, yv), vmulf(sq.n, sq.z0))));
        pdf    = Pdf{ safe_div(1.0f, sq.s), 0 };
        weight = sq.s;
    }
    __device__ __forceinline__ Pdf pdf_direct(Vec3 from) const { return Pdf{ safe_div(1.0f, compute_sq(from).s), 0 }; }
};

// make_sphere_area_emitter::sample_direct, light/area.art:248-273: uniform on the sphere, mirrored through the centre onto
// the half facing `from`; pdf = 2 / area, weight = area / 2
static __device__ __noinline__ SpherePoint sphere_emitter_sample(const float* entities, int32_t entity, Vec3 origin, float radius, float u, float v,
                                                                 Vec3 from, float* inv_area)
{
    const Entity e{ entities + 36 * (size_t)entity };
    const Vec3   glb_org = xform_point(e.global_mat(), origin);
    SpherePoint  sp      = sphere_point_for_normal(equal_area_square_to_sphere(u, v), origin, radius, e);
    const Vec3   os = vsub(from, glb_org), ps = vsub(from, sp.point);
    if (!(dot3(ps, ps) <= dot3(os, os))) {
        const Vec3 op = vsub(sp.point, glb_org);
        const Vec3 np = vsub(sp.point, vmulf(op, 2.0f));
        sp            = sphere_point_for_normal(vnormalize(vsub(np, glb_org)), origin, radius, e);
    }
    *inv_area = __fdiv_rn(1.0f, ellipsoid_area(radius, e));
    return sp;
}
// make_shape_area_emitter::sample, light/area.art:48-58 (sample_triangle core/sampling.art:33-35)
__device__ __forceinline__ Surf shape_emitter_sample(const ShadeScene& sc, const igb200_light& l, float ux_, float uy, float& pdfv,
                                                    float& weight)
{
    const Entity  e = load_entity(sc, l.entity);
    const TriMesh m = load_trimesh(sc, e.shape_id());
    const float   ux = ux_ * (float)m.nf;
    const int32_t f  = min((int32_t)ux, m.nf - 1);
    float         u = ux - (float)f, v = uy;
    if (u + v > 1.0f) { u = 1.0f - u; v = 1.0f - v; }
    const Surf s = trimesh_surface_element_for_point(m, e, f, u, v);
    pdfv         = __fdiv_rn(s.inv_area, (float)m.nf);
    weight       = __fdiv_rn((float)m.nf, s.inv_area);
    return s;
}

// ---- core/cdf.art on the blob layout of cdf::make_cdf_2d_from_buffer (:156-160) ---------------------------------
struct Cdf1 { // make_cdf_1d_from_buffer :70-72 + make_cdf_1d :38-68
    const float* data;
    int32_t      n;
    __device__ __forceinline__ float get(int32_t i) const { return i == 0 ? 0.0f : __ldg(data + i - 1); }
    __device__ __forceinline__ float pdf_discrete(int32_t x) const { return get(x + 1) - get(x); }
    __device__ __forceinline__ int32_t sample_discrete(float u) const
    {
        int32_t first = 0, len_ = n + 1; // interval::binary_search, core/interval.art:7-23
        while (len_ > 0) {
            const int32_t half = len_ / 2, middle = first + half;
            if (get(middle) <= u) { first = middle + 1; len_ -= half + 1; } else { len_ = half; }
        }
        return min(min(max(first - 1, 0), n), n - 1);
    }
    __device__ __forceinline__ void sample_continuous(float u, int32_t& off, float& pos, float& pdf) const
    {
        off             = sample_discrete(u);
        const float pd  = pdf_discrete(off);
        const float rem = safe_div(u - get(off), pd);
        pos             = clampf(((float)off + rem) / (float)n, 0.0f, 1.0f);
        pdf             = pd * (float)n;
    }
    __device__ __forceinline__ void pdf_continuous(float x, int32_t& off, float& pdf) const
    {
        off = min(max((int32_t)(x * (float)n), 0), n - 1);
        pdf = pdf_discrete(off) * (float)n;
    }
};
struct Cdf2 {
    const float* base;
    int32_t      sx, sy;
    __device__ __forceinline__ Cdf1 marginal() const { return Cdf1{ base, sy }; }
    __device__ __forceinline__ Cdf1 conditional(int32_t i) const { return Cdf1{ base + sy + (size_t)i * sx, sx }; }
};

// light/env.art:13-21, core/warp.art:44-57
__device__ __forceinline__ Vec3 switch_env_up(Vec3 v) { return Vec3{ v.x, v.z, v.y }; }
__device__ __forceinline__ void map_env_uv(Vec3 dir, float& u, float& v)
{
    // acos argument clamped: a direction whose length rounds to 1 + 1 ulp gives acos(1.0000001) = NaN in the reference
    // (core/warp.art:45) and poisons the pixel for good; this is the only deviation, and only where that NaN would be
    const float theta = acosf(clampf(dir.z, -1.0f, 1.0f));
    float       phi   = atan2f(dir.y, dir.x);
    if (phi < 0.0f) phi += 2.0f * flt_pi;
    u = fractf(phi / (2.0f * flt_pi) + 0.25f);
    v = 1.0f - theta / flt_pi;
}
__device__ __forceinline__ Mat3 light_transform(const igb200_light& l)
{
    return Mat3{ Vec3{ l.v[0], l.v[1], l.v[2] }, Vec3{ l.v[3], l.v[4], l.v[5] }, Vec3{ l.v[6], l.v[7], l.v[8] } };
}
__device__ __forceinline__ Vec3 mat3_left_mul(const Mat3& m, Vec3 v) { return Vec3{ dot3(m.c0, v), dot3(m.c1, v), dot3(m.c2, v) }; }

// Light::emission of an infinite light along `dir` (light/env.art:93,146-151,163-166).  The __noinline__ helpers take
// the device pointers they need, never the ShadeScene (a reference to the kernel argument would copy it to the stack).
static __device__ __noinline__ Color env_tex_emission(const igb200_texture* textures, const uint8_t* blob, const igb200_light* l, Vec3 dir)
{
    float u, v;
    map_env_uv(switch_env_up(mat3_mul(light_transform(*l), dir)), u, v);
    return cmul(light_color(*l), tex_eval(textures, blob, l->tex, u, v));
}
// cie_wmean + make_cie_sky_light / make_cie_sunny_light, light/cie.art:1-42; `dir` in the light's own frame (Y up)
static __device__ __noinline__ Color sky_function(const uint8_t* blob, const igb200_light* lp, Vec3 dir)
{
    const igb200_light& l = *lp;
    const float* p = reinterpret_cast<const float*>(blob) + l.cdf;
    const bool   variant = l.flags & IGB200_SKYFLAG_VARIANT, has_ground = l.flags & IGB200_SKYFLAG_HAS_GROUND;
    const Color  scale{ __ldg(p + 0), __ldg(p + 1), __ldg(p + 2) }, zenith{ __ldg(p + 3), __ldg(p + 4), __ldg(p + 5) },
        ground{ __ldg(p + 6), __ldg(p + 7), __ldg(p + 8) };
    const float cos_theta = dir.y;
    if (!has_ground && cos_theta < 0.0f) return Color{ 0.0f, 0.0f, 0.0f };
    const float a  = powf(cos_theta + 1.01f, 10.0f);
    const float f1 = __fdiv_rn(a * a, a * a + 1.0f), f2 = __fdiv_rn(1.0f, a * a + 1.0f);
    if (l.tex == IGB200_SKY_CIE) {
        const float c1 = variant ? __fdiv_rn(1.0f + 2.0f * cos_theta, 3.0f) : 1.0f, c2 = variant ? 0.777777777f : 1.0f;
        return cadd(cmulf(cmulf(zenith, c1), f1), cmulf(cmulf(ground, __ldg(p + 10) * c2), f2));
    }
    const Vec3  sun{ __ldg(p + 12), __ldg(p + 13), __ldg(p + 14) };
    if (l.tex == IGB200_SKY_PEREZ) { // calc_perez + make_perez_light, light/cie.art:45-57 (zenith slot = lum)
        const float a = __ldg(p + 9), b = __ldg(p + 10), c_ = __ldg(p + 11), d = __ldg(p + 15), e = __ldg(p + 0);
        const float sun_c = clampf(dot3(dir, sun), -1.0f, 1.0f);
        const float sun_a = acosf(sun_c);
        const float A = 1.0f + a * expf(__fdiv_rn(b, fmaxf(0.01f, cos_theta)));
        const float B = 1.0f + c_ * expf(d * sun_a) + e * sun_c * sun_c;
        return cadd(cmulf(cmulf(zenith, A * B), f1), cmulf(ground, f2));
    }
    const float cos_gamma = dot3(dir, sun);
    const float gamma     = acosf(clampf(cos_gamma, -1.0f, 1.0f));
    float       c1;
    if (variant) {
        c1 = (0.91f + 10.0f * expf(-3.0f * gamma) + 0.45f * cos_gamma * cos_gamma)
             * (cos_theta >= 0.01f ? 1.0f - expf(__fdiv_rn(-0.32f, cos_theta)) : 1.0f);
    } else {
        const float theta = acosf(clampf(cos_theta, -1.0f, 1.0f)), stheta = acosf(clampf(sun.y, -1.0f, 1.0f));
        c1 = __fdiv_rn((1.35f * sinf(5.631f - 3.59f * theta) + 3.12f) * sinf(4.396f - 2.6f * stheta) + 6.37f - theta, 2.326f)
             * expf(gamma * (-0.563f) * ((2.629f - theta) * (1.562f - stheta) + 0.812f));
    }
    return cmul(scale, cadd(cmulf(cmulf(zenith, __ldg(p + 9) * c1), f1), cmulf(cmulf(ground, __ldg(p + 10) * __ldg(p + 11)), f2)));
}
// half = !has_ground for the CIE skies (light/cie.art:19,42), always false for the Perez sky (:57); light/env.art:104-108
__device__ __forceinline__ bool sky_is_hemi(const igb200_light& l) { return !(l.flags & IGB200_SKYFLAG_HAS_GROUND) && l.tex != IGB200_SKY_PEREZ; }
template <bool EXT>
__device__ __forceinline__ Color env_emission(const ShadeScene& sc, const igb200_light& l, Vec3 dir)
{
    if (EXT && l.type == IGB200_LIGHT_ENV_FUNCTION) { // light/env.art:49-57 (hemi), :94 (spherical)
        const Vec3 local = mat3_mul(light_transform(l), dir);
        if (sky_is_hemi(l) && !(local.y > flt_eps)) return Color{ 0.0f, 0.0f, 0.0f };
        return sky_function(sc.blob, &l, local);
    }
    return EXT && l.type == IGB200_LIGHT_ENV_TEXTURED ? env_tex_emission(sc.textures, sc.blob, &l, dir) : light_color(l);
}
__device__ __forceinline__ Cdf2 load_cdf2(const uint8_t* blob, uint32_t off4)
{
    const uint32_t* h = reinterpret_cast<const uint32_t*>(blob) + off4;
    return Cdf2{ reinterpret_cast<const float*>(h + 4), (int32_t)h[0], (int32_t)h[1] };
}
// Light::pdf_direct of an infinite light, solid angle (light/env.art:96,121-127,152)
static __device__ __noinline__ float env_tex_pdf(const uint8_t* blob, const igb200_light* l, Vec3 dir)
{
    const Vec3  ldir     = switch_env_up(mat3_mul(light_transform(*l), dir));
    const float sinTheta = safe_sqrt(1.0f - ldir.z * ldir.z);
    float       u, v;
    map_env_uv(ldir, u, v);
    const Cdf2 cdf = load_cdf2(blob, l->cdf);
    int32_t    oy, ox;
    float      p1, p2;
    cdf.marginal().pdf_continuous(v, oy, p1);
    cdf.conditional(oy).pdf_continuous(u, ox, p2);
    return safe_div(p1 * p2, sinTheta * flt_pi * flt_pi * 2.0f);
}
template <bool EXT>
__device__ __forceinline__ float env_pdf(const ShadeScene& sc, const igb200_light& l, Vec3 dir)
{
    if (EXT && l.type == IGB200_LIGHT_ENV_FUNCTION && sky_is_hemi(l)) { // light/env.art:58-61: cosine_hemisphere_pdf
        const float y = mat3_mul(light_transform(l), dir).y;
        return y > flt_eps ? __fdiv_rn(y, flt_pi) : 0.0f;
    }
    if (!EXT || l.type != IGB200_LIGHT_ENV_TEXTURED || l.cdf == IGB200_NO_CDF) return __fdiv_rn(1.0f, 4.0f * flt_pi);
    return env_tex_pdf(sc.blob, &l, dir);
}

struct EnvSample { Vec3 dir; Color intensity; float pdf; };
static __device__ __noinline__ EnvSample env_tex_sample_direct(const igb200_texture* textures, const uint8_t* blob, const igb200_light* lp,
                                                               float u0, float u1)
{
    const igb200_light& l = *lp;
    const Mat3 tr = light_transform(l);
    EnvSample  out;
    if (l.cdf == IGB200_NO_CDF) { // make_environment_light -> function_spherical, light/env.art:163-166,82-86
        const Vec3 d = equal_area_square_to_sphere(u0, u1);
        out.pdf      = __fdiv_rn(1.0f, 4.0f * flt_pi);
        float tu, tv;
        map_env_uv(switch_env_up(mat3_mul(tr, d)), tu, tv);
        out.intensity = cmulf(cmul(light_color(l), tex_eval(textures, blob, l.tex, tu, tv)), __fdiv_rn(1.0f, out.pdf));
        out.dir       = d;
    } else { // make_environment_light_textured::sample_dir, light/env.art:115-133 (radiance NOT scaled, as written)
        const Cdf2  cdf = load_cdf2(blob, l.cdf);
        int32_t     oy, ox;
        float       py, px, p1, p2;
        cdf.marginal().sample_continuous(u1, oy, py, p1);
        cdf.conditional(oy).sample_continuous(u0, ox, px, p2);
        Color       in    = tex_eval(textures, blob, l.tex, px, py);
        if (l.flags & IGB200_LIGHTFLAG_SCALED_NEE) in = cmul(light_color(l), in);
        const float theta = (1.0f - py) * flt_pi, phi = (px - 0.25f) * 2.0f * flt_pi;
        const float st = sinf(theta), ct = cosf(theta);
        const Vec3  d0{ st * cosf(phi), st * sinf(phi), ct }; // dir_from_spherical
        const float sinTheta = safe_sqrt(1.0f - d0.z * d0.z);
        out.pdf       = safe_div(p1 * p2, sinTheta * flt_pi * flt_pi * 2.0f);
        out.intensity = cmulf(in, 1.0f / out.pdf);
        out.dir       = mat3_left_mul(tr, switch_env_up(d0));
    }
    return out;
}

// Light::sample_direct of every supported light
template <bool EXT>
__device__ __forceinline__ DirectSample light_sample_direct(const ShadeScene& sc, const igb200_light& l, Rnd& rnd, const Surf& from)
{
    switch (l.type) {
    default:
    case IGB200_LIGHT_POINT: { // light/point.art:3-8
        const Vec3  pos{ l.v[0], l.v[1], l.v[2] };
        const Vec3  dir_ = vsub(pos, from.point);
        const float dist = vlen(dir_);
        return DirectSample{ pos, vmulf(dir_, safe_div(1.0f, dist)), light_color(l), Pdf{ 1.0f, 1 }, 1.0f, dist };
    }
    case IGB200_LIGHT_SPOT: { // light/spot.art:8-40
        const Vec3  pos{ l.v[0], l.v[1], l.v[2] }, dir{ l.v[3], l.v[4], l.v[5] };
        const float cosCutoff = cosf(l.v[6]), cosFalloff = cosf(l.v[7]);
        const float blendRange = cosFalloff - cosCutoff;
        const Vec3  out_dir_   = vsub(pos, from.point);
        const float dist       = vlen(out_dir_);
        const Vec3  out_dir    = vmulf(out_dir_, safe_div(1.0f, dist));
        const float cos_angle  = dot3(vneg(out_dir), dir);
        float       factor;
        if (blendRange <= flt_eps) {
            factor = cos_angle <= cosCutoff ? 0.0f : 1.0f;
        } else {
            const float x = clampf(__fdiv_rn(cos_angle - cosCutoff, blendRange), 0.0f, 1.0f);
            factor        = x * x * (3.0f - 2.0f * x);
        }
        const float cos = -dot3(out_dir, dir);
        return DirectSample{ pos, out_dir, cmulf(light_color(l), factor), Pdf{ cos_angle > cosCutoff ? 1.0f : 0.0f, 1 }, cos, dist };
    }
    case IGB200_LIGHT_DIRECTIONAL: { // light/directional.art:6
        const Vec3 dir{ l.v[0], l.v[1], l.v[2] };
        return DirectSample{ vadd(from.point, vmulf(dir, -sc.scene_radius)), vneg(dir), light_color(l), Pdf{ 1.0f, 0 }, 1.0f,
                             sc.scene_radius };
    }
    case IGB200_LIGHT_SUN: { // make_sun_light::sample_direct, light/sun.art:11-15 (sample_uniform_cone, core/sampling.art:106-116)
        if (!EXT) break;
        const Vec3  dir{ l.v[0], l.v[1], l.v[2] };
        const float cos_angle  = l.v[3];
        const float sun_radius = __fdiv_rn(__fsqrt_rn(1.0f - cos_angle * cos_angle), cos_angle), sun_area = flt_pi * sun_radius * sun_radius;
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        float       px, py; // square_to_concentric_disk, core/warp.art:2-22
        {
            const float a = 2.0f * u - 1.0f, b = 2.0f * v - 1.0f;
            if (a == 0.0f && b == 0.0f) { px = 0.0f; py = 0.0f; }
            else if (a * a > b * b) { const float phi = __fdiv_rn(flt_pi, 4.0f) * safe_div(b, a); px = cosf(phi) * a; py = sinf(phi) * a; }
            else { const float phi = __fdiv_rn(flt_pi, 2.0f) - __fdiv_rn(flt_pi, 4.0f) * safe_div(a, b); px = cosf(phi) * b; py = sinf(phi) * b; }
        }
        const float c1 = 1.0f - cos_angle, n2 = px * px + py * py;
        const float z  = cos_angle + c1 * (1.0f - n2), sc = safe_sqrt(c1 * (2.0f - c1 * n2));
        const float den = 2.0f * flt_pi * (1.0f - cos_angle);
        const float pdf = fabsf(den) <= flt_eps ? 1.0f : __fdiv_rn(1.0f, den); // uniform_cone_pdf: safe_div_one
        const Vec3  ndir = mat3_mul(make_orthonormal(dir), Vec3{ px * sc, py * sc, z });
        return DirectSample{ Vec3{ 0.0f, 0.0f, 0.0f }, vneg(ndir), cmulf(light_color(l), __fdiv_rn(1.0f, sun_area * pdf)), Pdf{ 1.0f, 0 }, z,
                             __int_as_float(0x7f800000) };
    }
    case IGB200_LIGHT_ENV: { // make_environment_light_function_spherical::sample_direct, light/env.art:82-86
        const float u   = rnd.next_f32();
        const float v   = rnd.next_f32();
        const Vec3  d   = equal_area_square_to_sphere(u, v);
        const float pdf = __fdiv_rn(1.0f, 4.0f * flt_pi);
        return DirectSample{ vadd(from.point, vmulf(d, sc.scene_radius)), d, cmulf(light_color(l), __fdiv_rn(1.0f, pdf)), Pdf{ pdf, 0 },
                             1.0f, sc.scene_radius };
    }
    case IGB200_LIGHT_ENV_FUNCTION: {
        if (!EXT) break;
        const Mat3  tr = light_transform(l);
        const float u  = rnd.next_f32();
        const float v  = rnd.next_f32();
        if (sky_is_hemi(l)) { // make_environment_light_function_hemi::sample_direct, light/env.art:32-38
            float       pdf;
            const Vec3  dir = switch_env_up(sample_cosine_hemisphere(u, v, pdf));
            const Color in  = cmulf(sky_function(sc.blob, &l, dir), __fdiv_rn(1.0f, pdf));
            const Vec3  gd  = mat3_left_mul(tr, dir);
            return DirectSample{ vadd(from.point, vmulf(gd, sc.scene_radius)), gd, in, Pdf{ pdf, 0 }, 1.0f, sc.scene_radius };
        }
        const Vec3  d   = equal_area_square_to_sphere(u, v); // _spherical::sample_direct, light/env.art:82-86
        const float pdf = __fdiv_rn(1.0f, 4.0f * flt_pi);
        return DirectSample{ vadd(from.point, vmulf(d, sc.scene_radius)), d, cmulf(sky_function(sc.blob, &l, mat3_mul(tr, d)), __fdiv_rn(1.0f, pdf)),
                             Pdf{ pdf, 0 }, 1.0f, sc.scene_radius };
    }
    case IGB200_LIGHT_ENV_TEXTURED: {
        if (!EXT) break;
        const float     u0 = rnd.next_f32();
        const float     u1 = rnd.next_f32();
        const EnvSample es = env_tex_sample_direct(sc.textures, sc.blob, &l, u0, u1);
        return DirectSample{ vadd(from.point, vmulf(es.dir, sc.scene_radius)), es.dir, es.intensity, Pdf{ es.pdf, 0 }, 1.0f, sc.scene_radius };
    }
    case IGB200_LIGHT_AREA_SHAPE:
    case IGB200_LIGHT_AREA_SPHERE:
    case IGB200_LIGHT_AREA_PLANE: { // make_area_light::sample_direct, light/area.art:12-26
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        Vec3  to_point, to_normal;
        Pdf   pdf{ 1.0f, 1 };
        float weight;
        if (l.type == IGB200_LIGHT_AREA_PLANE) {
            const PlaneEmitter pe(l);
            pe.sample_direct(u, v, from.point, to_point, pdf, weight);
            to_normal = pe.normal;
        } else if (EXT && l.type == IGB200_LIGHT_AREA_SPHERE) {
            float             inv_area;
            const SpherePoint sp = sphere_emitter_sample(sc.entities, l.entity, Vec3{ l.v[0], l.v[1], l.v[2] }, l.v[3], u, v, from.point, &inv_area);
            to_point  = sp.point;
            to_normal = sp.normal;
            pdf       = Pdf{ 2.0f * inv_area, 1 };
            weight    = __fdiv_rn(1.0f, 2.0f * inv_area);
        } else {
            float      pdfv;
            const Surf to = shape_emitter_sample(sc, l, u, v, pdfv, weight);
            to_point      = to.point;
            to_normal     = to.face_normal;
            pdf           = Pdf{ pdfv, 1 };
        }
        const Vec3  dir_ = vsub(to_point, from.point);
        const float dist = vlen(dir_);
        const Vec3  dir  = vmulf(dir_, safe_div(1.0f, dist));
        const float cos  = dot3(dir, to_normal) * (from.is_entering ? -1.0f : 1.0f);
        return DirectSample{ to_point, dir, cmulf(light_color(l), weight), pdf, cos, dist };
    }
    }
    return DirectSample{ from.point, Vec3{ 0.0f, 0.0f, 1.0f }, Color{ 0.0f, 0.0f, 0.0f }, Pdf{ 1.0f, 0 }, 0.0f, 0.0f }; // not reached
}

// Light::pdf_direct(ray, surf) for lights a path can hit (area) -- light/area.art:83,206-210,60-68
__device__ __forceinline__ Pdf light_pdf_direct(const ShadeScene& sc, const igb200_light& l, Vec3 ray_org, const Surf& surf)
{
    if (l.type == IGB200_LIGHT_AREA_PLANE) return PlaneEmitter(l).pdf_direct(ray_org);
    if (l.type == IGB200_LIGHT_AREA_SPHERE) // light/area.art:280-282
        return Pdf{ __fdiv_rn(2.0f, ellipsoid_area(l.v[3], load_entity(sc, l.entity))), 1 };
    if (l.type == IGB200_LIGHT_AREA_SHAPE) {
        float pdfv, w;
        shape_emitter_sample(sc, l, surf.pu, surf.pv, pdfv, w); // prim_coords reinterpreted as the sampling square
        return Pdf{ pdfv, 1 };
    }
    return Pdf{ 1.0f, 0 };
}

// make_uniform_light_selector / make_cdf_light_selector / make_hierarchy_light_selector, light/light_selector.art:19-110
struct HEntry { Vec3 pos, dir; float flux; int32_t id; bool has_dir, is_leaf; };
__device__ __forceinline__ const float* hier_nodes(const uint8_t* blob, uint32_t hierarchy, int32_t n_fin)
{
    return reinterpret_cast<const float*>(blob) + hierarchy + ((n_fin + 3) / 4) * 4;
}
__device__ __forceinline__ HEntry hier_load(const float* nodes, int32_t id) // light_hierarchy.art:24-38
{
    const float4  e1 = __ldg(reinterpret_cast<const float4*>(nodes) + 2 * id), e2 = __ldg(reinterpret_cast<const float4*>(nodes) + 2 * id + 1);
    const int32_t index = __float_as_int(e2.w);
    return HEntry{ Vec3{ e1.x, e1.y, e1.z }, Vec3{ e2.x, e2.y, e2.z }, fabsf(e1.w), index < 0 ? -index - 1 : index, !signbit(e1.w), index >= 0 };
}
__device__ __forceinline__ float hier_cost(const HEntry& e, Vec3 pos) // get_entry_cost :40-52
{
    const Vec3  cdir  = vsub(e.pos, pos);
    const float dist2 = dot3(cdir, cdir);
    const float cos_d = e.has_dir ? fabsf(dot3(e.dir, vnormalize(cdir))) : 1.0f;
    return safe_div(e.flux * cos_d, dist2);
}
__device__ __forceinline__ float hier_left_prop(const HEntry& l, const HEntry& r, Vec3 pos) // :54-58
{
    return 1.0f / (1.0f + hier_cost(r, pos) / hier_cost(l, pos));
}
struct HierPick { int32_t id; float pdf; uint32_t counter; };
static __device__ __noinline__ HierPick hier_sample(const uint8_t* blob, uint32_t hierarchy, int32_t n_fin, Rnd rnd, Vec3 pos) // :65-78
{
    float pdf = 1.0f;
    if (n_fin == 1) return HierPick{ 0, pdf, rnd.counter };
    const float* nodes = hier_nodes(blob, hierarchy, n_fin);
    HEntry       e     = hier_load(nodes, 0);
    while (!e.is_leaf) {
        const HEntry l = hier_load(nodes, e.id), r = hier_load(nodes, e.id + 1);
        const float  prop = hier_left_prop(l, r, pos);
        const bool   left = rnd.next_f32() < prop;
        e                 = left ? l : r;
        pdf *= left ? prop : 1.0f - prop;
    }
    return HierPick{ e.id, pdf, rnd.counter };
}
static __device__ __noinline__ float hier_pdf(const uint8_t* blob, uint32_t hierarchy, int32_t n_fin, int32_t light_id, Vec3 pos) // :80-96
{
    if (n_fin == 1) return 1.0f;
    const float* nodes = hier_nodes(blob, hierarchy, n_fin);
    uint32_t     code  = __ldg(reinterpret_cast<const uint32_t*>(blob) + hierarchy + light_id);
    float        pdf   = 1.0f;
    HEntry       e     = hier_load(nodes, 0);
    while (!e.is_leaf) {
        const HEntry l = hier_load(nodes, e.id), r = hier_load(nodes, e.id + 1);
        const float  prop = hier_left_prop(l, r, pos);
        const bool   left = (code & 0x1u) == 0u;
        e                 = left ? l : r;
        pdf *= left ? prop : 1.0f - prop;
        code >>= 1;
    }
    return pdf;
}
template <bool EXT>
__device__ __forceinline__ bool selector_hierarchical(const ShadeScene& sc) { return EXT && sc.selector == IGB200_SELECTOR_HIERARCHY && sc.n_fin != 0; }
template <bool EXT>
__device__ __forceinline__ bool selector_cdf(const ShadeScene& sc) { return EXT && sc.selector == IGB200_SELECTOR_CDF && sc.n_fin != 0; }
__device__ __forceinline__ Cdf1 selector_cdf_table(const ShadeScene& sc) // cdf::make_cdf_1d_from_buffer(buffer, finite_lights.count, 0)
{
    return Cdf1{ reinterpret_cast<const float*>(sc.blob) + sc.hierarchy, sc.n_fin };
}
__device__ __forceinline__ float selector_uniform_pdf(const ShadeScene& sc)
{
    const int32_t n = sc.n_inf + sc.n_fin;
    return n == 0 ? 1.0f : __fdiv_rn(1.0f, (float)n);
}
// pdf of the finite-light part of the non-uniform selectors (hierarchy: position dependent; cdf: flux table)
template <bool EXT>
__device__ __forceinline__ float selector_finite_pdf(const ShadeScene& sc, int32_t id, Vec3 pos)
{
    if (selector_cdf<EXT>(sc)) return selector_cdf_table(sc).pdf_discrete(id);
    return hier_pdf(sc.blob, sc.hierarchy, sc.n_fin, id, pos);
}
template <bool EXT>
__device__ __forceinline__ float selector_pdf(const ShadeScene& sc, bool infinite, int32_t id, Vec3 pos)
{
    if (!selector_hierarchical<EXT>(sc) && !selector_cdf<EXT>(sc)) return selector_uniform_pdf(sc);
    if (sc.n_inf == 0) return selector_finite_pdf<EXT>(sc, id, pos);
    return infinite ? (1.0f / (float)sc.n_inf) * 0.5f : selector_finite_pdf<EXT>(sc, id, pos) * (1.0f - 0.5f);
}
template <bool EXT>
__device__ __forceinline__ int32_t selector_finite_sample(const ShadeScene& sc, Rnd& rnd, Vec3 pos, float& pdf)
{
    if (selector_cdf<EXT>(sc)) { // make_cdf_light_selector: sampler.sample_discrete(rnd.next_f32()), light_selector.art:51,66
        const Cdf1    cdf = selector_cdf_table(sc);
        const int32_t off = cdf.sample_discrete(rnd.next_f32());
        pdf               = cdf.pdf_discrete(off);
        return off;
    }
    const HierPick hp = hier_sample(sc.blob, sc.hierarchy, sc.n_fin, rnd, pos);
    rnd.counter       = hp.counter;
    pdf               = hp.pdf;
    return hp.id;
}
template <bool EXT>
__device__ __forceinline__ const igb200_light& selector_sample(const ShadeScene& sc, Rnd& rnd, Vec3 pos, float& pdf)
{
    if (!selector_hierarchical<EXT>(sc) && !selector_cdf<EXT>(sc)) {
        pdf              = selector_uniform_pdf(sc);
        const int32_t n  = sc.n_inf + sc.n_fin;
        const int32_t id = n <= 1 ? 0 : rnd.next_i32(0, n - 1); // pick_light_id :19-25
        return id < sc.n_inf ? sc.inf_lights[id] : sc.fin_lights[id - sc.n_inf];
    }
    if (sc.n_inf == 0) return sc.fin_lights[selector_finite_sample<EXT>(sc, rnd, pos, pdf)];
    const float q = rnd.next_f32();
    if (q < 0.5f) {
        pdf = (1.0f / (float)sc.n_inf) * 0.5f;
        return sc.inf_lights[sc.n_inf <= 1 ? 0 : rnd.next_i32(0, sc.n_inf - 1)];
    }
    float               p;
    const igb200_light& l = sc.fin_lights[selector_finite_sample<EXT>(sc, rnd, pos, p)];
    pdf                   = p * (1.0f - 0.5f);
    return l;
}

__device__ __forceinline__ Color handle_color(const igb200_technique& t, Color c) // technique/pathtracer.art:46-50
{
    if (t.clamp > 0.0f) return Color{ fminf(c.r, t.clamp), fminf(c.g, t.clamp), fminf(c.b, t.clamp) };
    return c;
}

} // namespace shade
} // namespace igb
