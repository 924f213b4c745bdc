// shade_tech.cuh -- lights, the uniform light selector and the path / debug technique callbacks (sm_100a).
// Continuation of shade_lib.cuh; same citation and arithmetic conventions.
#pragma once
#include "shade_lib.cuh"

namespace igb {
namespace shade {

struct DirectSample { Vec3 pos, dir; Color intensity; Pdf pdf; float cos, dist; };

__device__ __forceinline__ bool light_delta(uint32_t type)
{
    return type == IGB200_LIGHT_POINT || type == IGB200_LIGHT_SPOT || type == IGB200_LIGHT_DIRECTIONAL;
}
__device__ __forceinline__ bool light_infinite(uint32_t type) { return type == IGB200_LIGHT_ENV || type == IGB200_LIGHT_DIRECTIONAL; }
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

// Light::sample_direct of every supported light
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
    case IGB200_LIGHT_ENV: { // make_environment_light_function_spherical::sample_direct, light/env.art:82-86
        const float u   = rnd.next_f32();
        const float v   = rnd.next_f32();
        const Vec3  d   = equal_area_square_to_sphere(u, v);
        const float pdf = __fdiv_rn(1.0f, 4.0f * flt_pi);
        return DirectSample{ vadd(from.point, vmulf(d, sc.scene_radius)), d, cmulf(light_color(l), __fdiv_rn(1.0f, pdf)), Pdf{ pdf, 0 },
                             1.0f, sc.scene_radius };
    }
    case IGB200_LIGHT_AREA_SHAPE:
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
}

// Light::pdf_direct(ray, surf) for lights a path can hit (area) -- light/area.art:83,206-210,60-68
__device__ __forceinline__ Pdf light_pdf_direct(const ShadeScene& sc, const igb200_light& l, Vec3 ray_org, const Surf& surf)
{
    if (l.type == IGB200_LIGHT_AREA_PLANE) return PlaneEmitter(l).pdf_direct(ray_org);
    if (l.type == IGB200_LIGHT_AREA_SHAPE) {
        float pdfv, w;
        shape_emitter_sample(sc, l, surf.pu, surf.pv, pdfv, w); // prim_coords reinterpreted as the sampling square
        return Pdf{ pdfv, 1 };
    }
    return Pdf{ 1.0f, 0 };
}

// make_uniform_light_selector, light/light_selector.art:26-45
__device__ __forceinline__ float selector_pdf(const ShadeScene& sc)
{
    const int32_t n = sc.n_inf + sc.n_fin;
    return n == 0 ? 1.0f : __fdiv_rn(1.0f, (float)n);
}
__device__ __forceinline__ const igb200_light& selector_sample(const ShadeScene& sc, Rnd& rnd)
{
    const int32_t n  = sc.n_inf + sc.n_fin;
    const int32_t id = n <= 1 ? 0 : rnd.next_i32(0, n - 1); // pick_light_id :19-25
    return id < sc.n_inf ? sc.inf_lights[id] : sc.fin_lights[id - sc.n_inf];
}

__device__ __forceinline__ Color handle_color(const igb200_technique& t, Color c) // technique/pathtracer.art:46-50
{
    if (t.clamp > 0.0f) return Color{ fminf(c.r, t.clamp), fminf(c.g, t.clamp), fminf(c.b, t.clamp) };
    return c;
}

} // namespace shade
} // namespace igb
