// shade_lib.cuh -- scene access, surface elements, BSDFs and lights of the wavefront path tracer (sm_100a).
// Continuation of shade.cuh; same citation and arithmetic conventions.
#pragma once
#include "shade.cuh"

namespace igb {
namespace shade {

// Everything the shading kernels read besides the ray streams.  Pointers are device pointers.
struct ShadeScene {
    const float*           entities;   // FixTable "entities": 36 floats per entity (LoaderEntity.cpp:150-176)
    const float*           shape_data; // DynTable "shapes" data, viewed as floats
    const uint2*           shape_info; // per shape: x = offset into shape_data in floats, y = IGB200_PRIM_*
    const igb200_bsdf*     bsdfs;
    const igb200_material* materials;
    const igb200_light*    inf_lights;
    const igb200_light*    fin_lights;
    int32_t                n_inf, n_fin, n_mat, n_ent;
    igb200_camera          cam;
    igb200_technique       tech;
    float                  scene_radius; // bbox_radius(scene_bbox) * 1.01 (light/env.art:76, light/directional.art:2)
};

// ---- driver/entity.art:12-29 --------------------------------------------------------------------------
struct Entity {
    const float* row;
    __device__ __forceinline__ const float* global_mat() const { return row + 12; }
    __device__ __forceinline__ Mat3 normal_mat() const
    {
        return Mat3{ Vec3{ row[24], row[25], row[26] }, Vec3{ row[27], row[28], row[29] }, Vec3{ row[30], row[31], row[32] } };
    }
    __device__ __forceinline__ int32_t shape_id() const { return __float_as_int(row[33]); }
    __device__ __forceinline__ int32_t mat_id() const { return __float_as_int(row[34]); }
};
__device__ __forceinline__ Entity load_entity(const ShadeScene& sc, int32_t id) { return Entity{ sc.entities + 36 * (size_t)id }; }

// ---- make_trimesh_from_buffer, shapes/trimesh.art:75-91 ---------------------------------------------------
struct TriMesh {
    const float* data;
    int32_t      nf, nv, nn;
    __device__ __forceinline__ Vec3 vertex(int32_t i) const
    {
        const float4 p = __ldg(reinterpret_cast<const float4*>(data + 12) + i);
        return Vec3{ p.x, p.y, p.z };
    }
    __device__ __forceinline__ Vec3 normal(int32_t i) const
    {
        const float4 p = __ldg(reinterpret_cast<const float4*>(data + 12 + 4 * (size_t)nv) + i);
        return Vec3{ p.x, p.y, p.z };
    }
    __device__ __forceinline__ int4 tri(int32_t i) const
    {
        return __ldg(reinterpret_cast<const int4*>(data + 12 + 4 * (size_t)nv + 4 * (size_t)nn) + i);
    }
    __device__ __forceinline__ float2 tex(int32_t i) const
    {
        return __ldg(reinterpret_cast<const float2*>(data + 12 + 4 * (size_t)nv + 4 * (size_t)nn + 4 * (size_t)nf) + i);
    }
};
__device__ __forceinline__ TriMesh load_trimesh(const ShadeScene& sc, int32_t shape_id)
{
    const float* d = sc.shape_data + sc.shape_info[shape_id].x;
    const int4   h = __ldg(reinterpret_cast<const int4*>(d));
    return TriMesh{ d, h.x, h.y, h.z };
}

// make_triangle, core/triangle.art:11-26: unit normal and area of (v0, v1, v2)
__device__ __forceinline__ void make_triangle(Vec3 v0, Vec3 v1, Vec3 v2, Vec3& n, float& area)
{
    const Vec3  nn = cross3(vsub(v1, v0), vsub(v2, v0));
    const float l  = vlen(nn);
    n              = vmulf(nn, __fdiv_rn(1.0f, l));
    area           = __fdiv_rn(l, 2.0f);
}

// make_trimesh_shape::surface_element, shapes/trimesh.art:14-39
__device__ __forceinline__ Surf trimesh_surface_element(const TriMesh& m, const Entity& e, Vec3 org, Vec3 dir, float t, float u,
                                                       float v, int32_t prim)
{
    const int4 idx = m.tri(prim);
    Vec3       fn;
    float      area;
    make_triangle(xform_point(e.global_mat(), m.vertex(idx.x)), xform_point(e.global_mat(), m.vertex(idx.y)),
                  xform_point(e.global_mat(), m.vertex(idx.z)), fn, area);
    const Vec3   normal   = vnormalize(mat3_mul(e.normal_mat(), lerp2v(m.normal(idx.x), m.normal(idx.y), m.normal(idx.z), u, v)));
    const bool   entering = dot3(dir, fn) <= 0.0f;
    const float2 t0 = m.tex(idx.x), t1 = m.tex(idx.y), t2 = m.tex(idx.z);
    Surf s;
    s.is_entering = entering;
    s.point       = vadd(org, vmulf(dir, t));
    s.face_normal = entering ? fn : vneg(fn);
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = u; s.pv = v;
    s.tu = lerp2(t0.x, t1.x, t2.x, u, v);
    s.tv = lerp2(t0.y, t1.y, t2.y, u, v);
    s.local = make_orthonormal(entering ? normal : vneg(normal));
    return s;
}
// make_trimesh_shape::surface_element_for_point, shapes/trimesh.art:40-68
__device__ __forceinline__ Surf trimesh_surface_element_for_point(const TriMesh& m, const Entity& e, int32_t prim, float u, float v)
{
    const int4 idx = m.tri(prim);
    const Vec3 g0 = xform_point(e.global_mat(), m.vertex(idx.x)), g1 = xform_point(e.global_mat(), m.vertex(idx.y)),
               g2 = xform_point(e.global_mat(), m.vertex(idx.z));
    Vec3  fn;
    float area;
    make_triangle(g0, g1, g2, fn, area);
    const Vec3   normal = vnormalize(mat3_mul(e.normal_mat(), lerp2v(m.normal(idx.x), m.normal(idx.y), m.normal(idx.z), u, v)));
    const float2 t0 = m.tex(idx.x), t1 = m.tex(idx.y), t2 = m.tex(idx.z);
    Surf s;
    s.is_entering = true;
    s.point       = lerp2v(g0, g1, g2, u, v);
    s.face_normal = fn;
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = u; s.pv = v;
    s.tu = lerp2(t0.x, t1.x, t2.x, u, v);
    s.tv = lerp2(t0.y, t1.y, t2.y, u, v);
    s.local = make_orthonormal(normal);
    return s;
}
// compute_ellipsoid_area + make_sphere_shape::surface_element, shapes/sphere.art:22-29,48-67
__device__ __forceinline__ Surf sphere_surface_element(float4 sph, const Entity& e, Vec3 org, Vec3 dir, float t, float u, float v)
{
    const float* g = e.global_mat();
    const Mat3   lin{ Vec3{ g[0], g[1], g[2] }, Vec3{ g[3], g[4], g[5] }, Vec3{ g[6], g[7], g[8] } };
    const float  l1 = vlen(mat3_mul(lin, Vec3{ sph.w, 0.0f, 0.0f })), l2 = vlen(mat3_mul(lin, Vec3{ 0.0f, sph.w, 0.0f })),
                l3 = vlen(mat3_mul(lin, Vec3{ 0.0f, 0.0f, sph.w }));
    const float P    = 1.6f;
    const float area = 4.0f * flt_pi * powf(__fdiv_rn(powf(l1 * l2, P) + powf(l1 * l3, P) + powf(l2 * l3, P), 3.0f), __fdiv_rn(1.0f, P));
    const Vec3  point  = vadd(org, vmulf(dir, t));
    const Vec3  d      = vsub(point, xform_point(e.global_mat(), Vec3{ sph.x, sph.y, sph.z }));
    const Vec3  normal = vmulf(d, __fdiv_rn(1.0f, vlen(d)));
    Surf s;
    s.is_entering = true;
    s.point       = point;
    s.face_normal = normal;
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = s.tu = u;
    s.pv = s.tv = v;
    s.local = make_orthonormal(normal);
    return s;
}

// ---- BSDFs ------------------------------------------------------------------------------------------------
// core/fresnel.art:7-36
__device__ __forceinline__ float fresnel_factor(float eta, float cos_i, float cos_t)
{
    const float R_s = safe_div(eta * cos_i - cos_t, eta * cos_i + cos_t);
    const float R_p = safe_div(cos_i - eta * cos_t, cos_i + eta * cos_t);
    return clampf((R_s * R_s + R_p * R_p) * 0.5f, 0.0f, 1.0f);
}
__device__ __forceinline__ bool fresnel(float eta, float cos_i, float& cos_t_out, float& factor)
{
    const float eta2   = cos_i < 0.0f ? __fdiv_rn(1.0f, eta) : eta;
    const float cos2_t = 1.0f - (1.0f - cos_i * cos_i) * eta2 * eta2;
    if (cos2_t <= 0.0f) return false;
    const float cos_t = __fsqrt_rn(cos2_t);
    cos_t_out         = cos_i < 0.0f ? -cos_t : cos_t;
    factor            = fresnel_factor(eta2, fabsf(cos_i), cos_t);
    return true;
}
__device__ __forceinline__ float conductor_factor(float n, float k, float cos_i)
{
    const float f   = n * n + k * k;
    const float d1  = f * cos_i * cos_i;
    const float d2  = 2.0f * n * cos_i;
    const float R_s = safe_div(d1 - d2, d1 + d2);
    const float R_p = safe_div(f - d2 + cos_i * cos_i, f + d2 + cos_i * cos_i);
    return clampf((R_s * R_s + R_p * R_p) * 0.5f, 0.0f, 1.0f);
}
__device__ __forceinline__ Vec3 reflect(Vec3 v, Vec3 n) { return vsub(vmulf(n, 2.0f * dot3(n, v)), v); }
__device__ __forceinline__ Vec3 refract(Vec3 v, Vec3 n, float eta, float cos_i, float cos_t)
{
    return vsub(vmulf(n, eta * cos_i - cos_t), vmulf(v, eta));
}

struct BsdfSample { Vec3 in_dir; float pdf; Color col; float eta; };

struct Bsdf {
    igb200_bsdf d;
    Vec3        n; // surf.local.col(2)
    bool        one_sided;
    __device__ __forceinline__ bool  is_specular() const { return d.type != IGB200_BSDF_DIFFUSE; }
    __device__ __forceinline__ Color kd() const { return Color{ d.color[0], d.color[1], d.color[2] }; }
    // bsdf/diffuse.art:3-4; specular BSDFs: black / 0 (dielectric.art:3-4, conductor.art:3-4,14-15)
    __device__ __forceinline__ Color eval(Vec3 in_dir) const
    {
        if (is_specular()) return Color{ 0.0f, 0.0f, 0.0f };
        const float c = dot3(in_dir, n);
        return cmulf(kd(), (one_sided ? (c >= 0.0f ? c : 0.0f) : fabsf(c)) * flt_inv_pi);
    }
    __device__ __forceinline__ float pdf(Vec3 in_dir) const
    {
        if (is_specular()) return 0.0f;
        const float c = dot3(in_dir, n);
        return __fdiv_rn(c >= 0.0f ? c : 0.0f, flt_pi);
    }
    __device__ __forceinline__ BsdfSample sample(Rnd& rnd, Vec3 out_dir, const Surf& surf) const
    {
        switch (d.type) {
        default:
        case IGB200_BSDF_DIFFUSE: { // bsdf/diffuse.art:5-9
            const float u = rnd.next_f32();
            const float v = rnd.next_f32();
            float       pdf;
            const Vec3  l = sample_cosine_hemisphere(u, v, pdf);
            return BsdfSample{ mat3_mul(surf.local, l), pdf, kd(), 1.0f };
        }
        case IGB200_BSDF_DIELECTRIC: { // bsdf/dielectric.art:5-21, adjoint = false
            const float n1 = d.p[0], n2 = d.p[1];
            const float k     = surf.is_entering ? __fdiv_rn(n1, n2) : __fdiv_rn(n2, n1);
            const float cos_o = dot3(out_dir, n);
            float       cos_t = 0.0f, factor = 1.0f;
            if (!fresnel(k, cos_o, cos_t, factor)) { cos_t = 0.0f; factor = 1.0f; }
            if (rnd.next_f32() > factor)
                return BsdfSample{ refract(out_dir, n, k, cos_o, cos_t), 1.0f, Color{ d.color2[0], d.color2[1], d.color2[2] }, k };
            return BsdfSample{ reflect(out_dir, n), 1.0f, kd(), 1.0f };
        }
        case IGB200_BSDF_MIRROR: // bsdf/conductor.art:5-7
            return BsdfSample{ reflect(out_dir, n), 1.0f, kd(), 1.0f };
        case IGB200_BSDF_CONDUCTOR: { // bsdf/conductor.art:16-28
            const float cos_i = dot3(out_dir, n);
            const Color f{ conductor_factor(d.color2[0], d.p[0], cos_i), conductor_factor(d.color2[1], d.p[1], cos_i),
                           conductor_factor(d.color2[2], d.p[2], cos_i) };
            return BsdfSample{ reflect(out_dir, n), 1.0f, cmul(kd(), f), 1.0f };
        }
        }
    }
};

} // namespace shade
} // namespace igb
