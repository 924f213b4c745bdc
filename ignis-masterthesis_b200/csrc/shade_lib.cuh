// shade_lib.cuh -- scene access, surface elements, BSDFs and lights of the wavefront path tracer (sm_100a).
// Continuation of shade.cuh; same citation and arithmetic conventions.
#pragma once
#include "shade_tex.cuh"

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
    const igb200_texture*  textures;
    const uint8_t*         blob;         // texels, CDFs, light hierarchy (igb200_shading::blob)
    uint32_t               selector, hierarchy;
    uint32_t               extended; // host-side: the scene uses textures / rough conductors / bump maps / textured
                                     // environments / the light hierarchy -> the shade kernel's EXT instantiation
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
    __device__ __forceinline__ const float2* tex_base() const
    {
        return reinterpret_cast<const float2*>(data + 12 + 4 * (size_t)nv + 4 * (size_t)nn + 4 * (size_t)nf);
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
    Surf s;
    s.is_entering = entering;
    s.point       = vadd(org, vmulf(dir, t));
    s.face_normal = entering ? fn : vneg(fn);
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = u; s.pv = v;
    s.tex = m.tex_base(); s.tix = idx.x; s.tiy = idx.y; s.tiz = idx.z;
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
    Surf s;
    s.is_entering = true;
    s.point       = lerp2v(g0, g1, g2, u, v);
    s.face_normal = fn;
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = u; s.pv = v;
    s.tex = m.tex_base(); s.tix = idx.x; s.tiy = idx.y; s.tiz = idx.z;
    s.local = make_orthonormal(normal);
    return s;
}
// compute_ellipsoid_area, shapes/sphere.art:22-29
__device__ __forceinline__ float ellipsoid_area(float radius, const Entity& e)
{
    const float* g = e.global_mat();
    const Mat3   lin{ Vec3{ g[0], g[1], g[2] }, Vec3{ g[3], g[4], g[5] }, Vec3{ g[6], g[7], g[8] } };
    const float  l1 = vlen(mat3_mul(lin, Vec3{ radius, 0.0f, 0.0f })), l2 = vlen(mat3_mul(lin, Vec3{ 0.0f, radius, 0.0f })),
                l3 = vlen(mat3_mul(lin, Vec3{ 0.0f, 0.0f, radius }));
    const float P    = 1.6f;
    return 4.0f * flt_pi * powf(__fdiv_rn(powf(l1 * l2, P) + powf(l1 * l3, P) + powf(l2 * l3, P), 3.0f), __fdiv_rn(1.0f, P));
}
// make_sphere_shape::surface_element, shapes/sphere.art:48-67
__device__ __forceinline__ Surf sphere_surface_element(float4 sph, const Entity& e, Vec3 org, Vec3 dir, float t, float u, float v)
{
    const float area   = ellipsoid_area(sph.w, e);
    const Vec3  point  = vadd(org, vmulf(dir, t));
    const Vec3  d      = vsub(point, xform_point(e.global_mat(), Vec3{ sph.x, sph.y, sph.z }));
    const Vec3  normal = vmulf(d, __fdiv_rn(1.0f, vlen(d)));
    Surf s;
    s.is_entering = true;
    s.point       = point;
    s.face_normal = normal;
    s.inv_area    = __fdiv_rn(1.0f, area);
    s.pu = u; s.pv = v;
    s.tex = nullptr; s.tix = s.tiy = s.tiz = 0;
    s.local = make_orthonormal(normal);
    return s;
}
// sphere_compute_surface_element_for_normal, shapes/sphere.art:31-45: the point and normal an emitter sample needs
// (prim / tex coords feed only textured emission, which this backend does not have)
struct SpherePoint { Vec3 point, normal; };
__device__ __forceinline__ SpherePoint sphere_point_for_normal(Vec3 normal, Vec3 origin, float radius, const Entity& e)
{
    return SpherePoint{ xform_point(e.global_mat(), vadd(origin, vmulf(normal, radius))), vnormalize(mat3_mul(e.normal_mat(), normal)) };
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
// conductor_factor (core/fresnel.art:29-36): shade_tex.cuh
__device__ __forceinline__ Vec3 reflect(Vec3 v, Vec3 n) { return vsub(vmulf(n, 2.0f * dot3(n, v)), v); }
__device__ __forceinline__ Vec3 refract(Vec3 v, Vec3 n, float eta, float cos_i, float cos_t)
{
    return vsub(vmulf(n, eta * cos_i - cos_t), vmulf(v, eta));
}

struct BsdfSample { Vec3 in_dir; float pdf; Color col; float eta; };

// EXT = false compiles the texture / microfacet / bump paths out (the kernel variant for scenes without them).
template <bool EXT>
struct Bsdf {
    igb200_bsdf d;
    Mat3        local; // surf.local, re-oriented by bump maps
    bool        one_sided, is_entering;
    Color       kd_;   // DIFFUSE reflectance (constant or texture), ks otherwise
    __device__ __forceinline__ float au() const { return d.q[0].f; }
    __device__ __forceinline__ float av() const { return d.q[1].f; }
    // check_if_delta_distribution, core/microfacet.art:304
    __device__ __forceinline__ bool  rough() const { return EXT && d.type == IGB200_BSDF_CONDUCTOR && !(au() <= 1e-4f || av() <= 1e-4f); }
    __device__ __forceinline__ bool  is_specular() const { return d.type != IGB200_BSDF_DIFFUSE && !rough(); }
    __device__ __forceinline__ Color kd() const { return kd_; }
    __device__ __forceinline__ Color fresnel_term(float c) const // make_rough_conductor_bsdf, bsdf/conductor.art:102-110
    {
        return Color{ conductor_factor(d.color2[0], d.p[0], c), conductor_factor(d.color2[1], d.p[1], c),
                      conductor_factor(d.color2[2], d.p[2], c) };
    }
    __device__ __forceinline__ RoughParams rough_params() const
    {
        return RoughParams{ local, au(), av(), kd_, Color{ d.color2[0], d.color2[1], d.color2[2] }, Color{ d.p[0], d.p[1], d.p[2] } };
    }
    // bsdf/diffuse.art:3-4; specular BSDFs: black / 0 (dielectric.art:3-4, conductor.art:3-4,14-15)
    __device__ __forceinline__ Color eval(Vec3 in_dir, Vec3 out_dir) const
    {
        if (rough()) return rough_eval(rough_params(), in_dir, out_dir);
        if (is_specular()) return Color{ 0.0f, 0.0f, 0.0f };
        const float c = dot3(in_dir, local.c2);
        return cmulf(kd(), (one_sided ? (c >= 0.0f ? c : 0.0f) : fabsf(c)) * flt_inv_pi);
    }
    __device__ __forceinline__ float pdf(Vec3 in_dir, Vec3 out_dir) const
    {
        if (rough()) return rough_pdf(local, au(), av(), in_dir, out_dir);
        if (is_specular()) return 0.0f;
        const float c = dot3(in_dir, local.c2);
        return __fdiv_rn(c >= 0.0f ? c : 0.0f, flt_pi);
    }
    // false = reject_bsdf_sample() (the path ends)
    __device__ __forceinline__ bool sample(Rnd& rnd, Vec3 out_dir, BsdfSample& out) const
    {
        const Vec3 n = local.c2;
        switch (d.type) {
        default:
        case IGB200_BSDF_DIFFUSE: { // bsdf/diffuse.art:5-9
            const float u = rnd.next_f32();
            const float v = rnd.next_f32();
            float       pdf;
            const Vec3  l = sample_cosine_hemisphere(u, v, pdf);
            out = BsdfSample{ mat3_mul(local, l), pdf, kd(), 1.0f };
            return true;
        }
        case IGB200_BSDF_DIELECTRIC: { // bsdf/dielectric.art:5-21, adjoint = false
            const float n1 = d.p[0], n2 = d.p[1];
            const float k     = is_entering ? __fdiv_rn(n1, n2) : __fdiv_rn(n2, n1);
            const float cos_o = dot3(out_dir, n);
            float       cos_t = 0.0f, factor = 1.0f;
            if (!fresnel(k, cos_o, cos_t, factor)) { cos_t = 0.0f; factor = 1.0f; }
            if (rnd.next_f32() > factor)
                out = BsdfSample{ refract(out_dir, n, k, cos_o, cos_t), 1.0f, Color{ d.color2[0], d.color2[1], d.color2[2] }, k };
            else
                out = BsdfSample{ reflect(out_dir, n), 1.0f, kd(), 1.0f };
            return true;
        }
        case IGB200_BSDF_THIN_DIELECTRIC: { // make_thin_dielectric_bsdf, bsdf/dielectric.art:30-45
            const float k     = __fdiv_rn(d.p[0], d.p[1]); // always from outside to inside
            const float cos_o = fabsf(dot3(out_dir, n));
            float       cos_t = 0.0f, fterm = 1.0f;
            if (!fresnel(k, cos_o, cos_t, fterm)) fterm = 1.0f; // math::fresnel_dielectric, core/math.art:120-123
            const float F = fterm + __fdiv_rn((1.0f - fterm) * fterm, fterm + 1.0f);
            if (rnd.next_f32() > F)
                out = BsdfSample{ vneg(out_dir), 1.0f, Color{ d.color2[0], d.color2[1], d.color2[2] }, 1.0f };
            else
                out = BsdfSample{ vnormalize(reflect(out_dir, n)), 1.0f, kd(), 1.0f };
            return true;
        }
        case IGB200_BSDF_MIRROR: // bsdf/conductor.art:5-7
            out = BsdfSample{ reflect(out_dir, n), 1.0f, kd(), 1.0f };
            return true;
        case IGB200_BSDF_CONDUCTOR: {
            if (rough()) { // make_rough_base_conductor_bsdf::sample; the two draws of sample_vndf_ggx_11 happen here
                const float cos_o = fabsf(dot3(out_dir, n));
                if (cos_o <= flt_eps) return false;
                const float       u0 = rnd.next_f32();
                const float       u1 = rnd.next_f32();
                const RoughSample rs = rough_sample(rough_params(), u0, u1, out_dir);
                out                  = BsdfSample{ rs.in_dir, rs.pdf, rs.col, 1.0f };
                return rs.ok;
            }
            const float cos_i = dot3(out_dir, n); // make_pure_conductor_bsdf, bsdf/conductor.art:16-28
            out = BsdfSample{ reflect(out_dir, n), 1.0f, cmul(kd(), fresnel_term(cos_i)), 1.0f };
            return true;
        }
        }
    }
};

// Resolves material -> BSDF: bump-map chains re-orient the frame, textures replace constant colours
// (what the generated `bsdf_<id>` shader closures do, runtime/bsdf/MapBSDF.cpp, DiffuseBSDF.cpp).
template <bool EXT>
__device__ __forceinline__ Bsdf<EXT> make_bsdf(const ShadeScene& sc, int32_t id, const Surf& surf, Vec3 ray_dir, bool one_sided)
{
    Bsdf<EXT> out;
    out.local = surf.local;
    out.d     = sc.bsdfs[id];
    for (int guard = 0; EXT && guard < 8 && out.d.type == IGB200_BSDF_BUMPMAP; ++guard) {
        float tu, tv;
        surf_tex_coords(surf, tu, tv);
        out.local = bump_frame(sc.textures, sc.blob, out.d.q[0].u - 1u, out.d.p[0], tu, tv, surf.face_normal, ray_dir, out.local);
        out.d = sc.bsdfs[(int32_t)out.d.q[1].u - 1];
    }
    out.one_sided   = one_sided;
    out.is_entering = surf.is_entering;
    out.kd_         = Color{ out.d.color[0], out.d.color[1], out.d.color[2] };
    if (EXT && out.d.type == IGB200_BSDF_DIFFUSE && out.d.q[0].u != 0u) {
        float tu, tv;
        surf_tex_coords(surf, tu, tv);
        out.kd_ = tex_eval(sc.textures, sc.blob, out.d.q[0].u - 1u, tu, tv);
    }
    return out;
}

} // namespace shade
} // namespace igb
