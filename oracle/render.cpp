// TEST INFRASTRUCTURE ONLY (see oracle.h).
// CPU restatement of the reference's path-traced render iteration: ray generation, hit / miss shading, next
// event estimation, bounce sampling and film accumulation, function by function with file:line citations
// (all paths relative to the reference's src/artic unless stated otherwise).
//
// The reference's CPU driver is a wavefront over 16x16 tiles (driver/mapping_cpu.art:694-836).  Every path is
// independent given its seed (create_random_seed(sample, iter, frame, x, y, user), core/random.art:34-43) and
// the RNG counter is carried in the stream between bounces (mapping_cpu.art:478, 530-531), so the wavefront
// order only changes the order of f32 additions into the film.  This file therefore walks one path at a time
// in the per-bounce order of cpu_hit_shade (mapping_cpu.art:446-538): on_hit -> on_shadow -> on_bounce.
//
// Arithmetic contract: as in oracle.h (explicit fma only where the reference has fmaf, no contraction).
// libm's sinf/cosf/acosf/atan2f/tanf stand in for AnyDSL's math_builtins -- not bit-pinned by anything.
#include "oracle.h"
#include "scene.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace igo {
namespace {

constexpr float flt_eps    = 1.1920928955e-07f; // core/common.art constants
constexpr float flt_max    = 3.4028234664e+38f;
constexpr float flt_pi     = 3.14159265359f;
constexpr float flt_inv_pi = 0.31830988618379067154f;

constexpr uint32_t ray_flag_camera = 0x1, ray_flag_bounce = 0x4, ray_flag_shadow = 0x8; // traversal/ray.art:21-25

struct Color { float r, g, b; };
static inline Color color(float r, float g, float b) { return Color{ r, g, b }; }
static inline Color cmul(Color a, Color b) { return color(a.r * b.r, a.g * b.g, a.b * b.b); }
static inline Color cmulf(Color a, float f) { return color(a.r * f, a.g * f, a.b * f); }
static inline Color cadd(Color a, Color b) { return color(a.r + b.r, a.g + b.g, a.b + b.b); }
static const Color  black = { 0, 0, 0 };

static inline V3    mulf(V3 a, float f) { return mul(a, v3(f, f, f)); } // vec3_mulf
static inline float len(V3 a) { return sqrtf(dot(a, a)); }              // vec3_len
static inline V3    normalize(V3 a) { return mulf(a, 1.0f / len(a)); }  // vec3_normalize
static inline float fminf_(float a, float b) { return fminf(a, b); }
static inline float clampf(float v, float l, float u) { return fminf(u, fmaxf(l, v)); } // core/common.art:167
static inline float safe_div(float a, float b) { return fabsf(b) <= flt_eps ? 0.0f : a / b; }
static inline float safe_sqrt(float a) { return sqrtf(fmaxf(0.0f, a)); }
static inline float lerp2(float a, float b, float c, float k1, float k2) { return (1 - k1 - k2) * a + k1 * b + k2 * c; }
static inline V3    lerp2v(V3 a, V3 b, V3 c, float u, float v)
{
    return v3(lerp2(a.x, b.x, c.x, u, v), lerp2(a.y, b.y, c.y, u, v), lerp2(a.z, b.z, c.z, u, v));
}
// core/common.art:140-145
static inline float sum_of_prod(float a, float b, float c, float d)
{
    const float cd  = c * d;
    const float sum = fmaf(a, b, cd);
    const float err = fmaf(c, d, -cd);
    return sum + err;
}

// ---- core/matrix.art ----------------------------------------------------------------------------
struct Mat3 { V3 c0, c1, c2; }; // columns
static inline V3 mat3_mul(const Mat3& m, V3 v) // mat3x3_mul: rows dot v
{
    return v3(dot(v3(m.c0.x, m.c1.x, m.c2.x), v), dot(v3(m.c0.y, m.c1.y, m.c2.y), v), dot(v3(m.c0.z, m.c1.z, m.c2.z), v));
}
static inline V3 mat3_left_mul(const Mat3& m, V3 v) { return v3(dot(m.c0, v), dot(m.c1, v), dot(m.c2, v)); }
// make_orthonormal_mat3x3, core/matrix.art:20-28
static inline Mat3 make_orthonormal(V3 n)
{
    const float sign = copysignf(1.0f, n.z);
    const float a    = -1.0f / (sign + n.z);
    const float b    = n.x * n.y * a;
    const V3    t    = v3(1 + sign * n.x * n.x * a, sign * b, -sign * n.x);
    const V3    bt   = v3(b, sign + n.y * n.y * a, -n.y);
    return Mat3{ t, bt, n };
}
// mat3x4_transform_point / direction over a column-major 3x4 (core/matrix.art:95-98,232-233)
static inline V3 xform_point(const float* m, V3 p)
{
    return v3(dot4(m[0], m[3], m[6], m[9], p.x, p.y, p.z, 1.0f), dot4(m[1], m[4], m[7], m[10], p.x, p.y, p.z, 1.0f),
              dot4(m[2], m[5], m[8], m[11], p.x, p.y, p.z, 1.0f));
}

// ---- core/random.art ----------------------------------------------------------------------------
static inline uint32_t hash_combine(uint32_t h, uint32_t d)
{
    h = (h * 16777619u) ^ (d & 0xFF);
    h = (h * 16777619u) ^ ((d >> 8) & 0xFF);
    h = (h * 16777619u) ^ ((d >> 16) & 0xFF);
    h = (h * 16777619u) ^ ((d >> 24) & 0xFF);
    return h;
}
static inline uint32_t sample_tea_u32(uint32_t v0, uint32_t v1)
{
    uint32_t sum = 0;
    for (int i = 0; i < 4; ++i) {
        sum += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + sum) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + sum) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v1;
}
static inline uint32_t create_random_seed(int32_t sample, int32_t iter, int32_t frame, int32_t x, int32_t y, int32_t user)
{
    uint32_t h = 0x811C9DC5u;
    h = hash_combine(h, (uint32_t)sample);
    h = hash_combine(h, (uint32_t)iter);
    h = hash_combine(h, (uint32_t)frame);
    h = hash_combine(h, (uint32_t)x);
    h = hash_combine(h, (uint32_t)y);
    h = hash_combine(h, (uint32_t)user);
    return h;
}
struct Rnd { // create_random_generator_live, core/random.art:82-87
    uint32_t seed, counter;
    uint32_t next_u32() { return sample_tea_u32(seed, counter++); }
    float    next_f32()
    {
        const uint32_t x = (next_u32() & 0x7FFFFFu) | 0x3F800000u;
        float          f;
        std::memcpy(&f, &x, 4);
        return f - 1.0f;
    }
    uint32_t next_u32_range(uint32_t range) // core/random.art:45-62
    {
        const uint32_t rng_range = 0xFFFFFFFFu;
        if (rng_range == range) return next_u32();
        const uint32_t erange = range + 1, scaling = rng_range / erange, past = erange * scaling;
        uint32_t       ret = next_u32();
        while (ret >= past) ret = next_u32();
        return ret / scaling;
    }
    int32_t next_i32(int32_t s, int32_t e) { return (int32_t)next_u32_range((uint32_t)(e - s)) + s; }
};

// ---- core/sampling.art, core/warp.art ------------------------------------------------------------
struct DirSample { V3 dir; float pdf; };
static inline DirSample sample_cosine_hemisphere(float u, float v) // core/sampling.art:62-74
{
    const float c = safe_sqrt(v), s = safe_sqrt(1 - v), phi = 2 * flt_pi * u;
    return DirSample{ v3(s * cosf(phi), s * sinf(phi), c), c / flt_pi };
}
static inline V3 equal_area_square_to_sphere(float px, float py) // core/warp.art:63-91
{
    const float u = 2 * px - 1, v = 2 * py - 1;
    const float au = fabsf(u), av = fabsf(v);
    const float signedDistance = 1 - (au + av);
    const float d = fabsf(signedDistance), r = 1 - d;
    const float phi      = (r == 0 ? 1.0f : (av - au) / r + 1) * flt_pi / 4;
    const float cosTheta = copysignf(1 - r * r, signedDistance);
    const float sinTheta = safe_sqrt(2 - r * r) * r;
    const float cosPhi = copysignf(cosf(phi), u), sinPhi = copysignf(sinf(phi), v);
    return v3(cosPhi * sinTheta, sinPhi * sinTheta, cosTheta);
}

// ---- driver/pdf.art ------------------------------------------------------------------------------
enum class Measure { Solid, Area, Delta };
struct Pdf {
    float   value;
    Measure m;
    float as_solid(float cos, float dist2) const { return m == Measure::Area ? value * dist2 / cos : value; }
};
static inline Pdf area_pdf(float v) { return Pdf{ v, Measure::Area }; }
static inline Pdf solid_pdf(float v) { return Pdf{ v, Measure::Solid }; }
static inline Pdf delta_pdf() { return Pdf{ 1.0f, Measure::Delta }; }

// ---- driver/surface_element.art, driver/entity.art -------------------------------------------------
struct Surf {
    bool  is_entering;
    V3    point, face_normal;
    float inv_area;
    float pu, pv; // prim_coords
    float tu, tv; // tex_coords
    Mat3  local;
};
struct Entity {
    const float* local_mat;  // 12
    const float* global_mat; // 12
    Mat3         normal_mat;
    int32_t      shape_id, mat_id;
};
static inline Entity load_entity(const igo_scene& s, int32_t id) // driver/entity.art:12-29
{
    const float* d = s.entities.data() + 36 * (size_t)id;
    Entity       e;
    e.local_mat  = d;
    e.global_mat = d + 12;
    e.normal_mat = Mat3{ v3(d[24], d[25], d[26]), v3(d[27], d[28], d[29]), v3(d[30], d[31], d[32]) };
    std::memcpy(&e.shape_id, d + 33, 4);
    std::memcpy(&e.mat_id, d + 34, 4);
    return e;
}

// make_trimesh_from_buffer, shapes/trimesh.art:75-91
struct TriMesh {
    const float* data;
    int32_t      nf, nv, nn, nt;
    V3 vertex(int i) const { const float* p = data + 12 + 4 * (size_t)i; return v3(p[0], p[1], p[2]); }
    V3 normal(int i) const { const float* p = data + 12 + 4 * (size_t)nv + 4 * (size_t)i; return v3(p[0], p[1], p[2]); }
    void tri(int i, int32_t& a, int32_t& b, int32_t& c) const
    {
        const float* p = data + 12 + 4 * (size_t)nv + 4 * (size_t)nn + 4 * (size_t)i;
        std::memcpy(&a, p, 4); std::memcpy(&b, p + 1, 4); std::memcpy(&c, p + 2, 4);
    }
    void tex(int i, float& u, float& v) const
    {
        const float* p = data + 12 + 4 * (size_t)nv + 4 * (size_t)nn + 4 * (size_t)nf + 2 * (size_t)i;
        u = p[0]; v = p[1];
    }
};
static inline TriMesh load_trimesh(const igo_scene& s, int32_t shape_id)
{
    TriMesh m;
    m.data = reinterpret_cast<const float*>(s.shape_data.data() + s.lookups[shape_id].offset);
    std::memcpy(&m.nf, m.data, 4); std::memcpy(&m.nv, m.data + 1, 4);
    std::memcpy(&m.nn, m.data + 2, 4); std::memcpy(&m.nt, m.data + 3, 4);
    return m;
}

// make_triangle, core/triangle.art:11-26
struct Triangle { V3 n; float area; };
static inline Triangle make_triangle(V3 v0, V3 v1, V3 v2)
{
    const V3    e1 = sub(v1, v0), e2 = sub(v2, v0);
    const V3    n  = cross(e1, e2);
    const float nn = len(n);
    return Triangle{ mulf(n, 1 / nn), nn / 2 };
}

// make_trimesh_shape::surface_element, shapes/trimesh.art:14-39
static Surf trimesh_surface_element(const TriMesh& m, const Entity& e, const Ray& ray, const Hit& hit)
{
    int32_t i0, i1, i2;
    m.tri(hit.prim_id, i0, i1, i2);
    const Triangle tri = make_triangle(xform_point(e.global_mat, m.vertex(i0)), xform_point(e.global_mat, m.vertex(i1)),
                                       xform_point(e.global_mat, m.vertex(i2)));
    const V3   face_normal = tri.n;
    const V3   normal = normalize(mat3_mul(e.normal_mat, lerp2v(m.normal(i0), m.normal(i1), m.normal(i2), hit.u, hit.v)));
    const bool entering = dot(ray.dir, face_normal) <= 0;
    float      t0u, t0v, t1u, t1v, t2u, t2v;
    m.tex(i0, t0u, t0v); m.tex(i1, t1u, t1v); m.tex(i2, t2u, t2v);
    Surf s;
    s.is_entering = entering;
    s.point       = add(ray.org, mulf(ray.dir, hit.distance));
    s.face_normal = entering ? face_normal : neg(face_normal);
    s.inv_area    = 1 / tri.area;
    s.pu = hit.u; s.pv = hit.v;
    s.tu = lerp2(t0u, t1u, t2u, hit.u, hit.v);
    s.tv = lerp2(t0v, t1v, t2v, hit.u, hit.v);
    s.local = make_orthonormal(entering ? normal : neg(normal));
    return s;
}
// make_trimesh_shape::surface_element_for_point, shapes/trimesh.art:40-68
static Surf trimesh_surface_element_for_point(const TriMesh& m, const Entity& e, int32_t prim_id, float u, float v)
{
    int32_t i0, i1, i2;
    m.tri(prim_id, i0, i1, i2);
    const V3       gv0 = xform_point(e.global_mat, m.vertex(i0)), gv1 = xform_point(e.global_mat, m.vertex(i1)),
             gv2 = xform_point(e.global_mat, m.vertex(i2));
    const Triangle tri    = make_triangle(gv0, gv1, gv2);
    const V3       normal = normalize(mat3_mul(e.normal_mat, lerp2v(m.normal(i0), m.normal(i1), m.normal(i2), u, v)));
    float          t0u, t0v, t1u, t1v, t2u, t2v;
    m.tex(i0, t0u, t0v); m.tex(i1, t1u, t1v); m.tex(i2, t2u, t2v);
    Surf s;
    s.is_entering = true;
    s.point       = lerp2v(gv0, gv1, gv2, u, v);
    s.face_normal = tri.n;
    s.inv_area    = 1 / tri.area;
    s.pu = u; s.pv = v;
    s.tu = lerp2(t0u, t1u, t2u, u, v);
    s.tv = lerp2(t0v, t1v, t2v, u, v);
    s.local = make_orthonormal(normal);
    return s;
}
// compute_ellipsoid_area + make_sphere_shape::surface_element, shapes/sphere.art:22-29,48-67
static Surf sphere_surface_element(const Sphere& sp, const Entity& e, const Ray& ray, const Hit& hit)
{
    const float* g  = e.global_mat;
    auto to_gdir    = [&](V3 p) { return mat3_mul(Mat3{ v3(g[0], g[1], g[2]), v3(g[3], g[4], g[5]), v3(g[6], g[7], g[8]) }, p); };
    const float l1 = len(to_gdir(v3(sp.radius, 0, 0))), l2 = len(to_gdir(v3(0, sp.radius, 0))), l3 = len(to_gdir(v3(0, 0, sp.radius)));
    const float P    = 1.6f;
    const float area = 4 * flt_pi * powf((powf(l1 * l2, P) + powf(l1 * l3, P) + powf(l2 * l3, P)) / 3, 1 / P);
    const V3    point  = add(ray.org, mulf(ray.dir, hit.distance));
    const V3    dir    = sub(point, xform_point(e.global_mat, sp.origin));
    const V3    normal = mulf(dir, 1 / len(dir));
    Surf s;
    s.is_entering = true;
    s.point       = point;
    s.face_normal = normal;
    s.inv_area    = 1 / area;
    s.pu = s.tu = hit.u;
    s.pv = s.tv = hit.v;
    s.local = make_orthonormal(normal);
    return s;
}

// ---- BSDFs ---------------------------------------------------------------------------------------
struct BsdfSample { bool valid; V3 in_dir; float pdf; Color col; float eta; };

// core/fresnel.art:7-27
static inline float fresnel_factor(float eta, float cos_i, float cos_t)
{
    const float R_s = safe_div(eta * cos_i - cos_t, eta * cos_i + cos_t);
    const float R_p = safe_div(cos_i - eta * cos_t, cos_i + eta * cos_t);
    return clampf((R_s * R_s + R_p * R_p) * 0.5f, 0, 1);
}
static inline bool fresnel(float eta, float cos_i, float& cos_t_out, float& factor)
{
    const float eta2   = cos_i < 0 ? 1 / eta : eta;
    const float cos2_t = 1 - (1 - cos_i * cos_i) * eta2 * eta2; // snell
    if (cos2_t <= 0.0f) return false;
    const float cos_t = sqrtf(cos2_t);
    cos_t_out         = cos_i < 0 ? -cos_t : cos_t;
    factor            = fresnel_factor(eta2, fabsf(cos_i), cos_t);
    return true;
}
// core/fresnel.art:29-36
static inline float conductor_factor(float n, float k, float cos_i)
{
    const float f   = n * n + k * k;
    const float d1  = f * cos_i * cos_i;
    const float d2  = 2.0f * n * cos_i;
    const float R_s = safe_div(d1 - d2, d1 + d2);
    const float R_p = safe_div(f - d2 + cos_i * cos_i, f + d2 + cos_i * cos_i);
    return clampf((R_s * R_s + R_p * R_p) * 0.5f, 0, 1);
}
static inline V3 reflect(V3 v, V3 n) { return sub(mulf(n, 2 * dot(n, v)), v); } // vec3_reflect, core/vector.art:127
static inline V3 refract(V3 v, V3 n, float eta, float cos_i, float cos_t)      // vec3_refract, core/vector.art:129
{
    return sub(mulf(n, eta * cos_i - cos_t), mulf(v, eta));
}

struct Bsdf {
    const igb200_bsdf* d;
    const Surf*        surf;
    bool               one_sided; // IGB200_TECHFLAG_ONE_SIDED_DIFFUSE (not in the reference)
    bool is_specular() const { return d->type != IGB200_BSDF_DIFFUSE; }
    Color kd() const { return color(d->color[0], d->color[1], d->color[2]); }
    // eval / pdf: bsdf/diffuse.art:3-4; specular BSDFs return black / 0 (dielectric.art:3-4, conductor.art:3-4,14-15)
    Color eval(V3 in_dir, V3) const
    {
        if (is_specular()) return black;
        const float c = dot(in_dir, surf->local.c2);
        return cmulf(kd(), (one_sided ? (c >= 0 ? c : 0.0f) : fabsf(c)) * flt_inv_pi); // absolute_cos unless flagged
    }
    float pdf(V3 in_dir, V3) const
    {
        if (is_specular()) return 0.0f;
        const float c = dot(in_dir, surf->local.c2);
        return (c >= 0 ? c : 0.0f) / flt_pi; // cosine_hemisphere_pdf(positive_cos(..))
    }
    BsdfSample sample(Rnd& rnd, V3 out_dir) const
    {
        const V3 n = surf->local.c2;
        switch (d->type) {
        default:
        case IGB200_BSDF_DIFFUSE: { // bsdf/diffuse.art:5-9
            const float     u = rnd.next_f32();
            const float     v = rnd.next_f32();
            const DirSample s = sample_cosine_hemisphere(u, v);
            return BsdfSample{ true, mat3_mul(surf->local, s.dir), s.pdf, kd(), 1.0f };
        }
        case IGB200_BSDF_DIELECTRIC: { // bsdf/dielectric.art:5-21 (adjoint = false)
            const float n1 = d->p[0], n2 = d->p[1];
            const float k     = surf->is_entering ? n1 / n2 : n2 / n1;
            const float cos_o = dot(out_dir, n);
            float       cos_t = 0, factor = 1;
            if (!fresnel(k, cos_o, cos_t, factor)) { cos_t = 0; factor = 1; }
            if (rnd.next_f32() > factor) {
                const V3 t = refract(out_dir, n, k, cos_o, cos_t);
                return BsdfSample{ true, t, 1.0f, cmulf(color(d->color2[0], d->color2[1], d->color2[2]), 1.0f), k };
            }
            return BsdfSample{ true, reflect(out_dir, n), 1.0f, kd(), 1.0f };
        }
        case IGB200_BSDF_MIRROR: // bsdf/conductor.art:5-7
            return BsdfSample{ true, reflect(out_dir, n), 1.0f, kd(), 1.0f };
        case IGB200_BSDF_CONDUCTOR: { // bsdf/conductor.art:16-28
            const float cos_i = dot(out_dir, n);
            const Color f = color(conductor_factor(d->color2[0], d->p[0], cos_i), conductor_factor(d->color2[1], d->p[1], cos_i),
                                  conductor_factor(d->color2[2], d->p[2], cos_i));
            return BsdfSample{ true, reflect(out_dir, n), 1.0f, cmul(kd(), f), 1.0f };
        }
        }
    }
};

// ---- lights --------------------------------------------------------------------------------------
struct DirectSample { V3 pos, dir; Color intensity; Pdf pdf; float cos, dist; };

struct Ctx { // the slice of the render state lights and the technique need
    const igo_scene*      scene;
    const igb200_shading* sh;
    float                 scene_radius; // bbox_radius(scene_bbox) * 1.01, light/env.art:76, directional.art:2
};

static inline bool light_delta(const igb200_light& l)
{
    return l.type == IGB200_LIGHT_POINT || l.type == IGB200_LIGHT_SPOT || l.type == IGB200_LIGHT_DIRECTIONAL;
}
static inline bool light_infinite(const igb200_light& l) { return l.type == IGB200_LIGHT_ENV || l.type == IGB200_LIGHT_DIRECTIONAL; }
static inline Color light_color(const igb200_light& l) { return color(l.color[0], l.color[1], l.color[2]); }

// make_plane_area_emitter::compute_sq, light/area.art:117-160
struct SQ { V3 o, n; float x0, y0, z0, x1, y1, b0, b1, k, s; };
struct PlaneEmitter {
    V3    origin, x_axis, y_axis, normal, ex, ey;
    float area, inv_area, width, height;
    explicit PlaneEmitter(const igb200_light& l)
    {
        origin = v3(l.v[0], l.v[1], l.v[2]); x_axis = v3(l.v[3], l.v[4], l.v[5]); y_axis = v3(l.v[6], l.v[7], l.v[8]);
        normal = v3(l.v[9], l.v[10], l.v[11]); area = l.v[12];
        inv_area = 1 / area; width = len(x_axis); height = len(y_axis);
        ex = mulf(x_axis, 1 / width); ey = mulf(y_axis, 1 / height);
    }
    SQ compute_sq(V3 from) const
    {
        const V3    dir = sub(origin, from);
        const float x0 = dot(dir, ex), y0 = dot(dir, ey), z0_ = dot(dir, normal);
        const float x1 = x0 + width, y1 = y0 + height;
        const bool  pos = !std::signbit(z0_);
        const float z0  = pos ? -z0_ : z0_;
        const V3    n   = pos ? neg(normal) : normal;
        const float diff[4] = { x0 - x1, y1 - y0, x1 - x0, y0 - y1 };
        const float nz_[4]  = { y0 * diff[0], x1 * diff[1], y1 * diff[2], x0 * diff[3] };
        float       nz[4];
        for (int i = 0; i < 4; ++i) nz[i] = nz_[i] / sqrtf(diff[i] * diff[i] * (z0 * z0) + nz_[i] * nz_[i]);
        auto safe_acos = [](float a) { return acosf(clampf(a, -1, 1)); };
        const float g0 = safe_acos(-nz[0] * nz[1]), g1 = safe_acos(-nz[1] * nz[2]);
        const float g2 = safe_acos(-nz[2] * nz[3]), g3 = safe_acos(-nz[3] * nz[0]);
        const float k = 2 * flt_pi - g2 - g3, s = g0 + g1 - k;
        return SQ{ from, n, x0, y0, z0, x1, y1, nz[0], nz[2], k, s };
    }
    // sample_direct, light/area.art:162-204
    void sample_direct(float ux, float uy, V3 from, Surf& surf, Pdf& pdf, float& weight) const
    {
        const SQ    sq = compute_sq(from);
        const float au = fmaf(ux, sq.s, sq.k);
        const float fu = fmaf(cosf(au), sq.b0, -sq.b1) / sinf(au);
        const float cu = clampf(copysignf(1.0f, fu) / sqrtf(sum_of_prod(fu, fu, sq.b0, sq.b0)), -1, 1);
        const float xu = clampf(-(cu * sq.z0) / sqrtf(fmaf(-cu, cu, 1.0f)), sq.x0, sq.x1);
        const float d   = sqrtf(sum_of_prod(xu, xu, sq.z0, sq.z0));
        const float h0  = sq.y0 / sqrtf(sum_of_prod(d, d, sq.y0, sq.y0));
        const float h1  = sq.y1 / sqrtf(sum_of_prod(d, d, sq.y1, sq.y1));
        const float hv  = fmaf(uy, h1 - h0, h0);
        const float hv2 = hv * hv;
        const float yv  = hv2 < 1 - 1e-6f ? (hv * d) / sqrtf(1 - hv2) : sq.y1;
        const V3    p     = add(sq.o, add(mulf(ex, xu), add(mulf(ey, yv), mulf(sq.n, sq.z0))));
        const float pdf_s = safe_div(1, sq.s);
        const float tx = dot(sub(p, origin), ex) / width, ty = dot(sub(p, origin), ey) / height;
        surf.is_entering = true;
        surf.point       = p;
        surf.face_normal = normal;
        surf.inv_area    = inv_area;
        surf.pu = tx; surf.pv = ty;
        surf.tu = tx; surf.tv = ty; // constant-colour emitters never read tex_coords
        surf.local = make_orthonormal(normal);
        pdf    = solid_pdf(pdf_s);
        weight = sq.s;
    }
    Pdf pdf_direct(V3 from) const { return solid_pdf(safe_div(1, compute_sq(from).s)); }
};

// make_shape_area_emitter::sample / pdf, light/area.art:48-68 (sample_triangle core/sampling.art:33-35)
static void shape_emitter_sample(const Ctx& c, const igb200_light& l, float ux_, float uy, Surf& surf, float& pdfv, float& weight)
{
    const Entity  e = load_entity(*c.scene, l.entity);
    const TriMesh m = load_trimesh(*c.scene, e.shape_id);
    const float   ux = ux_ * (float)m.nf;
    const int32_t f  = std::min((int32_t)ux, m.nf - 1);
    float         u = ux - (float)f, v = uy;
    if (u + v > 1) { u = 1 - u; v = 1 - v; }
    surf   = trimesh_surface_element_for_point(m, e, f, u, v);
    pdfv   = surf.inv_area / (float)m.nf;
    weight = (float)m.nf / surf.inv_area;
}

static DirectSample light_sample_direct(const Ctx& c, const igb200_light& l, Rnd& rnd, const Surf& from)
{
    switch (l.type) {
    default:
    case IGB200_LIGHT_POINT: { // light/point.art:3-8
        const V3    pos  = v3(l.v[0], l.v[1], l.v[2]);
        const V3    dir_ = sub(pos, from.point);
        const float dist = len(dir_);
        return DirectSample{ pos, mulf(dir_, safe_div(1, dist)), light_color(l), area_pdf(1), 1.0f, dist };
    }
    case IGB200_LIGHT_SPOT: { // light/spot.art:8-40
        const V3    pos = v3(l.v[0], l.v[1], l.v[2]), dir = v3(l.v[3], l.v[4], l.v[5]);
        const float cosCutoff = cosf(l.v[6]), cosFalloff = cosf(l.v[7]);
        const float blendRange = cosFalloff - cosCutoff;
        const V3    out_dir_ = sub(pos, from.point);
        const float dist     = len(out_dir_);
        const V3    out_dir  = mulf(out_dir_, safe_div(1, dist));
        const float cos_angle = dot(neg(out_dir), dir);
        float       factor;
        if (blendRange <= flt_eps) {
            factor = cos_angle <= cosCutoff ? 0.0f : 1.0f;
        } else {
            const float x = clampf((cos_angle - cosCutoff) / blendRange, 0, 1);
            factor        = x * x * (3 - 2 * x); // smoothstep
        }
        const float cos      = -dot(out_dir, dir);
        const float area_pdfv = dot(neg(out_dir), dir) > cosCutoff ? 1.0f : 0.0f;
        return DirectSample{ pos, out_dir, cmulf(light_color(l), factor), area_pdf(area_pdfv), cos, dist };
    }
    case IGB200_LIGHT_DIRECTIONAL: { // light/directional.art:6
        const V3 dir = v3(l.v[0], l.v[1], l.v[2]);
        return DirectSample{ add(from.point, mulf(dir, -c.scene_radius)), neg(dir), light_color(l), delta_pdf(), 1.0f, c.scene_radius };
    }
    case IGB200_LIGHT_ENV: { // make_environment_light_function_spherical::sample_direct, light/env.art:82-86
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        const V3    d = equal_area_square_to_sphere(u, v);
        const float pdf = 1 / (4 * flt_pi);
        return DirectSample{ add(from.point, mulf(d, c.scene_radius)), d, cmulf(light_color(l), 1 / pdf), solid_pdf(pdf), 1.0f, c.scene_radius };
    }
    case IGB200_LIGHT_AREA_SHAPE:
    case IGB200_LIGHT_AREA_PLANE: { // make_area_light::sample_direct, light/area.art:12-26
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        Surf  to;
        Pdf   pdf = area_pdf(1);
        float weight;
        if (l.type == IGB200_LIGHT_AREA_PLANE) {
            PlaneEmitter(l).sample_direct(u, v, from.point, to, pdf, weight);
        } else {
            float pdfv;
            shape_emitter_sample(c, l, u, v, to, pdfv, weight);
            pdf = area_pdf(pdfv);
        }
        const V3    dir_ = sub(to.point, from.point);
        const float dist = len(dir_);
        const V3    dir  = mulf(dir_, safe_div(1, dist));
        const float cos  = dot(dir, to.face_normal) * (from.is_entering ? -1.0f : 1.0f);
        return DirectSample{ to.point, dir, cmulf(light_color(l), weight), pdf, cos, dist };
    }
    }
}

// Light::pdf_direct(ray, surf) for the lights a path can HIT or MISS into (area, env)
static Pdf light_pdf_direct(const Ctx& c, const igb200_light& l, const Ray& ray, const Surf& surf)
{
    switch (l.type) {
    case IGB200_LIGHT_ENV: return solid_pdf(1 / (4 * flt_pi)); // light/env.art:96
    case IGB200_LIGHT_AREA_PLANE: return PlaneEmitter(l).pdf_direct(ray.org); // light/area.art:206-210
    case IGB200_LIGHT_AREA_SHAPE: { // light/area.art:60-68,83: prim_coords reinterpreted as the sampling square
        Surf  s;
        float pdfv, w;
        shape_emitter_sample(c, l, surf.pu, surf.pv, s, pdfv, w);
        return area_pdf(pdfv);
    }
    default: return delta_pdf();
    }
}

// make_uniform_light_selector, light/light_selector.art:26-45 (pick_light_id :19-25)
struct Selector {
    const Ctx* c;
    int32_t    n_inf, n_fin;
    int32_t count() const { return n_inf + n_fin; }
    float   pdf() const { return count() == 0 ? 1.0f : 1 / (float)count(); }
    const igb200_light& sample(Rnd& rnd) const
    {
        const int32_t n  = count();
        const int32_t id = n <= 1 ? 0 : rnd.next_i32(0, n - 1);
        return id < n_inf ? c->sh->infinite_lights[id] : c->sh->finite_lights[id - n_inf];
    }
};

// ---- technique/pathtracer.art ---------------------------------------------------------------------
struct Payload { float inv_pdf; Color contrib; int32_t depth; float eta; };

struct Film {
    float*   px;
    uint32_t w, h;
    float    inv_spi;
    void splat(int32_t pixel, Color c) const // make_standard_accumulator, driver/accumulator.art:23-30
    {
        px[3 * (size_t)pixel + 0] += c.r * inv_spi;
        px[3 * (size_t)pixel + 1] += c.g * inv_spi;
        px[3 * (size_t)pixel + 2] += c.b * inv_spi;
    }
};

struct Stats { uint64_t camera = 0, bounce = 0, shadow = 0; };

static inline Color handle_color(const igb200_technique& t, Color c) // pathtracer.art:46-50
{
    if (t.clamp > 0) return color(fminf_(c.r, t.clamp), fminf_(c.g, t.clamp), fminf_(c.b, t.clamp));
    return c;
}

static void trace_path(const Ctx& c, const igb200_render_settings& rs, int32_t x, int32_t y, int32_t sample, Ray ray,
                       Rnd rnd, const Film& film, Stats& st)
{
    const igb200_technique& tech   = c.sh->technique;
    const Selector          sel{ &c, (int32_t)c.sh->infinite_light_count, (int32_t)c.sh->finite_light_count };
    const float             offset = 0.001f;
    const int32_t           pixel  = y * (int32_t)rs.width + x;
    Payload                 pt{ 0.0f, color(1, 1, 1), 1, 1.0f }; // init_pt_raypayload, pathtracer.art:33-38
    (void)sample;

    for (;;) {
        const Hit hit = trace_one(*c.scene, ray, false);
        if (hit.prim_id < 0 || hit.ent_id < 0) {
            // on_miss, pathtracer.art:134-160
            if (tech.type == IGB200_TECH_PATH) {
                int   inflights = 0;
                Color col       = black;
                for (uint64_t i = 0; i < c.sh->infinite_light_count; ++i) {
                    const igb200_light& l = c.sh->infinite_lights[i];
                    if (light_infinite(l) && !light_delta(l)) {
                        ++inflights;
                        const Color emit  = light_color(l); // env emission: func(transform * dir) with constant func
                        const float pdf_s = 1 / (4 * flt_pi);
                        const float mis   = tech.enable_nee ? 1 / (1 + pt.inv_pdf * sel.pdf() * pdf_s) : 1.0f;
                        col               = cadd(col, handle_color(tech, cmulf(cmul(pt.contrib, emit), mis)));
                    }
                }
                if (inflights > 0) film.splat(pixel, col);
            }
            return;
        }

        const Entity ent = load_entity(*c.scene, hit.ent_id);
        Surf         surf;
        if (c.scene->shape_type[ent.shape_id] == IGB200_PRIM_SPHERE)
            surf = sphere_surface_element(c.scene->spheres[ent.shape_id], ent, ray, hit);
        else
            surf = trimesh_surface_element(load_trimesh(*c.scene, ent.shape_id), ent, ray, hit);
        const igb200_material& mat = c.sh->materials[ent.mat_id];
        const Bsdf             bsdf{ &c.sh->bsdfs[mat.bsdf], &surf, (tech.flags & IGB200_TECHFLAG_ONE_SIDED_DIFFUSE) != 0 };
        const bool             emissive = mat.light >= 0;

        if (tech.type == IGB200_TECH_DEBUG) { // technique/debugtracer.art:29-177
            Color   out;
            const V3 n = surf.local.c2;
            auto    absv = [](V3 a) { return color(fabsf(a.x), fabsf(a.y), fabsf(a.z)); };
            switch (tech.debug_mode) {
            case 1: out = absv(surf.local.c0); break;
            case 2: out = absv(surf.local.c1); break;
            case 3: out = absv(surf.face_normal); break;
            case 8: out = color(fabsf(surf.tu), fabsf(surf.tv), 0); break;
            case 9: out = color(fabsf(hit.u), fabsf(hit.v), 0); break;
            case 10: out = color(surf.point.x, surf.point.y, surf.point.z); break;
            case 13: out = color(hit.distance, hit.distance, hit.distance); break;
            case 14: { const float a = safe_div(1, surf.inv_area); out = color(a, a, a); break; }
            case 15: { const float v = (float)hit.prim_id; out = color(v, v, v); break; }
            case 17: { const float v = (float)hit.ent_id; out = color(v, v, v); break; }
            case 19: { const float v = (float)ent.mat_id; out = color(v, v, v); break; }
            case 21: out = emissive ? color(0, 0, 1) : color(1, 0, 0); break;
            case 22: out = bsdf.is_specular() ? color(0, 0, 1) : color(1, 0, 0); break;
            case 23: out = surf.is_entering ? color(0, 0, 1) : color(1, 0, 0); break;
            default: out = absv(n); break;
            }
            film.splat(pixel, out);
            return; // TechniqueNoBounceFunction
        }

        // on_hit, pathtracer.art:112-132
        if (emissive && surf.is_entering) {
            const float d = -dot(ray.dir, surf.local.c2);
            if (d > flt_eps) {
                const igb200_light& l     = c.sh->finite_lights[mat.light];
                const Color         emit  = light_color(l);
                const Pdf           pdf   = light_pdf_direct(c, l, ray, surf);
                const float         pdf_s = pdf.as_solid(d, hit.distance * hit.distance);
                const float         mis   = tech.enable_nee ? 1 / (1 + pt.inv_pdf * sel.pdf() * pdf_s) : 1.0f;
                film.splat(pixel, handle_color(tech, cmulf(cmul(pt.contrib, emit), mis)));
            }
        }

        // on_shadow, pathtracer.art:52-110
        const V3 out_dir = neg(ray.dir);
        if (tech.enable_nee && !bsdf.is_specular() && sel.count() != 0 && pt.depth + 1 <= tech.max_depth) {
            const igb200_light& l  = sel.sample(rnd);
            const DirectSample  ls = light_sample_direct(c, l, rnd, surf);
            const float pdf_l_s    = ls.pdf.as_solid(ls.cos, ls.dist * ls.dist) * sel.pdf();
            if (!(pdf_l_s <= flt_eps) && ls.cos > flt_eps) {
                const V3    in_dir = ls.dir;
                const float mis    = light_delta(l) ? 1.0f : 1 / (1 + bsdf.pdf(in_dir, out_dir) / pdf_l_s);
                const float factor = ls.pdf.value / pdf_l_s;
                const Color contrib =
                    handle_color(tech, cmulf(cmul(ls.intensity, cmul(pt.contrib, bsdf.eval(in_dir, out_dir))), mis * factor));
                const Ray shadow = light_infinite(l)
                                       ? make_ray(surf.point, in_dir, offset, flt_max, ray_flag_shadow)
                                       : make_ray(surf.point, sub(ls.pos, surf.point), offset, 1 - offset, ray_flag_shadow);
                st.shadow++;
                // cpu_traverse_secondary + the simple-shadow splat loop (driver/mapping_cpu.art:395-424,816-829)
                const Hit sh = trace_one(*c.scene, shadow, true);
                if (sh.prim_id < 0) film.splat(pixel, contrib);
            }
        }

        // on_bounce, pathtracer.art:162-200
        if (pt.depth + 1 > tech.max_depth) return;
        const BsdfSample ms = bsdf.sample(rnd, out_dir);
        if (!ms.valid) return;
        const Color contrib = cmul(pt.contrib, ms.col);
        float       rr_prob = 1.0f;
        if (pt.depth + 1 > tech.min_depth) {
            const Color cc = cmulf(contrib, pt.eta * pt.eta);
            rr_prob        = clampf(fmaxf(fmaxf(cc.r, cc.g), cc.b), 0.05f, 0.95f); // russian_roulette_pbrt
        }
        if (rnd.next_f32() >= rr_prob) return;
        pt.inv_pdf = bsdf.is_specular() ? 0.0f : 1 / ms.pdf;
        pt.contrib = cmulf(contrib, 1 / rr_prob);
        pt.depth   = pt.depth + 1;
        pt.eta     = pt.eta * ms.eta;
        ray        = make_ray(surf.point, ms.in_dir, offset, flt_max, ray_flag_bounce);
        st.bounce++;
    }
}

} // namespace
} // namespace igo

using namespace igo;

// One render iteration (Device::render, src/runtime/device/Device.cpp:1588-1612 -> cpu_trace).
// film: w*h*3 floats, accumulated into (the caller keeps the iteration count).  stats3 = camera / bounce / shadow rays.
extern "C" int igo_render(const igo_scene* scene, const igb200_shading* sh, const igb200_render_settings* rs, float* film,
                          uint64_t* stats3, int threads)
{
    if (!scene || !sh || !rs || !film) return -1;
    const bool     list = rs->rays != nullptr;
    const uint32_t w = list ? (uint32_t)rs->ray_count : rs->width, h = list ? 1u : rs->height;
    if (w == 0 || h == 0 || rs->spi == 0) return -1;
    for (uint64_t i = 0; i < sh->material_count; ++i)
        if (sh->materials[i].bsdf < 0 || (uint64_t)sh->materials[i].bsdf >= sh->bsdf_count ||
            sh->materials[i].light >= (int64_t)sh->finite_light_count)
            return -1;

    Ctx c{ scene, sh, 0.0f };
    {
        const V3 size = v3(scene->bbox[3] - scene->bbox[0], scene->bbox[4] - scene->bbox[1], scene->bbox[5] - scene->bbox[2]);
        c.scene_radius = len(size) / 2 * 1.01f; // bbox_radius(scene_bbox) * 1.01
    }
    igb200_render_settings s = *rs;
    s.width = w; s.height = h;
    const Film film_{ film, w, h, 1 / (float)rs->spi };

    // make_perspective_camera, camera/perspective.art:29-41
    const igb200_camera& cam = sh->camera;
    const V3   eye = v3(cam.eye[0], cam.eye[1], cam.eye[2]), dir = v3(cam.dir[0], cam.dir[1], cam.dir[2]),
             up    = v3(cam.up[0], cam.up[1], cam.up[2]);
    const V3   right = normalize(cross(dir, up));
    const Mat3 view{ right, up, dir };

    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<Stats>       per(threads);
    std::atomic<uint32_t>    next_row{ 0 };
    const uint32_t           rows_per_job = list ? 1 : 4;
    const uint32_t           cols_per_job = list ? 4096 : w;
    const uint32_t           jobs_x = (w + cols_per_job - 1) / cols_per_job, jobs_y = (h + rows_per_job - 1) / rows_per_job;
    auto worker = [&](int tid) {
        Stats st;
        for (;;) {
            const uint32_t job = next_row.fetch_add(1);
            if (job >= jobs_x * jobs_y) break;
            const uint32_t jy = job / jobs_x, jx = job % jobs_x;
            for (uint32_t y = jy * rows_per_job; y < std::min(h, (jy + 1) * rows_per_job); ++y)
                for (uint32_t x = jx * cols_per_job; x < std::min(w, (jx + 1) * cols_per_job); ++x)
                    for (uint32_t sample = 0; sample < rs->spi; ++sample) {
                        // make_camera_emitter / make_list_emitter, driver/emitter.art:6-31
                        Rnd rnd{ create_random_seed((int32_t)sample, (int32_t)rs->iteration, (int32_t)rs->frame, (int32_t)x,
                                                    (int32_t)y, (int32_t)rs->user_seed),
                                 1u };
                        Ray ray;
                        if (list) {
                            igb200_ray r = rs->rays[x];
                            igo_normalize_rays(&r, 1);
                            ray = make_ray(v3(r.org[0], r.org[1], r.org[2]), v3(r.dir[0], r.dir[1], r.dir[2]), r.tmin, r.tmax, 0);
                        } else {
                            const float rx = rnd.next_f32(); // make_uniform_pixel_sampler, sampler/pixel_sampler.art:4-10
                            const float ry = rnd.next_f32();
                            // make_pixelcoord_from_xy, driver/camera.art:21-29
                            const float nx = 2 * ((float)x + rx) / (float)w - 1;
                            const float ny = 1 - 2 * ((float)y + ry) / (float)h;
                            const V3    d  = normalize(mat3_mul(view, v3(cam.scale[0] * nx, cam.scale[1] * ny, 1)));
                            ray            = make_ray(eye, d, cam.tmin, cam.tmax, ray_flag_camera);
                        }
                        st.camera++;
                        trace_path(c, s, (int32_t)x, (int32_t)y, (int32_t)sample, ray, rnd, film_, st);
                    }
        }
        per[tid] = st;
    };
    if (threads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(worker, t);
        for (auto& t : pool) t.join();
    }
    if (stats3) {
        stats3[0] = stats3[1] = stats3[2] = 0;
        for (auto& p : per) { stats3[0] += p.camera; stats3[1] += p.bounce; stats3[2] += p.shadow; }
    }
    return 0;
}

// Known-answer hooks for the integer RNG (bit-exact on every platform): seed and the first `n` raw draws.
extern "C" void igo_kat_rng(int32_t sample, int32_t iter, int32_t frame, int32_t x, int32_t y, int32_t user, uint32_t n,
                            uint32_t* out_seed, uint32_t* out_raw, float* out_f32)
{
    Rnd a{ create_random_seed(sample, iter, frame, x, y, user), 1u };
    Rnd b = a;
    if (out_seed) *out_seed = a.seed;
    for (uint32_t i = 0; i < n; ++i) {
        if (out_raw) out_raw[i] = a.next_u32();
        if (out_f32) out_f32[i] = b.next_f32();
    }
}
