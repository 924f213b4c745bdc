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
#include <limits>
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
// compute_ellipsoid_area, shapes/sphere.art:22-29
static float ellipsoid_area(float radius, const Entity& e)
{
    const float* g  = e.global_mat;
    auto to_gdir    = [&](V3 p) { return mat3_mul(Mat3{ v3(g[0], g[1], g[2]), v3(g[3], g[4], g[5]), v3(g[6], g[7], g[8]) }, p); };
    const float l1 = len(to_gdir(v3(radius, 0, 0))), l2 = len(to_gdir(v3(0, radius, 0))), l3 = len(to_gdir(v3(0, 0, radius)));
    const float P    = 1.6f;
    return 4 * flt_pi * powf((powf(l1 * l2, P) + powf(l1 * l3, P) + powf(l2 * l3, P)) / 3, 1 / P);
}
// make_sphere_shape::surface_element, shapes/sphere.art:48-67
static Surf sphere_surface_element(const Sphere& sp, const Entity& e, const Ray& ray, const Hit& hit)
{
    const float area   = ellipsoid_area(sp.radius, e);
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
// sphere_map_uv + sphere_compute_surface_element_for_normal, shapes/sphere.art:1-6,31-45
static Surf sphere_surface_element_for_normal(V3 normal, const Sphere& sp, const Entity& e)
{
    const V3    d     = v3(normal.y, -normal.x, normal.z);
    const float theta = acosf(d.z); // spherical_from_dir, core/common.art
    float       phi   = atan2f(d.y, d.x);
    if (phi < 0) phi += 2 * flt_pi;
    const V3 gn = normalize(mat3_mul(e.normal_mat, normal));
    Surf s;
    s.is_entering = true;
    s.point       = xform_point(e.global_mat, add(sp.origin, mulf(normal, sp.radius)));
    s.face_normal = gn;
    s.inv_area    = 1 / ellipsoid_area(sp.radius, e);
    s.pu = s.tu = phi / (2 * flt_pi);
    s.pv = s.tv = theta / flt_pi;
    s.local = make_orthonormal(gn);
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

// ---- textures: texture/checkerboard.art, texture/image.art, driver/image.art -------------------------
struct Tex {
    const igb200_shading* sh;
    static float wrapf(float v, float mn, float mx) // math::wrap, core/math.art:88-91
    {
        const float range = mx - mn;
        return range <= flt_eps ? mn : v - (range * floorf((v - mn) / range));
    }
    static float fract(float x) { return x - floorf(x); } // core/math.art:74
    static int32_t border(uint32_t mode, int32_t x, int32_t w) // texture/image.art:9-40
    {
        if (mode == IGB200_WRAP_CLAMP) return std::min(std::max(x, 0), w - 1);
        if (mode == IGB200_WRAP_MIRROR) {
            const int32_t t = x < 0 ? -1 - x : x;
            const int32_t i = t / w, k = t - i * w;
            return (i & 1) == 0 ? w - 1 - k : k;
        }
        const int32_t t = x % w;
        return t < 0 ? t + w : t;
    }
    void texel(const igb200_texture& t, int32_t x, int32_t y, float out[4]) const // img.pixels(x, y)
    {
        const size_t i = (size_t)y * t.width + (size_t)x;
        if (t.format == IGB200_IMAGE_RGBA32F) {
            const float* px = reinterpret_cast<const float*>(sh->blob + t.offset) + 4 * i;
            out[0] = px[0]; out[1] = px[1]; out[2] = px[2]; out[3] = px[3];
        } else { // image_rgba_unpack, driver/image.art:8-14
            const uint32_t q = reinterpret_cast<const uint32_t*>(sh->blob + t.offset)[i];
            out[0] = (float)(q & 0xFFu) / 255;
            out[1] = (float)((q & 0xFF00u) >> 8) / 255;
            out[2] = (float)((q & 0xFF0000u) >> 16) / 255;
            out[3] = t.opaque ? 1.0f : (float)((q & 0xFF000000u) >> 24) / 255;
        }
    }
    Color eval(uint32_t id, float u, float v) const
    {
        const igb200_texture& t = sh->textures[id];
        // mat3x3_transform_point_affine(transform, uv), core/matrix.art:223-226 (rows dot (u, v, 1))
        const float* m  = t.transform;
        const float  tu = fmaf(m[0], u, fmaf(m[3], v, m[6] * 1.0f));
        const float  tv = fmaf(m[1], u, fmaf(m[4], v, m[7] * 1.0f));
        if (t.type == IGB200_TEX_CHECKERBOARD) { // texture/checkerboard.art:4-13
            const float sx = tu * t.scale[0], sy = tv * t.scale[1];
            const bool  px = ((int32_t)wrapf(sx, 0, 2) % 2) == 0;
            const bool  py = ((int32_t)wrapf(sy, 0, 2) % 2) == 0;
            return (px ^ py) ? color(t.color0[0], t.color0[1], t.color0[2]) : color(t.color1[0], t.color1[1], t.color1[2]);
        }
        const int32_t w = (int32_t)t.width, h = (int32_t)t.height;
        float         p[4][4];
        if (t.filter == IGB200_FILTER_NEAREST) { // map_uv_to_image_pixel_nearest + make_nearest_filter
            const float fu = tu * (float)w, fv = tv * (float)h;
            texel(t, border(t.wrap_u, (int32_t)floorf(fu), w), border(t.wrap_v, (int32_t)floorf(fv), h), p[0]);
            return color(p[0][0], p[0][1], p[0][2]);
        }
        const float   fu = tu * (float)w - 0.5f, fv = tv * (float)h - 0.5f; // map_uv_to_image_pixel :70-79
        const int32_t ix = (int32_t)floorf(fu), iy = (int32_t)floorf(fv);
        const float   fx = fract(fu), fy = fract(fv);
        if (t.filter == IGB200_FILTER_BILINEAR) { // :91-106
            const int32_t x0 = border(t.wrap_u, ix, w), y0 = border(t.wrap_v, iy, h);
            const int32_t x1 = border(t.wrap_u, ix + 1, w), y1 = border(t.wrap_v, iy + 1, h);
            texel(t, x0, y0, p[0]); texel(t, x1, y0, p[1]); texel(t, x0, y1, p[2]); texel(t, x1, y1, p[3]);
            float o[3];
            for (int k = 0; k < 3; ++k) {
                const float a = (1 - fx) * p[0][k] + fx * p[1][k], b = (1 - fx) * p[2][k] + fx * p[3][k]; // color_lerp
                o[k]          = (1 - fy) * a + fy * b;
            }
            return color(o[0], o[1], o[2]);
        }
        // make_bicubic_filter :108-146
        auto w0 = [](float a) { return (a * (a * (-a + 3) - 3) + 1) / 6; };
        auto w1 = [](float a) { return (a * a * (3 * a - 6) + 4) / 6; };
        auto w2 = [](float a) { return (a * (a * (-3 * a + 3) + 3) + 1) / 6; };
        auto w3 = [](float a) { return (a * a * a) / 6; };
        auto g0 = [&](float a) { return w0(a) + w1(a); };
        auto g1 = [&](float a) { return w2(a) + w3(a); };
        auto h0 = [&](float a) { return (w1(a) / g0(a)) - 1; };
        auto h1 = [&](float a) { return (w3(a) / g1(a)) + 1; };
        const float   g0x = g0(fx), g0y = g0(fy), g1x = g1(fx), g1y = g1(fy);
        const int32_t ix0 = (int32_t)floorf((float)ix + h0(fx) + 0.5f), iy0 = (int32_t)floorf((float)iy + h0(fy) + 0.5f);
        const int32_t ix1 = (int32_t)floorf((float)ix + h1(fx) + 0.5f), iy1 = (int32_t)floorf((float)iy + h1(fy) + 0.5f);
        const int32_t x0 = border(t.wrap_u, ix0, w), y0 = border(t.wrap_v, iy0, h);
        const int32_t x1 = border(t.wrap_u, ix1, w), y1 = border(t.wrap_v, iy1, h);
        texel(t, x0, y0, p[0]); texel(t, x1, y0, p[1]); texel(t, x0, y1, p[2]); texel(t, x1, y1, p[3]);
        float o[3];
        for (int k = 0; k < 3; ++k)
            o[k] = (p[0][k] * (g0x * g0y) + p[1][k] * (g1x * g0y)) + (p[2][k] * (g0x * g1y) + p[3][k] * (g1x * g1y));
        return color(o[0], o[1], o[2]);
    }
};

// ---- core/microfacet.art: GGX distribution of visible normals (the loader's default, runtime/bsdf/BSDF.cpp:75) ----
static inline V3 to_local(const Mat3& l, V3 v) { return v3(dot(l.c0, v), dot(l.c1, v), dot(l.c2, v)); }                // shading.art:9
static inline V3 to_world(const Mat3& l, V3 v) { return add(add(mulf(l.c0, v.x), mulf(l.c1, v.y)), mulf(l.c2, v.z)); } // :8
static inline float diff_of_prod(float a, float b, float c, float d) // core/common.art:132-137
{
    const float cd = c * d, diff = fmaf(a, b, -cd), err = fmaf(-c, d, cd);
    return diff + err;
}
static float g_1_smith(const Mat3& l, V3 w, float au, float av) // microfacet.art:159-177
{
    const float cosZ = dot(l.c2, w);
    if (fabsf(cosZ) <= flt_eps) return 0;
    const float kx = au * dot(l.c0, w), ky = av * dot(l.c1, w);
    const float a2 = kx * kx + ky * ky;
    if (a2 <= flt_eps) return 1;
    const float k2 = a2 / (cosZ * cosZ);
    return 2 / (1 + sqrtf(1 + k2));
}
static float ndf_ggx(const Mat3& l, V3 m, float au, float av) // :191-201
{
    const float cosZ = dot(l.c2, m), kx = dot(l.c0, m) / au, ky = dot(l.c1, m) / av;
    const float k = kx * kx + ky * ky + cosZ * cosZ;
    return safe_div(1, flt_pi * au * av * k * k);
}
static float pdf_vndf_ggx(const Mat3& l, V3 w, V3 h, float au, float av) // :378-381
{
    const float cosZ = fabsf(dot(l.c2, w));
    return safe_div(g_1_smith(l, w, au, av) * fabsf(dot(w, h)) * ndf_ggx(l, h, au, av), cosZ);
}
static V3 sample_vndf_ggx(Rnd& rnd, const Mat3& l, V3 vN, float au, float av) // :324-376
{
    const V3 vL = to_local(l, vN);
    const V3 sL = normalize(v3(au * vL.x, av * vL.y, vL.z));
    // shading::sin_cos_phi
    const float sin_t = safe_sqrt(1 - sL.z * sL.z);
    float       sin_phi = 0, cos_phi = 1;
    if (fabsf(sin_t) > flt_eps) { sin_phi = sL.y / sin_t; cos_phi = sL.x / sin_t; }
    // sample_vndf_ggx_11
    const float cos_theta = fabsf(sL.z);
    const float u0 = rnd.next_f32();
    const float u1 = rnd.next_f32();
    float       pcx, pcy; // square_to_concentric_disk, core/warp.art:2-22
    {
        const float a = 2 * u0 - 1, b = 2 * u1 - 1;
        if (a == 0 && b == 0) { pcx = 0; pcy = 0; }
        else if (a * a > b * b) { const float phi = (flt_pi / 4) * safe_div(b, a); pcx = cosf(phi) * a; pcy = sinf(phi) * a; }
        else { const float phi = (flt_pi / 2) - (flt_pi / 4) * safe_div(a, b); pcx = cosf(phi) * b; pcy = sinf(phi) * b; }
    }
    const float s  = 0.5f * (1 + cos_theta);
    const float y  = (1 - s) * safe_sqrt(1 - pcx * pcx) + s * pcy;
    const float z  = safe_sqrt(1 - y * y - pcx * pcx);
    const float st = safe_sqrt(1 - cos_theta * cos_theta);
    const float nm = safe_div(1, sum_of_prod(st, y, cos_theta, z));
    const float slx = diff_of_prod(cos_theta, y, st, z) * nm, sly = pcx * nm;
    const float s2x = (cos_phi * slx - sin_phi * sly) * au, s2y = (sin_phi * slx + cos_phi * sly) * av;
    const V3    Nh  = !std::isfinite(s2x) ? v3(0, 0, 0) : normalize(v3(-s2x, -s2y, 1));
    return to_world(l, Nh);
}

// core/sampling.art:120-165
static V3 ensure_valid_reflection(V3 Ng, V3 I, V3 N)
{
    const V3    R         = reflect(I, N);
    const float threshold = fminf(0.9f * dot(Ng, I), 0.01f);
    if (dot(Ng, R) >= threshold) return N;
    const float NdotNg = dot(N, Ng);
    const V3    X      = normalize(sub(N, mulf(Ng, NdotNg)));
    const float Ix = dot(I, X), Iz = dot(I, Ng), Ix2 = Ix * Ix, Iz2 = Iz * Iz, a = Ix2 + Iz2;
    const float b = safe_sqrt(Ix2 * (a - threshold * threshold)), c = Iz * threshold + a;
    const float fac = 0.5f / a, N1_z2 = fac * (b + c), N2_z2 = fac * (-b + c);
    const bool  valid1 = (N1_z2 > 1e-5f) && (N1_z2 <= (1.0f + 1e-5f)), valid2 = (N2_z2 > 1e-5f) && (N2_z2 <= (1.0f + 1e-5f));
    float nx, ny;
    if (valid1 && valid2) {
        const float n1x = safe_sqrt(1 - N1_z2), n1y = safe_sqrt(N1_z2), n2x = safe_sqrt(1 - N2_z2), n2y = safe_sqrt(N2_z2);
        const float R1 = 2 * (n1x * Ix + n1y * Iz) * n1y - Iz, R2 = 2 * (n2x * Ix + n2y * Iz) * n2y - Iz;
        const bool  valid3 = R1 >= 1e-5f, valid4 = R2 >= 1e-5f;
        const bool  first  = (valid3 && valid4) ? (R1 < R2) : (R1 > R2);
        nx = first ? n1x : n2x; ny = first ? n1y : n2y;
    } else if (valid1 || valid2) {
        const float Nz2 = valid1 ? N1_z2 : N2_z2;
        nx = safe_sqrt(1 - Nz2); ny = safe_sqrt(Nz2);
    } else { nx = 0; ny = 1; }
    return add(mulf(X, nx), mulf(Ng, ny));
}
// core/matrix.art:247-270 and :110-113
static Mat3 align_vectors(V3 a, V3 b)
{
    const V3    ax   = cross(b, a);
    const float cosA = dot(a, b);
    if (cosA <= -1) return Mat3{ v3(-1, 0, 0), v3(0, -1, 0), v3(0, 0, -1) };
    const float k = 1 / (1 + cosA);
    return Mat3{ v3((ax.x * ax.x * k) + cosA, (ax.y * ax.x * k) - ax.z, (ax.z * ax.x * k) + ax.y),
                 v3((ax.x * ax.y * k) + ax.z, (ax.y * ax.y * k) + cosA, (ax.z * ax.y * k) - ax.x),
                 v3((ax.x * ax.z * k) - ax.y, (ax.y * ax.z * k) + ax.x, (ax.z * ax.z * k) + cosA) };
}
static Mat3 matmul(const Mat3& a, const Mat3& b) { return Mat3{ mat3_mul(a, b.c0), mat3_mul(a, b.c1), mat3_mul(a, b.c2) }; }

struct Bsdf {
    const igb200_bsdf* d;
    Mat3               local;       // surf.local, re-oriented by bump maps
    V3                 face_normal;
    bool               is_entering;
    bool               one_sided;   // IGB200_TECHFLAG_ONE_SIDED_DIFFUSE (not in the reference)
    Color              kd_;         // DIFFUSE reflectance (constant or texture), ks otherwise
    float              au, av;      // CONDUCTOR roughness
    bool rough() const { return d->type == IGB200_BSDF_CONDUCTOR && !(au <= 1e-4f || av <= 1e-4f); } // check_if_delta_distribution
    bool is_specular() const { return d->type != IGB200_BSDF_DIFFUSE && !rough(); }
    Color kd() const { return kd_; }
    Color fresnel_term(float c) const // make_rough_conductor_bsdf, bsdf/conductor.art:102-110
    {
        return color(conductor_factor(d->color2[0], d->p[0], c), conductor_factor(d->color2[1], d->p[1], c),
                     conductor_factor(d->color2[2], d->p[2], c));
    }
    // eval / pdf: bsdf/diffuse.art:3-4; specular BSDFs return black / 0 (dielectric.art:3-4, conductor.art:3-4,14-15)
    Color eval(V3 in_dir, V3 out_dir) const
    {
        if (rough()) { // make_rough_base_conductor_bsdf::eval (kd = black), bsdf/conductor.art:53-66
            const V3    N     = local.c2;
            const float cos_o = fabsf(dot(out_dir, N)), cos_i = fabsf(dot(in_dir, N));
            if (cos_o <= flt_eps || cos_i <= flt_eps) return black;
            const V3    H = normalize(add(in_dir, out_dir));
            const float D = ndf_ggx(local, H, au, av);
            const float G = g_1_smith(local, in_dir, au, av) * g_1_smith(local, out_dir, au, av);
            const Color F = fresnel_term(fabsf(dot(out_dir, H)));
            const Color IF = color(1 - F.r, 1 - F.g, 1 - F.b);
            return cmulf(cadd(cmul(black, IF), cmul(kd_, F)), D * G / (4 * cos_o));
        }
        if (is_specular()) return black;
        const float c = dot(in_dir, local.c2);
        return cmulf(kd(), (one_sided ? (c >= 0 ? c : 0.0f) : fabsf(c)) * flt_inv_pi); // absolute_cos unless flagged
    }
    float pdf(V3 in_dir, V3 out_dir) const
    {
        if (rough()) { // bsdf/conductor.art:70-75
            const V3    H       = normalize(add(in_dir, out_dir));
            const float cos_h_o = fabsf(dot(out_dir, H));
            return pdf_vndf_ggx(local, out_dir, H, au, av) * safe_div(1, 4 * cos_h_o);
        }
        if (is_specular()) return 0.0f;
        const float c = dot(in_dir, local.c2);
        return (c >= 0 ? c : 0.0f) / flt_pi; // cosine_hemisphere_pdf(positive_cos(..))
    }
    BsdfSample sample(Rnd& rnd, V3 out_dir) const
    {
        const V3         n = local.c2;
        const BsdfSample reject{ false, v3(0, 0, 0), 0, black, 1 };
        switch (d->type) {
        default:
        case IGB200_BSDF_DIFFUSE: { // bsdf/diffuse.art:5-9
            const float     u = rnd.next_f32();
            const float     v = rnd.next_f32();
            const DirSample s = sample_cosine_hemisphere(u, v);
            return BsdfSample{ true, mat3_mul(local, s.dir), s.pdf, kd(), 1.0f };
        }
        case IGB200_BSDF_DIELECTRIC: { // bsdf/dielectric.art:5-21 (adjoint = false)
            const float n1 = d->p[0], n2 = d->p[1];
            const float k     = is_entering ? n1 / n2 : n2 / n1;
            const float cos_o = dot(out_dir, n);
            float       cos_t = 0, factor = 1;
            if (!fresnel(k, cos_o, cos_t, factor)) { cos_t = 0; factor = 1; }
            if (rnd.next_f32() > factor) {
                const V3 t = refract(out_dir, n, k, cos_o, cos_t);
                return BsdfSample{ true, t, 1.0f, cmulf(color(d->color2[0], d->color2[1], d->color2[2]), 1.0f), k };
            }
            return BsdfSample{ true, reflect(out_dir, n), 1.0f, kd(), 1.0f };
        }
        case IGB200_BSDF_THIN_DIELECTRIC: { // make_thin_dielectric_bsdf, bsdf/dielectric.art:30-45
            const float k     = d->p[0] / d->p[1]; // always from outside to inside
            const float cos_o = fabsf(dot(out_dir, n));
            float       cos_t = 0, fterm = 1;
            if (!fresnel(k, cos_o, cos_t, fterm)) fterm = 1; // math::fresnel_dielectric, core/math.art:120-123
            const float F = fterm + (1 - fterm) * fterm / (fterm + 1);
            if (rnd.next_f32() > F) return BsdfSample{ true, neg(out_dir), 1.0f, color(d->color2[0], d->color2[1], d->color2[2]), 1.0f };
            return BsdfSample{ true, normalize(reflect(out_dir, n)), 1.0f, kd(), 1.0f };
        }
        case IGB200_BSDF_MIRROR: // bsdf/conductor.art:5-7
            return BsdfSample{ true, reflect(out_dir, n), 1.0f, kd(), 1.0f };
        case IGB200_BSDF_CONDUCTOR: {
            if (!rough()) { // make_pure_conductor_bsdf, bsdf/conductor.art:16-28
                const float cos_i = dot(out_dir, n);
                return BsdfSample{ true, reflect(out_dir, n), 1.0f, cmul(kd(), fresnel_term(cos_i)), 1.0f };
            }
            // make_rough_base_conductor_bsdf::sample, bsdf/conductor.art:76-97
            const float cos_o = fabsf(dot(out_dir, n));
            if (cos_o <= flt_eps) return reject;
            const V3    m    = sample_vndf_ggx(rnd, local, out_dir, au, av);
            const float spdf = pdf_vndf_ggx(local, out_dir, m, au, av);
            if (dot(m, m) <= flt_eps) return reject;
            const V3 oH     = normalize(m);
            const V3 H      = std::signbit(dot(oH, out_dir)) ? neg(oH) : oH;
            const V3 in_dir = reflect(out_dir, H);
            const float cos_i = fabsf(dot(in_dir, n));
            if (cos_i <= flt_eps) return reject;
            const float cos_h_o = fabsf(dot(out_dir, H));
            const float jacob   = 1 / (4 * cos_h_o);
            const float pdfv    = spdf * jacob;
            return BsdfSample{ true, in_dir, pdfv, cmulf(eval(in_dir, out_dir), safe_div(1, pdfv)), 1.0f };
        }
        }
    }
};

struct Ctx;
static Bsdf make_bsdf(const igb200_shading& sh, int32_t id, const Surf& surf, V3 ray_dir, bool one_sided)
{
    const Tex tex{ &sh };
    Mat3      local = surf.local;
    // make_bumpmap -> make_normal_set (bsdf/map.art:36-44,64-68): only the shading frame changes (adjoint = false)
    for (int guard = 0; guard < 8 && sh.bsdfs[id].type == IGB200_BSDF_BUMPMAP; ++guard) {
        const igb200_bsdf& b     = sh.bsdfs[id];
        const uint32_t     map   = b.q[0].u - 1u;
        const float        delta = 0.001f; // texture_dx / texture_dy, texture/common.art:28-38
        const Color        t0 = tex.eval(map, surf.tu, surf.tv);
        const float        dx = (tex.eval(map, surf.tu + delta, surf.tv).r - t0.r) * (1 / delta);
        const float        dy = (tex.eval(map, surf.tu, surf.tv + delta).r - t0.r) * (1 / delta);
        const V3 N     = normalize(sub(local.c2, mulf(add(mulf(local.c0, dx), mulf(local.c1, dy)), b.p[0])));
        const V3 nn    = ensure_valid_reflection(surf.face_normal, neg(ray_dir), normalize(N));
        const Mat3 trans = align_vectors(local.c2, nn);
        local          = matmul(trans, local);
        id             = (int32_t)b.q[1].u - 1;
    }
    const igb200_bsdf& d = sh.bsdfs[id];
    Bsdf               out{ &d, local, surf.face_normal, surf.is_entering, one_sided, color(d.color[0], d.color[1], d.color[2]), 0, 0 };
    if (d.type == IGB200_BSDF_DIFFUSE && d.q[0].u != 0) out.kd_ = tex.eval(d.q[0].u - 1u, surf.tu, surf.tv);
    if (d.type == IGB200_BSDF_CONDUCTOR) { out.au = d.q[0].f; out.av = d.q[1].f; }
    return out;
}

// ---- lights --------------------------------------------------------------------------------------
struct DirectSample { V3 pos, dir; Color intensity; Pdf pdf; float cos, dist; };

struct Ctx { // the slice of the render state lights and the technique need
    const igo_scene*      scene;
    const igb200_shading* sh;
    float                 scene_radius; // bbox_radius(scene_bbox) * 1.01, light/env.art:76, directional.art:2
};

static inline bool light_delta(const igb200_light& l)
{
    return l.type == IGB200_LIGHT_POINT || l.type == IGB200_LIGHT_SPOT || l.type == IGB200_LIGHT_DIRECTIONAL || l.type == IGB200_LIGHT_SUN;
}
static inline bool light_infinite(const igb200_light& l)
{
    return l.type == IGB200_LIGHT_ENV || l.type == IGB200_LIGHT_DIRECTIONAL || l.type == IGB200_LIGHT_ENV_TEXTURED ||
           l.type == IGB200_LIGHT_ENV_FUNCTION || l.type == IGB200_LIGHT_SUN;
}
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

// make_sphere_area_emitter::sample_direct, light/area.art:248-273: uniform on the sphere, mirrored through the centre
// onto the half facing `from`; pdf = 2 / area
static void sphere_emitter_sample(const Ctx& c, const igb200_light& l, float u, float v, V3 from, Surf& surf, float& pdfv, float& weight)
{
    const Entity e = load_entity(*c.scene, l.entity);
    const Sphere sp{ v3(l.v[0], l.v[1], l.v[2]), l.v[3] };
    const V3     glb_org = xform_point(e.global_mat, sp.origin);
    surf = sphere_surface_element_for_normal(equal_area_square_to_sphere(u, v), sp, e);
    const V3 p  = surf.point;
    const V3 os = sub(from, glb_org), ps = sub(from, p);
    pdfv   = 2 * surf.inv_area;
    weight = 1 / (2 * surf.inv_area);
    if (!(dot(ps, ps) <= dot(os, os))) {
        const V3 op   = sub(p, glb_org);
        const V3 np   = sub(p, mulf(op, 2));
        const V3 norm = normalize(sub(np, glb_org));
        surf          = sphere_surface_element_for_normal(norm, sp, e);
    }
}
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

// ---- core/cdf.art on the blob layout of cdf::make_cdf_2d_from_buffer (:156-160) ---------------------------
struct Cdf1 { // make_cdf_1d_from_buffer :70-72 + make_cdf_1d :38-68
    const float* data;
    int32_t      n;
    float get(int32_t i) const { return i == 0 ? 0.0f : data[i - 1]; }
    float pdf_discrete(int32_t x) const { return get(x + 1) - get(x); }
    int32_t sample_discrete(float u) const
    {
        int32_t first = 0, len_ = n + 1; // interval::binary_search(func_size + 1, get(i) <= u), core/interval.art:7-23
        while (len_ > 0) {
            const int32_t half = len_ / 2, middle = first + half;
            if (get(middle) <= u) { first = middle + 1; len_ -= half + 1; } else { len_ = half; }
        }
        return std::min(std::min(std::max(first - 1, 0), n), n - 1);
    }
    void sample_continuous(float u, int32_t& off, float& pos, float& pdf) const
    {
        off              = sample_discrete(u);
        const float pd   = pdf_discrete(off);
        const float rem  = safe_div(u - get(off), pd);
        pos              = clampf(((float)off + rem) / (float)n, 0, 1);
        pdf              = pd * (float)n;
    }
    void pdf_continuous(float x, int32_t& off, float& pdf) const
    {
        off = std::min(std::max((int32_t)(x * (float)n), 0), n - 1);
        pdf = pdf_discrete(off) * (float)n;
    }
};
struct Cdf2 {
    const float* base;
    int32_t      sx, sy;
    Cdf1 marginal() const { return Cdf1{ base, sy }; }
    Cdf1 conditional(int32_t i) const { return Cdf1{ base + sy + (size_t)i * sx, sx }; }
};
static inline Cdf2 load_cdf2(const igb200_shading& sh, uint32_t off4)
{
    const uint32_t* h = reinterpret_cast<const uint32_t*>(sh.blob) + off4;
    return Cdf2{ reinterpret_cast<const float*>(h + 4), (int32_t)h[0], (int32_t)h[1] };
}

// light/env.art:13-21, core/warp.art:44-57
static inline V3 switch_env_up(V3 v) { return v3(v.x, v.z, v.y); }
static inline void map_env_uv(V3 dir, float& u, float& v)
{
    // acos argument clamped: |dir.z| can round to 1 + 1 ulp, for which the reference's acos (core/warp.art:45) is NaN;
    // the CUDA path clamps the same way (note 3)
    const float theta = acosf(clampf(dir.z, -1, 1));
    float       phi   = atan2f(dir.y, dir.x);
    if (phi < 0) phi += 2 * flt_pi;
    u = Tex::fract(phi / (2 * flt_pi) + 0.25f);
    v = 1 - theta / flt_pi;
}
static inline Mat3 light_transform(const igb200_light& l) { return Mat3{ v3(l.v[0], l.v[1], l.v[2]), v3(l.v[3], l.v[4], l.v[5]), v3(l.v[6], l.v[7], l.v[8]) }; }

// Light::emission of an infinite light along `dir` (light/env.art:93,146-151,163-166)
// cie_wmean + make_cie_sky_light / make_cie_sunny_light, light/cie.art:1-42; `dir` in the light's own frame (Y up)
static Color sky_function(const igb200_shading& sh, const igb200_light& l, V3 dir)
{
    const float* p = reinterpret_cast<const float*>(sh.blob) + l.cdf;
    const bool   variant = l.flags & IGB200_SKYFLAG_VARIANT, has_ground = l.flags & IGB200_SKYFLAG_HAS_GROUND;
    const Color  scale = color(p[0], p[1], p[2]), zenith = color(p[3], p[4], p[5]), ground = color(p[6], p[7], p[8]);
    const float  cos_theta = dir.y;
    if (!has_ground && cos_theta < 0) return black;
    auto wmean = [&](Color c1, Color c2) {
        const float a  = powf(cos_theta + 1.01f, 10);
        const float f1 = a * a / (a * a + 1), f2 = 1 / (a * a + 1);
        return cadd(cmulf(c1, f1), cmulf(c2, f2));
    };
    if (l.tex == IGB200_SKY_CIE) {
        const float c1 = variant ? (1 + 2 * cos_theta) / 3 : 1.0f, c2 = variant ? 0.777777777f : 1.0f;
        return wmean(cmulf(zenith, c1), cmulf(ground, p[10] * c2));
    }
    const V3    sun = v3(p[12], p[13], p[14]);
    if (l.tex == IGB200_SKY_PEREZ) { // calc_perez + make_perez_light, light/cie.art:45-57 (zenith slot = lum)
        const float a = p[9], b = p[10], c_ = p[11], d = p[15], e = p[0];
        const float sun_c = clampf(dot(dir, sun), -1, 1);
        const float sun_a = acosf(sun_c);
        const float A = 1 + a * expf(b / fmaxf(0.01f, cos_theta));
        const float B = 1 + c_ * expf(d * sun_a) + e * sun_c * sun_c;
        return wmean(cmulf(zenith, A * B), ground);
    }
    const float cos_gamma = dot(dir, sun);
    const float gamma     = acosf(clampf(cos_gamma, -1, 1));
    float       c1;
    if (variant) {
        c1 = (0.91f + 10 * expf(-3 * gamma) + 0.45f * cos_gamma * cos_gamma) * (cos_theta >= 0.01f ? 1 - expf(-0.32f / cos_theta) : 1.0f);
    } else {
        const float theta = acosf(clampf(cos_theta, -1, 1)), stheta = acosf(clampf(sun.y, -1, 1));
        c1 = ((1.35f * sinf(5.631f - 3.59f * theta) + 3.12f) * sinf(4.396f - 2.6f * stheta) + 6.37f - theta) / 2.326f
             * expf(gamma * (-0.563f) * ((2.629f - theta) * (1.562f - stheta) + 0.812f));
    }
    return cmul(scale, wmean(cmulf(zenith, p[9] * c1), cmulf(ground, p[10] * p[11])));
}
// half = !has_ground for the CIE skies (light/cie.art:19,42), always false for the Perez sky (:57); light/env.art:104-108
static bool sky_is_hemi(const igb200_light& l) { return !(l.flags & IGB200_SKYFLAG_HAS_GROUND) && l.tex != IGB200_SKY_PEREZ; }

static Color env_emission(const igb200_shading& sh, const igb200_light& l, V3 dir)
{
    if (l.type == IGB200_LIGHT_ENV_FUNCTION) { // light/env.art:49-57 (hemi), :94 (spherical)
        const V3 local = mat3_mul(light_transform(l), dir);
        if (sky_is_hemi(l) && !(local.y > flt_eps)) return black;
        return sky_function(sh, l, local);
    }
    if (l.type != IGB200_LIGHT_ENV_TEXTURED) return light_color(l);
    float u, v;
    map_env_uv(switch_env_up(mat3_mul(light_transform(l), dir)), u, v);
    return cmul(light_color(l), Tex{ &sh }.eval(l.tex, u, v));
}
// Light::pdf_direct of an infinite light, solid angle (light/env.art:96,121-127,152)
static float env_pdf(const igb200_shading& sh, const igb200_light& l, V3 dir)
{
    if (l.type == IGB200_LIGHT_ENV_FUNCTION && sky_is_hemi(l)) { // light/env.art:58-61
        const float y = mat3_mul(light_transform(l), dir).y;
        return y > flt_eps ? y / flt_pi : 0.0f; // cosine_hemisphere_pdf
    }
    if (l.type != IGB200_LIGHT_ENV_TEXTURED || l.cdf == IGB200_NO_CDF) return 1 / (4 * flt_pi);
    const V3    ldir     = switch_env_up(mat3_mul(light_transform(l), dir));
    const float sinTheta = safe_sqrt(1 - ldir.z * ldir.z);
    float       u, v;
    map_env_uv(ldir, u, v);
    const Cdf2 cdf = load_cdf2(sh, l.cdf);
    int32_t    oy, ox;
    float      p1, p2;
    cdf.marginal().pdf_continuous(v, oy, p1);
    cdf.conditional(oy).pdf_continuous(u, ox, p2);
    return safe_div(p1 * p2, sinTheta * flt_pi * flt_pi * 2);
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
    case IGB200_LIGHT_SUN: { // make_sun_light::sample_direct, light/sun.art:11-15 (sample_uniform_cone, core/sampling.art:106-116)
        const V3    dir = v3(l.v[0], l.v[1], l.v[2]);
        const float cos_angle = l.v[3];
        const float sun_radius = sqrtf(1 - cos_angle * cos_angle) / cos_angle, sun_area = flt_pi * sun_radius * sun_radius;
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        float       px, py; // square_to_concentric_disk, core/warp.art:2-22
        {
            const float a = 2 * u - 1, b = 2 * v - 1;
            if (a == 0 && b == 0) { px = 0; py = 0; }
            else if (a * a > b * b) { const float phi = (flt_pi / 4) * safe_div(b, a); px = cosf(phi) * a; py = sinf(phi) * a; }
            else { const float phi = (flt_pi / 2) - (flt_pi / 4) * safe_div(a, b); px = cosf(phi) * b; py = sinf(phi) * b; }
        }
        const float c1 = 1 - cos_angle, n2 = px * px + py * py;
        const float z  = cos_angle + c1 * (1 - n2), sc = safe_sqrt(c1 * (2 - c1 * n2));
        const float den = 2 * flt_pi * (1 - cos_angle);
        const float pdf = fabsf(den) <= flt_eps ? 1.0f : 1 / den; // uniform_cone_pdf: safe_div_one
        const V3    ndir = mat3_mul(make_orthonormal(dir), v3(px * sc, py * sc, z));
        return DirectSample{ v3(0, 0, 0), neg(ndir), cmulf(light_color(l), 1 / (sun_area * pdf)), delta_pdf(), z, std::numeric_limits<float>::infinity() };
    }
    case IGB200_LIGHT_ENV: { // make_environment_light_function_spherical::sample_direct, light/env.art:82-86
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        const V3    d = equal_area_square_to_sphere(u, v);
        const float pdf = 1 / (4 * flt_pi);
        return DirectSample{ add(from.point, mulf(d, c.scene_radius)), d, cmulf(light_color(l), 1 / pdf), solid_pdf(pdf), 1.0f, c.scene_radius };
    }
    case IGB200_LIGHT_ENV_FUNCTION: {
        const Mat3  tr = light_transform(l);
        const float u  = rnd.next_f32();
        const float v  = rnd.next_f32();
        if (sky_is_hemi(l)) { // make_environment_light_function_hemi::sample_direct, light/env.art:32-38
            const DirSample sm  = sample_cosine_hemisphere(u, v);
            const V3        dir = switch_env_up(sm.dir);
            const Color     in  = cmulf(sky_function(*c.sh, l, dir), 1 / sm.pdf);
            const V3        gd  = mat3_left_mul(tr, dir);
            return DirectSample{ add(from.point, mulf(gd, c.scene_radius)), gd, in, solid_pdf(sm.pdf), 1.0f, c.scene_radius };
        }
        const V3    d   = equal_area_square_to_sphere(u, v); // _spherical::sample_direct, light/env.art:82-86
        const float pdf = 1 / (4 * flt_pi);
        return DirectSample{ add(from.point, mulf(d, c.scene_radius)), d, cmulf(sky_function(*c.sh, l, mat3_mul(tr, d)), 1 / pdf), solid_pdf(pdf),
                             1.0f, c.scene_radius };
    }
    case IGB200_LIGHT_ENV_TEXTURED: {
        const Mat3 tr = light_transform(l);
        if (l.cdf == IGB200_NO_CDF) { // make_environment_light -> function_spherical, light/env.art:163-166,82-86
            const float u = rnd.next_f32();
            const float v = rnd.next_f32();
            const V3    d = equal_area_square_to_sphere(u, v);
            const float pdf = 1 / (4 * flt_pi);
            float       tu, tv;
            map_env_uv(switch_env_up(mat3_mul(tr, d)), tu, tv);
            const Color in = cmul(light_color(l), Tex{ c.sh }.eval(l.tex, tu, tv));
            return DirectSample{ add(from.point, mulf(d, c.scene_radius)), d, cmulf(in, 1 / pdf), solid_pdf(pdf), 1.0f, c.scene_radius };
        }
        // make_environment_light_textured::sample_dir / sample_direct, light/env.art:115-133 (the radiance is NOT scaled here)
        const float u0 = rnd.next_f32();
        const float u1 = rnd.next_f32();
        const Cdf2  cdf = load_cdf2(*c.sh, l.cdf);
        int32_t     oy, ox;
        float       py, px, p1, p2;
        cdf.marginal().sample_continuous(u1, oy, py, p1);
        cdf.conditional(oy).sample_continuous(u0, ox, px, p2);
        Color       in    = Tex{ c.sh }.eval(l.tex, px, py);
        if (l.flags & IGB200_LIGHTFLAG_SCALED_NEE) in = cmul(light_color(l), in);
        const float theta = (1 - py) * flt_pi, phi = (px - 0.25f) * 2 * flt_pi;
        const float st = sinf(theta), ct = cosf(theta);
        const V3    dir = v3(st * cosf(phi), st * sinf(phi), ct); // dir_from_spherical
        const float sinTheta = safe_sqrt(1 - dir.z * dir.z);
        const float pdf_dir  = safe_div(p1 * p2, sinTheta * flt_pi * flt_pi * 2);
        const V3    d        = mat3_left_mul(tr, switch_env_up(dir));
        return DirectSample{ add(from.point, mulf(d, c.scene_radius)), d, cmulf(in, 1 / pdf_dir), solid_pdf(pdf_dir), 1.0f, c.scene_radius };
    }
    case IGB200_LIGHT_AREA_SHAPE:
    case IGB200_LIGHT_AREA_SPHERE:
    case IGB200_LIGHT_AREA_PLANE: { // make_area_light::sample_direct, light/area.art:12-26
        const float u = rnd.next_f32();
        const float v = rnd.next_f32();
        Surf  to;
        Pdf   pdf = area_pdf(1);
        float weight;
        if (l.type == IGB200_LIGHT_AREA_PLANE) {
            PlaneEmitter(l).sample_direct(u, v, from.point, to, pdf, weight);
        } else if (l.type == IGB200_LIGHT_AREA_SPHERE) {
            float pdfv;
            sphere_emitter_sample(c, l, u, v, from.point, to, pdfv, weight);
            pdf = area_pdf(pdfv);
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
    case IGB200_LIGHT_AREA_SPHERE: // light/area.art:280-282: 2 / area
        return area_pdf(2 / ellipsoid_area(l.v[3], load_entity(*c.scene, l.entity)));
    case IGB200_LIGHT_AREA_SHAPE: { // light/area.art:60-68,83: prim_coords reinterpreted as the sampling square
        Surf  s;
        float pdfv, w;
        shape_emitter_sample(c, l, surf.pu, surf.pv, s, pdfv, w);
        return area_pdf(pdfv);
    }
    default: return delta_pdf();
    }
}

// make_uniform_light_selector / make_cdf_light_selector / make_hierarchy_light_selector, light/light_selector.art:19-110
struct HEntry { V3 pos, dir; float flux; int32_t id; bool has_dir, is_leaf; };
struct Selector {
    const Ctx* c;
    int32_t    n_inf, n_fin;
    int32_t count() const { return n_inf + n_fin; }
    bool    hierarchical() const { return c->sh->selector == IGB200_SELECTOR_HIERARCHY && n_fin != 0; }
    bool    by_cdf() const { return c->sh->selector == IGB200_SELECTOR_CDF && n_fin != 0; } // make_cdf_light_selector :47-74
    Cdf1    cdf() const { return Cdf1{ reinterpret_cast<const float*>(c->sh->blob) + c->sh->hierarchy, n_fin }; }
    // the finite-light half of the two non-uniform selectors
    int32_t finite_sample(Rnd& rnd, V3 pos, float& pdf) const
    {
        if (by_cdf()) {
            const int32_t off = cdf().sample_discrete(rnd.next_f32());
            pdf               = cdf().pdf_discrete(off);
            return off;
        }
        return hierarchy_sample(rnd, pos, pdf);
    }
    float finite_pdf(int32_t id, V3 pos) const { return by_cdf() ? cdf().pdf_discrete(id) : hierarchy_pdf(id, pos); }
    float   uniform_pdf() const { return count() == 0 ? 1.0f : 1 / (float)count(); }
    int32_t pick(Rnd& rnd, int32_t n) const { return n <= 1 ? 0 : rnd.next_i32(0, n - 1); }

    // ---- light/light_hierarchy.art ----
    const uint32_t* codes() const { return reinterpret_cast<const uint32_t*>(c->sh->blob) + c->sh->hierarchy; }
    const float*    nodes() const { return reinterpret_cast<const float*>(codes() + ((n_fin + 3) / 4) * 4); }
    HEntry load(int32_t id) const // load_entry :24-38
    {
        const float* e = nodes() + 8 * (size_t)id;
        int32_t      index;
        memcpy(&index, e + 7, 4);
        return HEntry{ v3(e[0], e[1], e[2]), v3(e[4], e[5], e[6]), fabsf(e[3]), index < 0 ? -index - 1 : index, !std::signbit(e[3]), index >= 0 };
    }
    static float cost(const HEntry& e, V3 pos) // get_entry_cost :40-52
    {
        const V3    cdir  = sub(e.pos, pos);
        const float dist2 = dot(cdir, cdir);
        const float cos_d = e.has_dir ? fabsf(dot(e.dir, normalize(cdir))) : 1.0f;
        return safe_div(e.flux * cos_d, dist2);
    }
    static float left_prop(const HEntry& l, const HEntry& r, V3 pos) { return 1 / (1 + cost(r, pos) / cost(l, pos)); } // :54-58
    int32_t hierarchy_sample(Rnd& rnd, V3 pos, float& pdf) const // sample_light_id :65-78, make_light_hierarchy :105-112
    {
        pdf = 1.0f;
        if (n_fin == 1) return 0;
        HEntry e = load(0);
        while (!e.is_leaf) {
            const HEntry l = load(e.id), r = load(e.id + 1);
            const float  prop = left_prop(l, r, pos);
            const bool   left = rnd.next_f32() < prop;
            e                 = left ? l : r;
            pdf *= left ? prop : 1 - prop;
        }
        return e.id;
    }
    float hierarchy_pdf(int32_t light_id, V3 pos) const // compute_pdf :80-96
    {
        if (n_fin == 1) return 1.0f;
        uint32_t code = codes()[light_id];
        float    pdf  = 1.0f;
        HEntry   e    = load(0);
        while (!e.is_leaf) {
            const HEntry l = load(e.id), r = load(e.id + 1);
            const float  prop = left_prop(l, r, pos);
            const bool   left = (code & 0x1u) == 0;
            e                 = left ? l : r;
            pdf *= left ? prop : 1 - prop;
            code >>= 1;
        }
        return pdf;
    }

    const igb200_light& sample(Rnd& rnd, V3 pos, float& pdf) const
    {
        if (!hierarchical() && !by_cdf()) {
            pdf              = uniform_pdf();
            const int32_t id = pick(rnd, count());
            return id < n_inf ? c->sh->infinite_lights[id] : c->sh->finite_lights[id - n_inf];
        }
        if (n_inf == 0) return c->sh->finite_lights[finite_sample(rnd, pos, pdf)];
        const float q = rnd.next_f32();
        if (q < 0.5f) {
            pdf = (1 / (float)n_inf) * 0.5f;
            return c->sh->infinite_lights[pick(rnd, n_inf)];
        }
        float               p;
        const igb200_light& l = c->sh->finite_lights[finite_sample(rnd, pos, p)];
        pdf                   = p * (1 - 0.5f);
        return l;
    }
    float pdf(bool infinite, int32_t id, V3 pos) const
    {
        if (!hierarchical() && !by_cdf()) return uniform_pdf();
        if (n_inf == 0) return finite_pdf(id, pos);
        return infinite ? (1 / (float)n_inf) * 0.5f : finite_pdf(id, pos) * (1 - 0.5f);
    }
};

// ---- technique/pathtracer.art ---------------------------------------------------------------------
struct Payload { float inv_pdf; Color contrib; int32_t depth; float eta; };

struct Film {
    float*   px;
    uint32_t w, h;
    float    inv_spi;
    float*   aov[3]; // path + aov_mis: Direct Weights, NEE Weights; infobuffer: Normals, Albedo, Depth (else null)
    static void add(float* img, int32_t pixel, Color c, float inv_spi) // make_standard_accumulator, driver/accumulator.art:23-30
    {
        // non-finite contributions (0 * inf in the reference's MIS weights) are dropped, as in the CUDA path:
        // DESIGN.md section 4.1, note 6
        if (!img || !std::isfinite(c.r + c.g + c.b)) return;
        img[3 * (size_t)pixel + 0] += c.r * inv_spi;
        img[3 * (size_t)pixel + 1] += c.g * inv_spi;
        img[3 * (size_t)pixel + 2] += c.b * inv_spi;
    }
    void splat(int32_t pixel, Color c) const { add(px, pixel, c, inv_spi); }
    void splat_aov(int k, int32_t pixel, Color c) const { add(aov[k], pixel, c, inv_spi); } // make_atomic_aov_image, driver/aov.art:22-36
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
    float                   ib_distance = 0.0f;                  // init_ib_raypayload, infobuffer.art:19-22 (depth = pt.depth)
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
                        const Color emit  = env_emission(*c.sh, l, ray.dir);
                        const float pdf_s = env_pdf(*c.sh, l, ray.dir);
                        const float mis   = tech.enable_nee ? 1 / (1 + pt.inv_pdf * sel.pdf(true, (int32_t)i, ray.org) * pdf_s) : 1.0f;
                        col               = cadd(col, handle_color(tech, cmulf(cmul(pt.contrib, emit), mis)));
                    }
                }
                if (inflights > 0) {
                    film.splat_aov(0, pixel, col); // aov_di, pathtracer.art:158
                    film.splat(pixel, col);
                }
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
        const Bsdf bsdf = make_bsdf(*c.sh, mat.bsdf, surf, ray.dir, (tech.flags & IGB200_TECHFLAG_ONE_SIDED_DIFFUSE) != 0);
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

        if (tech.type == IGB200_TECH_INFOBUFFER) { // make_infobuffer_renderer, technique/infobuffer.art:28-104
            const bool handle_specular = (tech.flags & IGB200_TECHFLAG_IB_HANDLE_SPECULAR) != 0;
            const V3   out_dir         = neg(ray.dir);
            if (!handle_specular || !bsdf.is_specular()) { // on_hit :36-63
                Color     albedo  = black;
                const int samples = 16;
                Rnd       arnd{ hash_combine(0x811C9DC5u, (uint32_t)pixel), 1u }; // create_random_generator(hash_combine(hash_init(), pixel))
                for (int i = 0; i < samples; ++i) {
                    const BsdfSample ms = bsdf.sample(arnd, out_dir);
                    if (ms.valid) {
                        const float cos = fabsf(dot(ms.in_dir, surf.local.c2));
                        albedo          = cadd(albedo, cmulf(ms.col, safe_div(ms.pdf, (float)samples * cos))); // without the cosine term
                    }
                }
                const float d = hit.distance + (handle_specular ? ib_distance : 0.0f);
                film.splat_aov(0, pixel, color(surf.local.c2.x, surf.local.c2.y, surf.local.c2.z));
                film.splat_aov(1, pixel, color(fminf_(albedo.r, 1), fminf_(albedo.g, 1), fminf_(albedo.b, 1))); // color_saturate
                film.splat_aov(2, pixel, color(d, d, d));
            }
            // on_bounce :65-91
            if (!(handle_specular && bsdf.is_specular()) || !(pt.depth <= tech.max_depth)) return;
            const BsdfSample ms = bsdf.sample(rnd, out_dir);
            if (!ms.valid) return;
            pt.depth    = pt.depth + 1;
            ib_distance = ib_distance + hit.distance;
            ray         = make_ray(surf.point, ms.in_dir, offset, flt_max, ray_flag_bounce);
            st.bounce++;
            continue;
        }

        // on_hit, pathtracer.art:112-132
        if (emissive && surf.is_entering) {
            const float d = -dot(ray.dir, surf.local.c2);
            if (d > flt_eps) {
                const igb200_light& l     = c.sh->finite_lights[mat.light];
                const Color         emit  = light_color(l);
                const Pdf           pdf   = light_pdf_direct(c, l, ray, surf);
                const float         pdf_s = pdf.as_solid(d, hit.distance * hit.distance);
                const float         mis   = tech.enable_nee ? 1 / (1 + pt.inv_pdf * sel.pdf(false, mat.light, ray.org) * pdf_s) : 1.0f;
                const Color         contrib = handle_color(tech, cmulf(cmul(pt.contrib, emit), mis));
                film.splat_aov(0, pixel, contrib); // aov_di, pathtracer.art:128
                film.splat(pixel, contrib);
            }
        }

        // on_shadow, pathtracer.art:52-110
        const V3 out_dir = neg(ray.dir);
        if (tech.enable_nee && !bsdf.is_specular() && sel.count() != 0 && pt.depth + 1 <= tech.max_depth) {
            float               sel_pdf;
            const igb200_light& l  = sel.sample(rnd, surf.point, sel_pdf);
            const DirectSample  ls = light_sample_direct(c, l, rnd, surf);
            const float pdf_l_s    = ls.pdf.as_solid(ls.cos, ls.dist * ls.dist) * sel_pdf;
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
                if (sh.prim_id < 0) {
                    film.splat_aov(1, pixel, contrib); // on_shadow_miss: aov_nee, pathtracer.art:202-208
                    film.splat(pixel, contrib);
                }
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

// Texture lookups on their own (tests pin them against an independent numpy restatement).
extern "C" int igo_texture_eval(const igb200_shading* sh, uint32_t id, const float* uv, uint64_t n, float* rgb)
{
    if (!sh || !uv || !rgb || id >= sh->texture_count) return -1;
    const Tex tex{ sh };
    for (uint64_t i = 0; i < n; ++i) {
        const Color c = tex.eval(id, uv[2 * i], uv[2 * i + 1]);
        rgb[3 * i] = c.r; rgb[3 * i + 1] = c.g; rgb[3 * i + 2] = c.b;
    }
    return 0;
}

// One render iteration (Device::render, src/runtime/device/Device.cpp:1588-1612 -> cpu_trace).
// make_fishlens_camera::compute_d, camera/fishlens.art:38-50 (fov = pi)
static V3 fishlens_dir(const Mat3& view, float xasp, float yasp, float x, float y)
{
    const float nx = x * xasp, ny = y * yasp;
    const float r     = sqrtf(nx * nx + ny * ny);
    const float theta = r * flt_pi / 2;
    const float sT = sinf(theta), cT = cosf(theta);
    const float sP = r < flt_eps ? 0.0f : ny / r, cP = r < flt_eps ? 0.0f : nx / r;
    return normalize(mat3_mul(view, v3(sT * cP, sT * sP, cT)));
}

// film: w*h*3 floats, accumulated into (the caller keeps the iteration count).  stats3 = camera / bounce / shadow rays.
extern "C" int igo_render(const igo_scene* scene, const igb200_shading* sh, const igb200_render_settings* rs, float* film,
                          uint64_t* stats3, int threads)
{
    return igo_render_aov(scene, sh, rs, film, nullptr, nullptr, nullptr, stats3, threads);
}

// The same with the technique's AOV images (each w*h*3 floats or NULL): path + IGB200_TECHFLAG_AOV_MIS writes
// aov0 = "Direct Weights", aov1 = "NEE Weights"; IGB200_TECH_INFOBUFFER writes aov0..2 = "Normals", "Albedo", "Depth"
// and leaves the film alone (LockFramebuffer).
extern "C" int igo_render_aov(const igo_scene* scene, const igb200_shading* sh, const igb200_render_settings* rs, float* film,
                              float* aov0, float* aov1, float* aov2, uint64_t* stats3, int threads)
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
    Film film_{ film, w, h, 1 / (float)rs->spi, { nullptr, nullptr, nullptr } };
    if (sh->technique.type == IGB200_TECH_INFOBUFFER) {
        film_.aov[0] = aov0; film_.aov[1] = aov1; film_.aov[2] = aov2;
    } else if (sh->technique.type == IGB200_TECH_PATH && (sh->technique.flags & IGB200_TECHFLAG_AOV_MIS)) {
        film_.aov[0] = aov0; film_.aov[1] = aov1;
    }

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
                            const V3    d  = cam.type == IGB200_CAMERA_FISHLENS ? fishlens_dir(view, cam.scale[0], cam.scale[1], nx, ny)
                                                                                : normalize(mat3_mul(view, v3(cam.scale[0] * nx, cam.scale[1] * ny, 1)));
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

// ig_tonemap_pipeline (src/artic/entrypoints/tonemap.art:1-70) over n pixels of RGB sums; `scale` already carries
// 1 / IterationCount (Device::tonemap, Device.cpp:1662-1687).  Packed colour = a << 24 | r << 16 | g << 8 | b.
extern "C" void igo_tonemap(const float* in, uint64_t n, uint32_t method, int use_gamma, float scale, float exposure_factor,
                            float exposure_offset, uint32_t* out)
{
    auto pack = [](uint32_t r, uint32_t g, uint32_t b, uint32_t a) { return (a << 24) | (r << 16) | (g << 8) | b; };
    auto byte = [](float c) { return (uint32_t)(clampf(c, 0.0f, 1.0f) * 255.0f) & 0xFFu; }; // make_byte_color: `as u8` truncates
    auto gamma = [](float x) { return x <= 0.0031308f ? 12.92f * x : 1.055f * powf(x, 0.416666667f) - 0.055f; }; // core/color.art:136-142
    auto curve = [&](float L) -> float { // mod tonemapping, core/color.art:101-134
        switch (method) {
        case 0: return L;
        case 1: return safe_div(L, 1.0f + L);
        case 2: return safe_div(L * (1.0f + L / (4.0f * 4.0f)), 1.0f + L);
        case 3: return safe_div(L * (2.51f * L + 0.03f), L * (2.43f * L + 0.59f) + 0.14f);
        default: {
            auto f = [](float x) {
                const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
                return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
            };
            return f(L) / f(11.2f);
        }
        }
    };
    for (uint64_t i = 0; i < n; ++i) {
        const float r = in[3 * i] * scale, g = in[3 * i + 1] * scale, b = in[3 * i + 2] * scale;
        const float X = 0.4124564f * r + 0.3575761f * g + 0.1804375f * b; // srgb_to_xyY, core/color.art:84-99
        const float Y = 0.2126729f * r + 0.7151522f * g + 0.0721750f * b;
        const float Z = 0.0193339f * r + 0.1191920f * g + 0.9503041f * b;
        const float s = X + Y + Z;
        float       cx = 0, cy = 0, cY = 0;
        if (!(s <= flt_eps)) { cx = X / s; cy = Y / s; cY = Y; }
        if (std::isnan(cY)) out[i] = pack(0, 255, 255, 255);
        else if (!std::isfinite(cY)) out[i] = pack(255, 0, 150, 255);
        else if (cx < 0 || cy < 0 || cY < 0) out[i] = pack(255, 255, 0, 255);
        else {
            const float nL = curve(exposure_factor * cY + exposure_offset);
            float       R = 0, G = 0, B = 0; // xyY_to_srgb, core/color.art:77-93
            if (!(cy <= flt_eps)) {
                const float x2 = cx * nL / cy, y2 = nL, z2 = (1 - cx - cy) * nL / cy;
                R = 3.2404542f * x2 - 1.5371385f * y2 - 0.4985314f * z2;
                G = -0.9692660f * x2 + 1.8760108f * y2 + 0.0415560f * z2;
                B = 0.0556434f * x2 - 0.2040259f * y2 + 1.0572252f * z2;
            }
            if (use_gamma) { R = gamma(R); G = gamma(G); B = gamma(B); }
            out[i] = pack(byte(R), byte(G), byte(B), 255);
        }
    }
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
