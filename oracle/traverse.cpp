// TEST INFRASTRUCTURE ONLY (see oracle.h).
// The reference's scalar CPU traversal, restated function by function.
#include "oracle.h"
#include "scene.h"

#include <algorithm>
#include <cfloat>
#include <thread>

namespace igo {
namespace {

constexpr float FltEps = 1.1920928955078125e-7f; // flt_eps
constexpr float FltMax = 3.402823466e+38f;

// make_default_min_max: src/artic/traversal/intersection.art:46-50
static inline float fminf_(float x, float y) { return x < y ? x : y; }
static inline float fmaxf_(float x, float y) { return x > y ? x : y; }

struct BBox { V3 min, max; };

// intersect_ray_box, unordered variant: src/artic/traversal/intersection.art:170-181
static inline void intersect_ray_box(const Ray& ray, const BBox& bbox, float& tentry, float& texit)
{
    const V3 t0 = add(mul(ray.inv_dir, bbox.min), ray.inv_org);
    const V3 t1 = add(mul(ray.inv_dir, bbox.max), ray.inv_org);
    // fmaxmaxf(a,b,c) = fmax(fmax(a,b),c); fminmaxf(a,b,c) = fmax(fmin(a,b),c)
    tentry = fmaxf_(fmaxf_(fminf_(t0.x, t1.x), fminf_(t0.y, t1.y)), fmaxf_(fminf_(t0.z, t1.z), ray.tmin));
    // fminminf(a,b,c) = fmin(fmin(a,b),c); fmaxminf(a,b,c) = fmin(fmax(a,b),c)
    texit = fminf_(fminf_(fmaxf_(t0.x, t1.x), fmaxf_(t0.y, t1.y)), fminf_(fmaxf_(t0.z, t1.z), ray.tmax));
}

// intersect_ray_box_single_section: intersection.art:194-203
static inline bool intersect_ray_box_single_section(const Ray& ray, const BBox& bbox, float& entry, float& exit)
{
    intersect_ray_box(ray, bbox, entry, exit);
    return (entry <= exit) & (exit >= 0);
}

// intersect_ray_tri_mt_gen, no back-face culling: intersection.art:71-101
static inline bool intersect_ray_tri(const Ray& ray, V3 v0, V3 e1, V3 e2, V3 n, float& t, float& u, float& v)
{
    const float tolerance = -FltEps;
    const V3    c         = sub(v0, ray.org);
    const V3    r         = cross(ray.dir, c);
    const float det       = dot(n, ray.dir);
    const float inv_det   = 1.0f / det;
    const float uu        = dot(r, e2) * inv_det;
    const float vv        = dot(r, e1) * inv_det;
    const float w         = 1.0f - uu - vv;
    bool mask = uu >= tolerance;
    mask &= vv >= tolerance;
    mask &= w >= tolerance;
    if (!mask) return false;
    const float tt = dot(c, n) * inv_det;
    mask &= tt >= ray.tmin;
    mask &= tt <= ray.tmax;
    if (!mask) return false;
    t = tt;
    u = uu > 0 ? uu : 0;
    v = vv > 0 ? vv : 0;
    return true;
}

// intersect_sphere: src/artic/shapes/sphere.art:104-132 (+ sphere_map_uv :1-6, spherical_from_dir core/warp.art:44-48)
static inline bool intersect_sphere(const Sphere& sphere, const Ray& ray, Hit& out)
{
    const float pi = 3.14159265359f;
    const V3    L  = sub(ray.org, sphere.origin);
    const float S  = -dot(L, ray.dir);
    const float D2 = dot(ray.dir, ray.dir);
    const float L2 = dot(L, L);
    const float R2 = sphere.radius * sphere.radius * D2;
    const float M2 = L2 * D2 - S * S;
    if ((S < 0) || (M2 > R2)) return false;
    const float Q   = sqrtf(R2 - M2);
    const float t0_ = (S - Q) / D2;
    const float t1_ = (S + Q) / D2;
    const float t0  = t0_ > t1_ ? t1_ : t0_;
    const float t1  = t0_ > t1_ ? t0_ : t1_;
    const float tm  = t0 < ray.tmin ? t1 : t0;
    if (tm >= ray.tmin && tm <= ray.tmax) {
        const float ir = 1.0f / sphere.radius;
        const V3    d  = v3((L.x + ray.dir.x * tm) * ir, (L.y + ray.dir.y * tm) * ir, (L.z + ray.dir.z * tm) * ir);
        // spherical_from_dir(make_vec3(dir.y, -dir.x, dir.z))
        const float theta = acosf(d.z);
        float       phi   = atan2f(-d.x, d.y);
        if (phi < 0) phi += 2 * pi;
        out = Hit{ tm, phi / (2 * pi), theta / pi, 0, -1 };
        return true;
    }
    return false;
}

// alloc_stack: src/artic/traversal/stack.art:52-123 (top cached outside the arrays)
struct Stack {
    int32_t nodes[64];
    float   tmins[64];
    int32_t node = 0;
    float   tmin = FltMax;
    int     ptr  = -1;
    void push(int32_t n, float t) { ++ptr; nodes[ptr] = node; tmins[ptr] = tmin; node = n; tmin = t; }
    void push_after(int32_t n, float t) { ++ptr; nodes[ptr] = n; tmins[ptr] = t; }
    void pop(int32_t& n, float& t) { n = node; t = tmin; node = nodes[ptr]; tmin = tmins[ptr]; --ptr; }
    bool is_empty() const { return node == 0; }
};

static inline BBox node_bbox(const Node4& n, int i)
{
    return BBox{ v3(n.bounds[0][i], n.bounds[2][i], n.bounds[4][i]), v3(n.bounds[1][i], n.bounds[3][i], n.bounds[5][i]) };
}

// cpu_traverse_helper_prim with vector_width = 1, single = false:
// src/artic/traversal/mapping_cpu.art:261-389
static Hit traverse_prim(Ray ray, const PrimBvh& bvh, bool any_hit, igo_counters& cnt)
{
    Hit   hit        = empty_hit(ray.tmax);
    bool  terminated = false;
    Stack stack;
    stack.push(1, ray.tmin);

    while (true) {
        // Cull nodes
        bool exit = false;
        while (true) {
            if (stack.is_empty()) { exit = true; break; }
            const bool active = (stack.tmin <= ray.tmax) & !terminated;
            if (active) break;
            int32_t n; float t;
            stack.pop(n, t);
        }
        if (exit) break;

        // Intersect inner nodes
        bool cull = false;
        while (stack.node > 0) {
            int32_t id; float tt;
            stack.pop(id, tt);
            const Node4& node = bvh.nodes[id - 1];
            cnt.prim_nodes++;
            bool pushed = false;
            for (int i = 0; i < 4; ++i) {
                const int32_t child_id = node.child[i];
                if (child_id == 0) break;
                float tentry, texit;
                intersect_ray_box(ray, node_bbox(node, i), tentry, texit);
                const bool miss = texit < tentry;
                if (!miss) {
                    const float thit = tentry;
                    if (any_hit || stack.tmin > thit)
                        stack.push(child_id, thit);
                    else
                        stack.push_after(child_id, thit);
                    pushed = true;
                }
            }
            if (!pushed) { cull = true; break; }
        }
        if (cull) continue;

        if (stack.node < 0) {
            bool    active = (stack.tmin <= ray.tmax) & !terminated;
            int32_t id; float tt;
            stack.pop(id, tt);
            int32_t prim_id = ~id;
            while (true) {
                const Tri4& tri = bvh.tris[prim_id++];
                cnt.tri_packets++;
                for (int i = 0; i < 4; ++i) {
                    if (tri.prim_id[i] == -1) break;
                    if (active) {
                        cnt.tri_tests++;
                        float t, u, v;
                        if (intersect_ray_tri(ray, v3(tri.v0[0][i], tri.v0[1][i], tri.v0[2][i]),
                                              v3(tri.e1[0][i], tri.e1[1][i], tri.e1[2][i]),
                                              v3(tri.e2[0][i], tri.e2[1][i], tri.e2[2][i]),
                                              v3(tri.n[0][i], tri.n[1][i], tri.n[2][i]), t, u, v)) {
                            hit      = Hit{ t, u, v, tri.prim_id[i] & 0x7FFFFFFF, -1 };
                            ray.tmax = t;
                            if (any_hit) { terminated = true; active = false; }
                        }
                    }
                    if (any_hit && terminated) return hit;
                }
                if (tri.prim_id[3] < 0) break;
            }
        }
    }
    return hit;
}

// cpu_traverse_helper: src/artic/traversal/mapping_cpu.art:398-493
static Hit traverse_scene(Ray ray, const igo_scene& s, int type, bool any_hit, Hit hit, igo_counters& cnt)
{
    const SceneBvh& bvh        = s.scene[type];
    bool            terminated = false;
    Stack           stack;
    stack.push(1, ray.tmin);

    while (true) {
        bool exit = false;
        while (true) {
            if (stack.is_empty()) { exit = true; break; }
            const bool active = (stack.tmin <= ray.tmax) & !terminated;
            if (active) break;
            int32_t n; float t;
            stack.pop(n, t);
        }
        if (exit) break;

        bool cull = false;
        while (stack.node > 0) {
            int32_t id; float tt;
            stack.pop(id, tt);
            const Node4& node = bvh.nodes[id - 1];
            cnt.scene_nodes++;
            bool pushed = false;
            for (int i = 0; i < 4; ++i) {
                const int32_t child_id = node.child[i];
                if (child_id == 0) break;
                float tentry, texit;
                intersect_ray_box(ray, node_bbox(node, i), tentry, texit);
                const bool miss = texit < tentry;
                if (!miss) {
                    const float thit = tentry;
                    if (any_hit || stack.tmin > thit)
                        stack.push(child_id, thit);
                    else
                        stack.push_after(child_id, thit);
                    pushed = true;
                }
            }
            if (!pushed) { cull = true; break; }
        }
        if (cull) continue;

        if (stack.node < 0) {
            bool    active = (stack.tmin <= ray.tmax) & !terminated;
            int32_t id; float tt;
            stack.pop(id, tt);
            int32_t ref_id = ~id;
            while (true) {
                const igb200_entity_leaf& leaf = bvh.leaves[ref_id++];
                if (check_ray_visibility(ray, leaf.flags)) {
                    const BBox lb{ v3(leaf.bbox_min[0], leaf.bbox_min[1], leaf.bbox_min[2]),
                                   v3(leaf.bbox_max[0], leaf.bbox_max[1], leaf.bbox_max[2]) };
                    float tmin, tmax;
                    if (intersect_ray_box_single_section(ray, lb, tmin, tmax)) {
                        if (tmin <= hit.distance) {
                            const Ray local_ray = transform_ray(ray, leaf.local);
                            cnt.entities++;
                            Hit local_hit;
                            if (type == (int)IGB200_PRIM_TRIMESH) {
                                local_hit = traverse_prim(local_ray, s.shapes[leaf.shape_id], any_hit, cnt);
                            } else {
                                // make_scene_local_handler_sphere: src/artic/shapes/sphere.art:134-143
                                if (!intersect_sphere(s.spheres[leaf.shape_id], local_ray, local_hit))
                                    local_hit = empty_hit(local_ray.tmax);
                            }
                            if (active) {
                                if (local_hit.prim_id != -1) {
                                    if (local_hit.distance <= hit.distance) {
                                        hit = Hit{ local_hit.distance, local_hit.u, local_hit.v, local_hit.prim_id,
                                                   leaf.entity_id & 0x7FFFFFFF };
                                        ray.tmax = hit.distance;
                                        if (any_hit) { terminated = true; active = false; }
                                    }
                                }
                            }
                        }
                    }
                }
                if (any_hit && terminated) return hit;
                if (leaf.entity_id < 0) break;
            }
        }
    }
    return hit;
}

template <typename F>
void parallel_chunks(uint64_t n, int threads, F&& f)
{
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    threads = (int)std::min<uint64_t>((uint64_t)threads, std::max<uint64_t>(1, n));
    if (threads == 1) { f(0, 0, n); return; }
    std::vector<std::thread> pool;
    const uint64_t           chunk = (n + threads - 1) / threads;
    for (int k = 0; k < threads; ++k) {
        const uint64_t b = std::min(n, chunk * k), e = std::min(n, chunk * (k + 1));
        pool.emplace_back([&f, k, b, e]() { f(k, b, e); });
    }
    for (auto& t : pool) t.join();
}

} // namespace

// cpu_traverse_primary / cpu_traverse_secondary for one ray: loop over the primitive types carrying the
// best hit (src/artic/driver/mapping_cpu.art:366-424).
Hit trace_one(const igo_scene& s, const Ray& ray, bool any_hit)
{
    igo_counters cnt{ 0, 0, 0, 0, 0 };
    Hit          hit = empty_hit(ray.tmax);
    for (int type = 0; type < 2; ++type) {
        if (s.scene[type].nodes.empty()) continue;
        hit = traverse_scene(ray, s, type, any_hit, hit, cnt);
    }
    return hit;
}

} // namespace igo

using namespace igo;

// cpu_traverse_primary / cpu_traverse_secondary: src/artic/driver/mapping_cpu.art:366-424
extern "C" void igo_trace(const igo_scene* s, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                          int threads, int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v,
                          igo_counters* counters)
{
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<igo_counters> per(threads, igo_counters{ 0, 0, 0, 0, 0 });
    parallel_chunks(n, threads, [&](int k, uint64_t b, uint64_t e) {
        igo_counters cnt{ 0, 0, 0, 0, 0 };
        for (uint64_t i = b; i < e; ++i) {
            const igb200_ray& sr  = rays[i];
            const Ray         ray = make_ray(v3(sr.org[0], sr.org[1], sr.org[2]), v3(sr.dir[0], sr.dir[1], sr.dir[2]),
                                             sr.tmin, sr.tmax, ray_flags);
            Hit hit = empty_hit(ray.tmax);
            for (int type = 0; type < 2; ++type) {
                if (s->scene[type].nodes.empty()) continue;
                hit = traverse_scene(ray, *s, type, any_hit != 0, hit, cnt);
            }
            if (ent_id) ent_id[i] = hit.ent_id;
            if (prim_id) prim_id[i] = hit.prim_id;
            if (t) t[i] = hit.distance;
            if (u) u[i] = hit.u;
            if (v) v[i] = hit.v;
        }
        per[k] = cnt;
    });
    if (counters) {
        *counters = igo_counters{ 0, 0, 0, 0, 0 };
        for (auto& c : per) {
            counters->scene_nodes += c.scene_nodes;
            counters->prim_nodes += c.prim_nodes;
            counters->tri_packets += c.tri_packets;
            counters->tri_tests += c.tri_tests;
            counters->entities += c.entities;
        }
    }
}

extern "C" void igo_brute(const igo_scene* s, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int threads,
                          int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v, uint8_t* tie,
                          float* second_t)
{
    parallel_chunks(n, threads, [&](int, uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            const igb200_ray& sr  = rays[i];
            const Ray         ray = make_ray(v3(sr.org[0], sr.org[1], sr.org[2]), v3(sr.dir[0], sr.dir[1], sr.dir[2]),
                                             sr.tmin, sr.tmax, ray_flags);
            Hit   best   = empty_hit(ray.tmax);
            float second = FltMax;
            bool  is_tie = false;
            auto  offer  = [&](const Hit& h) {
                if (best.prim_id == -1 || h.distance < best.distance) {
                    if (best.prim_id != -1) second = fminf_(second, best.distance);
                    best   = h;
                    is_tie = false;
                } else {
                    if (h.distance == best.distance) is_tie = true;
                    second = fminf_(second, h.distance);
                }
            };
            for (int type = 0; type < 2; ++type) {
                for (const igb200_entity_leaf& leaf : s->flat_leaves[type]) {
                    if (!check_ray_visibility(ray, leaf.flags)) continue;
                    const Ray local = transform_ray(ray, leaf.local);
                    if (type == (int)IGB200_PRIM_TRIMESH) {
                        for (const FlatTri& tr : s->shapes[leaf.shape_id].flat) {
                            float tt, uu, vv;
                            if (intersect_ray_tri(local, tr.v0, tr.e1, tr.e2, tr.n, tt, uu, vv))
                                offer(Hit{ tt, uu, vv, tr.prim_id, leaf.entity_id });
                        }
                    } else {
                        Hit h;
                        if (intersect_sphere(s->spheres[leaf.shape_id], local, h)) {
                            h.ent_id = leaf.entity_id;
                            offer(h);
                        }
                    }
                }
            }
            if (ent_id) ent_id[i] = best.ent_id;
            if (prim_id) prim_id[i] = best.prim_id;
            if (t) t[i] = best.distance;
            if (u) u[i] = best.u;
            if (v) v[i] = best.v;
            if (tie) tie[i] = is_tie ? 1 : 0;
            if (second_t) second_t[i] = second;
        }
    });
}

// Interface::loadRayList: src/runtime/device/Device.cpp:585-607
extern "C" void igo_normalize_rays(igb200_ray* rays, uint64_t n)
{
    for (uint64_t i = 0; i < n; ++i) {
        float* d    = rays[i].dir;
        float  norm = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); // Eigen norm(): plain sum of squares
        if (norm < FLT_EPSILON) norm = 1;
        d[0] /= norm;
        d[1] /= norm;
        d[2] /= norm;
    }
}

extern "C" int igo_kat_ray_triangle(const float* tri, const float* org, const float* dir, float tmin, float tmax,
                                    float* out_tuv)
{
    const V3  p0 = v3(tri[0], tri[1], tri[2]), p1 = v3(tri[3], tri[4], tri[5]), p2 = v3(tri[6], tri[7], tri[8]);
    const V3  e1 = sub(p0, p1), e2 = sub(p2, p0);
    const Ray ray = make_ray(v3(org[0], org[1], org[2]), v3(dir[0], dir[1], dir[2]), tmin, tmax, 0);
    float     t, u, v;
    if (!intersect_ray_tri(ray, p0, e1, e2, cross(e1, e2), t, u, v)) return 0;
    if (out_tuv) { out_tuv[0] = t; out_tuv[1] = u; out_tuv[2] = v; }
    return 1;
}

// intersect_ray_box_single: src/artic/traversal/intersection.art:183-192
extern "C" int igo_kat_ray_box(const float* box, const float* org, const float* dir, float tmin, float tmax, float* out_t)
{
    const Ray  ray = make_ray(v3(org[0], org[1], org[2]), v3(dir[0], dir[1], dir[2]), tmin, tmax, 0);
    const BBox bb{ v3(box[0], box[1], box[2]), v3(box[3], box[4], box[5]) };
    float      entry, exit;
    intersect_ray_box(ray, bb, entry, exit);
    if (!((entry <= exit) & (exit >= 0))) return 0;
    if (out_t) *out_t = entry < 1e-5f ? exit : entry;
    return 1;
}
