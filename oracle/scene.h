// TEST INFRASTRUCTURE ONLY (see oracle.h).
// Reference CPU data layouts (restated) and the oracle scene container.
#pragma once
#include "../include/ignis_b200.h"
#include "bvh_build.h"

#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace igo {

// ---- scalar math with the reference's contraction pattern -----------------------------------
struct V3 { float x, y, z; };
static inline V3 v3(float x, float y, float z) { return V3{ x, y, z }; }
static inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline V3 neg(V3 a) { return v3(-a.x, -a.y, -a.z); }
// src/artic/core/vector.art:99  vec3_dot = fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z))
static inline float dot(V3 a, V3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
// src/artic/core/vector.art:106-109 (no fma)
static inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// src/artic/core/vector.art:100 vec4_dot with w = 1 / 0, as used by mat3x4_mul (core/matrix.art:95-98,232-233)
static inline float dot4(float ax, float ay, float az, float aw, float bx, float by, float bz, float bw)
{
    return fmaf(ax, bx, fmaf(ay, by, fmaf(az, bz, aw * bw)));
}

// src/artic/core/common.art:93-98
static inline float prodsign(float x, float y)
{
    uint32_t a, b;
    std::memcpy(&a, &x, 4);
    std::memcpy(&b, &y, 4);
    a ^= (b & 0x80000000u);
    float r;
    std::memcpy(&r, &a, 4);
    return r;
}
static inline float safe_rcp(float x)
{
    const float min_rcp = 1e-8f;
    return ((x > 0 ? x : -x) < min_rcp) ? prodsign(3.402823466e+38f, x) : 1.0f / x;
}

// src/artic/traversal/ray.art:9-39
struct Ray {
    V3       org, dir, inv_dir, inv_org;
    float    tmin, tmax;
    uint32_t flags;
};
static inline Ray make_ray(V3 org, V3 dir, float tmin, float tmax, uint32_t flags)
{
    Ray r;
    r.org     = org;
    r.dir     = dir;
    r.inv_dir = v3(safe_rcp(dir.x), safe_rcp(dir.y), safe_rcp(dir.z));
    r.inv_org = neg(mul(org, r.inv_dir));
    r.tmin    = tmin;
    r.tmax    = tmax;
    r.flags   = flags;
    return r;
}
// src/artic/traversal/ray.art:51
static inline bool check_ray_visibility(const Ray& ray, uint32_t flags)
{
    const uint32_t mask = 0xFu;
    return (ray.flags & mask) == ((ray.flags & flags) & mask);
}
// src/artic/traversal/ray.art:56-59; matrix is column major (col[4] of Vec3)
static inline Ray transform_ray(const Ray& ray, const float* m)
{
    const V3 o = v3(dot4(m[0], m[3], m[6], m[9], ray.org.x, ray.org.y, ray.org.z, 1.0f),
                    dot4(m[1], m[4], m[7], m[10], ray.org.x, ray.org.y, ray.org.z, 1.0f),
                    dot4(m[2], m[5], m[8], m[11], ray.org.x, ray.org.y, ray.org.z, 1.0f));
    const V3 d = v3(dot4(m[0], m[3], m[6], m[9], ray.dir.x, ray.dir.y, ray.dir.z, 0.0f),
                    dot4(m[1], m[4], m[7], m[10], ray.dir.x, ray.dir.y, ray.dir.z, 0.0f),
                    dot4(m[2], m[5], m[8], m[11], ray.dir.x, ray.dir.y, ray.dir.z, 0.0f));
    return make_ray(o, d, ray.tmin, ray.tmax, ray.flags);
}

// src/artic/traversal/intersection.art:1-8
struct Hit {
    float   distance;
    float   u, v;
    int32_t prim_id;
    int32_t ent_id;
};
static inline Hit empty_hit(float tmax) { return Hit{ tmax, 0.0f, 0.0f, -1, -1 }; }

// ---- reference CPU layouts -------------------------------------------------------------------
// src/artic/traversal/mapping_cpu.art:3-7
struct Node4 {
    float   bounds[6][4]; // lo_x[4], hi_x[4], lo_y[4], hi_y[4], lo_z[4], hi_z[4]
    int32_t child[4];     // > 0 inner (id + 1), < 0 ~leaf_index, 0 unused
    int32_t pad[4];
};
static_assert(sizeof(Node4) == 128, "Node4 must be 128 bytes");
// src/artic/shapes/trimesh.art:116-122
struct Tri4 {
    float   v0[3][4], e1[3][4], e2[3][4], n[3][4];
    int32_t prim_id[4]; // -1 invalid lane; sign bit of lane 3 = last packet of the leaf
};
static_assert(sizeof(Tri4) == 208, "Tri4 must be 208 bytes");
static_assert(sizeof(igb200_entity_leaf) == 96, "EntityLeaf1 must be 96 bytes");

struct FlatTri { V3 v0, e1, e2, n; int32_t prim_id; };

struct PrimBvh {
    std::vector<Node4>   nodes;
    std::vector<Tri4>    tris;
    std::vector<FlatTri> flat; // prim-id order, for the brute-force check
};

struct Sphere { V3 origin; float radius; };

struct SceneBvh {
    std::vector<Node4>              nodes;
    std::vector<igb200_entity_leaf> leaves; // in BVH leaf order, with sentinel bits
};

} // namespace igo

struct igo_scene {
    std::vector<igo::PrimBvh>       shapes;  // by shape id (empty for spheres)
    std::vector<igo::Sphere>        spheres; // by shape id (radius 0 for meshes)
    std::vector<uint32_t>           shape_type;
    igo::SceneBvh                   scene[2];
    std::vector<igb200_entity_leaf> flat_leaves[2]; // input order, sentinel stripped
    // copies of the tables the shading path reads (render.cpp)
    std::vector<igb200_lookup_entry> lookups;
    std::vector<uint8_t>             shape_data;
    std::vector<float>               entities; // 36 floats per entity
    float                            bbox[6] = { 0, 0, 0, 0, 0, 0 };
};

namespace igo {
// One ray through cpu_traverse_primary / cpu_traverse_secondary (traverse.cpp); used by render.cpp.
Hit trace_one(const igo_scene& s, const Ray& ray, bool any_hit);
} // namespace igo
