/* oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's (SLebailly/Ignis-MasterThesis) CPU ray-traversal path in its
 * default configuration: scalar, vector width 1, BVH4 + Tri4 (src/runtime/device/Target.cpp:309-313,
 * src/runtime/shader/ShaderUtils.cpp:21-38).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may link or call anything in this directory.  The product
 * (ignis-masterthesis_b200/csrc) never does.
 *
 * Parity status: the reference itself cannot be compiled here (Artic/AnyDSL, TBB, madmann91/bvh absent),
 * so the restatement is pinned by (1) the reference's own known-answer tests
 * (src/tests/artic/test_intersection.art:1-134) and (2) an O(N*T) brute-force intersector that shares
 * only the Moeller-Trumbore arithmetic.  Tree SHAPE (madmann91/bvh v1 SBVH/PLOC) is NOT reproduced:
 * "tree shape: parity unpinned" -- closest-hit results do not depend on it except on exact-t ties.
 *
 * Arithmetic contract (the reference is compiled -ffast-math, so it defines no bit-exact result; this
 * file does): every `vec3_dot` is the explicit fma chain of src/artic/core/vector.art:99, `vec3_cross`
 * is un-fused (vector.art:106-109), everything else is plain IEEE f32 with contraction OFF
 * (compile with -ffp-contract=off).  The CUDA kernels follow the same contract.
 */
#ifndef IGNIS_ORACLE_H
#define IGNIS_ORACLE_H

#include "../include/ignis_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct igo_scene igo_scene;

typedef struct igo_counters {
    uint64_t scene_nodes;   /* Node4 of the scene BVH popped */
    uint64_t prim_nodes;    /* Node4 of prim BVHs popped */
    uint64_t tri_packets;   /* Tri4 packets fetched */
    uint64_t tri_tests;     /* valid triangle lanes tested */
    uint64_t entities;      /* entity leaves whose local BVH was entered */
} igo_counters;

/* Build BVH4/Tri4 for every trimesh shape and a BVH4 over the entity leaves of each primitive type.
 * Takes exactly the tables the product ABI takes. Returns NULL on malformed input. */
igo_scene* igo_scene_create(const igb200_scene* scene);
void       igo_scene_destroy(igo_scene* s);

/* Restated cpu_traverse_primary / cpu_traverse_secondary
 * (src/artic/driver/mapping_cpu.art:366-424 -> src/artic/traversal/mapping_cpu.art:261-512).
 * Rays are StreamRay records whose directions are ALREADY normalised (igo_normalize_rays).
 * threads <= 0 -> hardware_concurrency. counters may be NULL. */
void igo_trace(const igo_scene* s, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
               int threads, int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v,
               igo_counters* counters);

/* BVH-free check: every triangle/sphere of every visible entity, same MT arithmetic, strict-minimum t.
 * tie[i] = 1 when two different (entity, prim) pairs yield the bit-identical minimum t (the winner
 * then depends on tree order in the reference: src/artic/traversal/mapping_cpu.art:469 `<=`),
 * second_t[i] = smallest t of any OTHER primitive (FLT_MAX if none): degenerate-ray classification. */
void igo_brute(const igo_scene* s, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int threads,
               int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v, uint8_t* tie,
               float* second_t);

/* Interface::loadRayList normalisation (src/runtime/device/Device.cpp:585-607), in place. */
void igo_normalize_rays(igb200_ray* rays, uint64_t n);

/* Known-answer hooks for src/tests/artic/test_intersection.art. Return 1 on hit.
 * tri = v0,v1,v2 (9 floats); out = t,u,v. */
int igo_kat_ray_triangle(const float* tri, const float* org, const float* dir, float tmin, float tmax,
                         float* out_tuv);
/* box = min xyz, max xyz; out = entry (or exit when entry < 1e-5), as intersect_ray_box_single. */
int igo_kat_ray_box(const float* box, const float* org, const float* dir, float tmin, float tmax,
                    float* out_t);

/* Restated render iteration (Device::render, src/runtime/device/Device.cpp:1588-1612 -> cpu_trace,
 * src/artic/driver/mapping_cpu.art:694-836) for the path / debug techniques and the shading subset of
 * include/ignis_b200.h: adds one iteration (spi samples per pixel) into `film` (w*h*3 f32, RGB sums).
 * stats3 = camera, bounce, shadow ray counts (src/runtime/Statistics.h:56-63).  Returns 0 on success.
 * Pinned by the reference's closed-form integrator tests (src/tests/integrator/test_lights.py:5-44) and the
 * evaluation images (scenes/evaluation/references/ EXRs, scripts/RunEvaluations.py:79-113). */
int igo_render(const igo_scene* scene, const igb200_shading* shading, const igb200_render_settings* settings,
               float* film, uint64_t* stats3, int threads);

/* igo_render plus the technique's AOV images (each w*h*3 f32 sums, or NULL): path with IGB200_TECHFLAG_AOV_MIS adds into
 * aov0 = "Direct Weights" and aov1 = "NEE Weights" (pathtracer.art:128,158,206); IGB200_TECH_INFOBUFFER adds into
 * aov0..2 = "Normals", "Albedo", "Depth" (technique/infobuffer.art:28-104) and does not touch the film. */
int igo_render_aov(const igo_scene* scene, const igb200_shading* shading, const igb200_render_settings* settings,
                   float* film, float* aov0, float* aov1, float* aov2, uint64_t* stats3, int threads);

/* Restated tonemap pipeline (src/artic/entrypoints/tonemap.art, core/color.art:77-142): n pixels of RGB sums ->
 * packed 8-bit ARGB; `scale` carries 1 / IterationCount as Device::tonemap computes it (Device.cpp:1668,1677). */
void igo_tonemap(const float* in, uint64_t n, uint32_t method, int use_gamma, float scale, float exposure_factor,
                 float exposure_offset, uint32_t* out);

/* Integer RNG known answers (src/artic/core/random.art): seed = create_random_seed(...), then n draws of
 * sample_tea_u32(seed, counter++) from counter 1, raw and as next_f32. */
void igo_kat_rng(int32_t sample, int32_t iter, int32_t frame, int32_t x, int32_t y, int32_t user, uint32_t n,
                 uint32_t* out_seed, uint32_t* out_raw, float* out_f32);

/* Texture lookups on their own: n (u, v) pairs -> n rgb triples; -1 if the id is out of range. */
int igo_texture_eval(const igb200_shading* shading, uint32_t id, const float* uv, uint64_t n, float* rgb);

/* Structure statistics for reports. */
void igo_scene_info(const igo_scene* s, uint64_t* node4_count, uint64_t* tri4_count, uint64_t* tri_count);

#ifdef __cplusplus
}
#endif
#endif
