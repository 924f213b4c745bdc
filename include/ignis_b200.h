/* ignis_b200.h -- C ABI of the B200-native ray-traversal / wavefront backend for Ignis.
 *
 * This is the drop-in boundary.  Every entry point replaces one member of the
 * reference's `IG::Device` class (src/runtime/device/Device.h:14-74), which is the only seam the
 * reference's `IG::Runtime` talks to.  All arguments are plain pointers and sizes; inputs are
 * borrowed for the duration of the call only (the reference borrows them for the Device lifetime,
 * Device.h:25-30 -- we copy to HBM in `igb200_assign_scene`, so the caller may free afterwards).
 *
 * Error behaviour: the reference aborts on OOM and logs on bad input (Device.cpp:271-274,1303-1305).
 * Here every call returns 0 on success or a negative IGB200_E_* code; `igb200_last_error()` gives
 * the message.  There is NO CPU fallback: without a CUDA device `igb200_create` fails.
 */
#ifndef IGNIS_B200_H
#define IGNIS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGB200_ABI_VERSION 7

enum {
    IGB200_OK            = 0,
    IGB200_E_NODEVICE    = -1, /* no CUDA device / wrong architecture */
    IGB200_E_INVALID     = -2, /* malformed argument or table */
    IGB200_E_OOM         = -3, /* cudaMalloc failed (reference: std::abort, Device.cpp:271-274) */
    IGB200_E_CUDA        = -4, /* a CUDA call or kernel failed */
    IGB200_E_STATE       = -5, /* call order violated (e.g. trace before assign_scene) */
    IGB200_E_UNSUPPORTED = -6
};

/* Ray visibility bits; identical to src/artic/traversal/ray.art:21-25 and the entity flags
 * written by src/runtime/loader/LoaderEntity.cpp:122-131. */
#define IGB200_RAY_CAMERA 0x1u
#define IGB200_RAY_LIGHT  0x2u
#define IGB200_RAY_BOUNCE 0x4u
#define IGB200_RAY_SHADOW 0x8u

/* One igtrace input ray; byte-identical to `StreamRay` (src/artic/traversal/ray.art:2-7) and to
 * `IG::Ray` (src/runtime/RuntimeStructs.h:71-75).  32 bytes. */
typedef struct igb200_ray {
    float org[3];
    float dir[3];
    float tmin;
    float tmax;
} igb200_ray;

/* `LookupEntry` of a DynTable (src/runtime/table/DynTable.h:6-10). 16 bytes. */
typedef struct igb200_lookup_entry {
    uint32_t type_id;
    uint32_t flags;
    uint64_t offset; /* byte offset into the table data */
} igb200_lookup_entry;

/* Top-level BVH leaf; byte-identical to `EntityLeaf1` (src/artic/traversal/bvh.art:52-61, written by
 * src/runtime/bvh/SceneBVHAdapter.h:73-105).  96 bytes.  The sign bit of entity_id is the reference's
 * "last in leaf" sentinel and is ignored on input. */
typedef struct igb200_entity_leaf {
    float    bbox_min[3];
    int32_t  entity_id;
    float    bbox_max[3];
    int32_t  shape_id;
    float    local[12]; /* world->object 3x4, column major: col0 xyz, col1 xyz, col2 xyz, col3 xyz */
    uint32_t flags;     /* visibility bits IGB200_RAY_* */
    int32_t  mat_id;
    int32_t  user1;     /* reference: low/high words of the prim-BVH offset; unused by this backend */
    int32_t  user2;
} igb200_entity_leaf;

/* Primitive type of a shape table entry (`type_id` in the `shapes` DynTable lookups): the index of the
 * shape provider, src/runtime/loader/LoaderShape.cpp:20-60 ("trimesh" = 0, "sphere" = 1). */
#define IGB200_PRIM_TRIMESH 0u
#define IGB200_PRIM_SPHERE  1u

/* Scene description = the slices of `IG::SceneDatabase` (src/runtime/table/SceneDatabase.h:8-20) that the
 * hot path consumes.  Replaces `Device::SceneSettings` (Device.h:25-30).
 *  - shapes: DynTable "shapes" (TriMeshProvider.cpp:574-602): per trimesh
 *      [nf,nv,nn,nt : u32x4][bbox min xyz,0, max xyz,0][verts vec4-aligned][normals vec4-aligned]
 *      [indices 4 x u32 per face][texcoords 2 x f32]; per sphere (SphereProvider.cpp:33-45)
 *      [origin xyz, radius].
 *  - entities: FixTable "entities" (LoaderEntity.cpp:150-176): 36 floats per entity.
 *  - leaves[k]: `SceneBVHs[provider k].Leaves` -- the EntityLeaf1 array of primitive type k.
 *    The reference's own Node2/4/8 arrays are NOT consumed: the backend builds its compressed BVH8. */
typedef struct igb200_scene {
    const igb200_lookup_entry* shape_lookups;
    uint64_t                   shape_count;
    const uint8_t*             shape_data;
    uint64_t                   shape_bytes;
    const float*               entities;     /* entity_count * 36 floats, may be NULL for trace-only use */
    uint64_t                   entity_count;
    const igb200_entity_leaf*  leaves[2];    /* [IGB200_PRIM_TRIMESH], [IGB200_PRIM_SPHERE] */
    uint64_t                   leaf_count[2];
    float                      scene_bbox[6]; /* min xyz, max xyz */
    uint32_t                   material_count;
    uint32_t                   _pad;
} igb200_scene;

/* Replaces `Device::SetupSettings` (Device.h:18-23). */
typedef struct igb200_setup {
    int32_t device;        /* CUDA ordinal (Target::device) */
    int32_t acquire_stats; /* count nodes/triangles visited per trace call (slower kernels) */
    int32_t debug_trace;
    int32_t _pad;
} igb200_setup;

/* Per-call statistics (the algorithmic-byte terms of SURVEY.md 8(d)). */
typedef struct igb200_trace_stats {
    uint64_t rays;
    uint64_t nodes_visited;    /* BVH8 inner nodes fetched (80 B each) */
    uint64_t tris_tested;      /* triangle records fetched (48 B each) */
    uint64_t entities_entered; /* entity leaves entered (96 B each) */
    float    kernel_ms;        /* device time of the traversal kernel(s), CUDA events */
    float    _pad;
} igb200_trace_stats;

typedef struct igb200_device igb200_device;

/* ---- lifetime (Device::Device / ~Device, Device.cpp:1562-1573) ---- */
int         igb200_abi_version(void);
const char* igb200_last_error(void);
int         igb200_create(const igb200_setup* setup, igb200_device** out);
void        igb200_destroy(igb200_device* dev);

/* ---- scene (Device::assignScene, Device.h:52; BVH upload Device.cpp:663-706) ---- */
int igb200_assign_scene(igb200_device* dev, const igb200_scene* scene);

/* ---- igtrace ray list -> hit records.
 * Host-buffer form: the reference path is Runtime::trace (Runtime.cpp:375-407) ->
 * Interface::loadRayList (Device.cpp:573-611; directions are normalised here the same way) ->
 * K2 primary traversal (src/artic/driver/mapping_gpu.art:45-68).  Outputs are the five hit columns of
 * the primary stream (src/artic/driver/streams.art:80-100): ent_id, prim_id (-1 on miss), t (= tmax on
 * miss), u, v.  `any_hit` != 0 runs the shadow-ray variant (mapping_gpu.art:70-112): ent_id/prim_id are
 * then only meaningful as hit (>=0) / miss (-1).  Any output pointer may be NULL.
 * `ray_flags` is the ray-type visibility mask (0 = igtrace behaviour: visible to everything). */
int igb200_trace_rays(igb200_device* dev, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                      int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v);

/* Device-resident form of the same call: `d_rays` and the outputs are device pointers (e.g. torch
 * tensors' data_ptr) on this device; enqueued on `stream` (a cudaStream_t, 0 = default), not synchronised.
 * Directions must already be normalised.  Calls on DIFFERENT streams of one handle are ordered on the device (each
 * waits for the previous launch through an event): the traversal kernel's work counters belong to the handle, so two
 * launches never overlap; use one handle per stream for concurrent traversal.  Calls from different host threads on
 * one handle must be serialised by the caller, as for every igb200_* function. */
int igb200_trace_rays_device(igb200_device* dev, const igb200_ray* d_rays, uint64_t n, uint32_t ray_flags,
                             int any_hit, int32_t* d_ent_id, int32_t* d_prim_id, float* d_t, float* d_u,
                             float* d_v, void* stream);

/* Statistics of the most recent trace call (Device::getStatistics, Device.h:68). */
int igb200_get_trace_stats(igb200_device* dev, igb200_trace_stats* out);

/* Sizes of the device-side acceleration structure (for the roofline's working-set statement). */
int igb200_get_bvh_info(igb200_device* dev, uint64_t* node_bytes, uint64_t* tri_bytes, uint64_t* leaf_bytes,
                        uint32_t* max_depth);

/* =====================================================================================================
 * Wavefront path tracing.
 *
 * In the reference, cameras / BSDFs / lights / the technique reach the device as per-scene Artic SOURCE
 * that is JIT-compiled (src/runtime/shader/, , ShaderUtils.cpp:95-122, loader/LoaderLight.cpp:106-240,
 * technique/PathTechnique.cpp:36-79).  This backend has no JIT: the same parameters travel as the POD
 * tables below ("the one place the seam must be widened").  Every field names the Artic
 * constructor argument it feeds.
 * ===================================================================================================== */

/* make_perspective_camera(eye, dir, up, scale, w, h, tmin, tmax) (src/artic/camera/perspective.art:29-41);
 * scale = compute_scale_from_hfov / _vfov (perspective.art:1-11), chosen by the loader
 * (src/runtime/camera/PerspectiveCamera.cpp:40-54). */
#define IGB200_CAMERA_PERSPECTIVE 0u
#define IGB200_CAMERA_FISHLENS    1u /* make_fishlens_camera(eye, dir, up, w, h, mode, tmin, tmax) (src/artic/camera/fishlens.art:8-70):
                                        180-degree equidistant fisheye; scale = (xasp, yasp) of the aspect mode the loader picked
                                        (fishlens.art:14-36; runtime/camera/FishLensCamera.cpp uses Circular) */
typedef struct igb200_camera {
    float    eye[3];
    float    dir[3];
    float    up[3];
    float    scale[2];
    float    tmin; /* near clip */
    float    tmax; /* far clip */
    uint32_t type; /* IGB200_CAMERA_* (ABI v5) */
} igb200_camera;

#define IGB200_BSDF_DIFFUSE    0u /* make_lambertian_bsdf(surf, kd = color)           bsdf/diffuse.art:2-11 */
#define IGB200_BSDF_DIELECTRIC 1u /* make_pure_dielectric_bsdf(surf, n1 = p[0] (exterior), n2 = p[1] (interior),
                                     ks = color, kt = color2)                          bsdf/dielectric.art:2-23 */
#define IGB200_BSDF_MIRROR     2u /* make_mirror_bsdf(surf, ks = color)                bsdf/conductor.art:1-9 */
#define IGB200_BSDF_CONDUCTOR  3u /* make_conductor_bsdf(surf, eta = color2, k = p[0..2], ks = color, micro)
                                     bsdf/conductor.art:112-122.  micro = make_vndf_ggx_distribution(face_normal,
                                     local, alpha_u = q[0].f, alpha_v = q[1].f) (core/microfacet.art:383-407; the
                                     loader's default distribution, runtime/bsdf/BSDF.cpp:75); alpha <= 1e-4 is the
                                     delta distribution = make_pure_conductor_bsdf (conductor.art:12-30) */
#define IGB200_BSDF_BUMPMAP    4u /* make_bumpmap(ctx, inner = bsdfs[q[1].u - 1], texture_dx(map).r, texture_dy(map).r,
                                     strength = p[0]) with map = textures[q[0].u - 1]
                                     bsdf/map.art:64-68, texture/common.art:28-38, runtime/bsdf/MapBSDF.cpp */
#define IGB200_BSDF_THIN_DIELECTRIC 5u /* make_thin_dielectric_bsdf(surf, n1 = p[0], n2 = p[1], ks = color, kt = color2): the
                                     dielectric with "thin": true -- straight-through transmission, Fresnel term with the
                                     interreflection series              bsdf/dielectric.art:27-47,178-184 */
typedef union igb200_word {
    float    f;
    uint32_t u;
} igb200_word;
typedef struct igb200_bsdf {
    uint32_t    type;
    float       color[3];  /* DIFFUSE: reflectance unless q[0].u != 0, then textures[q[0].u - 1] replaces it */
    float       color2[3];
    float       p[3];
    igb200_word q[2];
} igb200_bsdf; /* 48 bytes */

/* Textures (src/artic/texture/, loader side src/runtime/pattern/).  uv is the surface texture coordinate. */
#define IGB200_TEX_CHECKERBOARD 0u /* make_checkerboard_texture(scale, color0, color1, transform)  texture/checkerboard.art:4-13 */
#define IGB200_TEX_IMAGE        1u /* make_image_texture(border, filter, image, transform)         texture/image.art:148-152 */
#define IGB200_FILTER_NEAREST  0u  /* texture/image.art:81-89 */
#define IGB200_FILTER_BILINEAR 1u  /* :91-106 */
#define IGB200_FILTER_BICUBIC  2u  /* :108-146; the loader's default, also for unknown names (pattern/ImagePattern.cpp:19-31) */
#define IGB200_WRAP_REPEAT 0u      /* texture/image.art:17-26 */
#define IGB200_WRAP_MIRROR 1u      /* :28-40 */
#define IGB200_WRAP_CLAMP  2u      /* :9-15 */
#define IGB200_IMAGE_RGBA8   0u    /* packed u32 per texel, image_rgba_unpack: byte / 255 (driver/image.art:8-14); the sRGB ->
                                      linear byte mapping was applied by the loader (runtime/Image.cpp:39-42,565-655) */
#define IGB200_IMAGE_RGBA32F 1u    /* four floats per texel (device.load_image_by_id, EXR / HDR / baked textures) */
typedef struct igb200_texture {
    uint32_t type, filter, wrap_u, wrap_v;
    float    transform[9]; /* 2-D affine transform of uv, column major (LoaderUtils::inlineTransformAs2d) */
    float    color0[3], color1[3];
    float    scale[2];
    uint32_t width, height, format;
    uint32_t opaque;       /* RGBA8: alpha reads as 1 (three-channel source, `hint_opaque`) */
    uint64_t offset;       /* byte offset of texel (0, 0) in igb200_shading::blob, 16-byte aligned; row-major, rows in the
                              order the reference holds them after loading (bottom row first, Image.cpp:559-560,575) */
    uint32_t _pad[2];
} igb200_texture; /* 120 bytes */

/* One row per material id (the ids written into the entity table, LoaderEntity.cpp:87-103,171):
 * make_material / make_emissive_material (src/artic/driver/material.art:10-28, ShaderUtils.cpp:95-122). */
typedef struct igb200_material {
    int32_t bsdf;  /* index into igb200_shading::bsdfs */
    int32_t light; /* index into the FINITE light table of the area light this material emits, or -1 */
} igb200_material;

#define IGB200_LIGHT_POINT      0u /* make_point_light(pos = v[0..2], intensity = color)   light/point.art:1-18 */
#define IGB200_LIGHT_ENV        1u /* make_environment_light(scene_bbox, scale*radiance = color, transform = v[0..8]
                                      column major): constant radiance, spherical sampling   light/env.art:75-101,163-166 */
#define IGB200_LIGHT_AREA_SHAPE 2u /* make_area_light(make_shape_area_emitter(entity, trimesh), radiance = color)
                                      light/area.art:10-88 */
#define IGB200_LIGHT_AREA_PLANE 3u /* make_area_light(make_plane_area_emitter(origin v[0..2], x_axis v[3..5],
                                      y_axis v[6..8], normal v[9..11], area v[12]), radiance = color)
                                      light/area.art:107-232, runtime/light/AreaLight.cpp:143-160 */
#define IGB200_LIGHT_SPOT       4u /* make_spot_light(pos v[0..2], dir v[3..5], cutoff v[6], falloff v[7] (radians),
                                      intensity = color)                                   light/spot.art */
#define IGB200_LIGHT_DIRECTIONAL 5u /* make_directional_light(dir v[0..2], scene_bbox, irradiance = color)
                                      light/directional.art */
#define IGB200_LIGHT_ENV_TEXTURED 6u /* make_environment_light_textured(scene_bbox, scale = color, tex = textures[tex],
                                      cdf, transform = v[0..8] column major) when cdf != IGB200_NO_CDF, else
                                      make_environment_light(scene_bbox, scale, tex, transform)   light/env.art:112-166 */
#define IGB200_LIGHT_AREA_SPHERE 7u /* make_area_light(make_sphere_area_emitter(entity, Sphere{origin v[0..2], radius v[3]}),
                                      radiance = color): analytic sphere entity as emitter, sampled on the side facing the
                                      shading point          light/area.art:241-296, runtime/light/AreaLight.cpp:73-82,153-160 */
#define IGB200_LIGHT_ENV_FUNCTION 8u /* make_environment_light_function(scene_bbox, func, transform = v[0..8], half = !has_ground)
                                      (light/env.art:24-107): an analytic sky.  tex = IGB200_SKY_* selects func, cdf = offset in
                                      blob (units of 4 bytes) of its 16 float parameters, flags = IGB200_SKYFLAG_*.
                                      Parameters: [0..2] scale, [3..5] zenith, [6..8] ground, [9] zenith brightness,
                                      [10] ground brightness, [11] c2, [12..14] sun direction (Y up), [15] 0 -- the arguments
                                      runtime/light/CIELight.cpp:41-110 passes to make_cie_sky_light / make_cie_sunny_light
                                      (light/cie.art:9-42) */
#define IGB200_LIGHT_SUN 9u /* make_sun_light(dir = v[0..2], scene_bbox, cos_sun_angle = v[3], color) (light/sun.art:4-28): a delta-treated
                              infinite light sampled uniformly over the sun's cone; the loader gives cos(angle / 2) or
                              sun_cos_angle_from_radius(radius) (runtime/light/SunLight.cpp:24-46) */
#define IGB200_SKY_CIE       0u /* cieuniform / ciecloudy (flag VARIANT = cloudy)          light/cie.art:10-19 */
#define IGB200_SKY_CIE_SUNNY 1u /* cieclear (flag VARIANT = is_clear) / cieintermediate    light/cie.art:22-42 */
#define IGB200_SKY_PEREZ     2u /* make_perez_light(sun_dir = [12..14], lum = [3..5], ground = [6..8], a b c d e = [9], [10], [11],
                                   [15], [0])                                        light/cie.art:45-57, runtime/light/PerezLight.cpp:58-110 */
#define IGB200_SKYFLAG_VARIANT    0x1u
#define IGB200_SKYFLAG_HAS_GROUND 0x2u
#define IGB200_NO_CDF 0xFFFFFFFFu
/* make_environment_light_textured::sample_direct returns the UNSCALED texture radiance (light/env.art:138-139 uses
 * `intensity` where emission, :150, uses scale * tex), so next event estimation under-weights a scaled environment map;
 * the reference's own Mitsuba golden for this light (scenes/evaluation/env.json, scale 100, ref-env-4096.exr) only holds
 * with the scale applied.  This light flag applies it; 0 restates the reference's code as written. */
#define IGB200_LIGHTFLAG_SCALED_NEE 0x1u
typedef struct igb200_light {
    uint32_t type;
    int32_t  entity; /* area lights: the emissive entity id */
    float    color[3];
    float    v[16];
    uint32_t tex;    /* ENV_TEXTURED: texture id */
    uint32_t cdf;    /* ENV_TEXTURED: offset of the 2-D CDF in igb200_shading::blob in units of 4 bytes: u32 size_x,
                        size_y, 0, 0, then size_y marginal floats and size_y rows of size_x conditional floats, each
                        without the leading 0 (cdf::make_cdf_2d_from_buffer, core/cdf.art:70-72,156-160) */
    uint32_t flags;  /* IGB200_LIGHTFLAG_*; 0 = the reference's behaviour */
} igb200_light; /* 96 bytes */

#define IGB200_TECH_PATH  0u /* make_path_renderer (src/artic/technique/pathtracer.art:40-218) */
#define IGB200_TECH_DEBUG 1u /* make_debug_renderer (src/artic/technique/debugtracer.art), mode = debug_mode */
#define IGB200_TECH_INFOBUFFER 2u /* make_infobuffer_renderer (src/artic/technique/infobuffer.art:28-104): the denoiser's
                                     AOVs "Normals", "Albedo", "Depth"; the film is locked (not written, iteration count
                                     unchanged: InfoBufferTechnique.cpp:36, Device.cpp:1463-1466); max_depth =
                                     __tech_ib_max_depth.  The reference runs it with spi 1 (OverrideSPI). */
/* The reference's lambertian eval is kd*|cos|/pi for ANY incoming direction (bsdf/diffuse.art:3), so next event
 * estimation lets light through one-sided geometry from behind (e.g. a full-sphere environment sample below a
 * ground plane, light/env.art:82-86).  The reference's own golden values that involve this case
 * (tests/integrator/test_lights.py:37-44 expects 1.0; scenes/evaluation plane-d6 vs Mitsuba) only hold with a
 * one-sided eval kd*max(cos,0)/pi; this flag selects it.  Default (0) restates the reference's code as written. */
#define IGB200_TECHFLAG_ONE_SIDED_DIFFUSE 0x1u
/* PathTechnique `aov_mis` (runtime/technique/PathTechnique.cpp:16,23-27,54-57): the AOVs "Direct Weights" (every
 * contribution a BSDF-sampled path collects: emission on hit, environment on miss; pathtracer.art:128,158) and
 * "NEE Weights" (unoccluded next-event contributions, pathtracer.art:202-208) are accumulated next to the film. */
#define IGB200_TECHFLAG_AOV_MIS 0x2u
/* IGB200_TECH_INFOBUFFER: `handle_specular` -- follow specular bounces and record the first non-specular hit
 * (runtime/technique/InfoBufferTechnique.cpp:11,77). */
#define IGB200_TECHFLAG_IB_HANDLE_SPECULAR 0x4u
typedef struct igb200_technique {
    uint32_t type;
    int32_t  max_depth;  /* __tech_max_depth, default 64 (runtime/technique/Technique.h:8) */
    int32_t  min_depth;  /* __tech_min_depth, default 2 */
    float    clamp;      /* __tech_clamp, <= 0 disables */
    int32_t  enable_nee; /* "nee", default true */
    int32_t  debug_mode; /* __debug_mode for IGB200_TECH_DEBUG */
    uint32_t flags;      /* IGB200_TECHFLAG_*; 0 = the reference's behaviour */
    int32_t  _pad;
} igb200_technique;

/* The light tables are split exactly like the reference's `infinite_lights` / `finite_lights`
 * (loader/LoaderLight.cpp:118-250); the light selector is make_uniform_light_selector
 * (src/artic/light/light_selector.art:26-45) unless `selector` says otherwise. */
typedef struct igb200_shading {
    igb200_camera          camera;
    igb200_technique       technique;
    const igb200_bsdf*     bsdfs;
    uint64_t               bsdf_count;
    const igb200_material* materials; /* material_count of the scene */
    uint64_t               material_count;
    const igb200_light*    infinite_lights;
    uint64_t               infinite_light_count;
    const igb200_light*    finite_lights;
    uint64_t               finite_light_count;
    /* ---- ABI v3 ---- */
    const igb200_texture*  textures;
    uint64_t               texture_count;
    const uint8_t*         blob;       /* texels, CDFs and the light hierarchy */
    uint64_t               blob_bytes;
    uint32_t               selector;   /* IGB200_SELECTOR_* */
    uint32_t               hierarchy;  /* IGB200_SELECTOR_HIERARCHY: offset of the buffer LightHierarchy::setup writes
                                          (runtime/light/LightHierarchy.cpp:84-127) in blob, in units of 4 bytes:
                                          round_up(finite_light_count, 4) u32 codes, then 8 floats per node */
} igb200_shading;
#define IGB200_SELECTOR_UNIFORM   0u /* make_uniform_light_selector    light/light_selector.art:26-45 */
#define IGB200_SELECTOR_HIERARCHY 1u /* make_hierarchy_light_selector  light/light_selector.art:76-110, light_hierarchy.art */
#define IGB200_SELECTOR_CDF       2u /* make_cdf_light_selector ("simple")  light/light_selector.art:47-74: `hierarchy` is then the
                                        offset (units of 4 bytes) in blob of finite_light_count floats, the running sum of the lights'
                                        estimated flux normalised to end at 1, without the leading 0 (CDF::computeForArray,
                                        runtime/CDF.cpp:11-40; loader/LoaderLight.cpp:454-478) */

/* Replaces Device::RenderSettings (Device.h:32-42).  One call = one ITERATION = spi samples per pixel
 * (Device::render, Device.cpp:1588-1612).  rays != NULL selects igtrace mode (list emitter,
 * src/artic/driver/emitter.art:18-31): width = ray_count, height = 1, ray i lands in pixel i. */
typedef struct igb200_render_settings {
    const igb200_ray* rays;
    uint64_t          ray_count;
    uint32_t          spi;
    uint32_t          width;
    uint32_t          height;
    uint32_t          iteration;
    uint32_t          frame;
    uint32_t          user_seed;
} igb200_render_settings;

/* Ray counters of src/runtime/Statistics.h:56-63 (the numerators of Msamples/s and Mrays/s). */
typedef struct igb200_render_stats {
    uint64_t camera_rays;
    uint64_t bounce_rays;
    uint64_t shadow_rays;
    uint64_t rounds;        /* wavefront rounds (generate/trace/sort/shade/shadow) of the last render call */
    uint64_t kernel_launches;
    float    device_ms;     /* device time of the last render call */
    float    _pad;
} igb200_render_stats;

/* Shading tables (no reference counterpart: replaces shader compilation, Runtime.cpp:542-626). Must follow
 * igb200_assign_scene; copies everything. */
int igb200_assign_shading(igb200_device* dev, const igb200_shading* shading);

/* Device::render (Device.h:53): synchronous; accumulates into the film and bumps the iteration count. */
int igb200_render(igb200_device* dev, const igb200_render_settings* settings);

/* `count` consecutive iterations (settings->iteration ... + count - 1) in ONE wavefront: the ray queue regenerates
 * across iteration boundaries, so the long low-occupancy tail of deep paths (dielectrics, max_depth 64) is paid once
 * per call instead of once per iteration.  The samples are exactly those of `count` igb200_render calls (the seed
 * hashes the iteration index, src/artic/core/random.art:34-43); only the order of the f32 film additions differs.
 * The frontends' loop `for i < spp/spi: runtime.step()` (src/frontend/cli/main.cpp:120-139) maps onto one call. */
int igb200_render_iterations(igb200_device* dev, const igb200_render_settings* settings, uint32_t count);

/* ABI v7.  One iteration like igb200_render, but the call returns as soon as the iteration's rays have been GENERATED: the paths
 * still in flight stay in the wavefront and continue in the rounds of the next igb200_render_pipelined call, so a frontend that
 * steps iteration by iteration (Runtime::step -> Device::render, src/runtime/Runtime.cpp:294-330; src/frontend/cli/main.cpp:120-139)
 * runs at the throughput of one long wavefront instead of paying the low-occupancy tail of deep paths on every step
 * (C4 diamond: 523 -> 1170 Msamples/s per step).  The film is complete again after igb200_flush; every entry that
 * reads, clears or re-allocates the film or changes what the paths depend on (igb200_framebuffer, igb200_tonemap,
 * igb200_clear_*, igb200_resize, igb200_set_framebuffer, igb200_assign_scene / _shading, a render call with another film size,
 * spi or partition, igb200_render / _iterations / _partition) flushes by itself, so the results are those of igb200_render.
 * The ray counters of a call cover the rounds that ran inside it; a flush adds what it traced to the last call's.
 * Ray-list mode and devices created with acquire_stats render synchronously as igb200_render does. */
int igb200_render_pipelined(igb200_device* dev, const igb200_render_settings* settings);
int igb200_flush(igb200_device* dev);

/* Device::getFramebufferForHost / ForDevice (Device.h:63-64, Device.cpp:1330-1443): RGB f32, w*h*3, SUM over
 * iterations; *iterations = IterationCount of that image.  name NULL, "" or "Color" = the film; "Direct Weights",
 * "NEE Weights" (path with IGB200_TECHFLAG_AOV_MIS) and "Normals", "Albedo", "Depth" (IGB200_TECH_INFOBUFFER) exist
 * once a technique that enables them has rendered; anything else is an error (Device.cpp:1303-1305).  The pointer is owned by
 * the device and valid until resize / destroy.  device != 0 returns the device pointer (no copy). */
int igb200_framebuffer(igb200_device* dev, const char* name, int device, float** data, uint64_t* iterations);

/* Device::clearAllFramebuffer / resize (Device.h:65-66, Device.cpp:306-313,1407-1443): the film and every AOV. */
int igb200_clear_framebuffer(igb200_device* dev);
int igb200_resize(igb200_device* dev, uint32_t width, uint32_t height);
/* Device::clearFramebuffer(name) (Device.h:65, Device.cpp:1413-1441): one image and its iteration count. */
int igb200_clear_aov(igb200_device* dev, const char* name);

/* Device::tonemap (Device.h:70, Device.cpp:1662-1687, src/artic/entrypoints/tonemap.art): the film or an AOV, divided
 * by its iteration count and multiplied by `scale`, to packed 8-bit ARGB (a << 24 | r << 16 | g << 8 | b), w*h
 * words written to the HOST buffer `out_pixels`.  NaN pixels come out cyan, infinite ones pink, negative ones
 * yellow, as in the reference.  This is the headless target for the viewer's display path. */
typedef struct igb200_tonemap_settings {
    const char* aov;             /* NULL, "" or "Color" = the film (RuntimeStructs.h:6-13) */
    uint32_t    method;          /* 0 none, 1 reinhard, 2 modified reinhard, 3 ACES, 4 uncharted2 */
    int32_t     use_gamma;       /* sRGB transfer curve on the output */
    float       scale;
    float       exposure_factor;
    float       exposure_offset;
} igb200_tonemap_settings;
int igb200_tonemap(igb200_device* dev, const igb200_tonemap_settings* settings, uint32_t* out_pixels);

/* Multi-GPU film combine: overwrites the film with `count` = w*h*3 floats from `d_src` (a device
 * pointer on this device, e.g. the result of an NCCL reduce over the per-GPU films) and sets the iteration count. */
int igb200_set_framebuffer(igb200_device* dev, const float* d_src, uint64_t count, uint64_t iterations);

int igb200_get_render_stats(igb200_device* dev, igb200_render_stats* out);

/* Per-kernel-kind device times, only gathered when the device was created with acquire_stats (Device::SetupSettings::
 * AcquireStats, `--stats`): the wavefront rounds then run un-graphed with a CUDA event between the kernels.  This is what
 * the reference's Statistics (src/runtime/Statistics.h:9-63, begin/endShaderLaunch + the GPUSort / GPUCompact sections
 * of src/artic/core/stats.art) reports; host/Statistics.h maps the slots onto its ShaderType / SectionType names:
 *   0 begin_round  -> ShaderType::Device            3 scatter (index sort) -> SectionType::GPUSortPrimary (+Sort)
 *   1 generate     -> ShaderType::RayGeneration     4 shade (+compaction)  -> ShaderType::Hit (Miss and GPUCompactPrimary are fused into it)
 *   2 primary traversal (+ sort histogram, scan) -> ShaderType::PrimaryTraversal (+ GPUSortPrimaryCount / Scan)
 *   5 shadow traversal (+ film splat)            -> ShaderType::SecondaryTraversal */
typedef struct igb200_section_stats {
    float    kernel_ms[6];
    uint64_t launches[6];
    float    framebuffer_host_update_ms; /* lazy D2H of the film, SectionType::FramebufferHostUpdate */
    uint32_t framebuffer_host_updates;
} igb200_section_stats;
int igb200_get_section_stats(igb200_device* dev, igb200_section_stats* out, int reset);

/* ---------------------------------------------------------------------------------------------------------------
 * Multi-GPU (ABI v6).  north_star: "work is partitioned by image tile and sample index, with the scene replicated per
 * GPU and accumulation buffers combined via NCCL reduce over NVLink".  The reference is single-device
 * (src/runtime/device/Device.cpp:232-243); its own partitioning analogue is the CPU driver's loop over 16 x 16 tiles
 * on TBB workers (src/artic/driver/mapping_cpu.art:694-706, src/artic/core/cpu_common.art:101-125).
 *
 * Partitioning of a request of `total_iterations` iterations (settings->iteration ... + total - 1):
 *   ITERATIONS  partition p renders a contiguous block of the iteration range.  The RNG seed hashes the iteration
 *               index (src/artic/core/random.art:34-43), so the union is EXACTLY the single-device sample set.
 *   TILE_ROWS   every partition renders all iterations of the 16-pixel tile rows t with t % parts == p (seeds hash
 *               the pixel, so again the same samples).  For spp < parts * spi, where there are not enough iterations.
 *   AUTO        ITERATIONS when total_iterations >= parts, else TILE_ROWS.
 * Either way the sum of the per-device films is the single-device film up to the order of the f32 additions.
 * ------------------------------------------------------------------------------------------------------------- */
enum { IGB200_PARTITION_AUTO = 0, IGB200_PARTITION_ITERATIONS = 1, IGB200_PARTITION_TILE_ROWS = 2 };

/* The share of partition `part` of `parts`: iterations [first, first + count) relative to settings->iteration and the
 * tile-row ownership (tile_rank of tile_count; tile_count == 1: whole film).  Pure host arithmetic, no device needed. */
int igb200_partition_share(uint32_t total_iterations, uint32_t part, uint32_t parts, uint32_t mode, uint32_t* first,
                           uint32_t* count, uint32_t* tile_rank, uint32_t* tile_count);
/* Renders this device's share in one wavefront (as igb200_render_iterations).  The device's film then holds a PARTIAL
 * sum; combine with igb200_comm_reduce_film / igb200_group_render.  Ray-list mode always splits by iteration. */
int igb200_render_partition(igb200_device* dev, const igb200_render_settings* settings, uint32_t total_iterations,
                            uint32_t part, uint32_t parts, uint32_t mode);
/* IterationCount of the film (name NULL / "" / "Color") or of an AOV, e.g. after a reduce (Device.cpp:1460-1472). */
int igb200_set_iteration_count(igb200_device* dev, const char* name, uint64_t iterations);
int igb200_device_ordinal(const igb200_device* dev);

/* ---- one process per GPU (torchrun / MPI launchers): an NCCL communicator next to a device handle ----
 * Rank 0 calls igb200_comm_unique_id and hands the 128 bytes to the other ranks by any means (the launcher's store, a
 * file, MPI_Bcast); then every rank calls igb200_comm_create.  NCCL is loaded at run time (libnccl.so.2). */
#define IGB200_COMM_ID_BYTES 128
typedef struct igb200_comm igb200_comm;
int  igb200_comm_unique_id(uint8_t id[IGB200_COMM_ID_BYTES]);
int  igb200_comm_create(igb200_device* dev, const uint8_t id[IGB200_COMM_ID_BYTES], uint32_t nranks, uint32_t rank, igb200_comm** out);
void igb200_comm_destroy(igb200_comm* comm);
/* This rank's share of the request (igb200_render_partition with part = rank, parts = nranks). */
int  igb200_comm_render(igb200_comm* comm, const igb200_render_settings* settings, uint32_t total_iterations, uint32_t mode);
/* ONE ncclReduce(sum, f32) per image (film + every AOV in use) onto `root` over NVLink / NVSwitch, timed with CUDA
 * events on the communicator's stream (*reduce_ms, may be NULL).  Afterwards root's images hold the sums and the
 * iteration counts of the whole request (`total_iterations` as passed to the matching igb200_comm_render, added to what
 * root held before); the other ranks' images are cleared so the next request starts from zero there.
 * warm != 0: additionally runs one untimed reduce of a scratch buffer of the film's size first (communicator
 * first-use cost: channel setup and buffer registration happen on the first large collective). */
int  igb200_comm_reduce_film(igb200_comm* comm, uint32_t root, uint32_t total_iterations, uint32_t mode, int warm, float* reduce_ms);

/* ---- one process, N devices (ncclCommInitAll): what igcli --gpus N and IG::Runtime use ----
 * Scene and shading are replicated; member 0 is the root that holds the combined film. */
typedef struct igb200_group igb200_group;
int            igb200_group_create(const int32_t* devices, uint32_t count, int acquire_stats, igb200_group** out);
void           igb200_group_destroy(igb200_group* group);
uint32_t       igb200_group_size(const igb200_group* group);
igb200_device* igb200_group_device(igb200_group* group, uint32_t index);
int            igb200_group_assign_scene(igb200_group* group, const igb200_scene* scene);
int            igb200_group_assign_shading(igb200_group* group, const igb200_shading* shading);
int            igb200_group_resize(igb200_group* group, uint32_t width, uint32_t height);
int            igb200_group_clear_framebuffer(igb200_group* group);
/* Partition, render on all members concurrently (one host thread and one stream per device), reduce onto member 0. */
int            igb200_group_render(igb200_group* group, const igb200_render_settings* settings, uint32_t total_iterations, uint32_t mode);
/* igtrace batches: contiguous ray ranges per member, results gathered by the D2H copies -- no collective. */
int            igb200_group_trace_rays(igb200_group* group, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                                       int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v);
/* Sum of the members' ray counters of the last igb200_group_render, max of their device times, and the reduce time. */
int            igb200_group_get_render_stats(igb200_group* group, igb200_render_stats* out, float* reduce_ms);

/* Host-only test hook (no GPU needed): builds the compressed BVH8 over `n` boxes (min xyz, max xyz) and
 * checks its invariants -- every primitive referenced exactly once, every dequantised child box contains
 * the exact bounds of its subtree, meta/imask encoding consistent.
 * out5 = {node_count, max_depth, violations, primitives_not_seen_exactly_once, leaf_count}. */
int igb200_debug_check_bvh8(const float* boxes, uint32_t n, uint32_t max_leaf, uint64_t* out5);
/* Host-only: builds the BVH8 and reports its quality and cost (row a25: the reference builds with madmann91/bvh's SBVH +
 * reinsertion, src/runtime/bvh/TriBVHAdapter.h:163-202, and collapses by area, NArityBvh.h:65-116).
 * out4 = {sum over wide nodes of A_n / A_root, sum over leaf children of A_l * N_l / A_root, node count, build ms}.
 * Builder switches (environment, read per build): IGB200_BVH_COLLAPSE = dp | greedy, IGB200_BVH_CNODE, IGB200_BVH_CPRIM,
 * IGB200_BVH_THREADS. */
int igb200_debug_bvh8_quality(const float* boxes, uint32_t n, uint32_t max_leaf, double* out4);

/* No counterpart in the reference.
 * igb200_trace_rays for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): same kernel, same host buffers in, same five host hit columns out (NULL = column not
 * wanted), same results -- but every byte of scratch (work counters, ray staging, hit columns) lives in ONE static device
 * array inside the library, so the call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limits and caveats:
 *   - n <= IGB200_NOALLOC_MAX_RAYS, otherwise IGB200_E_UNSUPPORTED (split the batch or use igb200_trace_rays);
 *   - every GPU that loads the library holds the array (256 B + 52 B per ray = 3.25 MiB) until the process exits;
 *   - the array is shared by EVERY call and every device handle on that GPU and keeps the last call's contents: the
 *     entry clears what must start at zero and is synchronous, but calls that overlap on the GPU (two host threads, two
 *     handles on one GPU, two processes do NOT share it) race on it -- serialise them;
 *   - no traversal statistics are gathered (igb200_get_trace_stats is unaffected by this call). */
#define IGB200_NOALLOC_MAX_RAYS 65536u
int igb200_trace_rays_noalloc(igb200_device* dev, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                              int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v);

#ifdef __cplusplus
}
#endif
#endif /* IGNIS_B200_H */
