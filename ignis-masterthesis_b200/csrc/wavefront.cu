// wavefront.cu -- kernels of the wavefront path tracer (see wavefront.cuh for the design and the reference rows).
#include "wavefront.cuh"

#include "shade_tech.cuh"

#include <algorithm>
#include <cstdio>

namespace igb {
namespace {

using namespace core;
using namespace shade;

constexpr uint32_t kRayCamera = 0x1u, kRayBounce = 0x4u, kRayShadow = 0x8u; // src/artic/traversal/ray.art:21-25

// ---------------------------------------------------------------------------------------------------------
// Traversal IO policies (trace_core.cuh)
// ---------------------------------------------------------------------------------------------------------

// The ray streams (hundreds of MB per round) are read and written once per kernel; the mesh, BVH and texture data are what
// the 126 MB L2 should keep.  IGB_WAVE_STREAM 1: evict-first loads / stores (ld/st.global.cs) for stream records in the
// generate and shade kernels (the traversal kernels' own switch is IGB_STREAM_IO).
#ifndef IGB_WAVE_STREAM
#define IGB_WAVE_STREAM 0
#endif
template <typename T>
__device__ __forceinline__ T ws_load(const T* p)
{
#if IGB_WAVE_STREAM
    return __ldcs(p);
#else
    return *p;
#endif
}
template <typename T>
__device__ __forceinline__ void ws_store(T* p, T v)
{
#if IGB_WAVE_STREAM
    __stcs(p, v);
#else
    *p = v;
#endif
}

// K2, gpu_traverse_primary (driver/mapping_gpu.art:45-68): closest hit over the current primary stream; the
// sort key (material id of the hit entity, n_mat for a miss) and its histogram are produced here
// (replaces the reset + count launches of gpu_sort_primary, mapping_gpu.art:431-446).
struct PrimaryIO {
    static constexpr bool kBins = true;
    PathStream     s;
    uint32_t*      keys;
    const int32_t* ent_mat;
    WaveControl*   ctl;
    uint32_t       n_bins; // n_mat + 1
    __device__ __forceinline__ uint32_t  count() const { return ctl->n_cur; }
    __device__ __forceinline__ uint32_t  bin_count() const { return n_bins; }
    __device__ __forceinline__ uint32_t* global_bins() const { return ctl->bins; }
#if IGB_RAY_PREFETCH
    __device__ __forceinline__ void prefetch(uint32_t first, uint32_t count, uint32_t lane) const
    {
        for (uint32_t k = lane; k < count; k += 32u) prefetch_line(s.rec + kRecVec * (size_t)(first + k)); // the ray sector of each record
    }
#endif
    __device__ __forceinline__ bool load(uint32_t i, Vec3& org, Vec3& dir, float& tmin, float& tmax, uint32_t& flags) const
    {
        const float4* r = s.rec + kRecVec * (size_t)i;
        const float4  o = stream_load(r, false), d = stream_load(r + 1, false);
        org   = Vec3{ o.x, o.y, o.z };
        dir   = Vec3{ d.x, d.y, d.z };
        tmin  = o.w;
        tmax  = d.w;
        flags = s.flags[i] & 0xFFu;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const HitRec& hit, uint32_t* s_bins) const
    {
        float4* r = s.rec + kRecVec * (size_t)i; // the whole third sector: no partial-sector write
        stream_store(r + 4, make_float4(__int_as_float(hit.ent), __int_as_float(hit.prim), hit.t, hit.u));
        stream_store(r + 5, make_float4(hit.v, 0.0f, 0.0f, 0.0f));
        const uint32_t key = hit.prim >= 0 ? (uint32_t)__ldg(ent_mat + hit.ent) : n_bins - 1u;
        keys[i]            = key;
        if (n_bins <= (uint32_t)kMaxSmemBins)
            atomicAdd(s_bins + key, 1u);
        else
            atomicAdd(ctl->bins + key, 1u);
    }
    // exclusive scan of the histogram -> scatter offsets (replaces the host-side scan between two blocking
    // copies, mapping_gpu.art:453-461)
    __device__ __forceinline__ void finalize_last_block() const
    {
        if (threadIdx.x < 32u) {
            uint32_t carry = 0u;
            for (uint32_t base = 0; base < n_bins; base += 32u) {
                const uint32_t b = base + threadIdx.x;
                const uint32_t v = b < n_bins ? __ldcg(ctl->bins + b) : 0u;
                uint32_t       x = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, o);
                    if ((int)threadIdx.x >= o) x += y;
                }
                if (b < n_bins) {
                    ctl->offsets[b] = carry + x - v;
                    ctl->cursors[b] = 0u;
                }
                carry += __shfl_sync(0xFFFFFFFFu, x, 31);
            }
        }
    }
};

// K3, gpu_traverse_secondary (driver/mapping_gpu.art:70-112): any hit over the shadow stream; an unoccluded ray
// splats its contribution (simple shadow handling), make_atomic_accumulator (driver/accumulator.art:4-21).
struct ShadowIO {
    static constexpr bool kBins = false;
    ShadowStream s;
    WaveControl* ctl;
    __device__ __forceinline__ uint32_t  count() const { return ctl->n_shadow; }
    __device__ __forceinline__ uint32_t  bin_count() const { return 0u; }
    __device__ __forceinline__ uint32_t* global_bins() const { return nullptr; }
    __device__ __forceinline__ void      finalize_last_block() const {}
#if IGB_RAY_PREFETCH
    __device__ __forceinline__ void prefetch(uint32_t first, uint32_t count, uint32_t lane) const
    {
        if (2u * lane < count) { // 32 B = two rays per lane and column
            prefetch_line(s.org_tmin + first + 2u * lane);
            prefetch_line(s.dir_tmax + first + 2u * lane);
        }
    }
#endif
    __device__ __forceinline__ bool load(uint32_t i, Vec3& org, Vec3& dir, float& tmin, float& tmax, uint32_t& flags) const
    {
        const float4 o = stream_load(s.org_tmin + i, false), d = stream_load(s.dir_tmax + i, false);
        org   = Vec3{ o.x, o.y, o.z };
        dir   = Vec3{ d.x, d.y, d.z };
        tmin  = o.w;
        tmax  = d.w;
        flags = kRayShadow;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const HitRec& hit, uint32_t*) const
    {
        if (hit.prim < 0) {
            const float4 c       = s.color_pixel[i];
            const float  inv_spi = ctl->set.inv_spi;
            float*       p       = ctl->set.film + 3 * (size_t)__float_as_uint(c.w);
            if (!isfinite(c.x + c.y + c.z)) return; // see splat()
            atomicAdd(p + 0, c.x * inv_spi);
            atomicAdd(p + 1, c.y * inv_spi);
            atomicAdd(p + 2, c.z * inv_spi);
            if (float* a = ctl->set.aov[1]) { // on_shadow_miss: aov_nee (pathtracer.art:202-208)
                a += 3 * (size_t)__float_as_uint(c.w);
                atomicAdd(a + 0, c.x * inv_spi);
                atomicAdd(a + 1, c.y * inv_spi);
                atomicAdd(a + 2, c.z * inv_spi);
            }
        }
    }
};

// ---------------------------------------------------------------------------------------------------------
// Round control + ray generation
// ---------------------------------------------------------------------------------------------------------

// Head of the loop body of gpu_trace (driver/mapping_gpu.art:756-767): how many rays to regenerate, stream sizes.
__global__ void begin_round_kernel(WaveControl* ctl, uint32_t n_bins, uint32_t* host_done)
{
    if (threadIdx.x == 0) {
        const uint32_t capacity = ctl->set.capacity;
        const uint32_t n_surv = ctl->n_next;
        ctl->bounce_rays += n_surv;
        ctl->shadow_rays += ctl->n_shadow;
        const uint32_t remaining = ctl->total - ctl->next_id;
        const uint32_t m         = min(remaining, capacity - n_surv);
        ctl->gen_base  = n_surv;
        ctl->gen_first = ctl->next_id;
        ctl->gen_count = m;
        ctl->next_id += m;
        ctl->camera_rays += m;
        ctl->n_cur    = n_surv + m;
        ctl->n_next   = 0u;
        ctl->n_shadow = 0u;
        volatile uint32_t* flags = host_done; // [0] nothing left at all, [1] every id of the call has been generated
        if (ctl->next_id == ctl->total && ctl->gen_done == 0u) {
            ctl->gen_done = 1u;
            flags[1]      = 1u;
            __threadfence_system();
        }
        if (n_surv + m == 0u) {
            ctl->done = 1u;
            flags[0]  = 1u;
            __threadfence_system();
        } else {
            ctl->rounds += 1u;
        }
    }
    for (uint32_t b = threadIdx.x; b < n_bins; b += blockDim.x) ctl->bins[b] = 0u;
}

// keep != 0: the survivors and the shadow-ray count of the previous call's last round stay (pipelined steps, igb200_render_pipelined:
// calls end on a round PAIR, so the survivors are in stream 0 where the next call's first round expects them)
__global__ void reset_kernel(WaveControl* ctl, const WaveSettings set, uint32_t total, uint32_t* host_done, int keep)
{
    ctl->set = set;
    ctl->n_cur = 0u;
    if (!keep) ctl->n_next = ctl->n_shadow = 0u;
    ctl->gen_base = ctl->gen_count = ctl->gen_first = 0u;
    ctl->next_id = 0u;
    ctl->total   = total;
    ctl->done    = 0u;
    ctl->gen_done = 0u;
    ctl->rounds  = 0u;
    ctl->work_primary[0] = ctl->work_primary[1] = ctl->work_shadow[0] = ctl->work_shadow[1] = 0u;
    ctl->camera_rays = ctl->bounce_rays = ctl->shadow_rays = 0ull;
    reinterpret_cast<volatile uint32_t*>(host_done)[0] = 0u;
    reinterpret_cast<volatile uint32_t*>(host_done)[1] = 0u;
    __threadfence_system();
}

// K1, gpu_generate_rays (driver/mapping_gpu.art:618-667) with make_camera_emitter / make_list_emitter
// (driver/emitter.art:6-31), make_perspective_camera (camera/perspective.art:29-41), make_uniform_pixel_sampler
// (sampler/pixel_sampler.art:4-10) and init_pt_raypayload (technique/pathtracer.art:33-38).
#ifndef IGB_GEN_MIN_BLOCKS
#define IGB_GEN_MIN_BLOCKS 1
#endif
#ifndef IGB_GEN_GRID_PER_SM
#define IGB_GEN_GRID_PER_SM 8
#endif
__global__ void __launch_bounds__(256, IGB_GEN_MIN_BLOCKS) generate_kernel(const WaveControl* __restrict__ ctl, const PathStream s, const ShadeScene sc)
{
    const WaveSettings set = ctl->set;
    const uint32_t count = ctl->gen_count, base = ctl->gen_base, first = ctl->gen_first;
    const Vec3     eye{ sc.cam.eye[0], sc.cam.eye[1], sc.cam.eye[2] }, cdir{ sc.cam.dir[0], sc.cam.dir[1], sc.cam.dir[2] },
        up{ sc.cam.up[0], sc.cam.up[1], sc.cam.up[2] };
    const Mat3 view{ vnormalize(cross3(cdir, up)), up, cdir };
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) {
        const uint32_t gid    = first + k;                 // id within the batch of iterations
        const uint32_t it     = gid / set.per_iter;        // iteration offset: the batch regenerates across iterations
        const uint32_t id     = gid - it * set.per_iter;   // ray id = (y * w + x) * spi + sample (mapping_gpu.art:655)
        const uint32_t sample = id % set.spi;
        uint32_t       pixel  = id / set.spi;
        uint32_t       y = pixel / set.width;
        const uint32_t x = pixel - y * set.width;
        if (set.tile_count > 1u) { // local row -> film row: owned tile rows are tile_rank, tile_rank + tile_count, ...
            y     = ((y / kTileSize) * set.tile_count + set.tile_rank) * kTileSize + (y % kTileSize);
            pixel = y * set.width + x;
        }
        Rnd            rnd{ create_random_seed(sample, set.iter + it, set.frame, x, y, set.seed), 1u };
        Vec3           org, dir;
        float          tmin, tmax;
        uint32_t       flags;
        if (set.list_rays) {
            const float4 r0 = __ldg(set.list_rays + 2 * (size_t)pixel), r1 = __ldg(set.list_rays + 2 * (size_t)pixel + 1);
            org   = Vec3{ r0.x, r0.y, r0.z };
            dir   = Vec3{ r0.w, r1.x, r1.y };
            tmin  = r1.z;
            tmax  = r1.w;
            flags = 0u;
        } else {
            const float rx = rnd.next_f32();
            const float ry = rnd.next_f32();
            // make_pixelcoord_from_xy (driver/camera.art:21-29)
            const float nx = __fdiv_rn(2.0f * ((float)x + rx), (float)set.width) - 1.0f;
            const float ny = 1.0f - __fdiv_rn(2.0f * ((float)y + ry), (float)set.height);
            org   = eye;
            if (sc.cam.type == IGB200_CAMERA_FISHLENS) { // make_fishlens_camera::compute_d, camera/fishlens.art:38-50 (fov = pi)
                const float fx = nx * sc.cam.scale[0], fy = ny * sc.cam.scale[1];
                const float r     = __fsqrt_rn(fx * fx + fy * fy);
                const float theta = __fdiv_rn(r * flt_pi, 2.0f);
                const float sT = sinf(theta), cT = cosf(theta);
                const float sP = r < flt_eps ? 0.0f : __fdiv_rn(fy, r), cP = r < flt_eps ? 0.0f : __fdiv_rn(fx, r);
                dir = vnormalize(mat3_mul(view, Vec3{ sT * cP, sT * sP, cT }));
            } else {
                dir = vnormalize(mat3_mul(view, Vec3{ sc.cam.scale[0] * nx, sc.cam.scale[1] * ny, 1.0f }));
            }
            tmin  = sc.cam.tmin;
            tmax  = sc.cam.tmax;
            flags = kRayCamera;
        }
        const uint32_t slot    = base + k;
        float4*        r       = s.rec + kRecVec * (size_t)slot;
        ws_store(r, make_float4(org.x, org.y, org.z, tmin));
        ws_store(r + 1, make_float4(dir.x, dir.y, dir.z, tmax));
        ws_store(r + 2, make_float4(__uint_as_float(pixel), __uint_as_float(rnd.counter), __uint_as_float(rnd.seed), 1.0f));
        ws_store(r + 3, make_float4(1.0f, 1.0f, 1.0f, 0.0f));
        s.flags[slot] = (uint16_t)(flags | (1u << 8)); // depth 1
    }
}

// Scatter step of the sort (gpu_sort_primary, driver/mapping_gpu.art:465-471), on indices only.  Each block ranks a
// tile of keys in shared memory (lanes with equal keys are grouped with match.any, one shared atomic per group),
// reserves one range per key with ONE global atomic per key and tile, then writes the slot indices.
constexpr int kScatterItems = 8; // keys per thread and tile
__global__ void __launch_bounds__(256) scatter_kernel(WaveControl* __restrict__ ctl, const uint32_t* __restrict__ keys,
                                                      uint32_t* __restrict__ sorted, const uint32_t n_bins, const bool tile_local)
{
    __shared__ uint32_t s_cnt[kMaxSmemBins], s_base[kMaxSmemBins];
    const uint32_t n    = ctl->n_cur;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t tile = blockDim.x * kScatterItems;
    if (n_bins > (uint32_t)kMaxSmemBins) { // many materials: warp-aggregated global cursors
        for (uint32_t i0 = blockIdx.x * blockDim.x; i0 < n; i0 += gridDim.x * blockDim.x) {
            const uint32_t i     = i0 + threadIdx.x;
            const bool     valid = i < n;
            const uint32_t key   = valid ? keys[i] : 0xFFFFFFFFu;
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
            const uint32_t lead  = __ffs(peers) - 1u;
            uint32_t       base  = 0u;
            if (valid && lane == lead) base = ctl->offsets[key] + atomicAdd(ctl->cursors + key, (uint32_t)__popc(peers));
            base = __shfl_sync(0xFFFFFFFFu, base, lead);
            if (valid) sorted[base + __popc(peers & ((1u << lane) - 1u))] = i;
        }
        return;
    }
    for (uint32_t t0 = blockIdx.x * tile; t0 < n; t0 += gridDim.x * tile) { // block-uniform trip count
        for (uint32_t b = threadIdx.x; b < n_bins; b += blockDim.x) s_cnt[b] = 0u;
        __syncthreads();
        uint32_t key[kScatterItems], rank[kScatterItems];
#pragma unroll
        for (int k = 0; k < kScatterItems; ++k) {
            const uint32_t i = t0 + k * blockDim.x + threadIdx.x;
            key[k]           = i < n ? keys[i] : 0xFFFFFFFFu;
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key[k]);
            const uint32_t lead  = __ffs(peers) - 1u;
            uint32_t       base  = 0u;
            if (i < n && lane == lead) base = atomicAdd(s_cnt + key[k], (uint32_t)__popc(peers));
            rank[k] = __shfl_sync(0xFFFFFFFFu, base, lead) + __popc(peers & ((1u << lane) - 1u));
        }
        __syncthreads();
        if (tile_local) {
            // sort inside the tile only: slots stay within a 2048-ray window, so the shade kernel's gather hits L2
            if (threadIdx.x < 32u) {
                uint32_t carry = t0;
                for (uint32_t b0 = 0; b0 < n_bins; b0 += 32u) {
                    const uint32_t b = b0 + lane;
                    const uint32_t c = b < n_bins ? s_cnt[b] : 0u;
                    uint32_t       x = c;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d);
                        if ((int)lane >= d) x += y;
                    }
                    if (b < n_bins) s_base[b] = carry + x - c;
                    carry += __shfl_sync(0xFFFFFFFFu, x, 31);
                }
            }
        } else {
            for (uint32_t b = threadIdx.x; b < n_bins; b += blockDim.x)
                if (s_cnt[b]) s_base[b] = ctl->offsets[b] + atomicAdd(ctl->cursors + b, s_cnt[b]);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kScatterItems; ++k) {
            const uint32_t i = t0 + k * blockDim.x + threadIdx.x;
            if (i < n) sorted[s_base[key[k]] + rank[k]] = i;
        }
        __syncthreads();
    }
}

__global__ void film_combine_kernel(float* film, const float* src, uint64_t count, bool overwrite)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) film[i] = overwrite ? src[i] : film[i] + src[i];
}

// ---------------------------------------------------------------------------------------------------------
// K7 + K8: hit and miss shading in ONE launch over the material-sorted index list
// (gpu_hit_shade, one launch per material, driver/mapping_gpu.art:114-205,787-795; gpu_miss_shade :229-266).
// Per ray: surface element -> material -> on_hit splat -> on_shadow (NEE) -> on_bounce, in the reference's order of
// RNG use (technique/pathtracer.art:52-200).  Survivors go to the OTHER primary stream at a warp-compacted
// position (this is the compaction pass, gpu_compact_primary :685-714); shadow rays go to the compacted
// secondary stream.
// ---------------------------------------------------------------------------------------------------------
struct NextState {
    float4   org_tmin, dir_tmax, meta, contrib_invpdf;
    uint16_t flags;
};
struct ShadowState {
    float4 org_tmin, dir_tmax, color_pixel;
};

__device__ __forceinline__ void splat(float* film, uint32_t pixel, Color c, float inv_spi)
{
    // The reference's estimator can evaluate 0 * inf (e.g. a BSDF sample of density 0 meeting a CDF cell of density 0
    // in the MIS weight, pathtracer.art:125,152) and leaves the NaN in the film for imageinfo to count
    // (entrypoints/imageinfo.art:81-86).  A NaN never leaves a pixel again, so non-finite contributions are dropped
    // here and in the oracle; results differ from the reference only where its pixel would be NaN / Inf.
    if (!isfinite(c.r + c.g + c.b)) return;
    float* p = film + 3 * (size_t)pixel;
    atomicAdd(p + 0, c.r * inv_spi);
    atomicAdd(p + 1, c.g * inv_spi);
    atomicAdd(p + 2, c.b * inv_spi);
}

template <bool EXT, bool IB>
__device__ __forceinline__ void shade_one(const ShadeScene& sc, const DeviceScene& geo, const WaveSettings& set, const PathStream& cur,
                                       uint32_t i, float* film, bool& alive, NextState& nx, bool& has_shadow, ShadowState& sh)
{
    const igb200_technique& tech = sc.tech;
    const float4* rec = cur.rec + kRecVec * (size_t)i;
    const float4  o4 = ws_load(rec), d4 = ws_load(rec + 1), meta = ws_load(rec + 2), cp = ws_load(rec + 3), h0 = ws_load(rec + 4);
    const Vec3   org{ o4.x, o4.y, o4.z }, dir{ d4.x, d4.y, d4.z };
    const int32_t ent = __float_as_int(h0.x), prim = __float_as_int(h0.y);
    const float   t = h0.z, u = h0.w;
    const uint32_t pixel = __float_as_uint(meta.x);
    const int32_t  depth = (int32_t)(cur.flags[i] >> 8);
    const Color    contrib0{ cp.x, cp.y, cp.z };
    const float    inv_pdf = cp.w;
    const float    offset  = 0.001f;

    if (prim < 0 || ent < 0) {
        // on_miss (technique/pathtracer.art:134-160): every infinite, non-delta light contributes its emission
        if (tech.type == IGB200_TECH_PATH) {
            int   inflights = 0;
            Color col{ 0.0f, 0.0f, 0.0f };
            for (int32_t k = 0; k < sc.n_inf; ++k) {
                const igb200_light& l = sc.inf_lights[k];
                if (light_infinite(l.type) && !light_delta(l.type)) {
                    ++inflights;
                    const float pdf_s = env_pdf<EXT>(sc, l, dir); // env.art:96,152
                    const float mis   = tech.enable_nee ? __fdiv_rn(1.0f, 1.0f + inv_pdf * selector_pdf<EXT>(sc, true, k, org) * pdf_s) : 1.0f;
                    col               = cadd(col, handle_color(tech, cmulf(cmul(contrib0, env_emission<EXT>(sc, l, dir)), mis)));
                }
            }
#ifdef IGB_DEBUG_NAN
            if (!isfinite(col.r + col.g + col.b))
                printf("NAN miss px %u depth %d contrib %g %g %g inv_pdf %g dir %g %g %g\n", pixel, depth, contrib0.r, contrib0.g, contrib0.b,
                       inv_pdf, dir.x, dir.y, dir.z);
#endif
            if (inflights > 0) {
                splat(film, pixel, col, set.inv_spi);
                if (set.aov[0]) splat(set.aov[0], pixel, col, set.inv_spi); // aov_di, pathtracer.art:158
            }
        }
        return;
    }

    const float  v = rec[5].x;
    const Entity e = load_entity(sc, ent);
    Surf         surf;
    const uint2  si = sc.shape_info[e.shape_id()];
    if (si.y == IGB200_PRIM_SPHERE)
        surf = sphere_surface_element(__ldg(reinterpret_cast<const float4*>(sc.shape_data + si.x)), e, org, dir, t, u, v);
    else
        surf = trimesh_surface_element(load_trimesh(sc, e.shape_id()), e, org, dir, t, u, v, prim);
    const igb200_material mat = sc.materials[e.mat_id()];
    const Bsdf<EXT> bsdf = make_bsdf<EXT>(sc, mat.bsdf, surf, dir, (tech.flags & IGB200_TECHFLAG_ONE_SIDED_DIFFUSE) != 0u);

    if constexpr (IB) { // make_infobuffer_renderer, technique/infobuffer.art:28-104; the payload's distance rides in cp.w
        const bool handle_specular = (tech.flags & IGB200_TECHFLAG_IB_HANDLE_SPECULAR) != 0u;
        const Vec3 out_dir         = vneg(dir);
        const bool specular        = bsdf.is_specular();
        if (!handle_specular || !specular) { // on_hit :36-63
            Color albedo{ 0.0f, 0.0f, 0.0f };
            Rnd   arnd{ hash_combine(0x811C9DC5u, pixel), 1u }; // create_random_generator(hash_combine(hash_init(), pixel.linear))
            for (int k = 0; k < 16; ++k) {
                BsdfSample ms;
                if (bsdf.sample(arnd, out_dir, ms)) {
                    const float c = fabsf(dot3(ms.in_dir, surf.local.c2));
                    albedo        = cadd(albedo, cmulf(ms.col, safe_div(ms.pdf, 16.0f * c))); // without the cosine term
                }
            }
            const float d = t + (handle_specular ? inv_pdf : 0.0f);
            splat(set.aov[0], pixel, Color{ surf.local.c2.x, surf.local.c2.y, surf.local.c2.z }, set.inv_spi);
            splat(set.aov[1], pixel, Color{ fminf(albedo.r, 1.0f), fminf(albedo.g, 1.0f), fminf(albedo.b, 1.0f) }, set.inv_spi);
            splat(set.aov[2], pixel, Color{ d, d, d }, set.inv_spi);
        }
        // on_bounce :65-91
        if (!(handle_specular && specular) || !(depth <= tech.max_depth)) return;
        Rnd        rnd{ __float_as_uint(meta.z), __float_as_uint(meta.y) };
        BsdfSample ms;
        if (!bsdf.sample(rnd, out_dir, ms)) return;
        alive             = true;
        nx.org_tmin       = make_float4(surf.point.x, surf.point.y, surf.point.z, offset);
        nx.dir_tmax       = make_float4(ms.in_dir.x, ms.in_dir.y, ms.in_dir.z, flt_max);
        nx.meta           = make_float4(meta.x, __uint_as_float(rnd.counter), meta.z, meta.w);
        nx.contrib_invpdf = make_float4(cp.x, cp.y, cp.z, inv_pdf + t);
        nx.flags          = (uint16_t)(kRayBounce | ((uint32_t)(depth + 1) << 8));
        return;
    }

    if (tech.type == IGB200_TECH_DEBUG) { // technique/debugtracer.art:29-177
        Color out;
        auto  absv = [](Vec3 a) { return Color{ fabsf(a.x), fabsf(a.y), fabsf(a.z) }; };
        switch (tech.debug_mode) {
        case 1: out = absv(surf.local.c0); break;
        case 2: out = absv(surf.local.c1); break;
        case 3: out = absv(surf.face_normal); break;
        case 8: { float tu, tv; surf_tex_coords(surf, tu, tv); out = Color{ fabsf(tu), fabsf(tv), 0.0f }; break; }
        case 9: out = Color{ fabsf(u), fabsf(v), 0.0f }; break;
        case 10: out = Color{ surf.point.x, surf.point.y, surf.point.z }; break;
        case 13: out = Color{ t, t, t }; break;
        case 14: { const float a = safe_div(1.0f, surf.inv_area); out = Color{ a, a, a }; break; }
        case 15: { const float f = (float)prim; out = Color{ f, f, f }; break; }
        case 17: { const float f = (float)ent; out = Color{ f, f, f }; break; }
        case 19: { const float f = (float)e.mat_id(); out = Color{ f, f, f }; break; }
        case 21: out = mat.light >= 0 ? Color{ 0.0f, 0.0f, 1.0f } : Color{ 1.0f, 0.0f, 0.0f }; break;
        case 22: out = bsdf.is_specular() ? Color{ 0.0f, 0.0f, 1.0f } : Color{ 1.0f, 0.0f, 0.0f }; break;
        case 23: out = surf.is_entering ? Color{ 0.0f, 0.0f, 1.0f } : Color{ 1.0f, 0.0f, 0.0f }; break;
        default: out = absv(surf.local.c2); break;
        }
        splat(film, pixel, out, set.inv_spi);
        return;
    }

    // on_hit (pathtracer.art:112-132): emission of area lights with MIS against NEE
    if (mat.light >= 0 && surf.is_entering) {
        const float d = -dot3(dir, surf.local.c2);
        if (d > flt_eps) {
            const igb200_light& l     = sc.fin_lights[mat.light];
            const Pdf           pdf   = light_pdf_direct(sc, l, org, surf);
            const float         pdf_s = pdf.as_solid(d, t * t);
            const float         mis   = tech.enable_nee ? __fdiv_rn(1.0f, 1.0f + inv_pdf * selector_pdf<EXT>(sc, false, mat.light, org) * pdf_s) : 1.0f;
            const Color         c     = handle_color(tech, cmulf(cmul(contrib0, light_color(l)), mis));
            splat(film, pixel, c, set.inv_spi);
            if (set.aov[0]) splat(set.aov[0], pixel, c, set.inv_spi); // aov_di, pathtracer.art:128
        }
    }

    Rnd            rnd{ __float_as_uint(meta.z), __float_as_uint(meta.y) }; // seed and counter travel with the ray
    const Vec3     out_dir = vneg(dir);

    // on_shadow (pathtracer.art:52-110): next event estimation
    if (tech.enable_nee && !bsdf.is_specular() && (sc.n_inf + sc.n_fin) != 0 && depth + 1 <= tech.max_depth) {
        float               sel_pdf;
        const igb200_light& l  = selector_sample<EXT>(sc, rnd, surf.point, sel_pdf);
        const DirectSample  ls = light_sample_direct<EXT>(sc, l, rnd, surf);
        const float pdf_l_s    = ls.pdf.as_solid(ls.cos, ls.dist * ls.dist) * sel_pdf;
        if (!(pdf_l_s <= flt_eps) && ls.cos > flt_eps) {
            const float mis    = light_delta(l.type) ? 1.0f : __fdiv_rn(1.0f, 1.0f + __fdiv_rn(bsdf.pdf(ls.dir, out_dir), pdf_l_s));
            const float factor = __fdiv_rn(ls.pdf.value, pdf_l_s);
            const Color c      = handle_color(tech, cmulf(cmul(ls.intensity, cmul(contrib0, bsdf.eval(ls.dir, out_dir))), mis * factor));
            has_shadow         = true;
            if (light_infinite(l.type)) {
                sh.org_tmin = make_float4(surf.point.x, surf.point.y, surf.point.z, offset);
                sh.dir_tmax = make_float4(ls.dir.x, ls.dir.y, ls.dir.z, flt_max);
            } else {
                const Vec3 dd = vsub(ls.pos, surf.point);
                sh.org_tmin   = make_float4(surf.point.x, surf.point.y, surf.point.z, offset);
                sh.dir_tmax   = make_float4(dd.x, dd.y, dd.z, 1.0f - offset);
            }
#ifdef IGB_DEBUG_NAN
            if (!isfinite(c.r + c.g + c.b))
                printf("NAN nee px %u depth %d contrib0 %g %g %g ls.int %g %g %g pdf_l_s %g lspdf %g mis %g eval %g bsdfpdf %g type %u au %g\n", pixel,
                       depth, contrib0.r, contrib0.g, contrib0.b, ls.intensity.r, ls.intensity.g, ls.intensity.b, pdf_l_s, ls.pdf.value, mis,
                       bsdf.eval(ls.dir, out_dir).r, bsdf.pdf(ls.dir, out_dir), bsdf.d.type, bsdf.au());
#endif
            sh.color_pixel = make_float4(c.r, c.g, c.b, __uint_as_float(pixel));
        }
    }

    // on_bounce (pathtracer.art:162-200)
    if (depth + 1 > tech.max_depth) return;
    BsdfSample ms;
    if (!bsdf.sample(rnd, out_dir, ms)) return;
#ifdef IGB_DEBUG_NAN
    if (!isfinite(ms.col.r + ms.col.g + ms.col.b + ms.pdf + ms.in_dir.x + ms.in_dir.y + ms.in_dir.z))
        printf("NAN sample px %u depth %d col %g %g %g pdf %g in %g %g %g out %g %g %g n %g %g %g type %u\n", pixel, depth, ms.col.r, ms.col.g,
               ms.col.b, ms.pdf, ms.in_dir.x, ms.in_dir.y, ms.in_dir.z, out_dir.x, out_dir.y, out_dir.z, bsdf.local.c2.x, bsdf.local.c2.y,
               bsdf.local.c2.z, bsdf.d.type);
#endif
    const Color      contrib = cmul(contrib0, ms.col);
    const float      eta0    = meta.w;
    float            rr_prob = 1.0f;
    if (depth + 1 > tech.min_depth) {
        const Color cc = cmulf(contrib, eta0 * eta0);
        rr_prob        = clampf(fmaxf(fmaxf(cc.r, cc.g), cc.b), 0.05f, 0.95f); // russian_roulette_pbrt
    }
    if (rnd.next_f32() >= rr_prob) return;
    const float new_inv_pdf = bsdf.is_specular() ? 0.0f : __fdiv_rn(1.0f, ms.pdf);
    const Color nc          = cmulf(contrib, __fdiv_rn(1.0f, rr_prob));
    alive             = true;
    nx.org_tmin       = make_float4(surf.point.x, surf.point.y, surf.point.z, offset);
    nx.dir_tmax       = make_float4(ms.in_dir.x, ms.in_dir.y, ms.in_dir.z, flt_max);
    nx.meta           = make_float4(meta.x, __uint_as_float(rnd.counter), meta.z, eta0 * ms.eta);
    nx.contrib_invpdf = make_float4(nc.r, nc.g, nc.b, new_inv_pdf);
    nx.flags          = (uint16_t)(kRayBounce | ((uint32_t)(depth + 1) << 8));
}

#ifndef IGB_SHADE_PREFETCH
#define IGB_SHADE_PREFETCH 0
#endif
#ifndef IGB_SHADE_MIN_BLOCKS
#define IGB_SHADE_MIN_BLOCKS 6 // 85 registers, 80 B of spills: +2 % over 4 blocks (112 registers) on B200
#endif
// Blocks of the grid-stride launch per SM.  Exactly the resident wave: with 16 per SM (2.67 waves of 6) the last partial wave
// cost 2-3 % on all three path-tracer workloads (profiles/r02_ab_shade_grid.txt: 6 / 12 / 16 / 18 / 24 per SM).
#ifndef IGB_SHADE_GRID_PER_SM
#define IGB_SHADE_GRID_PER_SM IGB_SHADE_MIN_BLOCKS
#endif
template <bool EXT, bool IB>
__global__ void __launch_bounds__(128, IGB_SHADE_MIN_BLOCKS) shade_kernel(WaveControl* __restrict__ ctl, const PathStream cur, const PathStream nxt,
                                                    const ShadowStream shq, const uint32_t* __restrict__ sorted, const ShadeScene sc,
                                                    const DeviceScene geo)
{
    const WaveSettings set  = ctl->set;
    float* const       film = set.film;
    const uint32_t     n    = ctl->n_cur;
    const uint32_t lane    = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t step    = gridDim.x * blockDim.x;
    for (uint32_t p0 = blockIdx.x * blockDim.x; p0 < n; p0 += step) { // block-uniform trip count
        const uint32_t p = p0 + threadIdx.x;
#if IGB_SHADE_PREFETCH
        // the record gather through the sorted index is a random 96-byte read: start the next trip's now (three 32-byte sectors)
        if (p + step < n) {
            const uint32_t nxi = sorted ? __ldg(sorted + p + step) : p + step;
            const float4*  nr  = cur.rec + kRecVec * (size_t)nxi;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nr));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nr + 2));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(nr + 4));
        }
#endif
        bool        alive = false, has_shadow = false;
        NextState   nx;
        ShadowState sh;
        if (p < n) shade_one<EXT, IB>(sc, geo, set, cur, sorted ? sorted[p] : p, film, alive, nx, has_shadow, sh);

        // stream compaction: one atomic per warp, ranks from the ballot (replaces gpu_compact_primary)
        const uint32_t am = __ballot_sync(0xFFFFFFFFu, alive);
        if (am) {
            uint32_t base = 0u;
            if (lane == (uint32_t)(__ffs(am) - 1)) base = atomicAdd(&ctl->n_next, (uint32_t)__popc(am));
            base = __shfl_sync(0xFFFFFFFFu, base, __ffs(am) - 1);
            if (alive) {
                const uint32_t slot      = base + __popc(am & lt_mask);
                float4*        r    = nxt.rec + kRecVec * (size_t)slot;
                ws_store(r, nx.org_tmin);
                ws_store(r + 1, nx.dir_tmax);
                ws_store(r + 2, nx.meta);
                ws_store(r + 3, nx.contrib_invpdf);
                nxt.flags[slot] = nx.flags;
            }
        }
        const uint32_t sm = __ballot_sync(0xFFFFFFFFu, has_shadow);
        if (sm) {
            uint32_t base = 0u;
            if (lane == (uint32_t)(__ffs(sm) - 1)) base = atomicAdd(&ctl->n_shadow, (uint32_t)__popc(sm));
            base = __shfl_sync(0xFFFFFFFFu, base, __ffs(sm) - 1);
            if (has_shadow) {
                const uint32_t slot    = base + __popc(sm & lt_mask);
                ws_store(shq.org_tmin + slot, sh.org_tmin);
                ws_store(shq.dir_tmax + slot, sh.dir_tmax);
                ws_store(shq.color_pixel + slot, sh.color_pixel);
            }
        }
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------------------
// Launchers
// ---------------------------------------------------------------------------------------------------------
cudaError_t launch_wave_reset(const WaveBuffers& buf, const WaveSettings& set, uint32_t total, cudaStream_t stream, bool keep_survivors)
{
    reset_kernel<<<1, 1, 0, stream>>>(buf.ctl, set, total, buf.host_done, keep_survivors ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_wave_round(const DeviceScene& scene, const shade::ShadeScene& sh, const WaveBuffers& buf, int parity,
                              int sort, int sm_count, cudaStream_t stream, cudaEvent_t* ev)
{
#define IGB_EV(k) do { if (ev) cudaEventRecord(ev[k], stream); } while (0)
    const PathStream& cur = buf.stream[parity & 1];
    const PathStream& nxt = buf.stream[(parity + 1) & 1];
    const uint32_t    n_bins = (uint32_t)sh.n_mat + 1u;
    if (n_bins > kWaveMaxBins) return cudaErrorInvalidValue;
    const int wide = sm_count * 8; // grid-stride kernels: a whole number of waves

    IGB_EV(0);
    begin_round_kernel<<<1, 256, 0, stream>>>(buf.ctl, n_bins, buf.host_done);
    IGB_EV(1);
    generate_kernel<<<sm_count * IGB_GEN_GRID_PER_SM, 256, 0, stream>>>(buf.ctl, cur, sh);
    IGB_EV(2);
    const PrimaryIO pio{ cur, buf.keys, buf.ent_mat, buf.ctl, n_bins };
    trace_kernel<PrimaryIO, false, false><<<sm_count * kBlocksPerSM, kBlock, 0, stream>>>(scene, pio, nullptr, buf.ctl->work_primary);
    IGB_EV(3);
    if (sort) scatter_kernel<<<wide / 2, 256, 0, stream>>>(buf.ctl, buf.keys, buf.sorted, n_bins, sort == 2 && n_bins <= kMaxSmemBins);
    IGB_EV(4);
    // four instantiations: {lean, extended material set} x {path / debug, infobuffer}, picked per scene and technique
#define SHADE_LAUNCH(EXT)                                                                                                           \
    do {                                                                                                                            \
        if (sh.tech.type == IGB200_TECH_INFOBUFFER)                                                                                 \
            shade_kernel<EXT, true><<<sm_count * IGB_SHADE_GRID_PER_SM, 128, 0, stream>>>(buf.ctl, cur, nxt, buf.shadow, sort ? buf.sorted : nullptr, sh, scene); \
        else                                                                                                                        \
            shade_kernel<EXT, false><<<sm_count * IGB_SHADE_GRID_PER_SM, 128, 0, stream>>>(buf.ctl, cur, nxt, buf.shadow, sort ? buf.sorted : nullptr, sh, scene); \
    } while (0)
    if (sh.extended)
        SHADE_LAUNCH(true);
    else
        SHADE_LAUNCH(false);
#undef SHADE_LAUNCH
    IGB_EV(5);
    const ShadowIO sio{ buf.shadow, buf.ctl };
    trace_kernel<ShadowIO, true, false><<<sm_count * kBlocksPerSM, kBlock, 0, stream>>>(scene, sio, nullptr, buf.ctl->work_shadow);
    IGB_EV(6);
#undef IGB_EV
    return cudaGetLastError();
}

cudaError_t launch_film_combine(float* film, const float* src, uint64_t count, bool overwrite, cudaStream_t stream)
{
    if (count == 0) return cudaSuccess;
    film_combine_kernel<<<(uint32_t)((count + 255) / 256), 256, 0, stream>>>(film, src, count, overwrite);
    return cudaGetLastError();
}

} // namespace igb
