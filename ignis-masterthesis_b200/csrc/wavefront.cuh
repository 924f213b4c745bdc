// wavefront.cuh -- ray streams, counters and launchers of the wavefront path tracer (sm_100a).
//
// Replaces the reference's GPU wavefront loop `gpu_trace` (src/artic/driver/mapping_gpu.art:728-870) and its
// streams (src/artic/driver/streams.art:1-44, sizing src/runtime/device/Device.cpp:474-514).  Differences by design:
//   * the primary stream is an array of 96-byte records (three full sectors per ray) instead of one float column
//     per component; the shadow stream is 16-byte-vector SoA (it is only ever read and written in order);
//   * survivors are compacted by the SHADE kernel itself (warp ballot + one atomic per warp) into the other
//     stream buffer, so there is no separate compaction pass (gpu_compact_primary, mapping_gpu.art:685-714);
//   * the sort is an index sort by MATERIAL id (histogram fused into the traversal kernel, scan by its last block,
//     warp-aggregated scatter) -- the reference moves whole 96-byte entries, sorts by entity and scans on the host
//     through two blocking copies (mapping_gpu.art:403-498);
//   * all counts live on the device; one wavefront round is a fixed sequence of six launches captured in a CUDA
//     graph, and the host only polls a mapped "done" flag (no blocking copies per bounce).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#include "../../include/ignis_b200.h"
#include "trace.cuh"

namespace igb {

// Primary stream (one of two ping-pong buffers): one 96-byte RECORD per ray = three 32-byte sectors, so that the
// shade kernel's gather through the material-sorted index list only ever fetches sectors it uses entirely
// (16-byte SoA columns made it fetch 210 B per ray where 92 B were needed, ncu r01_s7):
//   sector 0  [0] org.xyz, tmin        [1] dir.xyz, tmax                       read by traversal, written by generate/shade
//   sector 1  [2] meta                 [3] contrib.rgb, inv_pdf                written by generate/shade, read by shade
//   sector 2  [4] ent, prim, t, u      [5] v, -, -, -                          written by traversal, read by shade
// meta = { pixel, RNG counter (the `rnd` column), RNG seed = hash(sample, iteration, frame, x, y, user seed), eta };
// path payload: technique/pathtracer.art:7-12.  The seed travels with the ray so that a bounce neither re-hashes it
// nor divides the ray id back into pixel / sample / x / y.
// The ray visibility flags (low byte) and the path depth (high byte) live in their own 16-bit column: traversal
// needs the flags without touching sector 1.
constexpr int kRecVec = 6; // float4 per record
struct PathStream {
    float4*  rec;
    uint16_t* flags;
};

// Secondary (shadow) stream, compacted on write.  48 bytes per ray.
struct ShadowStream {
    float4* org_tmin;
    float4* dir_tmax;
    float4* color_pixel; // contribution rgb, pixel index (bit pattern)
};

constexpr uint32_t kWaveMaxBins = 4096; // material bins + miss bin supported by the sort

// Per-call settings; they live in the control block (device memory) so that the captured CUDA graph of a round
// pair stays valid from call to call.
struct WaveSettings {
    uint32_t      spi, width, height, iter, frame, seed; // iter = first iteration of the batch
    uint32_t      per_iter;                              // spi * width * height: ray ids per iteration
    float         inv_spi;
    uint32_t      capacity;
    const float4* list_rays; // igtrace mode: normalised StreamRays (2 x float4 each), else nullptr
    float*        film;      // w * h * 3, RGB sums
    float*        aov[3];    // technique AOVs, same layout, or nullptr: path + aov_mis: Direct Weights, NEE Weights;
                             // infobuffer: Normals, Albedo, Depth
    // Multi-GPU tile split (igb200_render_partition): this device renders the rows of the 16-pixel tile rows
    // t with t % tile_count == tile_rank (the reference's CPU driver walks the same 16 x 16 tiles,
    // mapping_cpu.art:694-706).  tile_count <= 1: the whole film.  per_iter counts the rays of the OWNED rows only.
    uint32_t      tile_rank, tile_count;
};
constexpr uint32_t kTileSize = 16; // rows per tile row of the multi-GPU split

// Device-resident control block of one render() call.
struct WaveControl {
    uint32_t n_cur;     // rays in the current primary stream (survivors + regenerated)
    uint32_t n_next;    // survivors appended to the other stream by the shade kernel
    uint32_t n_shadow;  // shadow rays appended by the shade kernel
    uint32_t gen_base;  // first slot the generator fills this round
    uint32_t gen_count; // rays the generator creates this round
    uint32_t gen_first; // id of the first generated ray
    uint32_t next_id;   // next ray id to generate
    uint32_t total;     // spi * width * height
    uint32_t done;      // set when a round starts with nothing to do
    uint32_t gen_done;  // set when every ray id of the call has been generated
    uint32_t rounds;
    uint32_t work_primary[2]; // persistent-kernel work counters (self-resetting)
    uint32_t work_shadow[2];
    WaveSettings set;
    unsigned long long camera_rays, bounce_rays, shadow_rays;
    uint32_t bins[kWaveMaxBins];    // histogram of sort keys (material id, miss = n_mat)
    uint32_t offsets[kWaveMaxBins]; // exclusive scan of bins
    uint32_t cursors[kWaveMaxBins]; // scatter cursors
};

struct WaveBuffers {
    PathStream     stream[2];
    ShadowStream   shadow;
    uint32_t*      keys;   // sort key per primary slot
    uint32_t*      sorted; // slot indices in key order
    WaveControl*   ctl;
    const int32_t* ent_mat; // material id per entity
    uint32_t*      host_done; // device alias of two mapped host flags: [0] drained, [1] all ids of the call generated
};

namespace shade { struct ShadeScene; }

// Enqueues ONE wavefront round (parity selects which stream buffer is current) on `stream`:
// begin_round, generate, trace_primary(+histogram+scan), scatter, shade, trace_shadow(+splat).
cudaError_t launch_wave_round(const DeviceScene& scene, const shade::ShadeScene& sh, const WaveBuffers& buf, int parity,
                              int sort, int sm_count, cudaStream_t stream, // sort: 0 off, 1 global, 2 tile-local
                              cudaEvent_t* events = nullptr); // stats mode: kWaveLaunchesPerRound + 1 events, one before each
                                                              // kernel and one after the last (not capturable into a graph)

// Resets the control block for a new call (total = spi * w * h * iterations) and stores its settings.  keep_survivors: the paths
// the previous call left in flight stay in the stream (pipelined steps).
cudaError_t launch_wave_reset(const WaveBuffers& buf, const WaveSettings& set, uint32_t total, cudaStream_t stream,
                              bool keep_survivors = false);

// film[i] += src[i] (overwrite = false) or film[i] = src[i]
// Device::tonemap (Device.cpp:1662-1687): scale already carries 1 / IterationCount
struct TonemapParams {
    uint32_t method;    // 0 none, 1 reinhard, 2 reinhard_modified, 3 aces, else uncharted2 (entrypoints/tonemap.art:31-37)
    int32_t  use_gamma;
    float    scale, exposure_factor, exposure_offset;
};
cudaError_t launch_tonemap(const float* in, uint32_t* out, uint64_t n, const TonemapParams& p, cudaStream_t stream);
cudaError_t launch_film_combine(float* film, const float* src, uint64_t count, bool overwrite, cudaStream_t stream);

constexpr int kWaveLaunchesPerRound = 6;

} // namespace igb
