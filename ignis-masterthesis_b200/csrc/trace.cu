// trace.cu -- igtrace ray-list front end of the traversal kernel (trace_core.cuh): AoS StreamRay in, SoA hit columns out.
//
// Replaces the reference's K2/K3 launches for Runtime::trace: gpu_traverse_primary / gpu_traverse_secondary
// (src/artic/driver/mapping_gpu.art:45-112) reading the list emitter's rays (src/artic/driver/emitter.art:18-31).
#include "trace_core.cuh"

#include <algorithm>

namespace igb {
namespace {

using namespace core;

// Ray list in (32 B StreamRay, two 128-bit loads), five hit columns out (src/artic/driver/streams.art:80-100).
// NORMALIZE: host-buffer path, Interface::loadRayList's normalisation (Device.cpp:588-592) fused into the load.
// ALL_OUT: all five columns requested (the usual case): no per-column null checks in the store (the checked store is ~40
// instructions per finished ray at ~7 live lanes, 5 % of the kernel's issue slots on C3).
template <bool NORMALIZE, bool ALL_OUT>
struct ListIO {
    static constexpr bool kBins = false;
    const float4* rays;
    uint32_t      n;
    uint32_t      ray_flags;
    TraceOutputs  out;
    __device__ __forceinline__ uint32_t count() const { return n; }
    __device__ __forceinline__ uint32_t bin_count() const { return 0u; }
    __device__ __forceinline__ uint32_t* global_bins() const { return nullptr; }
    __device__ __forceinline__ void finalize_last_block() const {}
#if IGB_RAY_PREFETCH
    __device__ __forceinline__ void prefetch(uint32_t first, uint32_t count, uint32_t lane) const
    {
        if (2u * lane < count) prefetch_line(rays + 2 * (size_t)(first + 2u * lane)); // 64 B = two rays per lane
    }
#endif
    __device__ __forceinline__ bool load(uint32_t i, Vec3& org, Vec3& dir, float& tmin, float& tmax, uint32_t& flags) const
    {
        const float4 r0 = stream_load(rays + 2 * (size_t)i, true);
        const float4 r1 = stream_load(rays + 2 * (size_t)i + 1, true);
        org   = Vec3{ r0.x, r0.y, r0.z };
        dir   = Vec3{ r0.w, r1.x, r1.y };
        if (NORMALIZE) { // same arithmetic as normalize_rays_kernel below
            float norm = __fsqrt_rn(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
            if (norm < kFltEps) norm = 1.0f;
            dir = Vec3{ __fdiv_rn(dir.x, norm), __fdiv_rn(dir.y, norm), __fdiv_rn(dir.z, norm) };
        }
        tmin  = r1.z;
        tmax  = r1.w;
        flags = ray_flags;
        return true;
    }
    __device__ __forceinline__ void store(uint32_t i, const HitRec& hit, uint32_t*) const
    {
        if (ALL_OUT || out.ent_id) stream_store(out.ent_id + i, hit.ent);
        if (ALL_OUT || out.prim_id) stream_store(out.prim_id + i, hit.prim);
        if (ALL_OUT || out.t) stream_store(out.t + i, hit.t);
        if (ALL_OUT || out.u) stream_store(out.u + i, hit.u);
        if (ALL_OUT || out.v) stream_store(out.v + i, hit.v);
    }
};

__global__ void normalize_rays_kernel(float4* rays, uint64_t n)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4      r0   = rays[2 * i];
    float4      r1   = rays[2 * i + 1];
    // Eigen's Vector3f::norm(): sqrt of the plain sum of squares; norm < eps -> 1 (Device.cpp:588-592)
    float norm = __fsqrt_rn(r0.w * r0.w + r1.x * r1.x + r1.y * r1.y);
    if (norm < kFltEps) norm = 1.0f;
    r0.w = __fdiv_rn(r0.w, norm);
    r1.x = __fdiv_rn(r1.x, norm);
    r1.y = __fdiv_rn(r1.y, norm);
    rays[2 * i]     = r0;
    rays[2 * i + 1] = r1;
}

template <typename IO>
void launch_list(const DeviceScene& scene, const IO& io, int any_hit, TraceCounters* d_counters, uint32_t* d_work, uint32_t grid, cudaStream_t stream)
{
    if (d_counters) {
        if (any_hit)
            trace_kernel<IO, true, true><<<grid, kBlock, 0, stream>>>(scene, io, d_counters, d_work);
        else
            trace_kernel<IO, false, true><<<grid, kBlock, 0, stream>>>(scene, io, d_counters, d_work);
    } else {
        if (any_hit)
            trace_kernel<IO, true, false><<<grid, kBlock, 0, stream>>>(scene, io, nullptr, d_work);
        else
            trace_kernel<IO, false, false><<<grid, kBlock, 0, stream>>>(scene, io, nullptr, d_work);
    }
}

} // namespace

cudaError_t launch_trace(const DeviceScene& scene, const float4* d_rays, uint64_t n, uint32_t ray_flags, int any_hit,
                         const TraceOutputs& out, TraceCounters* d_counters, uint32_t* d_work, int sm_count,
                         cudaStream_t stream, bool normalize)
{
    if (n == 0) return cudaSuccess;
    // every warp of the grid overshoots the work counter by up to two chunks: keep it well clear of the 32-bit wrap
    if (n > 0xFF000000ull) return cudaErrorInvalidValue;
    // persistent grid: one wave of resident CTAs (a multiple of the SM count), fewer for tiny batches
    const uint64_t want = (n + 31) / 32; // warps that can get at least one ray per lane (the kernel claims 32 rays at a time then)
    const uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)sm_count * kBlocksPerSM, (want + 3) / 4));
    const bool     all  = out.ent_id && out.prim_id && out.t && out.u && out.v;
    if (normalize) {
        if (all)
            launch_list(scene, ListIO<true, true>{ d_rays, (uint32_t)n, ray_flags, out }, any_hit, d_counters, d_work, grid, stream);
        else
            launch_list(scene, ListIO<true, false>{ d_rays, (uint32_t)n, ray_flags, out }, any_hit, d_counters, d_work, grid, stream);
    } else {
        if (all)
            launch_list(scene, ListIO<false, true>{ d_rays, (uint32_t)n, ray_flags, out }, any_hit, d_counters, d_work, grid, stream);
        else
            launch_list(scene, ListIO<false, false>{ d_rays, (uint32_t)n, ray_flags, out }, any_hit, d_counters, d_work, grid, stream);
    }
    return cudaGetLastError();
}

cudaError_t launch_normalize_rays(float4* d_rays, uint64_t n, cudaStream_t stream)
{
    if (n == 0) return cudaSuccess;
    const uint32_t grid = (uint32_t)((n + 255) / 256);
    normalize_rays_kernel<<<grid, 256, 0, stream>>>(d_rays, n);
    return cudaGetLastError();
}

} // namespace igb
