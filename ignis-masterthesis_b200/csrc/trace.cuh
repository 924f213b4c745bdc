// trace.cuh -- device-side scene view + launcher declarations of the sm_100a traversal kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace igb {

// Per-shape record (16 B): where the shape's BVH8 lives in the unified arrays.
struct ShapeRef {
    uint32_t root_node; // index into DeviceScene::nodes (child_base / prim_base already relocated)
    uint32_t type;      // IGB200_PRIM_*
    uint32_t sphere;    // index into DeviceScene::spheres for type == sphere
    uint32_t _pad;
};

// Everything the traversal kernels read. All pointers are device pointers, 16-byte aligned.
struct DeviceScene {
    const uint4*    nodes;   // 5 x uint4 per Node8; TLAS nodes first, then one block per trimesh shape
    const float4*   tris;    // 3 x float4 per triangle (Tri1 layout), all shapes, leaf-contiguous order
    const float4*   leaves;  // 6 x float4 per EntityLeaf1, in TLAS leaf order
    const ShapeRef* shapes;  // by shape id
    const float4*   spheres; // origin xyz, radius
    uint32_t        tlas_nodes; // number of TLAS nodes (staged in shared memory when small)
    uint32_t        leaf_count;
    // 0x47000000, set by the host.  It is a kernel PARAMETER on purpose: ptxas folds any in-code constant into the
    // immediate slot of PRMT and then has to move the byte selector into a register before each of the 48 PRMTs of a
    // node step; a constant-bank operand leaves the immediate slot to the selector (trace_core.cuh, biased_byte_to_float).
    uint32_t        byte_bias;
    uint32_t        half_bias; // 0x64006400 (half2 {1024, 1024}), a parameter for the same reason
    uint32_t        node_count, tri_count; // array sizes (debug bounds checks, IGB_DEBUG_CHECKS)
    uint32_t        tri_stream;            // != 0: triangle records bypass L1 allocation (set when the array dwarfs L1)
};

struct TraceOutputs {
    int32_t* ent_id;
    int32_t* prim_id;
    float*   t;
    float*   u;
    float*   v;
};

struct TraceCounters { // device-side, 3 x u64
    unsigned long long nodes, tris, ents;
};

#ifndef IGB_STACK
#define IGB_STACK 46
#endif
constexpr int kStackSize = IGB_STACK; // traversal stack entries per ray (uint2 each)

// Launches the closest-hit (any_hit = 0) or any-hit traversal of `n` rays on `stream`.
// counters may be nullptr (then the non-counting kernel variant is used).  d_work: two zero-initialised
// uint32 (work counter, retired-block counter) owned by the device handle; the kernel rewinds them itself,
// so only one trace launch per handle may be in flight at a time (device.cu orders launches on different streams by an event).
cudaError_t launch_trace(const DeviceScene& scene, const float4* d_rays, uint64_t n, uint32_t ray_flags, int any_hit,
                         const TraceOutputs& out, TraceCounters* d_counters, uint32_t* d_work, int sm_count,
                         cudaStream_t stream, bool normalize = false);

// In-place normalisation of ray directions (Interface::loadRayList, src/runtime/device/Device.cpp:585-607).
cudaError_t launch_normalize_rays(float4* d_rays, uint64_t n, cudaStream_t stream);

} // namespace igb
