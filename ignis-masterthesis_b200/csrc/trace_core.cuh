// trace_core.cuh -- closest-hit / any-hit ray traversal for sm_100a (kernel template shared by the igtrace
// ray-list path, trace.cu, and the wavefront path tracer, wavefront.cu).
//
// Replaces the reference's K2/K3 kernels: gpu_traverse_primary / gpu_traverse_secondary
// (src/artic/driver/mapping_gpu.art:45-112) -> gpu_traverse_scene / gpu_traverse_prim / gpu_traverse_helper
// (src/artic/traversal/mapping_gpu.art:73-226), which walk a binary Node2 tree with a 64-entry local stack.
// Design here: one ray per lane over a compressed 8-wide BVH (80 B nodes, five 16 B loads), octant-ordered
// child visiting driven by a bit mask, a small (node-group, hit-mask) stack, and ONE unified loop for the
// two-level scene (entity level and triangle level share the node step; only the leaf step differs).
//
// Arithmetic contract shared with oracle/ (bit-exact t,u,v for the same triangle):
//   * ray transform: mat3x4_mul = vec4_dot fma chains (src/artic/core/matrix.art:95-98, vector.art:100)
//   * Moeller-Trumbore: intersect_ray_tri_mt_gen (src/artic/traversal/intersection.art:71-101) with
//     vec3_dot = fma chain (vector.art:99) and vec3_cross un-fused (vector.art:106-109)
//   * this file is compiled with -fmad=false; every fused op is an explicit __fmaf_rn.
// Box tests are conservative (quantised boxes + ulp slack) and carry no parity requirement.
#pragma once
#include "trace.cuh"

#include <cstdio>
#include <cuda_fp16.h>

#include "../../include/ignis_b200.h"

namespace igb {
namespace core {

constexpr float    kFltEps   = 1.1920928955078125e-7f;
constexpr float    kFltMax   = 3.402823466e+38f;
constexpr uint32_t kSentinel = 0xFFFFFFFFu;
constexpr int      kBlock       = 128;
#ifndef IGB_BLOCKS_PER_SM
#define IGB_BLOCKS_PER_SM 8
#endif
constexpr int      kBlocksPerSM = IGB_BLOCKS_PER_SM;
constexpr int      kChunk       = 64; // rays a warp claims per atomic on the global work counter
#ifndef IGB_VOTE_TRI
#define IGB_VOTE_TRI 4
#endif
#ifndef IGB_VOTE_ENT
#define IGB_VOTE_ENT 6 // swept separately from VOTE_TRI in round 2 (profiles/r02_ab_vote_tri_ent_separately.txt): 6 is best
#endif
constexpr int      kVoteTri     = IGB_VOTE_TRI; // lanes that must wait for a triangle phase before it runs
constexpr int      kVoteEnt     = IGB_VOTE_ENT; // same for the entity-leaf phase
#ifndef IGB_VOTE_REFILL
#define IGB_VOTE_REFILL 1
#endif
constexpr int      kVoteRefill  = IGB_VOTE_REFILL; // idle lanes that must wait before the warp fetches new rays
#ifndef IGB_PARK_PRIMS
#define IGB_PARK_PRIMS 0 // 1: park triangle groups, 2: entity-leaf groups too (measured: see DESIGN.md section 7).  Experiment only:
                         // it pushes up to two entries per level, while the host's depth check (device.cu: stack_budget) counts one --
                         // run such a build with IGB200_STACK_BUDGET halved (the debug-checks build traps on an overflow)
#endif
#ifndef IGB_TRI_LOAD2
#define IGB_TRI_LOAD2 0
#endif
#ifndef IGB_ENT_ONE_WAIT
#define IGB_ENT_ONE_WAIT 1
#endif
#ifndef IGB_PARK_SMEM
#define IGB_PARK_SMEM 0 // 1: the world ray of a lane inside an instance is parked in shared memory instead of on the local-memory stack
#endif
#ifndef IGB_TRI_PREFETCH
#define IGB_TRI_PREFETCH 0
#endif
#ifndef IGB_ENT_BOX
#define IGB_ENT_BOX 1 // 0: skip the exact entity-box test of the leaf step (the quantised box of the TLAS node was tested already)
#endif
// Debug build (scripts/build_variant.sh dbg -DIGB_DEBUG_CHECKS=1): device-side bounds checks on every stack push and
// on every node / triangle / leaf index the traversal dereferences; a violation prints once and traps.  This is the
// stand-in for compute-sanitizer memcheck, which is closed on this GPU pool (profiles/r02_sanitizer_closed.txt).
#ifndef IGB_DEBUG_CHECKS
#define IGB_DEBUG_CHECKS 0
#endif
#if IGB_DEBUG_CHECKS
#define IGB_CHECK(cond, what)                                                                              \
    do {                                                                                                   \
        if (!(cond)) {                                                                                     \
            printf("IGB_CHECK failed: %s (%s:%d, block %u thread %u)\n", what, __FILE__, __LINE__, blockIdx.x, threadIdx.x); \
            __trap();                                                                                      \
        }                                                                                                  \
    } while (0)
#else
#define IGB_CHECK(cond, what) ((void)0)
#endif
// IGB_STACK_SMEM = N > 0: the first N stack entries of every thread live in shared memory (N x 8 B x 128 threads per CTA,
// lane-contiguous: conflict-free), deeper ones in local memory.  ncu on C3: the local-memory stack is 41 % of the kernel's
// L1 sectors and its loads miss L1 a third of the time.
#ifndef IGB_STACK_SMEM
#define IGB_STACK_SMEM 0
#endif
#ifndef IGB_TOS_CACHE
#define IGB_TOS_CACHE 0
#endif
#if IGB_STACK_SMEM
#define IGB_STACK_AT(i) (*((i) < IGB_STACK_SMEM ? &s_stack[(i) * kBlock + threadIdx.x] : &stack[(i) - IGB_STACK_SMEM]))
#define IGB_PUSH(v)                                             \
    do {                                                        \
        IGB_CHECK(sp < kStackSize, "traversal stack overflow"); \
        if (sp < IGB_STACK_SMEM)                                \
            s_stack[sp * kBlock + threadIdx.x] = (v);           \
        else                                                    \
            stack[sp - IGB_STACK_SMEM] = (v);                   \
        ++sp;                                                   \
    } while (0)
__device__ __forceinline__ uint2 stack_read(const uint2* s_stack, const uint2* stack, int i, int block, int tid)
{
    return i < IGB_STACK_SMEM ? s_stack[i * block + tid] : stack[i - IGB_STACK_SMEM];
}
#define IGB_TOP() stack_read(s_stack, stack, sp - 1, kBlock, threadIdx.x)
#define IGB_POP() stack_read(s_stack, stack, --sp, kBlock, threadIdx.x)
#elif IGB_TOS_CACHE
// The top of the stack lives in two registers (`top`); memory holds entries 0 .. sp - 2.  A pop hands out the registers and
// starts the load of the next top, which is not needed before the FOLLOWING pop: the local-memory latency (a third of the stack
// loads miss L1) leaves the critical path.  The reference's stack does the same (src/artic/traversal/stack.art:52-123).
#define IGB_PUSH(v)                                             \
    do {                                                        \
        IGB_CHECK(sp < kStackSize, "traversal stack overflow"); \
        if (sp > 0) stack[sp - 1] = top;                        \
        top = (v);                                              \
        ++sp;                                                   \
    } while (0)
#define IGB_TOP() top
#define IGB_POP() stack_pop(stack, top, sp)
__device__ __forceinline__ uint2 stack_pop(const uint2* stack, uint2& top, int& sp)
{
    const uint2 v = top;
    --sp;
    if (sp > 0) top = stack[sp - 1];
    return v;
}
#else
#define IGB_PUSH(v)                                        \
    do {                                                   \
        IGB_CHECK(sp < kStackSize, "traversal stack overflow"); \
        stack[sp++] = (v);                                 \
    } while (0)
#define IGB_TOP() stack[sp - 1]
#define IGB_POP() stack[--sp]
#endif
#ifndef IGB_NODE_REPS
#define IGB_NODE_REPS 1 // node steps per loop trip (lanes that can take another one do; the other phases run once per trip)
#endif
#if IGB_TOS_CACHE && (IGB_STACK_SMEM || IGB_NODE_REPS > 1)
#error "IGB_TOS_CACHE excludes IGB_STACK_SMEM and IGB_NODE_REPS > 1 (the quick pop between node steps reads the stack directly)"
#endif
#ifndef IGB_TRI_PER_TRIP
#define IGB_TRI_PER_TRIP 2
#endif
constexpr int      kTriPerTrip  = IGB_TRI_PER_TRIP; // triangles a lane tests per triangle phase (0 = all it has pending)

struct Vec3 { float x, y, z; };

__device__ __forceinline__ float dot3(Vec3 a, Vec3 b) { return __fmaf_rn(a.x, b.x, __fmaf_rn(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ Vec3  cross3(Vec3 a, Vec3 b)
{
    return Vec3{ a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x };
}
__device__ __forceinline__ float dot4(float ax, float ay, float az, float aw, float bx, float by, float bz, float bw)
{
    return __fmaf_rn(ax, bx, __fmaf_rn(ay, by, __fmaf_rn(az, bz, aw * bw)));
}

// Box-test reciprocal, clamped so that 0 * idir never produces NaN.  One MUFU.RCP (1 ulp): box tests
// carry no parity requirement and their exit distance has kBoxSlack of head room.
__device__ __forceinline__ float box_rcp(float d)
{
    // |1/d| clamped to 2^80 (the same values as clamping |d| to 2^-80 first, one instruction less: MUFU + two FMNMX)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return fminf(fmaxf(r, -1.2089258196146292e24f), 1.2089258196146292e24f);
}
constexpr float kBoxSlack = 1.000002f; // ~17 ulp on the exit distance

__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, 0x0, 0x0000BA98;" : "=r"(r) : "r"(x));
    return r;
}
// 32768 + byte j of w, as a float: the byte lands in mantissa bits 8..15 of 2^15 (one PRMT, no arithmetic).  The
// caller folds the -32768 into the addend of its FMA.
#ifndef IGB_TLAS_SMEM
#define IGB_TLAS_SMEM 0 // number of top-level nodes staged in shared memory (0 = off)
#endif
constexpr float kByteBias = 32768.0f;
// `bias` must hold 0x47000000 in a REGISTER (see byte_bias_reg): PRMT takes either its selector or its third operand as
// an immediate, and with the constant as the immediate nvcc moves the selector into a register before every one of the
// 48 PRMTs of a node step (IMAD.U32 R, RZ, RZ, UR: ~18 % of the node step's instructions in the round-1 SASS).
__device__ __forceinline__ float biased_byte_to_float(uint32_t w, int j, uint32_t bias)
{
    uint32_t r;
    switch (j) { // selector as an immediate: byte 3 = bias.3 (0x47), byte 2 = bias.2 (0), byte 1 = w.j, byte 0 = bias.0 (0)
    case 0: asm("prmt.b32 %0, %1, %2, 0x7604;" : "=r"(r) : "r"(w), "r"(bias)); break;
    case 1: asm("prmt.b32 %0, %1, %2, 0x7614;" : "=r"(r) : "r"(w), "r"(bias)); break;
    case 2: asm("prmt.b32 %0, %1, %2, 0x7624;" : "=r"(r) : "r"(w), "r"(bias)); break;
    default: asm("prmt.b32 %0, %1, %2, 0x7634;" : "=r"(r) : "r"(w), "r"(bias)); break;
    }
    return __uint_as_float(r);
}
__device__ __forceinline__ uint32_t extract_byte(uint32_t w, int j) { return (w >> (8 * j)) & 0xFFu; }

// Ray records and hit columns are touched once per launch; the BVH is what should stay in L1.
//   IGB_STREAM_IO 1: ld.global.cs / st.global.cs (evict-first) for ray input and hit output, 2: L1::no_allocate loads.
// Measured (profiles/r02_ab_stream_io_prefetch.txt): 7200 -> 7260 Mrays/s on C3, 1128 -> 1135 Msamples/s on C4 with 1; on.
#ifndef IGB_STREAM_IO
#define IGB_STREAM_IO 1
#endif
//   IGB_RAY_PREFETCH 1: a warp claims its ray chunks one ahead and prefetches the records into L2, 2: into L1.
#ifndef IGB_RAY_PREFETCH
#define IGB_RAY_PREFETCH 0
#endif
__device__ __forceinline__ void prefetch_line(const void* p)
{
#if IGB_RAY_PREFETCH == 2
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#else
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
#endif
}
__device__ __forceinline__ float4 stream_load(const float4* p, bool read_only)
{
#if IGB_STREAM_IO == 1
    return __ldcs(p);
#elif IGB_STREAM_IO == 2
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
#else
    return read_only ? __ldg(p) : *p;
#endif
}
template <typename T>
__device__ __forceinline__ void stream_store(T* p, T v)
{
#if IGB_STREAM_IO
    __stcs(p, v);
#else
    *p = v;
#endif
}

// Cache policy of the BVH fetches (L1 is the contended resource: nodes, triangles, ray records and the local-memory stack
// share it).  IGB_TRI_NOALLOC 1: triangle records bypass L1 allocation; IGB_NODE_EVICT_LAST 1: node fetches ask L1 to evict them last.
#ifndef IGB_TRI_NOALLOC
#define IGB_TRI_NOALLOC 0
#endif
#ifndef IGB_NODE_EVICT_LAST
#define IGB_NODE_EVICT_LAST 0
#endif
// `stream` is DeviceScene::tri_stream, a launch-uniform flag the host sets when the triangle array is much larger than L1:
// then a record is rarely reused by the same SM and caching it only evicts nodes and stack lines (C3: +1.2 %), whereas a small
// triangle set lives in L1 and must be cached (C4's 80 KB scene loses 41 % without).  IGB_TRI_NOALLOC 1 forces it on.
__device__ __forceinline__ float4 tri_load(const float4* p, bool stream)
{
    if (IGB_TRI_NOALLOC || stream) {
        float4 v;
        asm("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
        return v;
    }
    return __ldg(p);
}
__device__ __forceinline__ uint4 node_load(const uint4* p)
{
#if IGB_NODE_EVICT_LAST
    uint4 v;
    asm("ld.global.nc.L1::evict_last.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}

struct TravRay {
    Vec3     org, dir;    // current space (world or object); dir un-normalised in object space
    Vec3     idir;        // box-test reciprocals
    uint32_t octinv4;     // (7 - octant) replicated to four bytes
};

__device__ __forceinline__ void setup_trav_ray(TravRay& r)
{
    r.idir = Vec3{ box_rcp(r.dir.x), box_rcp(r.dir.y), box_rcp(r.dir.z) };
    const uint32_t oct = (r.dir.x < 0.0f ? 4u : 0u) | (r.dir.y < 0.0f ? 2u : 0u) | (r.dir.z < 0.0f ? 1u : 0u);
    r.octinv4          = (7u - oct) * 0x01010101u;
}

// ---- node-step instruction mix (round 2).  scripts/pipe_probe.cu measures the issue rates on B200: PRMT, FMNMX/FMNMX3,
// LOP3, SHF and FSETP share the ALU pipe, FFMA/FMUL/FADD/HADD2/IMAD the FMA pipe, and a scheduler can start one
// instruction per pipe every other cycle -- so a node step with 16 ALU-pipe and 7 FMA-pipe instructions per child is bound
// by the ALU pipe, not by issue slots.  The switches below move work across:
//   IGB_CVT_HALF   1: every byte -> float conversion goes through a half2 pair {1024 + b_j, 1024 + b_k} built by ONE PRMT and
//                     two HADD2.F32 (FMA pipe) instead of one PRMT per byte; 2: far planes only.
//   IGB_FFMA2      1: the plane distances of two children are computed by one packed fma.rn.f32x2.
//   IGB_SLACK_FOLD 1: the exit-distance slack is folded into the far-plane coefficients (6 FMUL per node instead of 8).
// Measured on B200 (profiles/r02_ab_node_step_mix.txt): C3 7021 -> 7125 (half) / 7130 (f32x2) / 7203 (both) Mrays/s,
// C4 1117 -> 1127 Msamples/s; the slack fold changes nothing.  Both are on.
#ifndef IGB_CVT_HALF
#define IGB_CVT_HALF 1
#endif
#ifndef IGB_FFMA2
#define IGB_FFMA2 1
#endif
#ifndef IGB_SLACK_FOLD
#define IGB_SLACK_FOLD 0
#endif
constexpr float kHalfBias = 1024.0f; // half 0x6400: the byte lands in mantissa bits 0..7

// bias + byte j of w for j = 0..3.  HALF: two PRMTs build the half2 pairs, four HADD2.F32 widen them.
template <bool HALF>
__device__ __forceinline__ void unpack4(uint32_t w, uint32_t bias, uint32_t hbias, float (&q)[4])
{
    if (HALF) {
        uint32_t h01, h23; // bytes {w.0, 0x64, w.1, 0x64} and {w.2, 0x64, w.3, 0x64}; hbias = 0x64006400 in a register
        asm("prmt.b32 %0, %1, %2, 0x7150;" : "=r"(h01) : "r"(w), "r"(hbias));
        asm("prmt.b32 %0, %1, %2, 0x7352;" : "=r"(h23) : "r"(w), "r"(hbias));
        const __half2 a = *reinterpret_cast<const __half2*>(&h01), b = *reinterpret_cast<const __half2*>(&h23);
        q[0] = __low2float(a);
        q[1] = __high2float(a);
        q[2] = __low2float(b);
        q[3] = __high2float(b);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) q[j] = biased_byte_to_float(w, j, bias);
    }
}

// t[j] = q[j] * s + o for four children
__device__ __forceinline__ void planes4(const float (&q)[4], float s, float o, float (&t)[4])
{
#if IGB_FFMA2
    unsigned long long q01, q23, ss, oo, t01, t23;
    asm("mov.b64 %0, {%1, %2};" : "=l"(q01) : "f"(q[0]), "f"(q[1]));
    asm("mov.b64 %0, {%1, %2};" : "=l"(q23) : "f"(q[2]), "f"(q[3]));
    asm("mov.b64 %0, {%1, %1};" : "=l"(ss) : "f"(s));
    asm("mov.b64 %0, {%1, %1};" : "=l"(oo) : "f"(o));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t01) : "l"(q01), "l"(ss), "l"(oo));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(t23) : "l"(q23), "l"(ss), "l"(oo));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(t[0]), "=f"(t[1]) : "l"(t01));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(t[2]), "=f"(t[3]) : "l"(t23));
#else
#pragma unroll
    for (int j = 0; j < 4; ++j) t[j] = __fmaf_rn(q[j], s, o);
#endif
}

// Tests the 8 quantised child boxes of one node; returns the (octant-ordered) hit mask:
// bits 31..24 inner children, bits 23..0 primitives of leaf children.
__device__ __forceinline__ uint32_t intersect_children(const uint4 n0, const uint4 n1, const uint4 n2, const uint4 n3,
                                                       const uint4 n4, const TravRay& r, float tmin, float tmax,
                                                       const uint32_t bias, const uint32_t hbias)
{
    constexpr bool  kHalfNear = IGB_CVT_HALF == 1, kHalfFar = IGB_CVT_HALF >= 1;
    constexpr float kBiasNear = kHalfNear ? kHalfBias : kByteBias, kBiasFar = kHalfFar ? kHalfBias : kByteBias;
    const uint32_t e  = n0.w;
    const float    sx = __uint_as_float((e & 0xFFu) << 23) * r.idir.x;
    const float    sy = __uint_as_float(((e >> 8) & 0xFFu) << 23) * r.idir.y;
    const float    sz = __uint_as_float(((e >> 16) & 0xFFu) << 23) * r.idir.z;
    const float    ox = (__uint_as_float(n0.x) - r.org.x) * r.idir.x;
    const float    oy = (__uint_as_float(n0.y) - r.org.y) * r.idir.y;
    const float    oz = (__uint_as_float(n0.z) - r.org.z) * r.idir.z;
    // plane distance = q * s + o = (bias + q) * s + (o - bias s).  Folding the bias costs one rounding of
    // magnitude <= |s| / 512 (s = one quantisation step in t), so entry planes move 1/256 step out, exit planes in.
    const float dx = fabsf(sx) * 0.00390625f, dy = fabsf(sy) * 0.00390625f, dz = fabsf(sz) * 0.00390625f;
    const float o0x = __fmaf_rn(-kBiasNear, sx, ox) - dx, o0y = __fmaf_rn(-kBiasNear, sy, oy) - dy, o0z = __fmaf_rn(-kBiasNear, sz, oz) - dz;
    float       o1x = __fmaf_rn(-kBiasFar, sx, ox) + dx, o1y = __fmaf_rn(-kBiasFar, sy, oy) + dy, o1z = __fmaf_rn(-kBiasFar, sz, oz) + dz;
#if IGB_SLACK_FOLD
    // min(t1x, t1y, t1z) * slack == min(t1x * slack, ...): the slack moves into the coefficients of the far planes
    const float fx = sx * kBoxSlack, fy = sy * kBoxSlack, fz = sz * kBoxSlack;
    o1x *= kBoxSlack, o1y *= kBoxSlack, o1z *= kBoxSlack;
#else
    const float fx = sx, fy = sy, fz = sz;
#endif
    const bool     nx = r.dir.x < 0.0f, ny = r.dir.y < 0.0f, nz = r.dir.z < 0.0f;

    uint32_t hitmask = 0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const uint32_t meta4       = i ? n1.w : n1.z;
        const uint32_t is_inner4   = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
        const uint32_t bit_index4  = (meta4 ^ (r.octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;

        const uint32_t qlox = i ? n2.y : n2.x, qloy = i ? n2.w : n2.z, qloz = i ? n3.y : n3.x;
        const uint32_t qhix = i ? n3.w : n3.z, qhiy = i ? n4.y : n4.x, qhiz = i ? n4.w : n4.z;
        const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
        const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
        const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
        float q[4], t0x[4], t0y[4], t0z[4], t1x[4], t1y[4], t1z[4];
        unpack4<kHalfNear>(nearx, bias, hbias, q), planes4(q, sx, o0x, t0x);
        unpack4<kHalfNear>(neary, bias, hbias, q), planes4(q, sy, o0y, t0y);
        unpack4<kHalfNear>(nearz, bias, hbias, q), planes4(q, sz, o0z, t0z);
        unpack4<kHalfFar>(farx, bias, hbias, q), planes4(q, fx, o1x, t1x);
        unpack4<kHalfFar>(fary, bias, hbias, q), planes4(q, fy, o1y, t1y);
        unpack4<kHalfFar>(farz, bias, hbias, q), planes4(q, fz, o1z, t1z);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float cmin = fmaxf(fmaxf(t0x[j], t0y[j]), fmaxf(t0z[j], tmin));
            // slack on the exit distance: the box test must never reject what the triangle test accepts
#if IGB_SLACK_FOLD
            const float cmax = fminf(fminf(fminf(t1x[j], t1y[j]), t1z[j]), tmax);
#else
            const float cmax = fminf(fminf(fminf(t1x[j], t1y[j]), t1z[j]) * kBoxSlack, tmax);
#endif
            if (cmin <= cmax) hitmask |= extract_byte(child_bits4, j) << extract_byte(bit_index4, j);
        }
    }
    return hitmask;
}

struct HitRec {
    float   t, u, v;
    int32_t prim, ent;
};

// Deterministic replacement rule.  The reference accepts `t <= tmax` (later equal-t hits replace earlier
// ones, src/artic/traversal/mapping_cpu.art:469), which makes exact ties depend on tree order; here ties
// resolve to the smallest (entity, primitive) so the result is independent of traversal order.
__device__ __forceinline__ bool accepts(const HitRec& h, float t, int32_t ent, int32_t prim)
{
    if (t < h.t) return true;
    if (t > h.t) return false;                 // also rejects NaN
    if (h.prim < 0) return true;               // t == initial tmax
    return (ent < h.ent) || (ent == h.ent && prim < h.prim);
}

// intersect_ray_tri_mt_gen (intersection.art:71-101), no back-face culling, Tri1 record (trimesh.art:146-170).
__device__ __forceinline__ void intersect_triangle(const float4 a, const float4 b, const float4 c4, const TravRay& r,
                                                   float tmin, int32_t ent, HitRec& hit)
{
    const Vec3  v0{ a.x, a.y, a.z }, e1{ b.x, b.y, b.z }, e2{ c4.x, c4.y, c4.z };
    const Vec3  n   = cross3(e1, e2);
    const Vec3  c   = Vec3{ v0.x - r.org.x, v0.y - r.org.y, v0.z - r.org.z };
    const Vec3  rr  = cross3(r.dir, c);
    const float det = dot3(n, r.dir);
    const float inv = __frcp_rn(det); // == IEEE 1.0f / det
    const float u   = dot3(rr, e2) * inv;
    const float v   = dot3(rr, e1) * inv;
    const float w   = 1.0f - u - v;
    if (u >= -kFltEps && v >= -kFltEps && w >= -kFltEps) {
        const float   t    = dot3(c, n) * inv;
        const int32_t prim = __float_as_int(c4.w) & 0x7FFFFFFF;
        if (t >= tmin && accepts(hit, t, ent, prim)) {
            hit.t    = t;
            hit.u    = u > 0.0f ? u : 0.0f;
            hit.v    = v > 0.0f ? v : 0.0f;
            hit.prim = prim;
            hit.ent  = ent;
        }
    }
}

// intersect_sphere (src/artic/shapes/sphere.art:104-132) on the object-space ray.
__device__ __forceinline__ void intersect_sphere(const float4 sph, const TravRay& r, float tmin, int32_t ent, HitRec& hit)
{
    const float pi = 3.14159265359f;
    const Vec3  L{ r.org.x - sph.x, r.org.y - sph.y, r.org.z - sph.z };
    const float S  = -dot3(L, r.dir);
    const float D2 = dot3(r.dir, r.dir);
    const float L2 = dot3(L, L);
    const float R2 = sph.w * sph.w * D2;
    const float M2 = L2 * D2 - S * S;
    if ((S < 0.0f) || (M2 > R2)) return;
    const float Q   = __fsqrt_rn(R2 - M2);
    const float t0_ = __fdiv_rn(S - Q, D2);
    const float t1_ = __fdiv_rn(S + Q, D2);
    const float t0  = t0_ > t1_ ? t1_ : t0_;
    const float t1  = t0_ > t1_ ? t0_ : t1_;
    const float tm  = t0 < tmin ? t1 : t0;
    if (tm >= tmin && tm <= hit.t && accepts(hit, tm, ent, 0)) {
        const float ir    = __frcp_rn(sph.w);
        const Vec3  d{ (L.x + r.dir.x * tm) * ir, (L.y + r.dir.y * tm) * ir, (L.z + r.dir.z * tm) * ir };
        const float theta = acosf(d.z);
        float       phi   = atan2f(-d.x, d.y);
        if (phi < 0.0f) phi += 2.0f * pi;
        hit.t    = tm;
        hit.u    = phi / (2.0f * pi);
        hit.v    = theta / pi;
        hit.prim = 0;
        hit.ent  = ent;
    }
}

// Persistent kernel: every lane owns one ray at a time and pulls the next ray from a warp-local pool
// (refilled in chunks from one global counter) the moment its ray finishes, so lanes never idle while the
// longest ray of the warp completes.  Each trip of the outer loop performs at most one node step, one
// primitive batch and one pop per lane, which keeps lanes in different phases converged on the same code.
//
// IO policy (how rays come in and results go out):
//   uint32_t count() const                                         number of rays (may read device memory)
//   bool     load(i, org, dir, tmin, tmax, flags) const            false = slot holds no ray
//   void     prefetch(first, count, lane) const                    the warp will load rays [first, first + count) soon
//   void     store(i, hit, s_bins)                                 ray i finished (hit.prim < 0: miss, hit.t = tmax)
//   static constexpr bool kBins                                    block-level key histogram in shared memory
//   uint32_t bin_count() const; uint32_t* global_bins() const; void finalize_last_block() const   (only when kBins)
constexpr int kMaxSmemBins = 512;

template <typename IO, bool ANY_HIT, bool STATS>
__global__ void __launch_bounds__(kBlock, kBlocksPerSM)
trace_kernel(const DeviceScene sc, const IO io, TraceCounters* __restrict__ counters, uint32_t* __restrict__ work)
{
    __shared__ uint32_t s_bins[IO::kBins ? kMaxSmemBins : 1];
    if (IO::kBins) {
        for (uint32_t b = threadIdx.x; b < min(io.bin_count(), (uint32_t)kMaxSmemBins); b += blockDim.x) s_bins[b] = 0u;
        __syncthreads();
    }
#if IGB_TLAS_SMEM
    // Experiment (north_star item b): top-level nodes staged in shared memory.  See DESIGN.md section 5.1 for the A/B.
    __shared__ uint4 s_top[IGB_TLAS_SMEM * 5];
    const uint32_t   n_top = min(sc.tlas_nodes, (uint32_t)IGB_TLAS_SMEM);
    for (uint32_t i = threadIdx.x; i < n_top * 5u; i += blockDim.x) s_top[i] = sc.nodes[i];
    __syncthreads();
#endif
#if IGB_PARK_SMEM
    __shared__ uint32_t s_park[10 * kBlock];
#endif
    const uint32_t n = io.count();
    uint32_t       ray_flags = 0;
    const uint32_t lane    = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t       c_nodes = 0, c_tris = 0, c_ents = 0;

    uint32_t pool_next = 0, pool_end = 0; // warp-uniform
#if IGB_RAY_PREFETCH
    constexpr uint32_t kNoChunk = 0xFFFFFFFFu;
    uint32_t           ahead    = kNoChunk; // warp-uniform: base of the chunk claimed one step ahead
#endif
    // warp-uniform.  A warp beyond the number of chunks never claims: in a nearly empty wavefront round (the deep-path tail) the
    // 4736 resident warps would otherwise serialise on one atomic just to learn that there is nothing to do
    // Rays per claim: 64, or 32 when the launch has no more than one ray per resident lane (nearly empty rounds: a warp then
    // traces its rays side by side instead of two after another, which halves the latency of the round).
#ifndef IGB_CHUNK32_LANES
#define IGB_CHUNK32_LANES 1 // claim 32 rays at a time while the launch has at most this many rays per resident lane
#endif
    const uint32_t chunk = (uint64_t)n <= (uint64_t)IGB_CHUNK32_LANES * gridDim.x * blockDim.x ? 32u : (uint32_t)kChunk;
    bool     exhausted = (uint64_t)((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * (uint64_t)chunk >= (uint64_t)n;
    // a lane is active while it owns a ray: ray_id != kNoRay (one register less than a separate flag, which spilled)
    constexpr uint32_t kNoRay = 0xFFFFFFFFu;
    uint32_t ray_id = kNoRay;
#define IGB_ACTIVE (ray_id != kNoRay)
    float    tmin   = 0.0f;
    TravRay  ray;
    HitRec   hit;
#if IGB_STACK_SMEM
    __shared__ uint2 s_stack[IGB_STACK_SMEM * kBlock];
    uint2            stack[kStackSize - IGB_STACK_SMEM];
#else
    uint2    stack[kStackSize];
#endif
#if IGB_TOS_CACHE
    uint2 top = make_uint2(0u, 0u);
#endif
    int      sp      = 0;
    uint2    ngroup  = make_uint2(0u, 0u);
    uint2    tgroup  = make_uint2(0u, 0u);
    bool     in_blas = false;
    int32_t  cur_ent = -1;

    for (;;) {
        const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !IGB_ACTIVE);
        if (idle != 0u && (kVoteRefill <= 1 || exhausted || __popc(idle) >= kVoteRefill)) {
            if (!exhausted) {
                if (pool_next == pool_end) {
#if IGB_RAY_PREFETCH
                    // One chunk ahead: the chunk this warp works on next was claimed (and its ray records prefetched)
                    // when the previous one was opened, so a refilled lane's first use of its ray does not stall the
                    // other 31 lanes of the warp for a DRAM round trip.
                    uint32_t base = ahead, next = 0;
                    if (lane == 0) {
                        if (base == kNoChunk) base = atomicAdd(work, chunk);
                        next = base < n ? atomicAdd(work, chunk) : base;
                    }
                    base  = __shfl_sync(0xFFFFFFFFu, base, 0);
                    ahead = __shfl_sync(0xFFFFFFFFu, next, 0);
                    if (ahead < n) io.prefetch(ahead, min(chunk, n - ahead), lane);
#else
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(work, chunk);
                    base = __shfl_sync(0xFFFFFFFFu, base, 0);
#endif
                    if (base >= n) {
                        exhausted = true;
                    } else {
                        pool_next = base;
                        pool_end  = min(base + chunk, n);
                    }
                }
                const uint32_t avail = pool_end - pool_next;
                const uint32_t rank  = __popc(idle & lt_mask);
                if (!IGB_ACTIVE && rank < avail) {
                    const uint32_t id = pool_next + rank;
                    float tmax = 0.0f;
                    if (io.load(id, ray.org, ray.dir, tmin, tmax, ray_flags)) {
                        ray_id  = id;
                        setup_trav_ray(ray);
                        hit     = HitRec{ tmax, 0.0f, 0.0f, -1, -1 };
                        sp      = 0;
                        ngroup  = make_uint2(0u, 0x80000000u);
                        tgroup  = make_uint2(0u, 0u);
                        in_blas = false;
                    }
                }
                pool_next += min(avail, (uint32_t)__popc(idle));
            } else if (idle == 0xFFFFFFFFu) {
                break;
            }
        }

#if IGB_PARK_PRIMS
        // Experiment: a lane with both inner children and primitives pending parks the primitive group on the stack and
        // keeps descending (it is tested when popped) instead of waiting for the primitive vote.
        if (IGB_ACTIVE && tgroup.y != 0u && ngroup.y > 0x00FFFFFFu && (IGB_PARK_PRIMS == 2 || in_blas)) {
            IGB_PUSH(tgroup);
            tgroup      = make_uint2(0u, 0u);
        }
#endif
        // ---- node phase: lanes with inner children pending and no primitives waiting
#if IGB_NODE_REPS > 1
#pragma unroll 1
        for (int rep = 0; rep < IGB_NODE_REPS; ++rep) {
        if (rep > 0 && IGB_ACTIVE && tgroup.y == 0u && ngroup.y <= 0x00FFFFFFu && sp > 0) {
            // quick pop between node steps: only a plain inner group (sentinels / parked primitive groups wait for the pop phase)
            const uint2 e = IGB_TOP();
            if (e.x != kSentinel && e.y > 0x00FFFFFFu) {
                ngroup = e;
                --sp;
            }
        }
#endif
        if (IGB_ACTIVE && tgroup.y == 0u && ngroup.y > 0x00FFFFFFu) {
            const uint32_t hits            = ngroup.y;
            const uint32_t imask           = ngroup.y & 0xFFu;
            const uint32_t child_bit_index = 31u - (uint32_t)__clz(hits);
            ngroup.y &= ~(1u << child_bit_index);
            if (ngroup.y > 0x00FFFFFFu) IGB_PUSH(ngroup);

            const uint32_t slot_index     = (child_bit_index - 24u) ^ (ray.octinv4 & 0xFFu);
            const uint32_t relative_index = __popc(imask & ~(0xFFFFFFFFu << slot_index));
#if IGB_TLAS_SMEM
            const uint32_t ni             = ngroup.x + relative_index;
            const uint4*   np             = ni < n_top ? s_top + 5 * ni : sc.nodes + 5 * (size_t)ni; // generic loads
            const uint4    n0 = np[0], n1 = np[1], n2 = np[2], n3 = np[3], n4 = np[4];
#else
            IGB_CHECK(ngroup.x + relative_index < sc.node_count, "node index out of range");
            const uint4*   np             = sc.nodes + 5 * (size_t)(ngroup.x + relative_index);
            const uint4    n0 = node_load(np + 0), n1 = node_load(np + 1), n2 = node_load(np + 2), n3 = node_load(np + 3), n4 = node_load(np + 4);
#endif
            if (STATS) c_nodes++;
            const uint32_t hitmask = intersect_children(n0, n1, n2, n3, n4, ray, tmin, hit.t, sc.byte_bias, sc.half_bias);
            ngroup                 = make_uint2(n1.x, (hitmask & 0xFF000000u) | (n0.w >> 24));
            tgroup                 = make_uint2(n1.y, hitmask & 0x00FFFFFFu);
#if IGB_TRI_PREFETCH
            // The triangle phase runs a few trips later (when enough lanes wait for it) and its first use of the record stalls the
            // whole warp; start the fetch of the first pending record now.  IGB_TRI_PREFETCH 1: into L2, 2: into L1.
            if (in_blas && tgroup.y != 0u) {
                const float4* tp = sc.tris + 3 * (size_t)(tgroup.x + (31u - (uint32_t)__clz(tgroup.y)));
#if IGB_TRI_PREFETCH == 2
                asm volatile("prefetch.global.L1 [%0];" ::"l"(tp));
#else
                asm volatile("prefetch.global.L2 [%0];" ::"l"(tp));
#endif
            }
#endif
        }
#if IGB_NODE_REPS > 1
        }
#endif

        // ---- primitive phases.  Triangle tests and entity-leaf steps are rare next to node steps (C3: 0.9 and 0.8 per
        // ray against 5 node steps), so run immediately they execute with ~2 and ~5 live lanes.  A lane with primitives
        // pending waits instead; a phase runs once kVoteTri / kVoteEnt lanes wait for it, or when no lane of the warp
        // can progress otherwise.
        const bool     want_p = IGB_ACTIVE && tgroup.y != 0u;
        const uint32_t tri_v  = __ballot_sync(0xFFFFFFFFu, want_p && in_blas);
        const uint32_t ent_v  = __ballot_sync(0xFFFFFFFFu, want_p && !in_blas);
        const uint32_t free_v = __ballot_sync(0xFFFFFFFFu, IGB_ACTIVE && tgroup.y == 0u);
        if (tri_v != 0u && (__popc(tri_v) >= kVoteTri || free_v == 0u)) {
            if (want_p && in_blas) {
#if IGB_TRI_LOAD2
                // two triangles per phase, both records in flight before the first test: one memory wait instead of two
                const uint32_t i0 = 31u - (uint32_t)__clz(tgroup.y);
                tgroup.y &= ~(1u << i0);
                const bool     two = tgroup.y != 0u;
                const uint32_t i1  = two ? 31u - (uint32_t)__clz(tgroup.y) : i0;
                tgroup.y &= ~(1u << i1);
                IGB_CHECK(tgroup.x + i0 < sc.tri_count, "triangle index out of range");
                const float4* tp0 = sc.tris + 3 * (size_t)(tgroup.x + i0);
                const float4* tp1 = sc.tris + 3 * (size_t)(tgroup.x + i1);
                const float4  a0 = __ldg(tp0 + 0), b0 = __ldg(tp0 + 1), c0 = __ldg(tp0 + 2);
                const float4  a1 = __ldg(tp1 + 0), b1 = __ldg(tp1 + 1), c1 = __ldg(tp1 + 2);
                if (STATS) c_tris += two ? 2u : 1u;
                intersect_triangle(a0, b0, c0, ray, tmin, cur_ent, hit);
                if (two && !(ANY_HIT && hit.prim >= 0)) intersect_triangle(a1, b1, c1, ray, tmin, cur_ent, hit);
#else
                int budget = kTriPerTrip > 0 ? kTriPerTrip : 32;
                while (tgroup.y != 0u && budget-- > 0) {
                    const uint32_t idx = 31u - (uint32_t)__clz(tgroup.y);
                    tgroup.y &= ~(1u << idx);
                    IGB_CHECK(tgroup.x + idx < sc.tri_count, "triangle index out of range");
                    const float4* tp = sc.tris + 3 * (size_t)(tgroup.x + idx);
                    const float4  a = tri_load(tp + 0, sc.tri_stream != 0u), b = tri_load(tp + 1, sc.tri_stream != 0u), c = tri_load(tp + 2, sc.tri_stream != 0u);
                    if (STATS) c_tris++;
                    intersect_triangle(a, b, c, ray, tmin, cur_ent, hit);
                    if (ANY_HIT && hit.prim >= 0) break;
                }
#endif
            }
        }
        if (ent_v != 0u && (__popc(ent_v) >= kVoteEnt || free_v == 0u)) {
            if (want_p && !in_blas) {
                while (tgroup.y != 0u) {
                    const uint32_t idx = 31u - (uint32_t)__clz(tgroup.y);
                    tgroup.y &= ~(1u << idx);
                    // Entity leaf step: gpu_traverse_scene::scene_handle (traversal/mapping_gpu.art:196-222)
                    IGB_CHECK(tgroup.x + idx < sc.leaf_count, "entity leaf index out of range");
                    const float4*  lp = sc.leaves + 6 * (size_t)(tgroup.x + idx);
                    const float4   l0 = __ldg(lp + 0), l1 = __ldg(lp + 1), l5 = __ldg(lp + 5);
#if IGB_ENT_ONE_WAIT
                    const float4   l2 = __ldg(lp + 2), l3 = __ldg(lp + 3), l4 = __ldg(lp + 4); // whole leaf in flight at once
#endif
                    const uint32_t leaf_flags = __float_as_uint(l5.x);
                    // check_ray_visibility (traversal/ray.art:51)
                    if ((ray_flags & 0xFu) != ((ray_flags & leaf_flags) & 0xFu)) continue;
#if IGB_ENT_BOX
                    // intersect_ray_box_single_section on the exact leaf box + `tmin <= hit.distance`
                    const float t0x = (l0.x - ray.org.x) * ray.idir.x, t1x = (l1.x - ray.org.x) * ray.idir.x;
                    const float t0y = (l0.y - ray.org.y) * ray.idir.y, t1y = (l1.y - ray.org.y) * ray.idir.y;
                    const float t0z = (l0.z - ray.org.z) * ray.idir.z, t1z = (l1.z - ray.org.z) * ray.idir.z;
                    const float entry = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), tmin));
                    const float exit_ = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z)) * kBoxSlack;
                    if (!(entry <= fminf(exit_, hit.t))) continue;
#endif

#if !IGB_ENT_ONE_WAIT
                    const float4 l2 = __ldg(lp + 2), l3 = __ldg(lp + 3), l4 = __ldg(lp + 4);
#endif
                    if (STATS) c_ents++;
                    cur_ent              = __float_as_int(l0.w) & 0x7FFFFFFF;
                    // the device leaf carries its shape's BVH root / type / sphere index in user1, user2 (device.cu)
                    ShapeRef shape;
                    shape.root_node = __float_as_uint(l5.z);
                    shape.type      = __float_as_uint(l5.w) & 0xFFu;
                    shape.sphere    = __float_as_uint(l5.w) >> 8;
                    // transform_ray (traversal/ray.art:56-59): matrix columns l2.xyz | l2.w l3.xy | l3.zw l4.x | l4.yzw
                    TravRay local;
                    local.org = Vec3{ dot4(l2.x, l2.w, l3.z, l4.y, ray.org.x, ray.org.y, ray.org.z, 1.0f),
                                      dot4(l2.y, l3.x, l3.w, l4.z, ray.org.x, ray.org.y, ray.org.z, 1.0f),
                                      dot4(l2.z, l3.y, l4.x, l4.w, ray.org.x, ray.org.y, ray.org.z, 1.0f) };
                    local.dir = Vec3{ dot4(l2.x, l2.w, l3.z, l4.y, ray.dir.x, ray.dir.y, ray.dir.z, 0.0f),
                                      dot4(l2.y, l3.x, l3.w, l4.z, ray.dir.x, ray.dir.y, ray.dir.z, 0.0f),
                                      dot4(l2.z, l3.y, l4.x, l4.w, ray.dir.x, ray.dir.y, ray.dir.z, 0.0f) };
                    if (shape.type == IGB200_PRIM_SPHERE) {
                        intersect_sphere(__ldg(sc.spheres + shape.sphere), local, tmin, cur_ent, hit);
                        if (ANY_HIT && hit.prim >= 0) break;
                        continue;
                    }
                    // Enter the shape's BVH: park the entity-level state and the world ray under a sentinel.
                    if (ngroup.y > 0x00FFFFFFu) IGB_PUSH(ngroup);
                    if (tgroup.y != 0u) IGB_PUSH(tgroup);
#if IGB_PARK_SMEM
                    // The scene has ONE level of instancing, so the world ray needs one parking slot per thread, not a
                    // stack: ten words in shared memory (lane-contiguous columns, conflict-free).  The local-memory stack
                    // carried 5 of the ~9 entries a C3 ray pushes for this, and its pops missed L1 a third of the time.
                    s_park[0 * kBlock + threadIdx.x] = __float_as_uint(ray.org.x);
                    s_park[1 * kBlock + threadIdx.x] = __float_as_uint(ray.org.y);
                    s_park[2 * kBlock + threadIdx.x] = __float_as_uint(ray.org.z);
                    s_park[3 * kBlock + threadIdx.x] = __float_as_uint(ray.dir.x);
                    s_park[4 * kBlock + threadIdx.x] = __float_as_uint(ray.dir.y);
                    s_park[5 * kBlock + threadIdx.x] = __float_as_uint(ray.dir.z);
                    s_park[6 * kBlock + threadIdx.x] = __float_as_uint(ray.idir.x);
                    s_park[7 * kBlock + threadIdx.x] = __float_as_uint(ray.idir.y);
                    s_park[8 * kBlock + threadIdx.x] = __float_as_uint(ray.idir.z);
                    s_park[9 * kBlock + threadIdx.x] = ray.octinv4;
#else
                    IGB_PUSH(make_uint2(__float_as_uint(ray.org.x), __float_as_uint(ray.org.y)));
                    IGB_PUSH(make_uint2(__float_as_uint(ray.org.z), __float_as_uint(ray.dir.x)));
                    IGB_PUSH(make_uint2(__float_as_uint(ray.dir.y), __float_as_uint(ray.dir.z)));
                    // the world ray's box-test reciprocals and octant travel too: leaving the shape costs loads, not three MUFUs
                    IGB_PUSH(make_uint2(__float_as_uint(ray.idir.x), __float_as_uint(ray.idir.y)));
                    IGB_PUSH(make_uint2(__float_as_uint(ray.idir.z), ray.octinv4));
#endif
                    IGB_PUSH(make_uint2(kSentinel, 0u));
                    ray         = local;
                    setup_trav_ray(ray);
                    in_blas = true;
                    ngroup  = make_uint2(shape.root_node, 0x80000000u);
                    tgroup  = make_uint2(0u, 0u);
                    break;
                }
            }
        }

        // ---- pop / finish: lanes with nothing pending at the current node
        if (IGB_ACTIVE) {
            bool finished = ANY_HIT && hit.prim >= 0;
            if (!finished && tgroup.y == 0u && ngroup.y <= 0x00FFFFFFu) {
                if (sp == 0) {
                    finished = true;
                } else {
                    uint2 e = IGB_POP();
                    IGB_CHECK(e.x != kSentinel || IGB_PARK_SMEM || sp >= 5, "sentinel without a parked world ray");
                    if (e.x == kSentinel) {
                        // Leave the shape: back to world space.
#if IGB_PARK_SMEM
                        ray.org     = Vec3{ __uint_as_float(s_park[0 * kBlock + threadIdx.x]), __uint_as_float(s_park[1 * kBlock + threadIdx.x]),
                                            __uint_as_float(s_park[2 * kBlock + threadIdx.x]) };
                        ray.dir     = Vec3{ __uint_as_float(s_park[3 * kBlock + threadIdx.x]), __uint_as_float(s_park[4 * kBlock + threadIdx.x]),
                                            __uint_as_float(s_park[5 * kBlock + threadIdx.x]) };
                        ray.idir    = Vec3{ __uint_as_float(s_park[6 * kBlock + threadIdx.x]), __uint_as_float(s_park[7 * kBlock + threadIdx.x]),
                                            __uint_as_float(s_park[8 * kBlock + threadIdx.x]) };
                        ray.octinv4 = s_park[9 * kBlock + threadIdx.x];
#else
                        const uint2 i1 = IGB_POP(), i0 = IGB_POP(), c = IGB_POP(), b = IGB_POP(), a = IGB_POP();
                        ray.org     = Vec3{ __uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(b.x) };
                        ray.dir     = Vec3{ __uint_as_float(b.y), __uint_as_float(c.x), __uint_as_float(c.y) };
                        ray.idir    = Vec3{ __uint_as_float(i0.x), __uint_as_float(i0.y), __uint_as_float(i1.x) };
                        ray.octinv4 = i1.y;
#endif
                        in_blas = false;
                        if (sp == 0)
                            finished = true;
                        else
                            e = IGB_POP();
                    }
                    if (!finished) {
                        if (e.y > 0x00FFFFFFu) {
                            ngroup = e;
                        } else { // a parked primitive group
                            tgroup = e;
                            ngroup = make_uint2(0u, 0u);
                        }
                    }
                }
            }
            if (finished) {
                IGB_CHECK(ray_id < n, "ray id out of range at store");
                io.store(ray_id, hit, s_bins);
                ray_id = kNoRay;
            }
        }
    }

    if (STATS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            c_nodes += __shfl_xor_sync(0xFFFFFFFFu, c_nodes, o);
            c_tris += __shfl_xor_sync(0xFFFFFFFFu, c_tris, o);
            c_ents += __shfl_xor_sync(0xFFFFFFFFu, c_ents, o);
        }
        if (lane == 0) {
            atomicAdd(&counters->nodes, (unsigned long long)c_nodes);
            atomicAdd(&counters->tris, (unsigned long long)c_tris);
            atomicAdd(&counters->ents, (unsigned long long)c_ents);
        }
    }

    // Self-resetting work counter: the last block to retire rewinds it for the next launch.
    __shared__ uint32_t s_last;
    __syncthreads();
    if (IO::kBins) {
        uint32_t* gb = io.global_bins();
        for (uint32_t b = threadIdx.x; b < min(io.bin_count(), (uint32_t)kMaxSmemBins); b += blockDim.x)
            if (s_bins[b]) atomicAdd(gb + b, s_bins[b]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        __threadfence();
        const uint32_t done = atomicAdd(work + 1, 1u);
        s_last              = done == gridDim.x - 1 ? 1u : 0u;
        if (s_last) {
            work[0] = 0u;
            work[1] = 0u;
            __threadfence();
        }
    }
    if (IO::kBins) {
        // the last block to retire sees every block's histogram: it turns it into scatter offsets
        __syncthreads();
        if (s_last) io.finalize_last_block();
    }
}

#undef IGB_ACTIVE
} // namespace core
} // namespace igb
