// device.cu -- the C ABI of include/ignis_b200.h: device lifetime, scene upload, igtrace-style ray tracing.
// Host-side counterpart of the reference's IG::Device / Interface (src/runtime/device/Device.cpp:136-1612).
#include "../../include/ignis_b200.h"
#include "bvh8.h"
#include "trace.cuh"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

thread_local std::string g_last_error;

int fail(int code, const char* fmt, ...)
{
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

#define IGB_CUDA(call)                                                                                          \
    do {                                                                                                        \
        cudaError_t err__ = (call);                                                                             \
        if (err__ != cudaSuccess)                                                                               \
            return fail(err__ == cudaErrorMemoryAllocation ? IGB200_E_OOM : IGB200_E_CUDA, "%s failed: %s", #call, \
                        cudaGetErrorString(err__));                                                             \
    } while (0)

template <typename T>
struct DeviceBuffer {
    T*     ptr   = nullptr;
    size_t count = 0;
    ~DeviceBuffer() { release(); }
    void release()
    {
        if (ptr) cudaFree(ptr);
        ptr   = nullptr;
        count = 0;
    }
    cudaError_t ensure(size_t n)
    {
        if (n <= count) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc(&ptr, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) count = n;
        return e;
    }
    cudaError_t upload(const T* src, size_t n, cudaStream_t s)
    {
        cudaError_t e = ensure(n);
        if (e != cudaSuccess || n == 0) return e;
        return cudaMemcpyAsync(ptr, src, n * sizeof(T), cudaMemcpyHostToDevice, s);
    }
};

} // namespace

struct igb200_device {
    int          ordinal = 0;
    bool         stats   = false;
    cudaStream_t stream  = nullptr;
    cudaEvent_t  ev0 = nullptr, ev1 = nullptr;
    int          sm_count = 0;
    static constexpr int      kPipe      = 3;         // streams of the host-buffer pipeline
    static constexpr uint64_t kChunkRays = 1u << 20;  // rays per pipeline chunk (32 MiB up, 20 MiB down)
    cudaStream_t             pipe[kPipe] = {};
    std::vector<cudaEvent_t> chunk_events;
    DeviceBuffer<uint32_t>   d_work; // persistent-kernel work counters (self-resetting): pair k for pipe[k], last pair for `stream`

    // scene
    bool                          has_scene = false;
    DeviceBuffer<uint4>           nodes;
    DeviceBuffer<float4>          tris;
    DeviceBuffer<float4>          leaves;
    DeviceBuffer<igb::ShapeRef>   shapes;
    DeviceBuffer<float4>          spheres;
    igb::DeviceScene              view{};
    uint64_t                      node_bytes = 0, tri_bytes = 0, leaf_bytes = 0;
    uint32_t                      max_depth = 0;

    // per-call staging (reused)
    DeviceBuffer<float4>             d_rays;
    DeviceBuffer<int32_t>            d_ent, d_prim;
    DeviceBuffer<float>              d_t, d_u, d_v;
    DeviceBuffer<igb::TraceCounters> d_counters;
    igb200_trace_stats               last_stats{};
};

extern "C" {

int igb200_abi_version(void) { return IGB200_ABI_VERSION; }

const char* igb200_last_error(void) { return g_last_error.c_str(); }

int igb200_create(const igb200_setup* setup, igb200_device** out)
{
    if (!out) return fail(IGB200_E_INVALID, "igb200_create: out is NULL");
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(IGB200_E_NODEVICE, "igb200_create: no CUDA device (this backend has no CPU fallback)");
    }
    const int ordinal = setup ? setup->device : 0;
    if (ordinal < 0 || ordinal >= count) return fail(IGB200_E_NODEVICE, "igb200_create: device %d of %d", ordinal, count);
    cudaDeviceProp prop;
    IGB_CUDA(cudaGetDeviceProperties(&prop, ordinal));
    if (prop.major != 10)
        return fail(IGB200_E_NODEVICE, "igb200_create: device %d is sm_%d%d; kernels are built for sm_100a only", ordinal,
                    prop.major, prop.minor);
    IGB_CUDA(cudaSetDevice(ordinal));
    igb200_device* dev = new igb200_device();
    dev->ordinal       = ordinal;
    dev->stats         = setup && setup->acquire_stats;
    dev->sm_count      = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&dev->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&dev->ev0) != cudaSuccess || cudaEventCreate(&dev->ev1) != cudaSuccess ||
        cudaStreamCreateWithFlags(&dev->pipe[0], cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&dev->pipe[1], cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&dev->pipe[2], cudaStreamNonBlocking) != cudaSuccess) {
        delete dev;
        return fail(IGB200_E_CUDA, "igb200_create: stream/event creation failed");
    }
    if (dev->d_work.ensure(2 * (igb200_device::kPipe + 1)) != cudaSuccess ||
        cudaMemset(dev->d_work.ptr, 0, 2 * (igb200_device::kPipe + 1) * sizeof(uint32_t)) != cudaSuccess) {
        delete dev;
        return fail(IGB200_E_OOM, "igb200_create: work counter allocation failed");
    }
    *out = dev;
    return IGB200_OK;
}

void igb200_destroy(igb200_device* dev)
{
    if (!dev) return;
    cudaSetDevice(dev->ordinal);
    cudaDeviceSynchronize();
    if (dev->ev0) cudaEventDestroy(dev->ev0);
    if (dev->ev1) cudaEventDestroy(dev->ev1);
    if (dev->stream) cudaStreamDestroy(dev->stream);
    for (cudaStream_t p : dev->pipe)
        if (p) cudaStreamDestroy(p);
    for (cudaEvent_t e : dev->chunk_events)
        if (e) cudaEventDestroy(e);
    delete dev; // DeviceBuffer destructors free the allocations (Device::releaseAll, Device.cpp:948-952)
}

int igb200_assign_scene(igb200_device* dev, const igb200_scene* sc)
{
    using namespace igb;
    if (!dev || !sc) return fail(IGB200_E_INVALID, "igb200_assign_scene: NULL argument");
    if (sc->shape_count && (!sc->shape_lookups || !sc->shape_data))
        return fail(IGB200_E_INVALID, "igb200_assign_scene: shape table pointers are NULL");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    dev->has_scene = false;

    // ---- entity leaves (both primitive types share one top-level BVH8) ----
    std::vector<igb200_entity_leaf> in_leaves;
    for (int k = 0; k < 2; ++k) {
        if (sc->leaf_count[k] && !sc->leaves[k]) return fail(IGB200_E_INVALID, "leaves[%d] is NULL", k);
        for (uint64_t i = 0; i < sc->leaf_count[k]; ++i) {
            igb200_entity_leaf l = sc->leaves[k][i];
            l.entity_id &= 0x7FFFFFFF; // drop the reference's sentinel bit (SceneBVHAdapter.h:103)
            if (l.shape_id < 0 || (uint64_t)l.shape_id >= sc->shape_count)
                return fail(IGB200_E_INVALID, "entity %d references shape %d of %llu", l.entity_id, l.shape_id,
                            (unsigned long long)sc->shape_count);
            if (sc->shape_lookups[l.shape_id].type_id != (uint32_t)k)
                return fail(IGB200_E_INVALID, "entity %d: shape %d is not of primitive type %d", l.entity_id, l.shape_id, k);
            in_leaves.push_back(l);
        }
    }

    std::vector<Node8>  all_nodes;
    std::vector<float4> all_tris;
    std::vector<float4> all_spheres;
    std::vector<ShapeRef> shape_refs(sc->shape_count);

    // TLAS first so that its root is node 0.
    Bvh8 tlas;
    {
        std::vector<float> boxes(in_leaves.size() * 6);
        for (size_t i = 0; i < in_leaves.size(); ++i)
            for (int k = 0; k < 3; ++k) {
                boxes[6 * i + k]     = in_leaves[i].bbox_min[k];
                boxes[6 * i + 3 + k] = in_leaves[i].bbox_max[k];
            }
        build_bvh8(boxes.data(), (uint32_t)in_leaves.size(), 1, tlas);
        all_nodes = tlas.nodes;
    }
    std::vector<igb200_entity_leaf> ordered(in_leaves.size());
    for (size_t i = 0; i < ordered.size(); ++i) ordered[i] = in_leaves[tlas.prim_indices[i]];

    // ---- shapes ----
    uint32_t blas_depth = 0;
    for (uint64_t s = 0; s < sc->shape_count; ++s) {
        const igb200_lookup_entry& e = sc->shape_lookups[s];
        if (e.offset > sc->shape_bytes) return fail(IGB200_E_INVALID, "shape %llu: offset out of range", (unsigned long long)s);
        const uint8_t* data  = sc->shape_data + e.offset;
        const uint64_t avail = sc->shape_bytes - e.offset;
        ShapeRef       ref{ 0, e.type_id, 0, 0 };
        if (e.type_id == IGB200_PRIM_SPHERE) {
            if (avail < 16) return fail(IGB200_E_INVALID, "shape %llu: truncated sphere", (unsigned long long)s);
            float4 sp;
            std::memcpy(&sp, data, 16);
            ref.sphere = (uint32_t)all_spheres.size();
            all_spheres.push_back(sp);
        } else if (e.type_id == IGB200_PRIM_TRIMESH) {
            // mesh table layout: src/runtime/shape/TriMeshProvider.cpp:577-598
            if (avail < 48) return fail(IGB200_E_INVALID, "shape %llu: truncated mesh header", (unsigned long long)s);
            uint32_t hdr[4];
            std::memcpy(hdr, data, 16);
            const uint64_t nf = hdr[0], nv = hdr[1], nn = hdr[2];
            if (avail < 4 * (12 + nv * 4 + nn * 4 + nf * 4))
                return fail(IGB200_E_INVALID, "shape %llu: truncated mesh data", (unsigned long long)s);
            const float*    verts = reinterpret_cast<const float*>(data) + 12;
            const uint32_t* idx   = reinterpret_cast<const uint32_t*>(data) + 12 + nv * 4 + nn * 4;
            std::vector<float> boxes(nf * 6);
            for (uint64_t i = 0; i < nf; ++i) {
                const uint32_t i0 = idx[4 * i], i1 = idx[4 * i + 1], i2 = idx[4 * i + 2];
                if (i0 >= nv || i1 >= nv || i2 >= nv)
                    return fail(IGB200_E_INVALID, "shape %llu: face %llu indexes past %llu vertices", (unsigned long long)s,
                                (unsigned long long)i, (unsigned long long)nv);
                for (int k = 0; k < 3; ++k) {
                    const float a = verts[4 * i0 + k], b = verts[4 * i1 + k], c = verts[4 * i2 + k];
                    boxes[6 * i + k]     = std::min(a, std::min(b, c));
                    boxes[6 * i + 3 + k] = std::max(a, std::max(b, c));
                }
            }
            Bvh8 blas;
            build_bvh8(boxes.data(), (uint32_t)nf, 3, blas);
            blas_depth = std::max(blas_depth, blas.max_depth);
            const uint32_t node_base = (uint32_t)all_nodes.size();
            const uint32_t tri_base  = (uint32_t)(all_tris.size() / 3);
            ref.root_node            = node_base;
            for (Node8 n : blas.nodes) {
                n.child_base += node_base;
                n.prim_base += tri_base;
                all_nodes.push_back(n);
            }
            for (uint32_t p : blas.prim_indices) {
                const float* v0 = verts + 4 * idx[4 * (size_t)p];
                const float* v1 = verts + 4 * idx[4 * (size_t)p + 1];
                const float* v2 = verts + 4 * idx[4 * (size_t)p + 2];
                // Tri1 record: v0, e1 = v0 - v1, e2 = v2 - v0, prim id (TriBVHAdapter.h:150-157; madmann91 Triangle)
                all_tris.push_back(make_float4(v0[0], v0[1], v0[2], 0.0f));
                all_tris.push_back(make_float4(v0[0] - v1[0], v0[1] - v1[1], v0[2] - v1[2], 0.0f));
                float4 c = make_float4(v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2], 0.0f);
                const int32_t pid = (int32_t)p;
                std::memcpy(&c.w, &pid, 4);
                all_tris.push_back(c);
            }
        } else {
            return fail(IGB200_E_UNSUPPORTED, "shape %llu: unknown primitive type %u", (unsigned long long)s, e.type_id);
        }
        shape_refs[s] = ref;
    }
    if (tlas.max_depth + blas_depth + 7 > (uint32_t)kStackSize)
        return fail(IGB200_E_UNSUPPORTED, "BVH too deep for the traversal stack (%u + %u levels)", tlas.max_depth, blas_depth);

    // ---- upload ----
    static_assert(sizeof(Node8) == 5 * sizeof(uint4), "node size");
    IGB_CUDA(dev->nodes.upload(reinterpret_cast<const uint4*>(all_nodes.data()), all_nodes.size() * 5, dev->stream));
    IGB_CUDA(dev->tris.upload(all_tris.data(), all_tris.size(), dev->stream));
    IGB_CUDA(dev->leaves.upload(reinterpret_cast<const float4*>(ordered.data()), ordered.size() * 6, dev->stream));
    IGB_CUDA(dev->shapes.upload(shape_refs.data(), shape_refs.size(), dev->stream));
    IGB_CUDA(dev->spheres.upload(all_spheres.data(), all_spheres.size(), dev->stream));
    IGB_CUDA(cudaStreamSynchronize(dev->stream));

    dev->view.nodes      = dev->nodes.ptr;
    dev->view.tris       = dev->tris.ptr;
    dev->view.leaves     = dev->leaves.ptr;
    dev->view.shapes     = dev->shapes.ptr;
    dev->view.spheres    = dev->spheres.ptr;
    dev->view.tlas_nodes = (uint32_t)tlas.nodes.size();
    dev->view.leaf_count = (uint32_t)ordered.size();
    dev->node_bytes      = all_nodes.size() * sizeof(Node8);
    dev->tri_bytes       = all_tris.size() * sizeof(float4);
    dev->leaf_bytes      = ordered.size() * sizeof(igb200_entity_leaf);
    dev->max_depth       = tlas.max_depth + blas_depth;
    dev->has_scene       = true;
    return IGB200_OK;
}

static int trace_on_stream(igb200_device* dev, const igb200_ray* d_rays, uint64_t n, uint32_t ray_flags, int any_hit,
                           const igb::TraceOutputs& out, cudaStream_t stream, bool timed)
{
    igb::TraceCounters* counters = nullptr;
    if (dev->stats) {
        IGB_CUDA(dev->d_counters.ensure(1));
        IGB_CUDA(cudaMemsetAsync(dev->d_counters.ptr, 0, sizeof(igb::TraceCounters), stream));
        counters = dev->d_counters.ptr;
    }
    if (timed) IGB_CUDA(cudaEventRecord(dev->ev0, stream));
    IGB_CUDA(igb::launch_trace(dev->view, reinterpret_cast<const float4*>(d_rays), n, ray_flags, any_hit, out, counters,
                               dev->d_work.ptr + 2 * igb200_device::kPipe, dev->sm_count, stream));
    if (timed) IGB_CUDA(cudaEventRecord(dev->ev1, stream));
    return IGB200_OK;
}

int igb200_trace_rays_device(igb200_device* dev, const igb200_ray* d_rays, uint64_t n, uint32_t ray_flags, int any_hit,
                             int32_t* d_ent_id, int32_t* d_prim_id, float* d_t, float* d_u, float* d_v, void* stream)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_trace_rays_device: NULL device");
    if (!dev->has_scene) return fail(IGB200_E_STATE, "igb200_trace_rays_device: no scene assigned");
    if (n && !d_rays) return fail(IGB200_E_INVALID, "igb200_trace_rays_device: rays is NULL");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    const igb::TraceOutputs out{ d_ent_id, d_prim_id, d_t, d_u, d_v };
    return trace_on_stream(dev, d_rays, n, ray_flags, any_hit, out, static_cast<cudaStream_t>(stream), false);
}

int igb200_trace_rays(igb200_device* dev, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                      int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_trace_rays: NULL device");
    if (!dev->has_scene) return fail(IGB200_E_STATE, "igb200_trace_rays: no scene assigned");
    if (n && !rays) return fail(IGB200_E_INVALID, "igb200_trace_rays: rays is NULL");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    IGB_CUDA(dev->d_rays.ensure(2 * n));
    if (ent_id) IGB_CUDA(dev->d_ent.ensure(n));
    if (prim_id) IGB_CUDA(dev->d_prim.ensure(n));
    if (t) IGB_CUDA(dev->d_t.ensure(n));
    if (u) IGB_CUDA(dev->d_u.ensure(n));
    if (v) IGB_CUDA(dev->d_v.ensure(n));

    igb::TraceCounters* counters = nullptr;
    if (dev->stats) {
        IGB_CUDA(dev->d_counters.ensure(1));
        IGB_CUDA(cudaMemsetAsync(dev->d_counters.ptr, 0, sizeof(igb::TraceCounters), dev->stream));
        IGB_CUDA(cudaStreamSynchronize(dev->stream));
        counters = dev->d_counters.ptr;
    }

    // Chunked software pipeline over kPipe streams: upload of chunk k+1, traversal of chunk k and download of
    // chunk k-1 overlap (PCIe is full duplex), instead of the reference's upload-all / trace / download-all
    // (Device.cpp:573-611, 1330-1363).  Chunks use disjoint slices of the staging buffers.
    const uint64_t chunk   = igb200_device::kChunkRays;
    const uint64_t nchunks = (n + chunk - 1) / chunk;
    if (dev->chunk_events.size() < 2 * nchunks) {
        const size_t old = dev->chunk_events.size();
        dev->chunk_events.resize(2 * nchunks, nullptr);
        for (size_t i = old; i < dev->chunk_events.size(); ++i) IGB_CUDA(cudaEventCreate(&dev->chunk_events[i]));
    }
    for (uint64_t c = 0; c < nchunks; ++c) {
        const int      slot = (int)(c % igb200_device::kPipe);
        cudaStream_t   s    = dev->pipe[slot];
        const uint64_t off = c * chunk, m = std::min(chunk, n - off);
        float4*        dr = dev->d_rays.ptr + 2 * off;
        IGB_CUDA(cudaMemcpyAsync(dr, rays + off, m * sizeof(igb200_ray), cudaMemcpyHostToDevice, s));
        IGB_CUDA(igb::launch_normalize_rays(dr, m, s));
        const igb::TraceOutputs out{ ent_id ? dev->d_ent.ptr + off : nullptr, prim_id ? dev->d_prim.ptr + off : nullptr,
                                     t ? dev->d_t.ptr + off : nullptr, u ? dev->d_u.ptr + off : nullptr,
                                     v ? dev->d_v.ptr + off : nullptr };
        IGB_CUDA(cudaEventRecord(dev->chunk_events[2 * c], s));
        IGB_CUDA(igb::launch_trace(dev->view, dr, m, ray_flags, any_hit, out, counters, dev->d_work.ptr + 2 * slot,
                                   dev->sm_count, s));
        IGB_CUDA(cudaEventRecord(dev->chunk_events[2 * c + 1], s));
        if (ent_id) IGB_CUDA(cudaMemcpyAsync(ent_id + off, out.ent_id, m * 4, cudaMemcpyDeviceToHost, s));
        if (prim_id) IGB_CUDA(cudaMemcpyAsync(prim_id + off, out.prim_id, m * 4, cudaMemcpyDeviceToHost, s));
        if (t) IGB_CUDA(cudaMemcpyAsync(t + off, out.t, m * 4, cudaMemcpyDeviceToHost, s));
        if (u) IGB_CUDA(cudaMemcpyAsync(u + off, out.u, m * 4, cudaMemcpyDeviceToHost, s));
        if (v) IGB_CUDA(cudaMemcpyAsync(v + off, out.v, m * 4, cudaMemcpyDeviceToHost, s));
    }
    for (int k = 0; k < igb200_device::kPipe; ++k) IGB_CUDA(cudaStreamSynchronize(dev->pipe[k]));

    dev->last_stats      = igb200_trace_stats{};
    dev->last_stats.rays = n;
    for (uint64_t c = 0; c < nchunks; ++c) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, dev->chunk_events[2 * c], dev->chunk_events[2 * c + 1]) == cudaSuccess)
            dev->last_stats.kernel_ms += ms;
    }
    if (dev->stats) {
        igb::TraceCounters c{};
        IGB_CUDA(cudaMemcpy(&c, dev->d_counters.ptr, sizeof(c), cudaMemcpyDeviceToHost));
        dev->last_stats.nodes_visited    = c.nodes;
        dev->last_stats.tris_tested      = c.tris;
        dev->last_stats.entities_entered = c.ents;
    }
    return IGB200_OK;
}

int igb200_get_trace_stats(igb200_device* dev, igb200_trace_stats* out)
{
    if (!dev || !out) return fail(IGB200_E_INVALID, "igb200_get_trace_stats: NULL argument");
    if (dev->stats && dev->d_counters.ptr && dev->has_scene) {
        // picks up counters of a preceding device-resident trace as well
        cudaSetDevice(dev->ordinal);
        igb::TraceCounters c{};
        IGB_CUDA(cudaDeviceSynchronize());
        IGB_CUDA(cudaMemcpy(&c, dev->d_counters.ptr, sizeof(c), cudaMemcpyDeviceToHost));
        dev->last_stats.nodes_visited    = c.nodes;
        dev->last_stats.tris_tested      = c.tris;
        dev->last_stats.entities_entered = c.ents;
    }
    *out = dev->last_stats;
    return IGB200_OK;
}

int igb200_get_bvh_info(igb200_device* dev, uint64_t* node_bytes, uint64_t* tri_bytes, uint64_t* leaf_bytes,
                        uint32_t* max_depth)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_get_bvh_info: NULL device");
    if (!dev->has_scene) return fail(IGB200_E_STATE, "igb200_get_bvh_info: no scene assigned");
    if (node_bytes) *node_bytes = dev->node_bytes;
    if (tri_bytes) *tri_bytes = dev->tri_bytes;
    if (leaf_bytes) *leaf_bytes = dev->leaf_bytes;
    if (max_depth) *max_depth = dev->max_depth;
    return IGB200_OK;
}

} // extern "C"
