// device.cu -- the C ABI of include/ignis_b200.h: device lifetime, scene upload, igtrace-style ray tracing.
// Host-side counterpart of the reference's IG::Device / Interface (src/runtime/device/Device.cpp:136-1612).
#include "../../include/ignis_b200.h"
#include "bvh8.h"
#include "trace.cuh"
#include "wavefront.cuh"
#include "shade_scene.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
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
    // igb200_trace_rays_device on caller streams: the kernel's work counters belong to the handle, so two launches must not
    // overlap.  ev_trace marks the end of the last one; a launch on ANOTHER stream waits for it (ordering on the device, no host wait).
    cudaEvent_t  ev_trace = nullptr;
    cudaStream_t last_trace_stream = nullptr;
    bool         trace_pending = false;
    int          sm_count = 0;
    static constexpr int      kPipe      = 3;         // streams of the host-buffer pipeline
    uint64_t                 chunk_rays = 1u << 20;   // rays per bulk chunk of the host-buffer pipeline (32 MiB up, 20 MiB down)
    bool                     pipe_align = true;       // downloads start at upload boundaries (IGB200_PIPE_ALIGN=0: right behind the traversal)
    int                      pipe_sms   = 0;          // SMs the pipeline's traversal launches occupy (0 = all; IGB200_PIPE_SMS)
    uint64_t                 chunk_min  = 1u << 18;   // first / last chunk of the ramp (IGB200_TRACE_CHUNK / _MIN override both)
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

    // ---- shading tables + wavefront state (igb200_assign_shading / igb200_render) ----
    DeviceBuffer<float>           d_entities;
    DeviceBuffer<float>           d_shape_data;
    DeviceBuffer<uint2>           d_shape_info;
    DeviceBuffer<int32_t>         d_ent_mat;
    DeviceBuffer<igb200_bsdf>     d_bsdfs;
    DeviceBuffer<igb200_material> d_materials;
    DeviceBuffer<igb200_light>    d_inf_lights, d_fin_lights;
    DeviceBuffer<igb200_texture>  d_textures;
    DeviceBuffer<uint8_t>         d_blob;
    igb::shade::ShadeScene        shade{};
    bool                          has_shading = false;
    uint64_t                      entity_count = 0;
    uint32_t                      scene_material_count = 0;
    float                         scene_bbox[6] = {};

    uint32_t                      wave_capacity = 0;
    int                           wave_sort     = 1; // 0 off, 1 global index sort, 2 tile-local index sort
    DeviceBuffer<float4>          w_rec[2], w_sh_f4[3];
    DeviceBuffer<uint16_t>        w_flags[2];
    DeviceBuffer<uint32_t>        w_keys, w_sorted;
    DeviceBuffer<igb::WaveControl> w_ctl;
    igb::WaveBuffers              wave{};
    uint32_t*                     h_done = nullptr; // two mapped pinned host flags: [0] wavefront drained, [1] all ids of the call generated
    bool                          wave_pending = false; // igb200_render_pipelined left paths in flight (survivors in stream 0)
    igb::WaveSettings             pending_set{};        // settings of the call that left them
    cudaEvent_t                   round_ev[4] = {};
    cudaGraphExec_t               round_graph = nullptr; // one round pair; invalidated by assign_scene / assign_shading
    cudaGraphExec_t               single_graph[2] = { nullptr, nullptr }; // one round of either parity (pipelined steps)
    int                           wave_parity = 0;       // stream buffer the next round reads (non-zero only between pipelined steps)
    DeviceBuffer<float4>          d_list_rays;

    DeviceBuffer<float>           d_film;
    float*                        h_film = nullptr; // pinned host copy
    uint32_t                      film_w = 0, film_h = 0;
    uint64_t                      iterations = 0;
    bool                          film_mapped = false; // host copy is current
    igb200_render_stats           render_stats{};
    igb200_section_stats          section{};   // accumulated since create / igb200_reset_section_stats (stats mode only)
    cudaEvent_t                   sec_ev[igb::kWaveLaunchesPerRound + 1] = {};

    // Named AOV images next to the film (Interface::aovs, Device.cpp:95-100,1300-1307): allocated at film size the first
    // time a technique that enables them renders, each with its own iteration count (markAOVAsUsed / present,
    // Device.cpp:1389-1403,1468-1471).
    struct Aov {
        DeviceBuffer<float> d;
        float*              h = nullptr; // pinned host copy, lazily filled
        uint64_t            iterations = 0;
        bool                mapped = false;
    };
    static constexpr int kAovCount = 5;
    Aov                  aov[kAovCount];
    DeviceBuffer<uint32_t> d_tonemap;
};
static const char* const kAovNames[igb200_device::kAovCount] = { "Direct Weights", "NEE Weights", "Normals", "Albedo", "Depth" };

extern "C" {

static int flush_wave(igb200_device* dev); // drains the paths left in flight by igb200_render_pipelined
static void drop_round_graphs(igb200_device* dev)
{
    if (dev->round_graph) cudaGraphExecDestroy(dev->round_graph);
    dev->round_graph = nullptr;
    for (cudaGraphExec_t& g : dev->single_graph) {
        if (g) cudaGraphExecDestroy(g);
        g = nullptr;
    }
}

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
        cudaEventCreateWithFlags(&dev->ev_trace, cudaEventDisableTiming) != cudaSuccess ||
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
    if (const char* e = std::getenv("IGB200_TRACE_CHUNK")) // experiment switch: rays per chunk of the host-buffer pipeline
        dev->chunk_rays = std::max<uint64_t>(1u << 12, std::min<uint64_t>(1u << 26, std::strtoull(e, nullptr, 10)));
    if (const char* e = std::getenv("IGB200_PIPE_ALIGN")) dev->pipe_align = std::atoi(e) != 0;
    if (const char* e = std::getenv("IGB200_PIPE_SMS")) // experiment switch: grid of the pipeline's launches = this many SMs x 8 CTAs
        dev->pipe_sms = std::atoi(e) > 0 ? std::min(dev->sm_count, std::atoi(e)) : 0;
    if (const char* e = std::getenv("IGB200_TRACE_CHUNK_MIN"))
        dev->chunk_min = std::max<uint64_t>(1u << 12, std::min<uint64_t>(1u << 26, std::strtoull(e, nullptr, 10)));
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
    if (dev->ev_trace) cudaEventDestroy(dev->ev_trace);
    if (dev->stream) cudaStreamDestroy(dev->stream);
    for (cudaStream_t p : dev->pipe)
        if (p) cudaStreamDestroy(p);
    for (cudaEvent_t e : dev->chunk_events)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : dev->round_ev)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : dev->sec_ev)
        if (e) cudaEventDestroy(e);
    drop_round_graphs(dev);
    if (dev->h_done) cudaFreeHost(dev->h_done);
    if (dev->h_film) cudaFreeHost(dev->h_film);
    for (auto& a : dev->aov)
        if (a.h) cudaFreeHost(a.h);
    delete dev; // DeviceBuffer destructors free the allocations (Device::releaseAll, Device.cpp:948-952)
}

// Levels the traversal stack has room for (kStackSize entries; an instance entry parks 6 + up to 2 of them, hence the + 9 at the
// call sites).  IGB200_STACK_BUDGET lowers it -- never raises it -- so that the balanced-rebuild path can be exercised on
// ordinary geometry (tests/test_limits_gpu.py) and deep trees can be avoided on purpose.
static uint32_t stack_budget()
{
    uint32_t budget = (uint32_t)igb::kStackSize;
    if (const char* e = std::getenv("IGB200_STACK_BUDGET")) budget = std::min<uint32_t>(budget, (uint32_t)std::max(14, std::atoi(e)));
    return budget;
}

// Disk cache of large mesh BVHs, the counterpart of setup_bvh's cache (src/runtime/shape/TriMeshProvider.cpp:329-353: meshes
// above 500 000 faces, keyed by a content hash, one file per mesh in the cache directory).  Opt-in through the environment
// variable IGB200_CACHE_DIR (the reference's CacheManager can be disabled likewise); a file that does not match the mesh
// (hash, face count, builder version) or is damaged is ignored and rewritten.  Key = FNV-1a over the triangle boxes the
// builder consumes, so it is independent of where the mesh sits in the shape table.
static constexpr uint64_t kBvhCacheMinFaces = 500000;
static constexpr uint32_t kBvhCacheVersion  = 2; // bump when bvh8_build.cpp changes the tree it produces
static void build_or_load_blas(const std::vector<float>& boxes, uint32_t nf, igb::Bvh8& blas)
{
    using igb::Node8;
    const char* dir = std::getenv("IGB200_CACHE_DIR");
    if (!dir || !*dir || nf <= kBvhCacheMinFaces) {
        igb::build_bvh8(boxes.data(), nf, 3, blas);
        return;
    }
    uint64_t h = 1469598103934665603ull;
    const uint8_t* bytes = reinterpret_cast<const uint8_t*>(boxes.data());
    for (size_t i = 0; i < boxes.size() * sizeof(float); ++i) h = (h ^ bytes[i]) * 1099511628211ull;
    char name[64];
    std::snprintf(name, sizeof(name), "/_bvh8_%016llx.bin", (unsigned long long)h);
    const std::string path = std::string(dir) + name;
    struct Head {
        char     magic[8];
        uint32_t version, faces;
        uint64_t hash, nodes, prims;
        uint32_t max_depth, pad;
        double   sah_node, sah_prim;
    };
    if (FILE* f = std::fopen(path.c_str(), "rb")) {
        Head hd{};
        bool ok = std::fread(&hd, sizeof(hd), 1, f) == 1 && std::memcmp(hd.magic, "IGBVH8\0\0", 8) == 0 && hd.version == kBvhCacheVersion &&
                  hd.faces == nf && hd.hash == h && hd.prims == nf && hd.nodes >= 1 && hd.nodes <= 2ull * nf + 1;
        if (ok) {
            blas.nodes.resize(hd.nodes);
            blas.prim_indices.resize(hd.prims);
            ok = std::fread(blas.nodes.data(), sizeof(Node8), hd.nodes, f) == hd.nodes &&
                 std::fread(blas.prim_indices.data(), sizeof(uint32_t), hd.prims, f) == hd.prims;
            for (size_t i = 0; ok && i < blas.prim_indices.size(); ++i) ok = blas.prim_indices[i] < nf; // never trust indices from disk
        }
        std::fclose(f);
        if (ok) {
            blas.max_depth     = hd.max_depth;
            blas.sah_node_term = hd.sah_node;
            blas.sah_prim_term = hd.sah_prim;
            return;
        }
        blas = igb::Bvh8{};
    }
    igb::build_bvh8(boxes.data(), nf, 3, blas);
    const std::string tmp = path + ".tmp";
    if (FILE* f = std::fopen(tmp.c_str(), "wb")) {
        Head hd{};
        std::memcpy(hd.magic, "IGBVH8\0\0", 8);
        hd.version   = kBvhCacheVersion;
        hd.faces     = nf;
        hd.hash      = h;
        hd.nodes     = blas.nodes.size();
        hd.prims     = blas.prim_indices.size();
        hd.max_depth = blas.max_depth;
        hd.sah_node  = blas.sah_node_term;
        hd.sah_prim  = blas.sah_prim_term;
        const bool ok = std::fwrite(&hd, sizeof(hd), 1, f) == 1 && std::fwrite(blas.nodes.data(), sizeof(Node8), blas.nodes.size(), f) == blas.nodes.size() &&
                        std::fwrite(blas.prim_indices.data(), sizeof(uint32_t), blas.prim_indices.size(), f) == blas.prim_indices.size();
        std::fclose(f);
        if (ok)
            std::rename(tmp.c_str(), path.c_str()); // atomic publish: readers never see a half-written file
        else
            std::remove(tmp.c_str());
    }
}

int igb200_assign_scene(igb200_device* dev, const igb200_scene* sc)
{
    using namespace igb;
    if (!dev || !sc) return fail(IGB200_E_INVALID, "igb200_assign_scene: NULL argument");
    if (sc->shape_count && (!sc->shape_lookups || !sc->shape_data))
        return fail(IGB200_E_INVALID, "igb200_assign_scene: shape table pointers are NULL");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
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
            // the shading kernels index ent_mat[], entities[36 * ent] and the shape tables with the hit's entity id unguarded
            if (sc->entities) {
                if (l.entity_id < 0 || (uint64_t)l.entity_id >= sc->entity_count)
                    return fail(IGB200_E_INVALID, "leaf %llu of type %d: entity id %d of %llu", (unsigned long long)i, k, l.entity_id,
                                (unsigned long long)sc->entity_count);
                int32_t row_shape;
                std::memcpy(&row_shape, sc->entities + 36 * (size_t)l.entity_id + 33, 4);
                if (row_shape != l.shape_id)
                    return fail(IGB200_E_INVALID, "entity %d: leaf says shape %d, entity row says shape %d", l.entity_id, l.shape_id, row_shape);
            }
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
        if (tlas.max_depth + 9 + 4 > stack_budget()) { // leave room for at least a shallow mesh below it
            tlas = Bvh8{};
            build_bvh8(boxes.data(), (uint32_t)in_leaves.size(), 1, tlas, true);
        }
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
            const uint64_t nf = hdr[0], nv = hdr[1], nn = hdr[2], nt = hdr[3];
            // TriMesh::normal(idx) and tex_base()[idx] (shade_lib.cuh) are indexed by VERTEX index: one normal and one
            // texture coordinate per vertex, and the texcoord block inside the table (TriMeshProvider.cpp:577-598)
            if (nn < nv || nt < nv)
                return fail(IGB200_E_INVALID, "shape %llu: %llu normals / %llu texcoords for %llu vertices", (unsigned long long)s,
                            (unsigned long long)nn, (unsigned long long)nt, (unsigned long long)nv);
            if (avail < 4 * (12 + nv * 4 + nn * 4 + nf * 4) + 8 * nt)
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
            build_or_load_blas(boxes, (uint32_t)nf, blas);
            if (tlas.max_depth + blas.max_depth + 9 > stack_budget()) {
                // SAH on sorted, sliver or exponentially spaced geometry can go deeper than the traversal stack: rebuild this
                // mesh with balanced splits (depth <= log2 n) instead of refusing the scene
                blas = Bvh8{};
                build_bvh8(boxes.data(), (uint32_t)nf, 3, blas, true);
            }
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
    if (tlas.max_depth + blas_depth + 9 > (uint32_t)kStackSize)
        return fail(IGB200_E_UNSUPPORTED, "BVH too deep for the traversal stack (%u + %u levels)", tlas.max_depth, blas_depth);

    // ---- upload ----
    static_assert(sizeof(Node8) == 5 * sizeof(uint4), "node size");
    IGB_CUDA(dev->nodes.upload(reinterpret_cast<const uint4*>(all_nodes.data()), all_nodes.size() * 5, dev->stream));
    IGB_CUDA(dev->tris.upload(all_tris.data(), all_tris.size(), dev->stream));
    // The device copy of a leaf carries its shape's BVH root and primitive type where the reference keeps the offset of
    // the shape's prim-BVH blob (user1 / user2, SceneBVHAdapter.h:73-105; that blob is not consumed here): the entity step
    // of the traversal then needs ONE dependent fetch instead of leaf -> shape record.
    for (igb200_entity_leaf& l : ordered) {
        const ShapeRef& r = shape_refs[l.shape_id];
        l.user1 = (int32_t)r.root_node;
        l.user2 = (int32_t)(r.type | (r.sphere << 8));
    }
    IGB_CUDA(dev->leaves.upload(reinterpret_cast<const float4*>(ordered.data()), ordered.size() * 6, dev->stream));
    IGB_CUDA(dev->shapes.upload(shape_refs.data(), shape_refs.size(), dev->stream));
    IGB_CUDA(dev->spheres.upload(all_spheres.data(), all_spheres.size(), dev->stream));
    // tables the shading kernels read (entities FixTable, shapes DynTable data)
    dev->has_shading = false;
    drop_round_graphs(dev);
    {
        std::vector<uint2>   info(sc->shape_count);
        for (uint64_t s = 0; s < sc->shape_count; ++s) {
            if (sc->shape_lookups[s].offset % 16 != 0) return fail(IGB200_E_INVALID, "shape %llu: offset is not 16-byte aligned", (unsigned long long)s);
            info[s] = make_uint2((uint32_t)(sc->shape_lookups[s].offset / 4), sc->shape_lookups[s].type_id);
        }
        std::vector<int32_t> ent_mat(sc->entities ? sc->entity_count : 0);
        for (size_t i = 0; i < ent_mat.size(); ++i) {
            std::memcpy(&ent_mat[i], sc->entities + 36 * i + 34, 4);
            int32_t shape_id;
            std::memcpy(&shape_id, sc->entities + 36 * i + 33, 4);
            if (shape_id < 0 || (uint64_t)shape_id >= sc->shape_count || ent_mat[i] < 0 || (uint32_t)ent_mat[i] >= sc->material_count)
                return fail(IGB200_E_INVALID, "entity %zu: shape %d / material %d out of range", i, shape_id, ent_mat[i]);
        }
        IGB_CUDA(dev->d_shape_info.upload(info.data(), info.size(), dev->stream));
        IGB_CUDA(dev->d_shape_data.upload(reinterpret_cast<const float*>(sc->shape_data), (sc->shape_bytes + 3) / 4, dev->stream));
        IGB_CUDA(dev->d_entities.upload(sc->entities, ent_mat.size() * 36, dev->stream));
        IGB_CUDA(dev->d_ent_mat.upload(ent_mat.data(), ent_mat.size(), dev->stream));
        dev->entity_count         = ent_mat.size();
        dev->scene_material_count = sc->material_count;
        std::memcpy(dev->scene_bbox, sc->scene_bbox, sizeof(dev->scene_bbox));
    }
    IGB_CUDA(cudaStreamSynchronize(dev->stream));

    dev->view.nodes      = dev->nodes.ptr;
    dev->view.tris       = dev->tris.ptr;
    dev->view.leaves     = dev->leaves.ptr;
    dev->view.shapes     = dev->shapes.ptr;
    dev->view.spheres    = dev->spheres.ptr;
    dev->view.tlas_nodes = (uint32_t)tlas.nodes.size();
    dev->view.leaf_count = (uint32_t)ordered.size();
    dev->view.byte_bias  = 0x47000000u;
    dev->view.half_bias  = 0x64006400u;
    dev->view.node_count = (uint32_t)all_nodes.size();
    dev->view.tri_count  = (uint32_t)(all_tris.size() / 3);
    {
        // Triangle records bypass L1 allocation when the NODE array fits L1 (<= 192 KiB; a B200 SM has ~200 KB of L1) and the
        // triangle array does not (> 256 KiB): the records then only evict nodes that would have stayed.  Measured
        // (profiles/r02_ab_tri_stream.txt): C3 instanced (138 KB nodes, 774 KB triangles) +1.2 %, its dense variant +1.7 %; with
        // everything large (C3-flat: 10 MB nodes) caching the records wins 1.5-2.8 % (neighbouring records share lines), and a
        // scene that fits L1 entirely loses 41-56 % without it (C4, office).  IGB200_TRI_STREAM_BYTES overrides the triangle
        // threshold (0 = always, huge = never), for A/Bs.
        uint64_t limit = 256u << 10, node_limit = 192u << 10;
        if (const char* e = std::getenv("IGB200_TRI_STREAM_BYTES")) {
            limit      = std::strtoull(e, nullptr, 10);
            node_limit = ~0ull;
        }
        dev->view.tri_stream = (all_tris.size() * sizeof(float4) > limit && all_nodes.size() * sizeof(Node8) <= node_limit) ? 1u : 0u;
    }
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
    if (dev->trace_pending && dev->last_trace_stream != stream) IGB_CUDA(cudaStreamWaitEvent(stream, dev->ev_trace, 0));
    if (timed) IGB_CUDA(cudaEventRecord(dev->ev0, stream));
    IGB_CUDA(igb::launch_trace(dev->view, reinterpret_cast<const float4*>(d_rays), n, ray_flags, any_hit, out, counters,
                               dev->d_work.ptr + 2 * igb200_device::kPipe, dev->sm_count, stream));
    if (timed) IGB_CUDA(cudaEventRecord(dev->ev1, stream));
    IGB_CUDA(cudaEventRecord(dev->ev_trace, stream));
    dev->last_trace_stream = stream;
    dev->trace_pending     = true;
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

    // Chunked software pipeline over three streams by ROLE -- upload, traversal, download -- chained by events: upload of
    // chunk k+1, traversal of chunk k and download of chunk k-1 overlap (PCIe is full duplex), instead of the
    // reference's upload-all / trace / download-all (Device.cpp:573-611, 1330-1363).  Chunks use disjoint slices of the
    // staging buffers; small chunks keep the un-overlapped head (first upload) and tail (last download) short.
    // Chunk schedule: sizes double from chunk_min up to chunk_rays and halve again at the end, so the un-overlapped
    // head (the first upload) and tail (the last download) are those of a small chunk while the bulk moves in large
    // ones (a chunk costs ~13 driver calls, ~0.1 ms of host time: small chunks throughout would be issue-bound).
    std::vector<uint64_t> sizes;
    {
        const uint64_t big = dev->chunk_rays, small = std::min(big, dev->chunk_min);
        std::vector<uint64_t> up;
        uint64_t              ramp = 0;
        for (uint64_t s = small; s < big; s *= 2) {
            up.push_back(s);
            ramp += s;
        }
        if (n > 2 * ramp + big) {
            sizes = up;
            uint64_t mid = n - 2 * ramp;
            for (; mid > big; mid -= big) sizes.push_back(big);
            sizes.push_back(mid);
            sizes.insert(sizes.end(), up.rbegin(), up.rend());
        } else {
            for (uint64_t left = n; left > 0;) {
                sizes.push_back(std::min(big, left));
                left -= sizes.back();
            }
        }
    }
    const uint64_t nchunks = sizes.size();
    // IGB200_PIPE_TIMELINE=1 (diagnostic): two more events per chunk and one at the start; the per-chunk times of upload end,
    // traversal start / end and download end go to stderr after the call
    static const bool timeline = std::getenv("IGB200_PIPE_TIMELINE") != nullptr;
    const size_t      n_events = 3 * nchunks + (timeline ? nchunks + 1 : 0);
    if (dev->chunk_events.size() < n_events) {
        const size_t old = dev->chunk_events.size();
        dev->chunk_events.resize(n_events, nullptr);
        for (size_t i = old; i < dev->chunk_events.size(); ++i) IGB_CUDA(cudaEventCreate(&dev->chunk_events[i]));
    }
    cudaStream_t s_up = dev->pipe[0], s_trace = dev->pipe[1], s_down = dev->pipe[2];
    uint64_t     off  = 0;
    if (timeline) IGB_CUDA(cudaEventRecord(dev->chunk_events[4 * nchunks], s_up));
    // Download schedule: the hit columns of chunk c-1 are queued behind BOTH its traversal and the upload of chunk c, so that a
    // download starts at an upload boundary.  Measured (scripts/pcie_chunk_probe.py, profiles/r02_ab_e2e_aligned_download.txt):
    // a download that starts in the middle of an upload costs the pair ~7 % of the link: 16 Mi rays end to end 1460 -> 1530 Mrays/s.
    // Exempting the ramp-down chunks changes nothing.  IGB200_PIPE_ALIGN=0 restores the round-1 schedule (download right behind
    // the traversal).
    auto deferred = [&](uint64_t c) { return dev->pipe_align && c + 1 < nchunks; }; // all but the last chunk
    auto download = [&](uint64_t c, uint64_t at, cudaEvent_t boundary) -> cudaError_t {
        const uint64_t m = sizes[c];
        cudaError_t    e = cudaStreamWaitEvent(s_down, dev->chunk_events[3 * c + 1], 0);
        if (e == cudaSuccess && boundary) e = cudaStreamWaitEvent(s_down, boundary, 0);
        if (e == cudaSuccess && ent_id) e = cudaMemcpyAsync(ent_id + at, dev->d_ent.ptr + at, m * 4, cudaMemcpyDeviceToHost, s_down);
        if (e == cudaSuccess && prim_id) e = cudaMemcpyAsync(prim_id + at, dev->d_prim.ptr + at, m * 4, cudaMemcpyDeviceToHost, s_down);
        if (e == cudaSuccess && t) e = cudaMemcpyAsync(t + at, dev->d_t.ptr + at, m * 4, cudaMemcpyDeviceToHost, s_down);
        if (e == cudaSuccess && u) e = cudaMemcpyAsync(u + at, dev->d_u.ptr + at, m * 4, cudaMemcpyDeviceToHost, s_down);
        if (e == cudaSuccess && v) e = cudaMemcpyAsync(v + at, dev->d_v.ptr + at, m * 4, cudaMemcpyDeviceToHost, s_down);
        if (e == cudaSuccess && timeline) e = cudaEventRecord(dev->chunk_events[3 * nchunks + c], s_down);
        return e;
    };
    uint64_t prev_off = 0;
    for (uint64_t c = 0; c < nchunks; prev_off = off, off += sizes[c], ++c) {
        const uint64_t m  = sizes[c];
        float4*        dr = dev->d_rays.ptr + 2 * off;
        IGB_CUDA(cudaMemcpyAsync(dr, rays + off, m * sizeof(igb200_ray), cudaMemcpyHostToDevice, s_up));
        IGB_CUDA(cudaEventRecord(dev->chunk_events[3 * c + 2], s_up));
        if (c > 0 && deferred(c - 1)) IGB_CUDA(download(c - 1, prev_off, dev->chunk_events[3 * c + 2]));
        IGB_CUDA(cudaStreamWaitEvent(s_trace, dev->chunk_events[3 * c + 2], 0));
        const igb::TraceOutputs out{ ent_id ? dev->d_ent.ptr + off : nullptr, prim_id ? dev->d_prim.ptr + off : nullptr,
                                     t ? dev->d_t.ptr + off : nullptr, u ? dev->d_u.ptr + off : nullptr,
                                     v ? dev->d_v.ptr + off : nullptr };
        IGB_CUDA(cudaEventRecord(dev->chunk_events[3 * c], s_trace));
        // Interface::loadRayList's direction normalisation (Device.cpp:588-592) happens in the kernel's ray load
        IGB_CUDA(igb::launch_trace(dev->view, dr, m, ray_flags, any_hit, out, counters, dev->d_work.ptr,
                                   dev->pipe_sms ? dev->pipe_sms : dev->sm_count, s_trace, true));
        IGB_CUDA(cudaEventRecord(dev->chunk_events[3 * c + 1], s_trace));
        if (!deferred(c)) IGB_CUDA(download(c, off, nullptr));
    }
    for (int k = 0; k < igb200_device::kPipe; ++k) IGB_CUDA(cudaStreamSynchronize(dev->pipe[k]));
    if (timeline) {
        std::fprintf(stderr, "# igb200_trace_rays timeline, ms since the first upload was queued: chunk rays up_end trace_start trace_end down_end\n");
        for (uint64_t c = 0; c < nchunks; ++c) {
            float a = 0, b = 0, d = 0, e = 0;
            cudaEventElapsedTime(&a, dev->chunk_events[4 * nchunks], dev->chunk_events[3 * c + 2]);
            cudaEventElapsedTime(&b, dev->chunk_events[4 * nchunks], dev->chunk_events[3 * c]);
            cudaEventElapsedTime(&d, dev->chunk_events[4 * nchunks], dev->chunk_events[3 * c + 1]);
            cudaEventElapsedTime(&e, dev->chunk_events[4 * nchunks], dev->chunk_events[3 * nchunks + c]);
            std::fprintf(stderr, "%3llu %8llu %7.3f %7.3f %7.3f %7.3f\n", (unsigned long long)c, (unsigned long long)sizes[c], a, b, d, e);
        }
    }

    dev->last_stats      = igb200_trace_stats{};
    dev->last_stats.rays = n;
    for (uint64_t c = 0; c < nchunks; ++c) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, dev->chunk_events[3 * c], dev->chunk_events[3 * c + 1]) == cudaSuccess)
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

// ---------------------------------------------------------------------------------------------------------
// Wavefront path tracing
// ---------------------------------------------------------------------------------------------------------
int igb200_assign_shading(igb200_device* dev, const igb200_shading* sh)
{
    if (dev)
        if (int rc = flush_wave(dev)) return rc;
    if (!dev || !sh) return fail(IGB200_E_INVALID, "igb200_assign_shading: NULL argument");
    if (!dev->has_scene) return fail(IGB200_E_STATE, "igb200_assign_shading: no scene assigned");
    if ((sh->bsdf_count && !sh->bsdfs) || (sh->material_count && !sh->materials) || (sh->infinite_light_count && !sh->infinite_lights) ||
        (sh->finite_light_count && !sh->finite_lights))
        return fail(IGB200_E_INVALID, "igb200_assign_shading: table pointer is NULL");
    if (sh->material_count < dev->scene_material_count)
        return fail(IGB200_E_INVALID, "igb200_assign_shading: %llu material rows for %u scene materials",
                    (unsigned long long)sh->material_count, dev->scene_material_count);
    if (sh->material_count + 1 > igb::kWaveMaxBins) return fail(IGB200_E_UNSUPPORTED, "more than %u materials", igb::kWaveMaxBins - 1);
    if (dev->entity_count == 0 && dev->view.leaf_count != 0)
        return fail(IGB200_E_STATE, "igb200_assign_shading: the scene was assigned without an entity table");
    for (uint64_t i = 0; i < sh->material_count; ++i) {
        const igb200_material& m = sh->materials[i];
        if (m.bsdf < 0 || (uint64_t)m.bsdf >= sh->bsdf_count) return fail(IGB200_E_INVALID, "material %llu: bsdf %d out of range", (unsigned long long)i, m.bsdf);
        if (m.light >= (int64_t)sh->finite_light_count) return fail(IGB200_E_INVALID, "material %llu: light %d out of range", (unsigned long long)i, m.light);
    }
    if ((sh->texture_count && !sh->textures) || (sh->blob_bytes && !sh->blob))
        return fail(IGB200_E_INVALID, "igb200_assign_shading: texture / blob pointer is NULL");
    for (uint64_t i = 0; i < sh->texture_count; ++i) {
        const igb200_texture& t = sh->textures[i];
        if (t.type > IGB200_TEX_IMAGE) return fail(IGB200_E_UNSUPPORTED, "texture %llu: unknown type %u", (unsigned long long)i, t.type);
        if (t.type != IGB200_TEX_IMAGE) continue;
        const uint64_t texel = t.format == IGB200_IMAGE_RGBA32F ? 16 : 4;
        if (t.format > IGB200_IMAGE_RGBA32F || t.filter > IGB200_FILTER_BICUBIC || t.wrap_u > IGB200_WRAP_CLAMP || t.wrap_v > IGB200_WRAP_CLAMP ||
            t.width == 0 || t.height == 0 || (t.offset & 15u) || t.offset + texel * t.width * t.height > sh->blob_bytes)
            return fail(IGB200_E_INVALID, "texture %llu: bad image description", (unsigned long long)i);
    }
    for (uint64_t i = 0; i < sh->bsdf_count; ++i) {
        const igb200_bsdf& b = sh->bsdfs[i];
        if (b.type > IGB200_BSDF_THIN_DIELECTRIC) return fail(IGB200_E_UNSUPPORTED, "bsdf %llu: unknown type %u", (unsigned long long)i, b.type);
        if (b.type == IGB200_BSDF_DIFFUSE && b.q[0].u > sh->texture_count)
            return fail(IGB200_E_INVALID, "bsdf %llu: texture %u out of range", (unsigned long long)i, b.q[0].u - 1);
        if (b.type == IGB200_BSDF_BUMPMAP &&
            (b.q[0].u == 0 || b.q[0].u > sh->texture_count || b.q[1].u == 0 || b.q[1].u > sh->bsdf_count || b.q[1].u - 1 == i))
            return fail(IGB200_E_INVALID, "bsdf %llu: bump map texture / inner bsdf out of range", (unsigned long long)i);
    }
    if (sh->selector > IGB200_SELECTOR_CDF) return fail(IGB200_E_UNSUPPORTED, "unknown light selector %u", sh->selector);
    if (sh->selector == IGB200_SELECTOR_CDF && sh->finite_light_count > 0 &&
        ((uint64_t)sh->hierarchy * 4 + sh->finite_light_count * 4 > sh->blob_bytes))
        return fail(IGB200_E_INVALID, "igb200_assign_shading: light CDF exceeds the blob");
    if (sh->selector == IGB200_SELECTOR_HIERARCHY && sh->finite_light_count > 1) {
        const uint64_t codes = (sh->finite_light_count + 3) / 4 * 4, nodes = 2 * sh->finite_light_count - 1;
        if ((sh->hierarchy & 3u) || (uint64_t)sh->hierarchy * 4 + codes * 4 + nodes * 32 > sh->blob_bytes)
            return fail(IGB200_E_INVALID, "igb200_assign_shading: light hierarchy exceeds the blob");
    }
    for (int t = 0; t < 2; ++t) {
        const igb200_light* ls = t ? sh->finite_lights : sh->infinite_lights;
        const uint64_t      n  = t ? sh->finite_light_count : sh->infinite_light_count;
        for (uint64_t i = 0; i < n; ++i) {
            const bool inf = ls[i].type == IGB200_LIGHT_ENV || ls[i].type == IGB200_LIGHT_DIRECTIONAL || ls[i].type == IGB200_LIGHT_ENV_TEXTURED ||
                             ls[i].type == IGB200_LIGHT_ENV_FUNCTION || ls[i].type == IGB200_LIGHT_SUN;
            if (ls[i].type == IGB200_LIGHT_ENV_FUNCTION &&
                (ls[i].tex > IGB200_SKY_PEREZ || (uint64_t)ls[i].cdf * 4 + 64 > sh->blob_bytes))
                return fail(IGB200_E_INVALID, "sky light %llu: unknown function or parameters outside the blob", (unsigned long long)i);
            if (ls[i].type == IGB200_LIGHT_ENV_TEXTURED) {
                if (ls[i].tex >= sh->texture_count) return fail(IGB200_E_INVALID, "environment light %llu: texture out of range", (unsigned long long)i);
                if (ls[i].cdf != IGB200_NO_CDF) {
                    if ((ls[i].cdf & 3u) || (uint64_t)ls[i].cdf * 4 + 16 > sh->blob_bytes)
                        return fail(IGB200_E_INVALID, "environment light %llu: CDF out of range", (unsigned long long)i);
                    const uint32_t* h = reinterpret_cast<const uint32_t*>(sh->blob) + ls[i].cdf;
                    if (h[0] == 0 || h[1] == 0 || (uint64_t)ls[i].cdf * 4 + 16 + 4ull * h[1] * (1ull + h[0]) > sh->blob_bytes)
                        return fail(IGB200_E_INVALID, "environment light %llu: CDF exceeds the blob", (unsigned long long)i);
                }
            }
            if (ls[i].type > IGB200_LIGHT_SUN || inf != (t == 0))
                return fail(IGB200_E_INVALID, "%s light %llu has type %u", t ? "finite" : "infinite", (unsigned long long)i, ls[i].type);
            if ((ls[i].type == IGB200_LIGHT_AREA_SHAPE || ls[i].type == IGB200_LIGHT_AREA_PLANE || ls[i].type == IGB200_LIGHT_AREA_SPHERE) &&
                (ls[i].entity < 0 || (uint64_t)ls[i].entity >= dev->entity_count))
                return fail(IGB200_E_INVALID, "area light %llu: entity %d out of range", (unsigned long long)i, ls[i].entity);
        }
    }
    if (sh->camera.type > IGB200_CAMERA_FISHLENS) return fail(IGB200_E_UNSUPPORTED, "unknown camera type %u", sh->camera.type);
    if (sh->technique.type > IGB200_TECH_INFOBUFFER) return fail(IGB200_E_UNSUPPORTED, "unknown technique %u", sh->technique.type);
    if (sh->technique.max_depth > 254) // the stream keeps the depth in 8 bits (reference default: 64, PathTechnique.cpp)
        return fail(IGB200_E_UNSUPPORTED, "technique.max_depth %d exceeds 254", sh->technique.max_depth);
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    IGB_CUDA(dev->d_bsdfs.upload(sh->bsdfs, sh->bsdf_count, dev->stream));
    IGB_CUDA(dev->d_materials.upload(sh->materials, sh->material_count, dev->stream));
    IGB_CUDA(dev->d_inf_lights.upload(sh->infinite_lights, sh->infinite_light_count, dev->stream));
    IGB_CUDA(dev->d_fin_lights.upload(sh->finite_lights, sh->finite_light_count, dev->stream));
    IGB_CUDA(dev->d_textures.upload(sh->textures, sh->texture_count, dev->stream));
    IGB_CUDA(dev->d_blob.upload(sh->blob, sh->blob_bytes, dev->stream));
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    igb::shade::ShadeScene& s = dev->shade;
    s.entities   = dev->d_entities.ptr;
    s.shape_data = dev->d_shape_data.ptr;
    s.shape_info = dev->d_shape_info.ptr;
    s.bsdfs      = dev->d_bsdfs.ptr;
    s.materials  = dev->d_materials.ptr;
    s.inf_lights = dev->d_inf_lights.ptr;
    s.fin_lights = dev->d_fin_lights.ptr;
    s.n_inf      = (int32_t)sh->infinite_light_count;
    s.n_fin      = (int32_t)sh->finite_light_count;
    s.n_mat      = (int32_t)sh->material_count;
    s.n_ent      = (int32_t)dev->entity_count;
    s.cam        = sh->camera;
    s.tech       = sh->technique;
    s.textures   = dev->d_textures.ptr;
    s.blob       = dev->d_blob.ptr;
    s.selector   = sh->selector;
    s.hierarchy  = sh->hierarchy;
    // The reference specialises its shaders per scene through the JIT; here the shade kernel has two instantiations
    // and the scene picks one: the extended one only when a feature it adds is actually used.
    s.extended = (sh->selector == IGB200_SELECTOR_HIERARCHY || sh->selector == IGB200_SELECTOR_CDF) && sh->finite_light_count > 0;
    for (uint64_t i = 0; i < sh->bsdf_count; ++i) {
        const igb200_bsdf& b = sh->bsdfs[i];
        if (b.type == IGB200_BSDF_BUMPMAP || (b.type == IGB200_BSDF_DIFFUSE && b.q[0].u != 0) ||
            (b.type == IGB200_BSDF_CONDUCTOR && !(b.q[0].f <= 1e-4f || b.q[1].f <= 1e-4f)))
            s.extended = 1;
    }
    for (uint64_t i = 0; i < sh->infinite_light_count; ++i)
        if (sh->infinite_lights[i].type == IGB200_LIGHT_ENV_TEXTURED || sh->infinite_lights[i].type == IGB200_LIGHT_ENV_FUNCTION ||
            sh->infinite_lights[i].type == IGB200_LIGHT_SUN)
            s.extended = 1;
    for (uint64_t i = 0; i < sh->finite_light_count; ++i)
        if (sh->finite_lights[i].type == IGB200_LIGHT_AREA_SPHERE) s.extended = 1; // the sphere emitter lives in the extended kernel
    {
        // bbox_radius(scene_bbox) * 1.01 with vec3_len = sqrt of the fma-chain dot (core/bbox.art:24, vector.art:99)
        const float* b = dev->scene_bbox;
        const float  x = b[3] - b[0], y = b[4] - b[1], z = b[5] - b[2];
        s.scene_radius = std::sqrt(std::fmaf(x, x, std::fmaf(y, y, z * z))) / 2 * 1.01f;
    }
    drop_round_graphs(dev);
    dev->has_shading = true;
    return IGB200_OK;
}

static int ensure_film(igb200_device* dev, uint32_t w, uint32_t h)
{
    if (dev->film_w == w && dev->film_h == h && dev->d_film.ptr) return IGB200_OK;
    const size_t n = (size_t)w * h * 3;
    IGB_CUDA(dev->d_film.ensure(n));
    if (dev->h_film) { cudaFreeHost(dev->h_film); dev->h_film = nullptr; }
    IGB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&dev->h_film), std::max<size_t>(n, 1) * sizeof(float), cudaHostAllocDefault));
    IGB_CUDA(cudaMemsetAsync(dev->d_film.ptr, 0, n * sizeof(float), dev->stream));
    dev->film_w = w; dev->film_h = h;
    dev->iterations  = 0;
    dev->film_mapped = false;
    for (auto& a : dev->aov) { // AOVs follow the film size (Device.cpp:306-321): dropped here, re-created on use
        a.d.release();
        if (a.h) { cudaFreeHost(a.h); a.h = nullptr; }
        a.iterations = 0;
        a.mapped     = false;
    }
    return IGB200_OK;
}

static int ensure_aov(igb200_device* dev, int k)
{
    igb200_device::Aov& a = dev->aov[k];
    if (a.d.ptr) return IGB200_OK;
    const size_t n = (size_t)dev->film_w * dev->film_h * 3;
    IGB_CUDA(a.d.ensure(n));
    IGB_CUDA(cudaMemsetAsync(a.d.ptr, 0, n * sizeof(float), dev->stream));
    a.iterations = 0;
    a.mapped     = false;
    return IGB200_OK;
}

// index into igb200_device::aov, or -1 for the film, or -2 if the name is not an AOV
static int find_image(const char* name)
{
    if (!name || name[0] == '\0' || std::strcmp(name, "Color") == 0) return -1;
    for (int k = 0; k < igb200_device::kAovCount; ++k)
        if (std::strcmp(name, kAovNames[k]) == 0) return k;
    return -2;
}

// Streams of the wavefront, sized to the request: `want` rays (all iterations of the call), rounded up to a power of two between
// 1 Mi (the reference's capacity, mapping_gpu.art:1125) and 32 Mi -- 232 B per ray, 7.8 GB at the cap.  Measured on B200
// (profiles/r02_ab_wave_capacity.txt): every round costs a fixed 6 launches, a 1-CTA kernel and the tails of two persistent
// traversals, so fuller rounds are cheaper per ray: 4 / 8 / 16 / 32 / 64 Mi = C4 1087 / 1162 / 1196 / 1210 / 1214 Msamples/s.
// Grows on demand (the captured graph holds the pointers: re-captured then); IGB200_WAVE_CAPACITY pins it.
static int ensure_wave(igb200_device* dev, uint64_t want)
{
    uint64_t cap = 1u << 20;
    while (cap < want && cap < (1u << 25)) cap <<= 1;
    if (const char* e = std::getenv("IGB200_WAVE_CAPACITY")) cap = (uint64_t)std::max(1024l, std::atol(e));
    if (const char* e = std::getenv("IGB200_WAVE_SORT")) dev->wave_sort = std::max(0, std::min(2, std::atoi(e)));
    if (cap <= dev->wave_capacity) return IGB200_OK;
    drop_round_graphs(dev);
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    for (int b = 0; b < 2; ++b) {
        IGB_CUDA(dev->w_rec[b].ensure((size_t)cap * igb::kRecVec));
        IGB_CUDA(dev->w_flags[b].ensure(cap));
        dev->wave.stream[b] = igb::PathStream{ dev->w_rec[b].ptr, dev->w_flags[b].ptr };
    }
    for (int k = 0; k < 3; ++k) IGB_CUDA(dev->w_sh_f4[k].ensure(cap));
    dev->wave.shadow = igb::ShadowStream{ dev->w_sh_f4[0].ptr, dev->w_sh_f4[1].ptr, dev->w_sh_f4[2].ptr };
    IGB_CUDA(dev->w_keys.ensure(cap));
    IGB_CUDA(dev->w_sorted.ensure(cap));
    dev->wave.keys   = dev->w_keys.ptr;
    dev->wave.sorted = dev->w_sorted.ptr;
    if (!dev->h_done) {
        IGB_CUDA(dev->w_ctl.ensure(1));
        IGB_CUDA(cudaMemset(dev->w_ctl.ptr, 0, sizeof(igb::WaveControl)));
        IGB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&dev->h_done), 2 * sizeof(uint32_t), cudaHostAllocMapped));
        dev->h_done[0] = dev->h_done[1] = 0;
        uint32_t* d_alias = nullptr;
        IGB_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&d_alias), dev->h_done, 0));
        dev->wave.host_done = d_alias;
        dev->wave.ctl       = dev->w_ctl.ptr;
        for (cudaEvent_t& e : dev->round_ev) IGB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    dev->wave_capacity = (uint32_t)cap;
    return IGB200_OK;
}

// rows of the film that tile-row partition `rank` of `parts` owns (16-pixel tile rows rank, rank + parts, ...)
static uint32_t owned_rows(uint32_t h, uint32_t rank, uint32_t parts)
{
    uint32_t rows = 0;
    for (uint32_t t = rank; t * igb::kTileSize < h; t += parts) rows += std::min(igb::kTileSize, h - t * igb::kTileSize);
    return rows;
}

// Launches wavefront rounds on dev->stream until the device reports the queue drained (until_generated = false: the host runs
// ahead by up to kAhead round pairs and polls the mapped flag; rounds enqueued after the drain are no-ops) or until every ray id
// of the call has been generated (pipelined steps: one host sync per pair, nothing is launched beyond what the ids need).
static int run_wave(igb200_device* dev, bool until_generated, uint64_t& rounds_launched)
{
    uint64_t pairs = 0, odd = 0;
    cudaStream_t       st        = dev->stream;
    constexpr int      kAhead    = 3;
    const uint64_t     max_pairs = 1ull << 24;
    volatile uint32_t* flags     = dev->h_done;
    // One round pair (12 launches) is captured once into a CUDA graph; per-call settings live in device memory, so
    // the graph only depends on the scene / shading tables and the stream buffers (re-captured when they change).
    auto capture = [&](cudaGraphExec_t& exec, int first_parity, int rounds) -> int {
        if (exec) return IGB200_OK;
        cudaGraph_t g  = nullptr;
        cudaError_t e1 = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal);
        for (int r = 0; r < rounds && e1 == cudaSuccess; ++r)
            e1 = igb::launch_wave_round(dev->view, dev->shade, dev->wave, (first_parity + r) & 1, dev->wave_sort, dev->sm_count, st);
        cudaError_t e3 = cudaStreamEndCapture(st, &g);
        if (e1 != cudaSuccess || e3 != cudaSuccess || !g)
            return fail(IGB200_E_CUDA, "igb200_render: graph capture failed: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e3));
        cudaError_t e4 = cudaGraphInstantiate(&exec, g, 0);
        cudaGraphDestroy(g);
        if (e4 != cudaSuccess) return fail(IGB200_E_CUDA, "igb200_render: cudaGraphInstantiate failed: %s", cudaGetErrorString(e4));
        return IGB200_OK;
    };
    if (int rc = capture(dev->round_graph, 0, 2)) return rc;
    if (dev->stats) {
        // --stats (Device::SetupSettings::AcquireStats): rounds are launched one by one with an event between the kernels
        // and a host sync per pair, so that the time of every kernel kind is known (Statistics.h:9-63 names on the host side).
        if (!dev->sec_ev[0])
            for (cudaEvent_t& e : dev->sec_ev) IGB_CUDA(cudaEventCreate(&e));
        for (; pairs < max_pairs; ++pairs) {
            for (int parity = 0; parity < 2; ++parity) {
                IGB_CUDA(igb::launch_wave_round(dev->view, dev->shade, dev->wave, parity, dev->wave_sort, dev->sm_count, st, dev->sec_ev));
                IGB_CUDA(cudaStreamSynchronize(st));
                for (int k = 0; k < igb::kWaveLaunchesPerRound; ++k) {
                    float ms = 0.0f;
                    if (cudaEventElapsedTime(&ms, dev->sec_ev[k], dev->sec_ev[k + 1]) == cudaSuccess) dev->section.kernel_ms[k] += ms;
                    dev->section.launches[k] += (k == 3 && !dev->wave_sort) ? 0 : 1;
                }
            }
            if (flags[0]) { ++pairs; break; }
        }
    } else if (until_generated) {
        // ONE round at a time, alternating the stream parity across calls: a step whose ids fit the stream is a single round of
        // survivors + a whole new iteration (`pairs` counts half pairs here; read_wave_stats is told so)
        for (; pairs < max_pairs; ++pairs) {
            if (int rc = capture(dev->single_graph[dev->wave_parity], dev->wave_parity, 1)) return rc;
            IGB_CUDA(cudaGraphLaunch(dev->single_graph[dev->wave_parity], st));
            dev->wave_parity ^= 1;
            IGB_CUDA(cudaEventRecord(dev->round_ev[0], st));
            IGB_CUDA(cudaEventSynchronize(dev->round_ev[0]));
            if (flags[0] || flags[1]) { ++pairs; break; }
        }
    } else {
        if (dev->wave_parity) { // a pipelined step ended on an odd round: one round of parity 1 brings the survivors back to stream 0
            if (int rc = capture(dev->single_graph[1], 1, 1)) return rc;
            IGB_CUDA(cudaGraphLaunch(dev->single_graph[1], st));
            dev->wave_parity = 0;
            odd              = 1;
        }
        for (; pairs < max_pairs; ++pairs) {
            IGB_CUDA(cudaGraphLaunch(dev->round_graph, st));
            IGB_CUDA(cudaEventRecord(dev->round_ev[pairs % 4], st));
            if (pairs >= (uint64_t)kAhead) IGB_CUDA(cudaEventSynchronize(dev->round_ev[(pairs - kAhead) % 4]));
            if (flags[0]) { ++pairs; break; }
        }
    }
    rounds_launched = (until_generated && !dev->stats) ? pairs : 2 * pairs + odd;
    return IGB200_OK;
}

// Copies the device-side ray counters of the call into (add = false) or onto (add = true) dev->render_stats.
static int read_wave_stats(igb200_device* dev, uint64_t rounds_launched, bool list, bool add)
{
    igb::WaveControl ctl;
    IGB_CUDA(cudaMemcpy(&ctl, dev->w_ctl.ptr, offsetof(igb::WaveControl, bins), cudaMemcpyDeviceToHost));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, dev->ev0, dev->ev1);
    igb200_render_stats s{};
    s.camera_rays     = ctl.camera_rays;
    s.bounce_rays     = ctl.bounce_rays;
    s.shadow_rays     = ctl.shadow_rays;
    s.rounds          = ctl.rounds;
    s.kernel_launches = 1 + rounds_launched * (igb::kWaveLaunchesPerRound - (dev->wave_sort ? 0 : 1)) + (list ? 1 : 0);
    s.device_ms       = ms;
    if (add) {
        s.camera_rays += dev->render_stats.camera_rays;
        s.bounce_rays += dev->render_stats.bounce_rays;
        s.shadow_rays += dev->render_stats.shadow_rays;
        s.rounds += dev->render_stats.rounds;
        s.kernel_launches += dev->render_stats.kernel_launches;
        s.device_ms += dev->render_stats.device_ms;
    }
    dev->render_stats = s;
    return IGB200_OK;
}

// Drains the paths igb200_render_pipelined left in flight (no new ids; same settings).  Every entry that reads, clears or
// re-allocates the film, or changes the scene, the shading tables or the streams, calls this first.  The rays it traces are
// added to the counters of the last call.
static int flush_wave(igb200_device* dev)
{
    if (!dev->wave_pending) return IGB200_OK;
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    cudaStream_t st = dev->stream;
    reinterpret_cast<volatile uint32_t*>(dev->h_done)[0] = 0u;
    reinterpret_cast<volatile uint32_t*>(dev->h_done)[1] = 0u;
    IGB_CUDA(cudaEventRecord(dev->ev0, st));
    IGB_CUDA(igb::launch_wave_reset(dev->wave, dev->pending_set, 0u, st, true));
    uint64_t pairs = 0;
    if (int rc = run_wave(dev, false, pairs)) return rc;
    IGB_CUDA(cudaEventRecord(dev->ev1, st));
    IGB_CUDA(cudaStreamSynchronize(st));
    dev->wave_pending = false;
    if (!dev->h_done[0]) return fail(IGB200_E_CUDA, "igb200_flush: wavefront did not drain");
    dev->film_mapped = false;
    for (igb200_device::Aov& a : dev->aov) a.mapped = false;
    return read_wave_stats(dev, pairs, false, true);
}

static int render_batch(igb200_device* dev, const igb200_render_settings* rs, uint32_t count, uint32_t tile_rank = 0, uint32_t tile_count = 1,
                        bool pipelined = false)
{
    if (!dev || !rs) return fail(IGB200_E_INVALID, "igb200_render: NULL argument");
    if (count == 0) return IGB200_OK;
    if (!dev->has_scene || !dev->has_shading) return fail(IGB200_E_STATE, "igb200_render: scene and shading must be assigned first");
    const bool     list = rs->rays != nullptr;
    const uint64_t w = list ? rs->ray_count : rs->width, h = list ? 1 : rs->height;
    if (w == 0 || h == 0 || rs->spi == 0) return fail(IGB200_E_INVALID, "igb200_render: empty film or spi == 0");
    if (w * h * rs->spi >= 0x7FFF0000ull) return fail(IGB200_E_UNSUPPORTED, "igb200_render: more than 2^31 samples per iteration");
    if (list || tile_count <= 1) { tile_rank = 0; tile_count = 1; }
    const uint64_t rows     = tile_count > 1 ? owned_rows((uint32_t)h, tile_rank, tile_count) : h;
    const uint64_t per_iter = w * rows * rs->spi; // ray ids per iteration on THIS device
    {
        // one wavefront may span several iterations (its queue regenerates across them); split the request so
        // that a batch has fewer than 2^31 ray ids and fewer than 2^24 iterations
        const uint64_t max_count = std::max<uint64_t>(1, std::min<uint64_t>(0x7FFF0000ull / std::max<uint64_t>(per_iter, 1), 1u << 23));
        if (count > max_count) {
            igb200_render_settings part = *rs;
            igb200_render_stats    sum{};
            for (uint32_t done = 0; done < count;) {
                const uint32_t c = (uint32_t)std::min<uint64_t>(max_count, count - done);
                part.iteration   = rs->iteration + done;
                if (int rc = render_batch(dev, &part, c, tile_rank, tile_count)) return rc;
                done += c;
                // the counters describe the whole call, not its last batch
                sum.camera_rays += dev->render_stats.camera_rays;
                sum.bounce_rays += dev->render_stats.bounce_rays;
                sum.shadow_rays += dev->render_stats.shadow_rays;
                sum.rounds += dev->render_stats.rounds;
                sum.kernel_launches += dev->render_stats.kernel_launches;
                sum.device_ms += dev->render_stats.device_ms;
            }
            dev->render_stats = sum;
            return IGB200_OK;
        }
    }
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    // Pipelined steps (igb200_render_pipelined): the call returns when its ray ids have been generated; the paths still in flight
    // stay in the stream and continue in the next call's rounds (or in flush_wave).  Streams four iterations deep, so that the
    // survivors of earlier steps and a whole new iteration fit one round.
    const bool pipe = pipelined && !list && !dev->stats;
    uint64_t   want = (uint64_t)per_iter * count * (pipe ? 4u : 1u);
    if (dev->wave_pending) {
        const igb::WaveSettings& p = dev->pending_set;
        uint64_t cap = 1u << 20;
        while (cap < want && cap < (1u << 25)) cap <<= 1;
        const bool same = pipe && p.spi == rs->spi && p.width == w && p.height == h && p.per_iter == per_iter &&
                          p.tile_rank == tile_rank && p.tile_count == tile_count && cap <= dev->wave_capacity;
        if (!same)
            if (int rc = flush_wave(dev)) return rc;
    }
    if (int rc = ensure_wave(dev, want)) return rc;
    if (int rc = ensure_film(dev, (uint32_t)w, (uint32_t)h)) return rc;
    cudaStream_t st = dev->stream;
    if (per_iter == 0) { // more partitions than tile rows: this device owns nothing, but its (zero) film takes part in the reduce
        dev->render_stats = igb200_render_stats{};
        if (dev->shade.tech.type != IGB200_TECH_INFOBUFFER) dev->iterations += count;
        dev->film_mapped = false;
        return IGB200_OK;
    }

    igb::WaveSettings set{};
    set.spi = rs->spi; set.width = (uint32_t)w; set.height = (uint32_t)h;
    set.iter = rs->iteration; set.frame = rs->frame; set.seed = rs->user_seed;
    set.per_iter = (uint32_t)per_iter;
    set.inv_spi  = 1.0f / (float)rs->spi;
    set.capacity = dev->wave_capacity;
    set.tile_rank = tile_rank; set.tile_count = tile_count;
    if (list) {
        // Interface::loadRayList (Device.cpp:573-611): upload + normalise
        IGB_CUDA(dev->d_list_rays.ensure(2 * rs->ray_count));
        IGB_CUDA(cudaMemcpyAsync(dev->d_list_rays.ptr, rs->rays, rs->ray_count * sizeof(igb200_ray), cudaMemcpyHostToDevice, st));
        IGB_CUDA(igb::launch_normalize_rays(dev->d_list_rays.ptr, rs->ray_count, st));
        set.list_rays = dev->d_list_rays.ptr;
    }
    set.film = dev->d_film.ptr;
    // the technique's AOVs (TechniqueInfo::EnabledAOVs: PathTechnique.cpp:23-27, InfoBufferTechnique.cpp:30-33)
    const igb200_technique& tech = dev->shade.tech;
    const bool infobuffer = tech.type == IGB200_TECH_INFOBUFFER;
    int        used[3]    = { -1, -1, -1 };
    if (infobuffer) {
        used[0] = 2; used[1] = 3; used[2] = 4;
    } else if (tech.type == IGB200_TECH_PATH && (tech.flags & IGB200_TECHFLAG_AOV_MIS)) {
        used[0] = 0; used[1] = 1;
    }
    for (int k = 0; k < 3; ++k) {
        set.aov[k] = nullptr;
        if (used[k] < 0) continue;
        if (int rc = ensure_aov(dev, used[k])) return rc;
        set.aov[k] = dev->aov[used[k]].d.ptr;
    }
    dev->wave.ent_mat = dev->d_ent_mat.ptr;

    // the mapped flags are cleared from the host before anything that could raise them is queued (no wavefront kernel is pending:
    // every call ends with a stream sync), so the host may start polling without waiting for the reset kernel
    reinterpret_cast<volatile uint32_t*>(dev->h_done)[0] = 0u;
    reinterpret_cast<volatile uint32_t*>(dev->h_done)[1] = 0u;
    IGB_CUDA(cudaEventRecord(dev->ev0, st));
    IGB_CUDA(igb::launch_wave_reset(dev->wave, set, (uint32_t)(per_iter * count), st, dev->wave_pending));
    uint64_t pairs = 0;
    if (int rc = run_wave(dev, pipe, pairs)) return rc;
    IGB_CUDA(cudaEventRecord(dev->ev1, st));
    IGB_CUDA(cudaStreamSynchronize(st));
    dev->wave_pending = pipe && !dev->h_done[0];
    if (!dev->wave_pending) dev->wave_parity = 0; // drained: the next call starts on stream 0 again
    dev->pending_set  = set;
    if (!pipe && !dev->h_done[0]) return fail(IGB200_E_CUDA, "igb200_render: wavefront did not drain");
    if (int rc = read_wave_stats(dev, pairs, list, false)) return rc;
    // Interface::present, Device.cpp:1460-1472: the film counts unless the technique locks it, every used AOV counts
    if (!infobuffer) {
        dev->iterations += count;
        dev->film_mapped = false;
    }
    for (int k = 0; k < 3; ++k)
        if (used[k] >= 0) {
            dev->aov[used[k]].iterations += count;
            dev->aov[used[k]].mapped = false;
        }
    return IGB200_OK;
}

int igb200_render(igb200_device* dev, const igb200_render_settings* rs) { return render_batch(dev, rs, 1); }

int igb200_render_iterations(igb200_device* dev, const igb200_render_settings* rs, uint32_t count)
{
    return render_batch(dev, rs, count);
}

int igb200_render_pipelined(igb200_device* dev, const igb200_render_settings* rs) { return render_batch(dev, rs, 1, 0, 1, true); }

int igb200_flush(igb200_device* dev)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_flush: NULL device");
    return flush_wave(dev);
}

int igb200_partition_share(uint32_t total_iterations, uint32_t part, uint32_t parts, uint32_t mode, uint32_t* first, uint32_t* count,
                           uint32_t* tile_rank, uint32_t* tile_count)
{
    if (parts == 0 || part >= parts || !first || !count || !tile_rank || !tile_count)
        return fail(IGB200_E_INVALID, "igb200_partition_share: partition %u of %u", part, parts);
    if (mode == IGB200_PARTITION_AUTO) mode = total_iterations >= parts ? IGB200_PARTITION_ITERATIONS : IGB200_PARTITION_TILE_ROWS;
    if (mode == IGB200_PARTITION_ITERATIONS) {
        // balanced contiguous blocks of the iteration range: the RNG seed hashes the iteration index
        // (src/artic/core/random.art:34-43), so the union over the partitions is exactly the single-device sample set
        const uint32_t base = total_iterations / parts, rem = total_iterations % parts;
        *first      = part * base + std::min(part, rem);
        *count      = base + (part < rem ? 1u : 0u);
        *tile_rank  = 0;
        *tile_count = 1;
    } else if (mode == IGB200_PARTITION_TILE_ROWS) {
        *first      = 0;
        *count      = total_iterations;
        *tile_rank  = part;
        *tile_count = parts;
    } else {
        return fail(IGB200_E_INVALID, "igb200_partition_share: unknown mode %u", mode);
    }
    return IGB200_OK;
}

int igb200_render_partition(igb200_device* dev, const igb200_render_settings* rs, uint32_t total_iterations, uint32_t part,
                            uint32_t parts, uint32_t mode)
{
    if (!dev || !rs) return fail(IGB200_E_INVALID, "igb200_render_partition: NULL argument");
    uint32_t first = 0, count = 0, tr = 0, tc = 1;
    if (int rc = igb200_partition_share(total_iterations, part, parts, rs->rays ? IGB200_PARTITION_ITERATIONS : mode, &first, &count, &tr, &tc))
        return rc;
    dev->render_stats = igb200_render_stats{};
    if (count == 0) return IGB200_OK;
    igb200_render_settings mine = *rs;
    mine.iteration += first;
    return render_batch(dev, &mine, count, tr, tc);
}

int igb200_set_iteration_count(igb200_device* dev, const char* name, uint64_t iterations)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_set_iteration_count: NULL device");
    const int img = find_image(name);
    if (img == -2 || (img >= 0 && !dev->aov[img].d.ptr)) return fail(IGB200_E_INVALID, "igb200_set_iteration_count: unknown AOV '%s'", name);
    if (img < 0) {
        dev->iterations  = iterations;
        dev->film_mapped = false;
    } else {
        dev->aov[img].iterations = iterations;
        dev->aov[img].mapped     = false;
    }
    return IGB200_OK;
}

int igb200_device_ordinal(const igb200_device* dev) { return dev ? dev->ordinal : -1; }

// hooks for multi_gpu.cpp (not part of the public header): error channel and film size
int igb200_internal_fail(int code, const char* message) { return fail(code, "%s", message ? message : ""); }
int igb200_internal_film_size(igb200_device* dev, uint32_t* w, uint32_t* h)
{
    if (!dev) return IGB200_E_INVALID;
    if (w) *w = dev->d_film.ptr ? dev->film_w : 0;
    if (h) *h = dev->d_film.ptr ? dev->film_h : 0;
    return IGB200_OK;
}

int igb200_framebuffer(igb200_device* dev, const char* name, int device, float** data, uint64_t* iterations)
{
    if (!dev || !data) return fail(IGB200_E_INVALID, "igb200_framebuffer: NULL argument");
    *data = nullptr;
    if (iterations) *iterations = 0;
    const int img = find_image(name);
    if (img == -2 || (img >= 0 && !dev->aov[img].d.ptr))
        return fail(IGB200_E_INVALID, "igb200_framebuffer: unknown AOV '%s'", name); // Device.cpp:1303-1305
    if (!dev->d_film.ptr) return fail(IGB200_E_STATE, "igb200_framebuffer: nothing rendered yet");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    if (img >= 0) {
        igb200_device::Aov& a = dev->aov[img];
        if (iterations) *iterations = a.iterations;
        if (device) {
            *data = a.d.ptr;
            return IGB200_OK;
        }
        const size_t bytes = (size_t)dev->film_w * dev->film_h * 3 * sizeof(float);
        if (!a.h) IGB_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&a.h), std::max<size_t>(bytes, 4), cudaHostAllocDefault));
        if (!a.mapped) {
            IGB_CUDA(cudaMemcpyAsync(a.h, a.d.ptr, bytes, cudaMemcpyDeviceToHost, dev->stream));
            IGB_CUDA(cudaStreamSynchronize(dev->stream));
            a.mapped = true;
        }
        *data = a.h;
        return IGB200_OK;
    }
    if (iterations) *iterations = dev->iterations;
    if (device) {
        *data = dev->d_film.ptr;
        return IGB200_OK;
    }
    if (!dev->film_mapped) { // lazy D2H once per iteration (the `Mapped` flag, Device.cpp:1330-1363)
        if (dev->stats) IGB_CUDA(cudaEventRecord(dev->ev0, dev->stream));
        IGB_CUDA(cudaMemcpyAsync(dev->h_film, dev->d_film.ptr, (size_t)dev->film_w * dev->film_h * 3 * sizeof(float),
                                 cudaMemcpyDeviceToHost, dev->stream));
        if (dev->stats) IGB_CUDA(cudaEventRecord(dev->ev1, dev->stream));
        IGB_CUDA(cudaStreamSynchronize(dev->stream));
        if (dev->stats) { // SectionType::FramebufferHostUpdate (Statistics.h:50)
            float ms = 0.0f;
            cudaEventElapsedTime(&ms, dev->ev0, dev->ev1);
            dev->section.framebuffer_host_update_ms += ms;
            dev->section.framebuffer_host_updates += 1;
        }
        dev->film_mapped = true;
    }
    *data = dev->h_film;
    return IGB200_OK;
}

static int clear_image(igb200_device* dev, int img)
{
    const size_t bytes = (size_t)dev->film_w * dev->film_h * 3 * sizeof(float);
    if (img < 0) {
        if (dev->d_film.ptr) IGB_CUDA(cudaMemsetAsync(dev->d_film.ptr, 0, bytes, dev->stream));
        dev->iterations  = 0;
        dev->film_mapped = false;
    } else {
        igb200_device::Aov& a = dev->aov[img];
        if (a.d.ptr) IGB_CUDA(cudaMemsetAsync(a.d.ptr, 0, bytes, dev->stream));
        a.iterations = 0;
        a.mapped     = false;
    }
    return IGB200_OK;
}

int igb200_clear_framebuffer(igb200_device* dev)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_clear_framebuffer: NULL device");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    for (int img = -1; img < igb200_device::kAovCount; ++img) // clearAllAOVs, Device.cpp:1406-1411
        if (int rc = clear_image(dev, img)) return rc;
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    return IGB200_OK;
}

int igb200_clear_aov(igb200_device* dev, const char* name)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_clear_aov: NULL device");
    const int img = find_image(name);
    if (img == -2 || (img >= 0 && !dev->aov[img].d.ptr)) return fail(IGB200_E_INVALID, "igb200_clear_aov: unknown AOV '%s'", name);
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    if (int rc = clear_image(dev, img)) return rc;
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    return IGB200_OK;
}

int igb200_tonemap(igb200_device* dev, const igb200_tonemap_settings* ts, uint32_t* out_pixels)
{
    if (!dev || !ts || !out_pixels) return fail(IGB200_E_INVALID, "igb200_tonemap: NULL argument");
    const int img = find_image(ts->aov);
    if (img == -2 || (img >= 0 && !dev->aov[img].d.ptr)) return fail(IGB200_E_INVALID, "igb200_tonemap: unknown AOV '%s'", ts->aov);
    if (!dev->d_film.ptr) return fail(IGB200_E_STATE, "igb200_tonemap: nothing rendered yet");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    const uint64_t n     = (uint64_t)dev->film_w * dev->film_h;
    const uint64_t iters = img < 0 ? dev->iterations : dev->aov[img].iterations;
    const float*   in    = img < 0 ? dev->d_film.ptr : dev->aov[img].d.ptr;
    igb::TonemapParams p{};
    p.method          = ts->method;
    p.use_gamma       = ts->use_gamma;
    p.scale           = ts->scale * (iters > 0 ? 1.0f / (float)iters : 0.0f); // Device.cpp:1668,1677
    p.exposure_factor = ts->exposure_factor;
    p.exposure_offset = ts->exposure_offset;
    IGB_CUDA(dev->d_tonemap.ensure(n));
    IGB_CUDA(igb::launch_tonemap(in, dev->d_tonemap.ptr, n, p, dev->stream));
    IGB_CUDA(cudaMemcpyAsync(out_pixels, dev->d_tonemap.ptr, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, dev->stream));
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    return IGB200_OK;
}

int igb200_resize(igb200_device* dev, uint32_t width, uint32_t height)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_resize: NULL device");
    if (width == 0 || height == 0) return fail(IGB200_E_INVALID, "igb200_resize: empty film");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    dev->film_w = dev->film_h = 0; // force re-allocation + zero (Device.cpp:306-313)
    return ensure_film(dev, width, height);
}

int igb200_set_framebuffer(igb200_device* dev, const float* d_src, uint64_t count, uint64_t iterations)
{
    if (!dev || !d_src) return fail(IGB200_E_INVALID, "igb200_set_framebuffer: NULL argument");
    if (!dev->d_film.ptr || count != (uint64_t)dev->film_w * dev->film_h * 3)
        return fail(IGB200_E_INVALID, "igb200_set_framebuffer: size does not match the film");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    if (int rc = flush_wave(dev)) return rc;
    IGB_CUDA(igb::launch_film_combine(dev->d_film.ptr, d_src, count, true, dev->stream));
    IGB_CUDA(cudaStreamSynchronize(dev->stream));
    dev->iterations  = iterations;
    dev->film_mapped = false;
    return IGB200_OK;
}

int igb200_get_section_stats(igb200_device* dev, igb200_section_stats* out, int reset)
{
    if (!dev || !out) return fail(IGB200_E_INVALID, "igb200_get_section_stats: NULL argument");
    *out = dev->section;
    if (reset) dev->section = igb200_section_stats{};
    return IGB200_OK;
}

int igb200_get_render_stats(igb200_device* dev, igb200_render_stats* out)
{
    if (!dev || !out) return fail(IGB200_E_INVALID, "igb200_get_render_stats: NULL argument");
    *out = dev->render_stats;
    return IGB200_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// igb200_trace_rays_noalloc -- no counterpart in the reference.
//
// The host-buffer trace (igb200_trace_rays above) stages rays and hit columns in cudaMalloc'd buffers and uses
// cudaMalloc'd work counters.  Callers that may not allocate (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc) get this variant: the same traversal kernel with ALL of its scratch -- work counters,
// ray staging, the five hit columns -- in one file-scope device array of fixed size.  Every GPU that loads the library
// holds the array (kNoAllocBytes, ~3.25 MiB) until the process exits.
//
// `g_noalloc_scratch` below is a HOST-SIDE STAND-IN in host code, not a device address; the device address is resolved
// once per device with cudaGetSymbolAddress (that device current) and only that pointer is handed to kernels / memcpys.
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct NoAllocScratch {
    uint32_t work[2];                 // persistent-kernel work counter + retired-block counter: must be 0 at launch
    uint32_t _pad[62];                // keeps the staging areas 256-byte aligned
    float4   rays[2 * IGB200_NOALLOC_MAX_RAYS];
    int32_t  ent[IGB200_NOALLOC_MAX_RAYS];
    int32_t  prim[IGB200_NOALLOC_MAX_RAYS];
    float    t[IGB200_NOALLOC_MAX_RAYS];
    float    u[IGB200_NOALLOC_MAX_RAYS];
    float    v[IGB200_NOALLOC_MAX_RAYS];
};
constexpr size_t kNoAllocBytes = sizeof(NoAllocScratch);
static_assert(kNoAllocBytes == 256 + (size_t)IGB200_NOALLOC_MAX_RAYS * 52, "scratch layout");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int   kNoAllocMaxDevices = 64;
NoAllocScratch* g_noalloc_ptr[kNoAllocMaxDevices] = {}; // device address of the array on each GPU, resolved on first use
std::mutex      g_noalloc_mutex;

// Device address of the scratch array on `ordinal`; that device must be current.  No allocation: the array is part of
// the loaded module.
int noalloc_scratch(int ordinal, NoAllocScratch** out)
{
    if (ordinal < 0 || ordinal >= kNoAllocMaxDevices) return fail(IGB200_E_UNSUPPORTED, "igb200_trace_rays_noalloc: device ordinal %d", ordinal);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[ordinal]) {
        void*             p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(IGB200_E_CUDA, "igb200_trace_rays_noalloc: cudaGetSymbolAddress failed on device %d: %s", ordinal, cudaGetErrorString(e));
        g_noalloc_ptr[ordinal] = static_cast<NoAllocScratch*>(p);
    }
    *out = g_noalloc_ptr[ordinal];
    return IGB200_OK;
}
} // namespace

extern "C" int igb200_trace_rays_noalloc(igb200_device* dev, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit,
                                         int32_t* ent_id, int32_t* prim_id, float* t, float* u, float* v)
{
    if (!dev) return fail(IGB200_E_INVALID, "igb200_trace_rays_noalloc: NULL device");
    if (!dev->has_scene) return fail(IGB200_E_STATE, "igb200_trace_rays_noalloc: no scene assigned");
    if (n > IGB200_NOALLOC_MAX_RAYS)
        return fail(IGB200_E_UNSUPPORTED, "igb200_trace_rays_noalloc: %llu rays exceed the static scratch of %u rays; split the batch or use igb200_trace_rays",
                    (unsigned long long)n, (unsigned)IGB200_NOALLOC_MAX_RAYS);
    if (n == 0) return IGB200_OK;
    if (!rays) return fail(IGB200_E_INVALID, "igb200_trace_rays_noalloc: rays is NULL");
    IGB_CUDA(cudaSetDevice(dev->ordinal));
    NoAllocScratch* s = nullptr;
    if (int rc = noalloc_scratch(dev->ordinal, &s)) return rc;
    cudaStream_t st = dev->stream;
    // The array keeps the previous call's contents.  What must start at zero is cleared here (the kernel rewinds its
    // counters itself, but a call that faulted half way would leave them behind); the staging areas are overwritten for
    // [0, n) before they are read, and nothing beyond n is ever read back.
    IGB_CUDA(cudaMemsetAsync(s->work, 0, sizeof(s->work), st));
    IGB_CUDA(cudaMemcpyAsync(s->rays, rays, n * sizeof(igb200_ray), cudaMemcpyHostToDevice, st));
    const igb::TraceOutputs out{ ent_id ? s->ent : nullptr, prim_id ? s->prim : nullptr, t ? s->t : nullptr, u ? s->u : nullptr,
                                 v ? s->v : nullptr };
    // same kernel and same in-kernel direction normalisation as igb200_trace_rays (Interface::loadRayList, Device.cpp:588-592)
    IGB_CUDA(igb::launch_trace(dev->view, s->rays, n, ray_flags, any_hit, out, nullptr, s->work, dev->sm_count, st, true));
    if (ent_id) IGB_CUDA(cudaMemcpyAsync(ent_id, s->ent, n * 4, cudaMemcpyDeviceToHost, st));
    if (prim_id) IGB_CUDA(cudaMemcpyAsync(prim_id, s->prim, n * 4, cudaMemcpyDeviceToHost, st));
    if (t) IGB_CUDA(cudaMemcpyAsync(t, s->t, n * 4, cudaMemcpyDeviceToHost, st));
    if (u) IGB_CUDA(cudaMemcpyAsync(u, s->u, n * 4, cudaMemcpyDeviceToHost, st));
    if (v) IGB_CUDA(cudaMemcpyAsync(v, s->v, n * 4, cudaMemcpyDeviceToHost, st));
    IGB_CUDA(cudaStreamSynchronize(st));
    return IGB200_OK;
}
