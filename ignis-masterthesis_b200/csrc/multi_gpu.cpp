// multi_gpu.cpp -- multi-GPU entries of the C ABI (include/ignis_b200.h, "Multi-GPU (ABI v6)").
//
// The reference is single-device (src/runtime/device/Device.cpp:232-243, one `Target::device`); the only partitioning
// it has is the CPU driver's tile loop on TBB workers (src/artic/driver/mapping_cpu.art:694-706,
// src/artic/core/cpu_common.art:101-125).  Here the same idea runs across the GPUs of one NVSwitch box:
//   * scene + shading tables replicated on every device,
//   * a request split by iteration index or by interleaved 16-pixel tile rows (igb200_partition_share, device.cu),
//   * ONE ncclReduce(sum) of each accumulation image onto the root, on the communicator's stream, over NVLink.
// Written on top of the single-device C ABI only; NCCL is resolved at run time (dlopen) so that a process which already
// carries an NCCL (torch's, under torchrun) shares it and a plain `igcli --gpus N` picks up the system one.
#include "../../include/ignis_b200.h"

#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

// ---- minimal NCCL surface (ABI-stable since 2.x; matches /usr/include/nccl.h) ----
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclSum = 0 };
enum { ncclFloat32 = 7 };
struct Nccl {
    void* lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId*)                                                                = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                                         = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*)                                                 = nullptr;
    int (*CommDestroy)(ncclComm_t)                                                                   = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t)               = nullptr;
    int (*GroupStart)()                                                                              = nullptr;
    int (*GroupEnd)()                                                                                = nullptr;
    const char* (*GetErrorString)(int)                                                               = nullptr;
    int (*GetVersion)(int*)                                                                          = nullptr;
    std::string error;
};

Nccl& nccl()
{
    static Nccl       n;
    static std::once_flag once;
    std::call_once(once, [] {
        // RTLD_NOLOAD first: reuse the NCCL the process already carries (torch's wheel under torchrun)
        for (const char* name : { "libnccl.so.2", "libnccl.so" }) {
            n.lib = dlopen(name, RTLD_NOW | RTLD_NOLOAD);
            if (n.lib) break;
        }
        if (!n.lib)
            for (const char* name : { "libnccl.so.2", "libnccl.so" }) {
                n.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
                if (n.lib) break;
            }
        if (!n.lib) {
            n.error = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?");
            return;
        }
        auto sym = [&](const char* s) {
            void* p = dlsym(n.lib, s);
            if (!p && n.error.empty()) n.error = std::string("NCCL symbol missing: ") + s;
            return p;
        };
        n.GetUniqueId    = reinterpret_cast<decltype(n.GetUniqueId)>(sym("ncclGetUniqueId"));
        n.CommInitRank   = reinterpret_cast<decltype(n.CommInitRank)>(sym("ncclCommInitRank"));
        n.CommInitAll    = reinterpret_cast<decltype(n.CommInitAll)>(sym("ncclCommInitAll"));
        n.CommDestroy    = reinterpret_cast<decltype(n.CommDestroy)>(sym("ncclCommDestroy"));
        n.Reduce         = reinterpret_cast<decltype(n.Reduce)>(sym("ncclReduce"));
        n.GroupStart     = reinterpret_cast<decltype(n.GroupStart)>(sym("ncclGroupStart"));
        n.GroupEnd       = reinterpret_cast<decltype(n.GroupEnd)>(sym("ncclGroupEnd"));
        n.GetErrorString = reinterpret_cast<decltype(n.GetErrorString)>(sym("ncclGetErrorString"));
        n.GetVersion     = reinterpret_cast<decltype(n.GetVersion)>(sym("ncclGetVersion"));
    });
    return n;
}

// igb200_last_error() is owned by device.cu; errors raised here go through this hook (device.cu)
} // namespace
extern "C" int igb200_internal_fail(int code, const char* message);
namespace {

int failf(int code, const char* fmt, ...)
{
    char    buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    return igb200_internal_fail(code, buf);
}

#define MG_CUDA(call)                                                                                    \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess) return failf(IGB200_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e__)); \
    } while (0)
#define MG_NCCL(call)                                                                                        \
    do {                                                                                                     \
        int r__ = (call);                                                                                    \
        if (r__ != ncclSuccess) return failf(IGB200_E_CUDA, "%s failed: %s", #call, nccl().GetErrorString(r__)); \
    } while (0)

const char* const kImages[] = { "Color", "Direct Weights", "NEE Weights", "Normals", "Albedo", "Depth" };
constexpr int     kImageCount = 6;

} // namespace

struct igb200_comm {
    igb200_device* dev     = nullptr;
    ncclComm_t     comm    = nullptr;
    uint32_t       nranks  = 1, rank = 0;
    cudaStream_t   stream  = nullptr;
    cudaEvent_t    ev0 = nullptr, ev1 = nullptr;
    float*         scratch = nullptr; // warm-up buffer
    size_t         scratch_count = 0;
    bool           warmed  = false;
};

namespace {

int comm_init_common(igb200_comm* c)
{
    MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
    MG_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    MG_CUDA(cudaEventCreate(&c->ev0));
    MG_CUDA(cudaEventCreate(&c->ev1));
    return IGB200_OK;
}

// device pointer of image k on this rank, or nullptr when the technique in use does not carry it
float* image_ptr(igb200_device* dev, int k, uint64_t* iterations)
{
    float* p = nullptr;
    if (igb200_framebuffer(dev, kImages[k], 1, &p, iterations) != IGB200_OK) return nullptr;
    return p;
}

// Enqueues the reduces of every image in use on c->stream (callers wrap several communicators in one NCCL group).
int enqueue_reduce(igb200_comm* c, uint32_t root, size_t count, bool timed)
{
    if (timed) MG_CUDA(cudaEventRecord(c->ev0, c->stream));
    for (int k = 0; k < kImageCount; ++k) {
        uint64_t it = 0;
        float*   p  = image_ptr(c->dev, k, &it);
        if (!p) continue;
        MG_NCCL(nccl().Reduce(p, p, count, ncclFloat32, ncclSum, (int)root, c->comm, c->stream));
    }
    if (timed) MG_CUDA(cudaEventRecord(c->ev1, c->stream));
    return IGB200_OK;
}

int enqueue_warm(igb200_comm* c, uint32_t root, size_t count)
{
    if (c->scratch_count < count) {
        if (c->scratch) cudaFree(c->scratch);
        c->scratch = nullptr;
        MG_CUDA(cudaMalloc(&c->scratch, count * sizeof(float)));
        c->scratch_count = count;
    }
    MG_CUDA(cudaMemsetAsync(c->scratch, 0, count * sizeof(float), c->stream));
    MG_NCCL(nccl().Reduce(c->scratch, c->scratch, count, ncclFloat32, ncclSum, (int)root, c->comm, c->stream));
    return IGB200_OK;
}

// After the reduce: root's iteration counts cover the whole request, the other ranks start from zero again.
int finish_reduce(igb200_comm* c, uint32_t root, const uint64_t before[kImageCount], const bool present[kImageCount],
                  uint32_t total_iterations, uint32_t own_count, bool film_locked)
{
    if (c->rank == root) {
        for (int k = 0; k < kImageCount; ++k) {
            if (!present[k]) continue;
            // this rank's render call added own_count (0 for a locked film); the request as a whole adds total_iterations
            const bool     counted = !(k == 0 && film_locked);
            const uint64_t base    = before[k] - (counted ? own_count : 0);
            if (int rc = igb200_set_iteration_count(c->dev, kImages[k], base + (counted ? total_iterations : 0))) return rc;
        }
        return IGB200_OK;
    }
    return igb200_clear_framebuffer(c->dev);
}

size_t film_count(igb200_device* dev, uint32_t* w_out = nullptr, uint32_t* h_out = nullptr);

} // namespace

extern "C" int igb200_internal_film_size(igb200_device* dev, uint32_t* w, uint32_t* h);

namespace {
size_t film_count(igb200_device* dev, uint32_t* w_out, uint32_t* h_out)
{
    uint32_t w = 0, h = 0;
    igb200_internal_film_size(dev, &w, &h);
    if (w_out) *w_out = w;
    if (h_out) *h_out = h;
    return (size_t)w * h * 3;
}
} // namespace

extern "C" {

int igb200_comm_unique_id(uint8_t id[IGB200_COMM_ID_BYTES])
{
    if (!id) return failf(IGB200_E_INVALID, "igb200_comm_unique_id: NULL");
    Nccl& n = nccl();
    if (!n.error.empty()) return failf(IGB200_E_UNSUPPORTED, "%s", n.error.c_str());
    ncclUniqueId uid;
    MG_NCCL(n.GetUniqueId(&uid));
    static_assert(sizeof(uid) == IGB200_COMM_ID_BYTES, "ncclUniqueId size");
    std::memcpy(id, &uid, sizeof(uid));
    return IGB200_OK;
}

int igb200_comm_create(igb200_device* dev, const uint8_t id[IGB200_COMM_ID_BYTES], uint32_t nranks, uint32_t rank, igb200_comm** out)
{
    if (!dev || !id || !out || nranks == 0 || rank >= nranks) return failf(IGB200_E_INVALID, "igb200_comm_create: bad argument");
    *out = nullptr;
    Nccl& n = nccl();
    if (!n.error.empty()) return failf(IGB200_E_UNSUPPORTED, "%s", n.error.c_str());
    igb200_comm* c = new igb200_comm();
    c->dev = dev; c->nranks = nranks; c->rank = rank;
    if (int rc = comm_init_common(c)) { igb200_comm_destroy(c); return rc; }
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof(uid));
    const int r = n.CommInitRank(&c->comm, (int)nranks, uid, (int)rank);
    if (r != ncclSuccess) {
        igb200_comm_destroy(c);
        return failf(IGB200_E_CUDA, "ncclCommInitRank failed: %s", n.GetErrorString(r));
    }
    *out = c;
    return IGB200_OK;
}

void igb200_comm_destroy(igb200_comm* c)
{
    if (!c) return;
    cudaSetDevice(igb200_device_ordinal(c->dev));
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm) nccl().CommDestroy(c->comm);
    if (c->scratch) cudaFree(c->scratch);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int igb200_comm_render(igb200_comm* c, const igb200_render_settings* settings, uint32_t total_iterations, uint32_t mode)
{
    if (!c) return failf(IGB200_E_INVALID, "igb200_comm_render: NULL communicator");
    return igb200_render_partition(c->dev, settings, total_iterations, c->rank, c->nranks, mode);
}

int igb200_comm_reduce_film(igb200_comm* c, uint32_t root, uint32_t total_iterations, uint32_t mode, int warm, float* reduce_ms)
{
    if (!c || root >= c->nranks) return failf(IGB200_E_INVALID, "igb200_comm_reduce_film: bad argument");
    if (reduce_ms) *reduce_ms = 0.0f;
    MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
    const size_t count = film_count(c->dev);
    if (count == 0) return failf(IGB200_E_STATE, "igb200_comm_reduce_film: no film (render or resize first)");
    uint32_t first = 0, own = 0, tr = 0, tc = 1;
    if (int rc = igb200_partition_share(total_iterations, c->rank, c->nranks, mode, &first, &own, &tr, &tc)) return rc;
    uint64_t before[kImageCount] = {};
    bool     present[kImageCount] = {};
    for (int k = 0; k < kImageCount; ++k) present[k] = image_ptr(c->dev, k, &before[k]) != nullptr;
    const bool film_locked = present[3]; // infobuffer technique: "Normals" exists and the film is locked (Device.cpp:1460-1472)
    if (c->nranks > 1) {
        if (warm && !c->warmed) {
            if (int rc = enqueue_warm(c, root, count)) return rc;
            MG_CUDA(cudaStreamSynchronize(c->stream));
            c->warmed = true;
        }
        if (int rc = enqueue_reduce(c, root, count, true)) return rc;
        MG_CUDA(cudaStreamSynchronize(c->stream));
        if (reduce_ms) cudaEventElapsedTime(reduce_ms, c->ev0, c->ev1);
    }
    return finish_reduce(c, root, before, present, total_iterations, own, film_locked);
}

} // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// One process, N devices
// ---------------------------------------------------------------------------------------------------------------
struct igb200_group {
    std::vector<igb200_comm*> members; // members[i]->rank == i
    igb200_render_stats       stats{};
    float                     reduce_ms = 0.0f;
};

namespace {

// Runs fn(i) for every member on its own host thread (member 0 on the caller's); returns the first error with its message.
template <typename F>
int for_each_member(igb200_group* g, F fn)
{
    const size_t             n = g->members.size();
    std::vector<int>         rc(n, IGB200_OK);
    std::vector<std::string> msg(n);
    auto run = [&](size_t i) {
        rc[i] = fn(i);
        if (rc[i] != IGB200_OK) msg[i] = igb200_last_error(); // thread-local in device.cu
    };
    std::vector<std::thread> th;
    for (size_t i = 1; i < n; ++i) th.emplace_back(run, i);
    run(0);
    for (auto& t : th) t.join();
    for (size_t i = 0; i < n; ++i)
        if (rc[i] != IGB200_OK) return igb200_internal_fail(rc[i], ("device group member " + std::to_string(i) + ": " + msg[i]).c_str());
    return IGB200_OK;
}

} // namespace

extern "C" {

int igb200_group_create(const int32_t* devices, uint32_t count, int acquire_stats, igb200_group** out)
{
    if (!out || count == 0) return failf(IGB200_E_INVALID, "igb200_group_create: bad argument");
    *out = nullptr;
    Nccl& n = nccl();
    if (count > 1 && !n.error.empty()) return failf(IGB200_E_UNSUPPORTED, "%s", n.error.c_str());
    igb200_group* g = new igb200_group();
    std::vector<int> ordinals(count);
    for (uint32_t i = 0; i < count; ++i) {
        ordinals[i] = devices ? devices[i] : (int)i;
        igb200_setup   setup{ ordinals[i], acquire_stats, 0, 0 };
        igb200_device* dev = nullptr;
        if (int rc = igb200_create(&setup, &dev)) { igb200_group_destroy(g); return rc; }
        igb200_comm* c = new igb200_comm();
        c->dev = dev; c->nranks = count; c->rank = i;
        g->members.push_back(c);
        if (int rc = comm_init_common(c)) { igb200_group_destroy(g); return rc; }
    }
    if (count > 1) {
        std::vector<ncclComm_t> comms(count, nullptr);
        const int r = n.CommInitAll(comms.data(), (int)count, ordinals.data());
        if (r != ncclSuccess) {
            igb200_group_destroy(g);
            return failf(IGB200_E_CUDA, "ncclCommInitAll failed: %s", n.GetErrorString(r));
        }
        for (uint32_t i = 0; i < count; ++i) g->members[i]->comm = comms[i];
    }
    *out = g;
    return IGB200_OK;
}

void igb200_group_destroy(igb200_group* g)
{
    if (!g) return;
    for (igb200_comm* c : g->members) {
        igb200_device* dev = c->dev;
        igb200_comm_destroy(c);
        igb200_destroy(dev);
    }
    delete g;
}

uint32_t igb200_group_size(const igb200_group* g) { return g ? (uint32_t)g->members.size() : 0; }

igb200_device* igb200_group_device(igb200_group* g, uint32_t index)
{
    return g && index < g->members.size() ? g->members[index]->dev : nullptr;
}

int igb200_group_assign_scene(igb200_group* g, const igb200_scene* scene)
{
    if (!g) return failf(IGB200_E_INVALID, "igb200_group_assign_scene: NULL group");
    // replicated: every member builds its own BVH8 copy (host build is the same work N times; the uploads overlap)
    return for_each_member(g, [&](size_t i) { return igb200_assign_scene(g->members[i]->dev, scene); });
}

int igb200_group_assign_shading(igb200_group* g, const igb200_shading* shading)
{
    if (!g) return failf(IGB200_E_INVALID, "igb200_group_assign_shading: NULL group");
    return for_each_member(g, [&](size_t i) { return igb200_assign_shading(g->members[i]->dev, shading); });
}

int igb200_group_resize(igb200_group* g, uint32_t w, uint32_t h)
{
    if (!g) return failf(IGB200_E_INVALID, "igb200_group_resize: NULL group");
    return for_each_member(g, [&](size_t i) { return igb200_resize(g->members[i]->dev, w, h); });
}

int igb200_group_clear_framebuffer(igb200_group* g)
{
    if (!g) return failf(IGB200_E_INVALID, "igb200_group_clear_framebuffer: NULL group");
    return for_each_member(g, [&](size_t i) { return igb200_clear_framebuffer(g->members[i]->dev); });
}

int igb200_group_render(igb200_group* g, const igb200_render_settings* settings, uint32_t total_iterations, uint32_t mode)
{
    if (!g || !settings) return failf(IGB200_E_INVALID, "igb200_group_render: NULL argument");
    const uint32_t n = (uint32_t)g->members.size();
    if (settings->rays) mode = IGB200_PARTITION_ITERATIONS;
    // resolve AUTO once so that render and reduce agree
    if (mode == IGB200_PARTITION_AUTO) mode = total_iterations >= n ? IGB200_PARTITION_ITERATIONS : IGB200_PARTITION_TILE_ROWS;
    if (int rc = for_each_member(g, [&](size_t i) { return igb200_comm_render(g->members[i], settings, total_iterations, mode); })) return rc;

    g->stats     = igb200_render_stats{};
    g->reduce_ms = 0.0f;
    for (igb200_comm* c : g->members) {
        igb200_render_stats st{};
        igb200_get_render_stats(c->dev, &st);
        g->stats.camera_rays += st.camera_rays;
        g->stats.bounce_rays += st.bounce_rays;
        g->stats.shadow_rays += st.shadow_rays;
        g->stats.kernel_launches += st.kernel_launches;
        g->stats.rounds    = std::max(g->stats.rounds, st.rounds);
        g->stats.device_ms = std::max(g->stats.device_ms, st.device_ms);
    }
    if (n == 1) return IGB200_OK;

    // members that own nothing (iterations < members, or more members than tile rows) may not have a film yet
    uint32_t w = 0, h = 0;
    for (igb200_comm* c : g->members) {
        uint32_t cw = 0, ch = 0;
        film_count(c->dev, &cw, &ch);
        w = std::max(w, cw); h = std::max(h, ch);
    }
    for (igb200_comm* c : g->members)
        if (film_count(c->dev) != (size_t)w * h * 3)
            if (int rc = igb200_resize(c->dev, w, h)) return rc;
    const size_t count = (size_t)w * h * 3;

    uint64_t before[kImageCount] = {};
    bool     present[kImageCount] = {};
    for (int k = 0; k < kImageCount; ++k) present[k] = image_ptr(g->members[0]->dev, k, &before[k]) != nullptr;
    uint32_t first = 0, own0 = 0, tr = 0, tc = 1;
    if (int rc = igb200_partition_share(total_iterations, 0, n, mode, &first, &own0, &tr, &tc)) return rc;

    Nccl& nc = nccl();
    if (!g->members[0]->warmed) { // communicator first use happens outside the timed reduce
        MG_NCCL(nc.GroupStart());
        for (igb200_comm* c : g->members) {
            MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
            if (int rc = enqueue_warm(c, 0, count)) { nc.GroupEnd(); return rc; }
        }
        MG_NCCL(nc.GroupEnd());
        for (igb200_comm* c : g->members) {
            MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
            MG_CUDA(cudaStreamSynchronize(c->stream));
            c->warmed = true;
        }
    }
    for (igb200_comm* c : g->members) {
        MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
        MG_CUDA(cudaEventRecord(c->ev0, c->stream));
    }
    MG_NCCL(nc.GroupStart());
    for (igb200_comm* c : g->members) {
        MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
        if (int rc = enqueue_reduce(c, 0, count, false)) { nc.GroupEnd(); return rc; }
    }
    MG_NCCL(nc.GroupEnd());
    for (igb200_comm* c : g->members) {
        MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
        MG_CUDA(cudaEventRecord(c->ev1, c->stream));
    }
    for (igb200_comm* c : g->members) {
        MG_CUDA(cudaSetDevice(igb200_device_ordinal(c->dev)));
        MG_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, c->ev0, c->ev1);
        g->reduce_ms = std::max(g->reduce_ms, ms);
    }
    const bool film_locked = present[3];
    for (igb200_comm* c : g->members)
        if (int rc = finish_reduce(c, 0, before, present, total_iterations, own0, film_locked)) return rc;
    MG_CUDA(cudaSetDevice(igb200_device_ordinal(g->members[0]->dev)));
    return IGB200_OK;
}

int igb200_group_trace_rays(igb200_group* g, const igb200_ray* rays, uint64_t n, uint32_t ray_flags, int any_hit, int32_t* ent_id,
                            int32_t* prim_id, float* t, float* u, float* v)
{
    if (!g) return failf(IGB200_E_INVALID, "igb200_group_trace_rays: NULL group");
    if (n && !rays) return failf(IGB200_E_INVALID, "igb200_group_trace_rays: rays is NULL");
    const uint64_t m = g->members.size();
    return for_each_member(g, [&](size_t i) {
        const uint64_t lo = n * i / m, hi = n * (i + 1) / m; // contiguous ray ranges; results land in place, no collective
        if (hi == lo) return (int)IGB200_OK;
        return igb200_trace_rays(g->members[i]->dev, rays + lo, hi - lo, ray_flags, any_hit, ent_id ? ent_id + lo : nullptr,
                                 prim_id ? prim_id + lo : nullptr, t ? t + lo : nullptr, u ? u + lo : nullptr, v ? v + lo : nullptr);
    });
}

int igb200_group_get_render_stats(igb200_group* g, igb200_render_stats* out, float* reduce_ms)
{
    if (!g || !out) return failf(IGB200_E_INVALID, "igb200_group_get_render_stats: NULL argument");
    *out = g->stats;
    if (reduce_ms) *reduce_ms = g->reduce_ms;
    return IGB200_OK;
}

} // extern "C"
