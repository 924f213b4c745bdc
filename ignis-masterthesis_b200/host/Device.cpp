// Device.cpp -- forwards IG::Device to libignis_b200.so.  In the reference this file is 2000 lines of AnyDSL runtime
// plumbing (src/runtime/device/Device.cpp); everything it does on the hot path now happens behind the C ABI.
#include "Device.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace IG {
// reference: `static Interface* sInterface` + IG_ASSERT(sInterface == nullptr, ...) (Device.cpp:1540,1564)
static Device* sInstance = nullptr;

[[noreturn]] static void fatal(const char* where)
{
    // reference: IG_LOG(L_FATAL) << ... ; std::abort()   (Device.cpp:271-274)
    std::fprintf(stderr, "[FATAL] %s: %s\n", where, igb200_last_error());
    std::abort();
}
static void error(const char* where) { std::fprintf(stderr, "[ERROR] %s: %s\n", where, igb200_last_error()); }

Device::Device(const SetupSettings& settings)
    : mSetup(settings)
{
    if (sInstance != nullptr) {
        std::fprintf(stderr, "[FATAL] Device: Only a single instance allowed!\n");
        std::abort();
    }
    if (mSetup.DeviceCount == 0) mSetup.DeviceCount = 1;
    std::vector<int32_t> ordinals(mSetup.DeviceCount);
    for (size_t i = 0; i < ordinals.size(); ++i) ordinals[i] = (int32_t)(settings.target.device() + i);
    if (igb200_group_create(ordinals.data(), (uint32_t)ordinals.size(), settings.AcquireStats ? 1 : 0, &mGroup) != IGB200_OK) fatal("Device");
    mHandle   = igb200_group_device(mGroup, 0);
    sInstance = this;
}

Device::~Device()
{
    igb200_group_destroy(mGroup);
    sInstance = nullptr;
}

void Device::releaseAll() {} // Device.cpp:948-952 frees host-side staging; the backend keeps none after assign_scene

void Device::assignScene(const SceneSettings& settings)
{
    const SceneDatabase& db = *settings.database;
    igb200_scene sc{};
    if (auto it = db.DynTables.find("shapes"); it != db.DynTables.end()) {
        const DynTable& shapes = it->second;
        sc.shape_lookups = reinterpret_cast<const igb200_lookup_entry*>(shapes.lookups().data());
        sc.shape_count   = shapes.entryCount();
        sc.shape_data    = shapes.data().data();
        sc.shape_bytes   = shapes.data().size();
    }
    if (auto it = db.FixTables.find("entities"); it != db.FixTables.end()) {
        sc.entities     = reinterpret_cast<const float*>(it->second.data().data());
        sc.entity_count = it->second.entryCount();
    }
    static const char* const kProviders[2] = { "trimesh", "sphere" }; // LoaderShape.cpp:20-38
    for (int k = 0; k < 2; ++k)
        if (auto it = db.SceneBVHs.find(kProviders[k]); it != db.SceneBVHs.end()) {
            sc.leaves[k]     = reinterpret_cast<const igb200_entity_leaf*>(it->second.Leaves.data());
            sc.leaf_count[k] = it->second.Leaves.size() / sizeof(igb200_entity_leaf);
        }
    std::memcpy(sc.scene_bbox + 0, db.SceneBBox.min.data(), 12);
    std::memcpy(sc.scene_bbox + 3, db.SceneBBox.max.data(), 12);
    sc.material_count = (uint32_t)db.MaterialCount;
    if (igb200_group_assign_scene(mGroup, &sc) != IGB200_OK) fatal("Device::assignScene");
}

void Device::assignShading(const igb200_shading& shading)
{
    if (igb200_group_assign_shading(mGroup, &shading) != IGB200_OK) fatal("Device::assignShading");
}

static igb200_render_settings to_abi(const Device::RenderSettings& rs)
{
    igb200_render_settings s{};
    s.rays      = reinterpret_cast<const igb200_ray*>(rs.rays);
    s.ray_count = rs.rays ? rs.width : 0;
    s.spi       = (uint32_t)rs.spi;
    s.width     = (uint32_t)rs.width;
    s.height    = (uint32_t)(rs.rays ? 1 : rs.height);
    s.iteration = (uint32_t)rs.iteration;
    s.frame     = (uint32_t)rs.frame;
    s.user_seed = (uint32_t)rs.user_seed;
    return s;
}

void Device::render(const TechniqueVariantShaderSet&, const RenderSettings& settings, const ParameterSet*)
{
    renderIterations(settings, 1);
}

void Device::renderIterations(const RenderSettings& settings, size_t count)
{
    const igb200_render_settings s = to_abi(settings);
    if (settings.rays) { // Interface::setupFramebuffer follows the ray count in tracer mode
        mWidth  = settings.width;
        mHeight = 1;
    }
    // one member: plain igb200_render_iterations on it; several: partition (AUTO: by iteration when there are enough of them,
    // else by tile rows), render concurrently, ncclReduce onto member 0
    if (igb200_group_render(mGroup, &s, (uint32_t)count, IGB200_PARTITION_AUTO) != IGB200_OK) fatal("Device::render");
}

void Device::resize(size_t width, size_t height)
{
    if (igb200_group_resize(mGroup, (uint32_t)width, (uint32_t)height) != IGB200_OK) fatal("Device::resize");
    mWidth  = width;
    mHeight = height;
}

Device::AOVAccessor Device::getFramebufferForHost(const std::string& name)
{
    float*   data  = nullptr;
    uint64_t iters = 0;
    if (igb200_framebuffer(mHandle, name.c_str(), 0, &data, &iters) != IGB200_OK) {
        error("Device::getFramebufferForHost"); // unknown AOV: Device.cpp:1303-1305
        return AOVAccessor{ nullptr, 0 };
    }
    return AOVAccessor{ data, (size_t)iters };
}

Device::AOVAccessor Device::getFramebufferForDevice(const std::string& name)
{
    float*   data  = nullptr;
    uint64_t iters = 0;
    if (igb200_framebuffer(mHandle, name.c_str(), 1, &data, &iters) != IGB200_OK) {
        error("Device::getFramebufferForDevice");
        return AOVAccessor{ nullptr, 0 };
    }
    return AOVAccessor{ data, (size_t)iters };
}

void Device::clearFramebuffer(const std::string& name)
{
    if (igb200_clear_aov(mHandle, name.c_str()) != IGB200_OK) error("Device::clearFramebuffer");
}
void Device::clearAllFramebuffer()
{
    if (igb200_group_clear_framebuffer(mGroup) != IGB200_OK) error("Device::clearAllFramebuffer");
}

igb200_render_stats Device::renderStats() const
{
    igb200_render_stats st{};
    igb200_group_get_render_stats(mGroup, &st, nullptr); // summed over the members
    return st;
}

float Device::lastFilmReduceMs() const
{
    igb200_render_stats st{};
    float               ms = 0.0f;
    igb200_group_get_render_stats(mGroup, &st, &ms);
    return ms;
}

void Device::tonemap(uint32_t* out, const TonemapSettings& s)
{
    igb200_tonemap_settings t{ s.AOV, (uint32_t)s.Method, s.UseGamma ? 1 : 0, s.Scale, s.ExposureFactor, s.ExposureOffset };
    if (igb200_tonemap(mHandle, &t, out) != IGB200_OK) error("Device::tonemap");
}
} // namespace IG
