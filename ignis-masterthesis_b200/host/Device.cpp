// Device.cpp -- forwards IG::Device to libignis_b200.so.  In the reference this file is 2000 lines of AnyDSL runtime
// plumbing (src/runtime/device/Device.cpp); everything it does on the hot path now happens behind the C ABI.
#include "Device.h"

#include <cstdio>
#include <limits>
#include <cstdlib>
#include <algorithm>
#include <cmath>
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
    igb200_render_stats st{};
    if (count == 1 && !settings.rays && igb200_group_size(mGroup) == 1) {
        // Runtime::step() on one GPU: pipelined steps -- the call returns when the iteration's rays have been generated, its deep
        // paths finish in the rounds of the next step (or in flush(); every framebuffer access flushes by itself)
        if (igb200_render_pipelined(mHandle, &s) != IGB200_OK) fatal("Device::render");
        igb200_get_render_stats(mHandle, &st);
    } else {
        // one member: plain igb200_render_iterations on it; several: partition (AUTO: by iteration when there are enough of them,
        // else by tile rows), render concurrently, ncclReduce onto member 0
        if (igb200_group_render(mGroup, &s, (uint32_t)count, IGB200_PARTITION_AUTO) != IGB200_OK) fatal("Device::render");
        st = renderStats();
    }
    // Quantity::{Camera,Bounce,Shadow}RayCount accumulate over the run (devices with AcquireStats render synchronously: exact)
    mRays[0] += st.camera_rays;
    mRays[1] += st.bounce_rays;
    mRays[2] += st.shadow_rays;
}

void Device::flush()
{
    if (igb200_flush(mHandle) != IGB200_OK) error("Device::flush");
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
    if (igb200_group_size(mGroup) == 1) igb200_get_render_stats(mHandle, &st); // also covers pipelined steps, which bypass the group
    else igb200_group_get_render_stats(mGroup, &st, nullptr);                  // summed over the members
    return st;
}

float Device::lastFilmReduceMs() const
{
    igb200_render_stats st{};
    float               ms = 0.0f;
    igb200_group_get_render_stats(mGroup, &st, &ms);
    return ms;
}

#ifndef IG_B200_IN_REFERENCE_TREE
const Statistics* Device::getStatistics()
{
    if (!mSetup.AcquireStats) return nullptr; // Runtime.cpp:445-448 only asks when AcquireStats is set
    mStats.reset();
    static const ShaderType kKind[6] = { ShaderType::Device, ShaderType::RayGeneration, ShaderType::PrimaryTraversal, ShaderType::Device,
                                         ShaderType::Hit, ShaderType::SecondaryTraversal };
    for (uint32_t i = 0; i < igb200_group_size(mGroup); ++i) { // summed over the devices, as the reference sums over threads
        igb200_section_stats sec{};
        if (igb200_get_section_stats(igb200_group_device(mGroup, i), &sec, 0) != IGB200_OK) continue;
        for (int k = 0; k < 6; ++k) {
            if (k == 3) mStats.addSection(SectionType::GPUSortPrimary, sec.kernel_ms[k], sec.launches[k]);
            else mStats.addShader(kKind[k], sec.kernel_ms[k], sec.launches[k]);
        }
        mStats.addSection(SectionType::FramebufferHostUpdate, sec.framebuffer_host_update_ms, sec.framebuffer_host_updates);
    }
    mStats.increase(Quantity::CameraRayCount, mRays[0]);
    mStats.increase(Quantity::BounceRayCount, mRays[1]);
    mStats.increase(Quantity::ShadowRayCount, mRays[2]);
    return &mStats;
}
#else
// Inside the reference tree its own Statistics class is used: its shader timers are wall-clock Timer objects driven by
// begin/endShaderLaunch, which have no meaning for kernels timed on the device, so only the ray counters are filled.
const Statistics* Device::getStatistics()
{
    if (!mSetup.AcquireStats) return nullptr;
    static Statistics stats;
    stats.reset();
    stats.increase(Quantity::CameraRayCount, mRays[0]);
    stats.increase(Quantity::BounceRayCount, mRays[1]);
    stats.increase(Quantity::ShadowRayCount, mRays[2]);
    return &stats;
}
#endif

// entrypoints/imageinfo.art:7-170 on the host copy of the image (a utility off the ray path; luminance = Y of linear sRGB)
ImageInfoOutput Device::imageinfo(const ImageInfoSettings& st)
{
    ImageInfoOutput   out{};
    const AOVAccessor acc = getFramebufferForHost(st.AOV ? st.AOV : "");
    const size_t      w = mWidth, h = mHeight, size = w * h;
    if (!acc.Data || size == 0) return out;
    auto clean = [](float a) { return std::isfinite(a) ? a : 0.0f; };
    auto lum   = [&](size_t i) {
        const float* p = acc.Data + 3 * i;
        return (0.2126729f * clean(p[0]) + 0.7151522f * clean(p[1]) + 0.0721750f * clean(p[2])) * st.Scale;
    };
    std::vector<float> L(size);
    double             sum = 0;
    out.Min = std::numeric_limits<float>::max();
    out.Max = -std::numeric_limits<float>::max();
    for (size_t i = 0; i < size; ++i) {
        L[i]    = lum(i);
        out.Min = std::min(out.Min, L[i]);
        out.Max = std::max(out.Max, L[i]);
        sum += L[i];
    }
    out.Average = (float)(sum / (double)size);
    if (w > 10 && h > 10) { // 3 x 3 window percentiles: 2nd smallest / median / 2nd largest, reduced over the image
        float  soft_min = std::numeric_limits<float>::max(), soft_max = 0;
        double median   = 0;
        for (size_t y = 1; y + 1 < h; ++y)
            for (size_t x = 1; x + 1 < w; ++x) {
                float win[9];
                for (int i = 0; i < 3; ++i)
                    for (int j = 0; j < 3; ++j) win[3 * i + j] = L[(y + i - 1) * w + (x + j - 1)];
                std::sort(win, win + 9);
                if (win[1] >= 0) soft_min = std::min(soft_min, win[1]);
                if (win[7] >= 0) soft_max = std::max(soft_max, win[7]);
                if (win[4] >= 0) median += win[4];
            }
        out.SoftMin = soft_min;
        out.SoftMax = soft_max;
        out.Median  = (float)(median / (double)((w - 2) * (h - 2)));
    } else {
        out.SoftMin = out.Min;
        out.SoftMax = out.Max;
        out.Median  = out.Average;
    }
    if (st.AcquireErrorStats)
        for (size_t i = 0; i < 3 * size; ++i) {
            const float c = acc.Data[i];
            if (std::isnan(c)) ++out.NaNCount;
            else if (std::isinf(c)) ++out.InfCount;
            if (std::signbit(c)) ++out.NegCount;
        }
    if (st.AcquireHistogram && st.Bins > 0) {
        const float start  = out.Min < out.Max ? out.Min : 0.0f;
        const float end    = out.Min < out.SoftMax ? out.SoftMax : (out.Min < out.Max ? out.Max : start + 1);
        const float range  = std::max(0.0f, end - start);
        const float factor = range > 0 ? (float)st.Bins / range : 0.0f;
        auto        bin    = [&](float v) { return (size_t)std::clamp((long)((v - start) * factor), 0l, (long)st.Bins - 1); };
        for (int* hst : { st.HistogramR, st.HistogramG, st.HistogramB, st.HistogramL })
            if (hst) std::fill(hst, hst + st.Bins, 0);
        for (size_t i = 0; i < size; ++i) {
            const float* p = acc.Data + 3 * i;
            if (st.HistogramR) ++st.HistogramR[bin(clean(p[0]) * st.Scale)];
            if (st.HistogramG) ++st.HistogramG[bin(clean(p[1]) * st.Scale)];
            if (st.HistogramB) ++st.HistogramB[bin(clean(p[2]) * st.Scale)];
            if (st.HistogramL) ++st.HistogramL[bin(L[i])];
        }
    }
    return out;
}

GlareOutput Device::evaluateGlare(uint32_t*, const GlareSettings&)
{
    static bool warned = false;
    if (!warned) std::fprintf(stderr, "[WARNING] Device::evaluateGlare: not available in the B200 backend (entrypoints/glare.art is off the ray path)\n");
    warned = true;
    return GlareOutput{};
}

void Device::bake(const ShaderOutput<void*>&, const std::vector<std::string>*, float* output)
{
    // The reference bakes by running a JIT-compiled texture shader on the device (Device.cpp:1690-1742); this backend has no JIT.
    // Environment maps and skies are baked by the loader side instead (host/SkyModel.cpp, sceneio.bake_image_texture).
    std::fprintf(stderr, "[ERROR] Device::bake: texture shaders are not executable in the B200 backend; output left untouched\n");
    (void)output;
}

void Device::tonemap(uint32_t* out, const TonemapSettings& s)
{
    igb200_tonemap_settings t{ s.AOV, (uint32_t)s.Method, s.UseGamma ? 1 : 0, s.Scale, s.ExposureFactor, s.ExposureOffset };
    if (igb200_tonemap(mHandle, &t, out) != IGB200_OK) error("Device::tonemap");
}
} // namespace IG
