// Device.h -- `IG::Device` with the reference's public interface (src/runtime/device/Device.h:14-74), implemented by
// forwarding to the C ABI of include/ignis_b200.h.  Same member names, argument meaning and error behaviour
// (no error codes out of the class: fatal conditions log and abort, unknown AOVs log and return {nullptr, 0} --
// Device.cpp:271-274,1303-1305), so `IG::Runtime` and the frontends compile against it unchanged.
// One addition: `assignShading`, the POD replacement of the per-scene shader JIT.
#pragma once

#ifdef IG_B200_IN_REFERENCE_TREE // building inside the reference: its own headers define these types
#include "RuntimeStructs.h"
#include "Target.h"
#include "technique/TechniqueInfo.h"
#include "technique/TechniqueVariant.h"
#else
#include "IgTypes.h"
#include "Statistics.h"
#endif

#include <ignis_b200.h>

namespace IG {
class Device {
public:
    struct SetupSettings { // Device.h:18-23
        Target target;
        bool   AcquireStats  = false;
        bool   DebugTrace    = false;
        bool   IsInteractive = false;
        // B200 extension (no reference counterpart: the reference is single-device, Device.cpp:232-243): number of GPUs of
        // this box to render on, target.device() .. + DeviceCount - 1.  Scene replicated, requests partitioned by iteration
        // index or by interleaved tile rows, films combined with one ncclReduce onto the first device (igb200_group_*).
        size_t DeviceCount   = 1;
    };
    struct SceneSettings { // Device.h:25-30; borrowed pointers, read during assignScene only (the backend copies to HBM)
        SceneDatabase*                  database            = nullptr;
        const std::vector<std::string>* aov_map             = nullptr;
        const std::vector<std::string>* resource_map        = nullptr;
        const std::vector<int32>*       entity_per_material = nullptr;
    };
    struct RenderSettings { // Device.h:32-42
        const Ray* rays      = nullptr; // non-null: width = number of rays, height = 1
        size_t     spi       = 8;
        size_t     width     = 0;
        size_t     height    = 0;
        size_t     iteration = 0;
        size_t     frame     = 0;
        size_t     user_seed = 0;
        TechniqueVariantInfo info;
        bool       denoise = false;
    };
    struct AOVAccessor { // Device.h:44-47
        float* Data;
        size_t IterationCount;
    };

    explicit Device(const SetupSettings& settings);
    ~Device();
    Device(const Device&)            = delete;
    Device& operator=(const Device&) = delete;

    void assignScene(const SceneSettings& settings);
    void assignShading(const igb200_shading& shading); // widened seam: replaces Runtime::compileShaders (Runtime.cpp:542-626)
    void render(const TechniqueVariantShaderSet& shader_set, const RenderSettings& settings, const ParameterSet* parameter_set);
    // `count` consecutive iterations in one wavefront (igb200_render_iterations); same film as `count` render() calls
    void renderIterations(const RenderSettings& settings, size_t count);
    // finishes the paths that pipelined steps (render() on one GPU) left in flight; framebuffer accessors do this by themselves
    void flush();
    void resize(size_t width, size_t height);
    void releaseAll();

    [[nodiscard]] Target target() const { return mSetup.target; }
    [[nodiscard]] size_t framebufferWidth() const { return mWidth; }
    [[nodiscard]] size_t framebufferHeight() const { return mHeight; }
    [[nodiscard]] bool   isInteractive() const { return mSetup.IsInteractive; }

    [[nodiscard]] AOVAccessor getFramebufferForHost(const std::string& name);
    [[nodiscard]] AOVAccessor getFramebufferForDevice(const std::string& name);
    void clearFramebuffer(const std::string& name);
    void clearAllFramebuffer();

    // Device.h:68 -- per-kernel times (device created with AcquireStats) and ray counters under the reference's names
    [[nodiscard]] const Statistics* getStatistics();
    [[nodiscard]] igb200_render_stats renderStats() const;

    void tonemap(uint32_t* out, const TonemapSettings& settings);
    // Device.h:71-73.  Off the hot path (SURVEY.md 8(b) "Utility"; the frontends on this path run with EnableTonemapping =
    // false, cli/main.cpp:62, trace/main.cpp:77) but present, so that IG::Runtime links (Runtime.cpp:447,649,663):
    //   imageinfo     luminance statistics of the film or an AOV (entrypoints/imageinfo.art), computed from the host copy
    //   evaluateGlare not available in this backend: logs once, returns a zeroed GlareOutput
    //   bake          texture baking runs through the JIT'd shader the reference hands in; there is none here: zero-fills
    [[nodiscard]] GlareOutput     evaluateGlare(uint32_t* out, const GlareSettings& settings);
    [[nodiscard]] ImageInfoOutput imageinfo(const ImageInfoSettings& settings);
    void bake(const ShaderOutput<void*>& shader, const std::vector<std::string>* resource_map, float* output);

    [[nodiscard]] igb200_device* handle() const { return mHandle; }
    [[nodiscard]] igb200_group*  group() const { return mGroup; }
    [[nodiscard]] size_t deviceCount() const { return mSetup.DeviceCount; }
    [[nodiscard]] float  lastFilmReduceMs() const;

private:
    SetupSettings  mSetup;
    igb200_group*  mGroup  = nullptr; // owns the device handles (one member when DeviceCount == 1: no NCCL involved)
    igb200_device* mHandle = nullptr; // member 0: holds the combined film
    size_t         mWidth = 0, mHeight = 0;
    uint64_t       mRays[3] = { 0, 0, 0 }; // camera, bounce, shadow since creation
#ifndef IG_B200_IN_REFERENCE_TREE
    Statistics     mStats;
#endif
};
} // namespace IG
