// Runtime.h -- the slice of `IG::Runtime` (src/runtime/Runtime.h, Runtime.cpp:121-448) that the igcli / igtrace
// frontends call, over IG::Device.  Scene loading reads the reference's scene JSON through SceneLoader.cpp (tables instead
// of generated shader source) or an `.igdb` (SceneFile.h: the loader's output, stored); everything after loading follows the
// reference's control flow.
#pragma once

#include "SceneFile.h"

#include <memory>
#include <optional>

namespace IG {
struct RuntimeOptions { // src/runtime/RuntimeSettings.h (the fields these frontends set)
    int    DeviceOrdinal = 0;
    int    DeviceCount   = 1;     // --gpus N: GPUs DeviceOrdinal .. + N - 1 (B200 extension, see Device::SetupSettings)
    bool   IsTracer      = false; // igtrace: list emitter, film = N x 1
    bool   AcquireStats  = false;
    size_t Seed          = 0;
    size_t SPI           = 0; // 0 = recommendSPI (Runtime.cpp:61-69)
    std::optional<std::pair<uint32_t, uint32_t>> OverrideFilmSize;
    std::string OverrideTechnique; // "", "path", "debug", "infobuffer"  (ProgramOptions --technique)
    int    DebugMode = 0;          // __debug_mode of the debug technique (debugtracer.art:27-177)
};

class Runtime {
public:
    explicit Runtime(const RuntimeOptions& opts);
    ~Runtime();

    bool loadFromFile(const std::string& path); // Runtime.cpp:157-166

    void step(bool ignoreDenoiser = false);                  // Runtime.cpp:294-349: one iteration
    void steps(size_t count);                                // `count` iterations in one wavefront (same film)
    void trace(const std::vector<Ray>& rays);                // Runtime.cpp:351-373
    void trace(const std::vector<Ray>& rays, std::vector<float>& data); // :375-383
    void reset();                                            // :437-443
    void resizeFramebuffer(size_t width, size_t height);     // :403-409

    [[nodiscard]] Device::AOVAccessor getFramebufferForHost(const std::string& name) const { return mDevice->getFramebufferForHost(name); }
    [[nodiscard]] size_t currentIterationCount() const { return mCurrentIteration; }
    [[nodiscard]] size_t currentSampleCount() const { return mCurrentSampleCount; }
    [[nodiscard]] size_t samplesPerIteration() const { return mSamplesPerIteration; }
    [[nodiscard]] size_t framebufferWidth() const { return mFilmWidth; }
    [[nodiscard]] size_t framebufferHeight() const { return mFilmHeight; }
    [[nodiscard]] Device& device() { return *mDevice; }

private:
    Device::RenderSettings renderSettings(const Ray* rays, size_t count) const;

    RuntimeOptions          mOptions;
    std::unique_ptr<Device> mDevice;
    SceneFile               mScene;
    TechniqueVariantInfo    mInfo;
    bool                    mLoaded              = false;
    size_t                  mSamplesPerIteration = 1;
    size_t                  mFilmWidth = 0, mFilmHeight = 0;
    size_t                  mCurrentIteration = 0, mCurrentSampleCount = 0, mCurrentFrame = 0;
};
} // namespace IG
