#include "Runtime.h"
#include "SceneLoader.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace IG {
// Runtime.cpp:61-69 (GPU target: spi_f = 8)
static size_t recommendSPI(size_t width, size_t height, bool interactive)
{
    size_t spi_f = 8;
    if (interactive) spi_f /= 2;
    const size_t spi = (size_t)std::ceil(spi_f / ((width / 1000.0f) * (height / 1000.0f)));
    return std::max<size_t>(1, std::min<size_t>(64, spi));
}

Runtime::Runtime(const RuntimeOptions& opts)
    : mOptions(opts)
{
    Device::SetupSettings setup;
    setup.target       = Target(opts.DeviceOrdinal);
    setup.AcquireStats = opts.AcquireStats;
    setup.DeviceCount  = (size_t)std::max(1, opts.DeviceCount);
    mDevice            = std::make_unique<Device>(setup);
}

Runtime::~Runtime() = default;

bool Runtime::loadFromFile(const std::string& path)
{
    std::string err;
    // scene JSON goes through the loader of this directory (SceneLoader.cpp); `.igdb` is the loader's output, stored
    const size_t ow = mOptions.OverrideFilmSize ? mOptions.OverrideFilmSize->first : 0, oh = mOptions.OverrideFilmSize ? mOptions.OverrideFilmSize->second : 0;
    if (!loadAnyScene(path, mScene, err, ow, oh)) {
        std::fprintf(stderr, "[ERROR] %s\n", err.c_str());
        return false;
    }
    igb200_technique& tech = mScene.shading.technique;
    if (mOptions.OverrideTechnique == "debug") {
        tech = igb200_technique{ IGB200_TECH_DEBUG, 1, 1, 0.0f, 0, mOptions.DebugMode, 0u, 0 };
    } else if (mOptions.OverrideTechnique == "infobuffer") {
        tech.type = IGB200_TECH_INFOBUFFER;
    } else if (mOptions.OverrideTechnique == "path") {
        tech.type = IGB200_TECH_PATH;
    } else if (!mOptions.OverrideTechnique.empty()) {
        std::fprintf(stderr, "[ERROR] unknown technique '%s'\n", mOptions.OverrideTechnique.c_str());
        return false;
    }
    // TechniqueInfo of the variant (PathTechnique.cpp:19-33, InfoBufferTechnique.cpp:28-70)
    mInfo = TechniqueVariantInfo{};
    if (tech.type == IGB200_TECH_INFOBUFFER) {
        mInfo.EnabledAOVs     = { "Normals", "Albedo", "Depth" };
        mInfo.LockFramebuffer = true;
        mInfo.OverrideSPI     = 1;
    } else if (tech.type == IGB200_TECH_PATH && (tech.flags & IGB200_TECHFLAG_AOV_MIS)) {
        mInfo.EnabledAOVs = { "Direct Weights", "NEE Weights" };
    }

    Device::SceneSettings settings;
    settings.database            = &mScene.database;
    settings.entity_per_material = &mScene.entity_per_material;
    mDevice->assignScene(settings);
    mDevice->assignShading(mScene.shading.view());

    mFilmWidth  = mOptions.OverrideFilmSize ? mOptions.OverrideFilmSize->first : mScene.width;
    mFilmHeight = mOptions.OverrideFilmSize ? mOptions.OverrideFilmSize->second : mScene.height;
    if (mFilmWidth != mScene.width || mFilmHeight != mScene.height) {
        // the camera's image-plane scale follows the film's aspect ratio (PerspectiveCamera.cpp:40-54); the .igdb holds
        // the scale of its own film, so an override keeps the horizontal field of view and rescales the vertical one
        igb200_camera& cam = mScene.shading.camera;
        const bool same_aspect = mScene.width * mFilmHeight == mScene.height * mFilmWidth;
        if (!mOptions.IsTracer && mScene.width && mScene.height && !same_aspect) {
            cam.scale[1] = cam.scale[0] * (float)mFilmHeight / (float)mFilmWidth;
            mDevice->assignShading(mScene.shading.view());
        }
    }
    mSamplesPerIteration = mOptions.SPI ? mOptions.SPI : recommendSPI(mFilmWidth, mFilmHeight, false);
    if (mOptions.IsTracer) mSamplesPerIteration = mOptions.SPI ? mOptions.SPI : 1; // trace/main.cpp:79-81
    mDevice->resize(mFilmWidth, mFilmHeight);
    mLoaded = true;
    reset();
    return true;
}

Device::RenderSettings Runtime::renderSettings(const Ray* rays, size_t count) const
{
    Device::RenderSettings s;
    s.rays      = rays;
    s.spi       = mInfo.GetSPI(mSamplesPerIteration);
    s.width     = rays ? count : mFilmWidth;
    s.height    = rays ? 1 : mFilmHeight;
    s.info      = mInfo;
    s.iteration = mCurrentIteration;
    s.frame     = mCurrentFrame;
    s.user_seed = mOptions.Seed;
    return s;
}

void Runtime::step(bool) { steps(1); }

void Runtime::steps(size_t count)
{
    if (mOptions.IsTracer) {
        std::fprintf(stderr, "[ERROR] Trying to use step() in a trace driver!\n"); // Runtime.cpp:296-299
        return;
    }
    if (!mLoaded) {
        std::fprintf(stderr, "[ERROR] No scene loaded!\n");
        return;
    }
    const Device::RenderSettings s = renderSettings(nullptr, 0);
    mDevice->renderIterations(s, count);
    mCurrentIteration += count;
    if (!mInfo.LockFramebuffer) mCurrentSampleCount += s.spi * count;
}

void Runtime::trace(const std::vector<Ray>& rays)
{
    if (!mOptions.IsTracer) {
        std::fprintf(stderr, "[ERROR] Trying to use trace() in a camera driver!\n"); // Runtime.cpp:353-356
        return;
    }
    if (!mLoaded) {
        std::fprintf(stderr, "[ERROR] No scene loaded!\n");
        return;
    }
    const Device::RenderSettings s = renderSettings(rays.data(), rays.size());
    mDevice->render(TechniqueVariantShaderSet{}, s, nullptr);
    if (!mInfo.LockFramebuffer) mCurrentSampleCount += s.spi;
    ++mCurrentIteration;
}

void Runtime::trace(const std::vector<Ray>& rays, std::vector<float>& data)
{
    trace(rays);
    const float* ptr = getFramebufferForHost({}).Data;
    data.resize(rays.size() * 3);
    if (ptr) std::memcpy(data.data(), ptr, sizeof(float) * data.size());
}

void Runtime::reset()
{
    mDevice->clearAllFramebuffer();
    mCurrentIteration   = 0;
    mCurrentSampleCount = 0;
}

void Runtime::resizeFramebuffer(size_t width, size_t height)
{
    mFilmWidth  = width;
    mFilmHeight = height;
    mDevice->resize(width, height);
    reset();
}
} // namespace IG
