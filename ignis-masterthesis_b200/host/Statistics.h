// Statistics.h -- the statistics object `IG::Device::getStatistics()` hands to the frontends (`--stats`), with the reference's
// category names (src/runtime/Statistics.h:9-63: ShaderType, SectionType, Quantity) so that what igcli prints keeps its
// vocabulary.  Standalone form only: inside the reference tree (IG_B200_IN_REFERENCE_TREE) the reference's own class is
// used and Device fills it through increase() / the public add().
//
// Source of the numbers: igb200_get_section_stats (CUDA events between the kernels of a wavefront round) and
// igb200_get_render_stats (the ray counters).  Where this backend fuses what the reference launches separately the fused
// kernel is reported under the first name and the others stay at zero launches:
//   Hit <- shade kernel (hit + miss shading + compaction);  PrimaryTraversal <- traversal + sort histogram + scan.
#pragma once

#include <array>
#include <cstdint>
#include <cstdio>
#include <string>

namespace IG {
enum class ShaderType { Device = 0, PrimaryTraversal, SecondaryTraversal, RayGeneration, Hit, Miss, AdvancedShadowHit,
                        AdvancedShadowMiss, Callback, Tonemap, Glare, ImageInfo, Bake, _COUNT };
enum class SectionType { GPUSortPrimary = 0, GPUSortSecondary = 1, GPUCompactPrimary = 2, GPUSortPrimaryReset = 3,
                         GPUSortPrimaryCount = 4, GPUSortPrimaryScan = 5, GPUSortPrimarySort = 6, GPUSortPrimaryCollapse = 7,
                         ImageInfoPercentile = 10, ImageInfoError = 11, ImageInfoHistogram = 12, ImageLoading, PackedImageLoading,
                         BufferLoading, BufferRequests, BufferReleases, FramebufferUpdate, AOVUpdate, TonemapUpdate,
                         FramebufferHostUpdate, AOVHostUpdate, _COUNT };
enum class Quantity { CameraRayCount = 0, ShadowRayCount, BounceRayCount, _COUNT };

class Statistics {
public:
    struct Entry {
        double   ms    = 0;
        uint64_t count = 0;
    };

    void reset() { *this = Statistics(); }
    void addShader(ShaderType t, double ms, uint64_t launches)
    {
        mShaders[(size_t)t].ms += ms;
        mShaders[(size_t)t].count += launches;
    }
    void addSection(SectionType t, double ms, uint64_t count)
    {
        mSections[(size_t)t].ms += ms;
        mSections[(size_t)t].count += count;
    }
    void increase(Quantity q, uint64_t value) { mQuantities[(size_t)q] += value; }

    [[nodiscard]] const Entry& shader(ShaderType t) const { return mShaders[(size_t)t]; }
    [[nodiscard]] const Entry& section(SectionType t) const { return mSections[(size_t)t]; }
    [[nodiscard]] uint64_t     quantity(Quantity q) const { return mQuantities[(size_t)q]; }

    /// Same shape as the reference's report: one line per shader kind / section that ran, then the ray counters.
    [[nodiscard]] std::string dump(size_t totalMS, size_t iter, bool verbose) const
    {
        static const char* const kShader[] = { "Device", "PrimaryTraversal", "SecondaryTraversal", "RayGeneration", "Hit", "Miss",
                                               "AdvancedShadowHit", "AdvancedShadowMiss", "Callback", "Tonemap", "Glare", "ImageInfo", "Bake" };
        std::string out = "Statistics:\n  Shader:\n";
        char        buf[256];
        const double total = totalMS > 0 ? (double)totalMS : 1.0;
        for (size_t i = 0; i < (size_t)ShaderType::_COUNT; ++i) {
            const Entry& e = mShaders[i];
            if (!verbose && e.count == 0) continue;
            std::snprintf(buf, sizeof(buf), "    %-20s> %10.3f ms (%5.1f %%) | %8llu launches | %9.4f ms per iteration\n", kShader[i], e.ms,
                          100.0 * e.ms / total, (unsigned long long)e.count, iter ? e.ms / (double)iter : 0.0);
            out += buf;
        }
        out += "  Sections:\n";
        const std::pair<SectionType, const char*> kSection[] = { { SectionType::GPUSortPrimary, "GPUSortPrimary" },
                                                                 { SectionType::GPUCompactPrimary, "GPUCompactPrimary" },
                                                                 { SectionType::FramebufferHostUpdate, "FramebufferHostUpdate" },
                                                                 { SectionType::AOVHostUpdate, "AOVHostUpdate" },
                                                                 { SectionType::TonemapUpdate, "TonemapUpdate" } };
        for (const auto& [t, name] : kSection) {
            const Entry& e = mSections[(size_t)t];
            if (!verbose && e.count == 0) continue;
            std::snprintf(buf, sizeof(buf), "    %-22s> %10.3f ms | %8llu times\n", name, e.ms, (unsigned long long)e.count);
            out += buf;
        }
        std::snprintf(buf, sizeof(buf), "  Quantities:\n    CameraRays> %llu\n    ShadowRays> %llu\n    BounceRays> %llu\n",
                      (unsigned long long)quantity(Quantity::CameraRayCount), (unsigned long long)quantity(Quantity::ShadowRayCount),
                      (unsigned long long)quantity(Quantity::BounceRayCount));
        out += buf;
        return out;
    }

private:
    std::array<Entry, (size_t)ShaderType::_COUNT>  mShaders{};
    std::array<Entry, (size_t)SectionType::_COUNT> mSections{};
    std::array<uint64_t, (size_t)Quantity::_COUNT> mQuantities{};
};
} // namespace IG
