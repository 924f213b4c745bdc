// Options.h -- the subset of the reference's ProgramOptions (src/frontend/common/ProgramOptions.cpp) that igcli and
// igtrace need on this path; same flag names.
#pragma once
#include "Runtime.h"

#include <cstdlib>
#include <cstring>
#include <iostream>
#include <optional>
#include <string>

namespace IG {
struct ProgramOptions {
    std::string           InputScene, InputRay, Output;
    std::optional<size_t> SPP, SPI, Width, Height;
    size_t                Seed   = 0;
    int                   Device = 0, DebugMode = 0, Gpus = 1;
    std::string           Technique;
    bool                  Quiet = false, ShouldExit = false, Stats = false;
    int                   ExitCode = EXIT_SUCCESS;

    ProgramOptions(int argc, char** argv, bool tracer, const char* title)
    {
        auto value = [&](int& i) -> const char* {
            if (i + 1 >= argc) {
                std::cerr << "Missing value for " << argv[i] << std::endl;
                ShouldExit = true;
                ExitCode   = EXIT_FAILURE;
                return "0";
            }
            return argv[++i];
        };
        for (int i = 1; i < argc; ++i) {
            const std::string a = argv[i];
            if (a == "-h" || a == "--help") {
                std::cout << title << "\nusage: " << argv[0] << " scene.igdb "
                          << (tracer ? "[-i rays.txt] " : "") << "[-o output] [--spp N] [--spi N] [--seed N] [--technique path|debug|infobuffer]"
                          << " [--debug-mode N] [--width W --height H] [--device N] [--gpus N] [--stats] [-q]" << std::endl;
                ShouldExit = true;
            } else if (a == "-o" || a == "--output") Output = value(i);
            else if (tracer && (a == "-i" || a == "--input")) InputRay = value(i);
            else if (a == "--spp") SPP = (size_t)std::strtoull(value(i), nullptr, 10);
            else if (a == "--spi") SPI = (size_t)std::strtoull(value(i), nullptr, 10);
            else if (a == "--seed") Seed = (size_t)std::strtoull(value(i), nullptr, 10);
            else if (a == "--width") Width = (size_t)std::strtoull(value(i), nullptr, 10);
            else if (a == "--height") Height = (size_t)std::strtoull(value(i), nullptr, 10);
            else if (a == "--device") Device = std::atoi(value(i));
            else if (a == "--gpus") Gpus = std::atoi(value(i));
            else if (a == "--technique") Technique = value(i);
            else if (a == "--debug-mode") DebugMode = std::atoi(value(i));
            else if (a == "--stats") Stats = true;
            else if (a == "-q" || a == "--quiet") Quiet = true;
            else if (!a.empty() && a[0] == '-') {
                std::cerr << "Unknown option " << a << std::endl;
                ShouldExit = true;
                ExitCode   = EXIT_FAILURE;
            } else InputScene = a;
        }
    }

    void populate(RuntimeOptions& opts) const
    {
        opts.DeviceOrdinal     = Device;
        opts.DeviceCount       = Gpus;
        opts.Seed              = Seed;
        opts.SPI               = SPI.value_or(0);
        opts.OverrideTechnique = Technique;
        opts.DebugMode         = DebugMode;
        opts.AcquireStats      = Stats;
        if (Width && Height) opts.OverrideFilmSize = { (uint32_t)*Width, (uint32_t)*Height };
    }
};
} // namespace IG
