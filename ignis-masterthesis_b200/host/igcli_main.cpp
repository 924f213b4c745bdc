// igcli -- scene in, path-traced image out; the control flow of the reference's src/frontend/cli/main.cpp:55-200:
// ceil(spp / spi) iterations of Runtime::step, samples/s per iteration = spi * w * h / seconds (main.cpp:135), the film
// divided by the iteration count saved as EXR (frontend/common/IO.cpp), min/med/max Msamples/s printed at the end.
// `--batch` renders all iterations in ONE wavefront (igb200_render_iterations; same film) and reports the average.
#include "ImageIO.h"
#include "Options.h"

#include <algorithm>
#include <chrono>
#include <cmath>

using namespace IG;

int main(int argc, char** argv)
{
    bool batch = false;
    std::vector<char*> args;
    for (int i = 0; i < argc; ++i) {
        if (std::string(argv[i]) == "--batch") batch = true;
        else args.push_back(argv[i]);
    }
    ProgramOptions cmd((int)args.size(), args.data(), false, "Command Line Interface (B200 backend)");
    if (cmd.ShouldExit) return cmd.ExitCode;
    if (cmd.InputScene.empty()) {
        std::cerr << "[ERROR] No input file given" << std::endl;
        return EXIT_FAILURE;
    }
    if (cmd.SPP.value_or(0) <= 0) {
        std::cerr << "[ERROR] No valid spp count given" << std::endl;
        return EXIT_FAILURE;
    }
    if (cmd.Output.empty()) {
        std::cerr << "[ERROR] No output file given" << std::endl;
        return EXIT_FAILURE;
    }
    using clock = std::chrono::high_resolution_clock;
    const auto t_all = clock::now();

    RuntimeOptions opts;
    cmd.populate(opts);
    Runtime runtime(opts);
    if (!runtime.loadFromFile(cmd.InputScene)) {
        std::cerr << "[ERROR] Could not load " << cmd.InputScene << std::endl;
        return EXIT_FAILURE;
    }
    const size_t SPI          = runtime.samplesPerIteration();
    const size_t desired_iter = (size_t)std::ceil(*cmd.SPP / (float)SPI);
    if (*cmd.SPP % SPI != 0)
        std::cerr << "[WARNING] Given spp " << *cmd.SPP << " is not a multiple of the spi " << SPI << ". Using spp " << desired_iter * SPI << " instead" << std::endl;

    const double per_iter = double(SPI * runtime.framebufferWidth() * runtime.framebufferHeight());
    std::vector<double> samples_sec;
    const auto t_render = clock::now();
    if (batch) {
        runtime.steps(desired_iter);
        const double s = std::chrono::duration<double>(clock::now() - t_render).count();
        samples_sec.push_back(per_iter * desired_iter / s);
    } else {
        while (samples_sec.size() < desired_iter) {
            const auto ticks = clock::now();
            runtime.step(samples_sec.size() != desired_iter - 1);
            const double s = std::chrono::duration<double>(clock::now() - ticks).count();
            samples_sec.push_back(per_iter / s);
        }
        runtime.device().flush(); // steps are pipelined on one GPU: the last iterations' deep paths finish here, inside the timed region
    }
    const double render_s = std::chrono::duration<double>(clock::now() - t_render).count();

    const Device::AOVAccessor film = runtime.getFramebufferForHost({});
    const bool saved = film.Data && film.IterationCount > 0
                       && saveImageRGB(cmd.Output, film.Data, runtime.framebufferWidth(), runtime.framebufferHeight(), 1.0f / film.IterationCount);
    if (!saved) std::cerr << "[ERROR] Failed to save " << cmd.Output << std::endl;

    if (!cmd.Quiet) {
        const igb200_render_stats st = runtime.device().renderStats();
        std::cout << "  Iterations: " << runtime.currentIterationCount() << "\n  SPP: " << runtime.currentSampleCount() << "\n  SPI: " << SPI
                  << "\n  Render>  " << render_s * 1e3 << " ms\n  Total>   " << std::chrono::duration<double>(clock::now() - t_all).count() * 1e3 << " ms" << std::endl;
        if (runtime.device().deviceCount() > 1)
            std::cout << "  GPUs: " << runtime.device().deviceCount() << "\n  Film reduce (last call)> " << runtime.device().lastFilmReduceMs() << " ms" << std::endl;
        if (cmd.Stats)
            if (const Statistics* stats = runtime.device().getStatistics()) // cli/main.cpp:155-166
                std::cout << stats->dump((size_t)(render_s * 1e3), runtime.currentIterationCount(), false);
        if (cmd.Stats)
            std::cout << "  Rays (last call)> camera " << st.camera_rays << " bounce " << st.bounce_rays << " shadow " << st.shadow_rays << std::endl;
        std::sort(samples_sec.begin(), samples_sec.end());
        std::cout << "# " << samples_sec.front() * 1e-6 << "/" << samples_sec[samples_sec.size() / 2] * 1e-6 << "/" << samples_sec.back() * 1e-6
                  << " (min/med/max Msamples/s)" << std::endl;
    }
    return saved ? EXIT_SUCCESS : EXIT_FAILURE;
}
