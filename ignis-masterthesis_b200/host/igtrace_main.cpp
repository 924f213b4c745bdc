// igtrace -- ray list in, contribution per ray out; the contract of the reference's src/frontend/trace/main.cpp:
//   input : one ray per line, `ox oy oz dx dy dz [tmin [tmax]]`; fewer than six numbers = line ignored; an empty line
//           or EOF ends the list; tmax <= tmin means FLT_MAX                                   (main.cpp:15-61)
//   run   : film = N x 1, samples per iteration fixed to 1, ceil(spp / spi) iterations, every iteration traces the
//           same rays                                                                          (main.cpp:79-81,124-137)
//   output: one line per ray, `r\tg\tb` in std::scientific, the film divided by the iteration count (main.cpp:63-67)
#include "Options.h"

#include <cmath>
#include <cstdio>
#include <fstream>
#include <limits>

using namespace IG;

// Text contract only (the format is the reference's, this parser is not): numbers separated by blanks, parsed with strtof
// until the first token that is not a number; six or more make a ray, fewer drop the line; an empty line ends the list.
static bool parseRayLine(const char* text, Ray& ray)
{
    float       v[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
    int         n    = 0;
    const char* p    = text;
    while (n < 8) {
        char*       end = nullptr;
        const float x   = std::strtof(p, &end);
        if (end == p) break;
        v[n++] = x;
        p      = end;
    }
    if (n < 6) return false;
    ray.Origin    = { v[0], v[1], v[2] };
    ray.Direction = { v[3], v[4], v[5] };
    const float tmin = v[6], tmax = v[7];
    ray.Range = { tmin, tmax > tmin ? tmax : std::numeric_limits<float>::max() }; // open range unless a valid one is given
    return true;
}

static std::vector<Ray> loadRayList(std::istream& in)
{
    std::vector<Ray> list;
    Ray              ray;
    for (std::string text; std::getline(in, text) && !text.empty();)
        if (parseRayLine(text.c_str(), ray)) list.push_back(ray);
    return list;
}

// film sums -> `r<TAB>g<TAB>b` per ray in %e (what std::scientific prints), divided by the number of iterations
static void saveContributions(std::FILE* out, const std::vector<float>& rgb, size_t iterations)
{
    const float inv = 1.0f / (float)iterations;
    for (size_t i = 0; i + 2 < rgb.size(); i += 3) std::fprintf(out, "%e\t%e\t%e\n", rgb[i] * inv, rgb[i + 1] * inv, rgb[i + 2] * inv);
    std::fflush(out);
}

int main(int argc, char** argv)
{
    ProgramOptions cmd(argc, argv, true, "Command Line Tracer (B200 backend)");
    if (cmd.ShouldExit) return cmd.ExitCode;
    if (cmd.InputScene.empty()) {
        std::cerr << "[ERROR] No input file given" << std::endl;
        return EXIT_FAILURE;
    }
    RuntimeOptions opts;
    cmd.populate(opts);
    opts.SPI      = 1; // "Fix samples per iteration to 1"
    opts.IsTracer = true;

    std::ifstream    file;
    if (!cmd.InputRay.empty()) file.open(cmd.InputRay);
    std::vector<Ray> rays = loadRayList(cmd.InputRay.empty() ? static_cast<std::istream&>(std::cin) : file);
    if (rays.empty()) {
        std::cerr << "[ERROR] No rays given" << std::endl;
        return EXIT_FAILURE;
    }
    opts.OverrideFilmSize = { (uint32_t)rays.size(), 1u };

    Runtime runtime(opts);
    if (!runtime.loadFromFile(cmd.InputScene)) {
        std::cerr << "[ERROR] Could not load " << cmd.InputScene << std::endl;
        return EXIT_FAILURE;
    }
    const size_t SPI          = runtime.samplesPerIteration();
    const size_t desired_iter = std::max<size_t>(1, (size_t)std::ceil(cmd.SPP.value_or(1) / (float)SPI));

    // every iteration traces the same rays into the same N x 1 film, which accumulates on the device: the read after the
    // last iteration is the sum over all of them
    std::vector<float> film;
    for (size_t it = 0; it < desired_iter; ++it) runtime.trace(rays, film);
    if (film.size() != rays.size() * 3) {
        std::cerr << "[FATAL] film holds " << film.size() << " values for " << rays.size() << " rays" << std::endl;
        return EXIT_FAILURE;
    }
    std::FILE* out = cmd.Output.empty() ? stdout : std::fopen(cmd.Output.c_str(), "w");
    if (!out) {
        std::cerr << "[ERROR] Could not open " << cmd.Output << std::endl;
        return EXIT_FAILURE;
    }
    saveContributions(out, film, runtime.currentIterationCount());
    if (out != stdout) std::fclose(out);
    return EXIT_SUCCESS;
}
