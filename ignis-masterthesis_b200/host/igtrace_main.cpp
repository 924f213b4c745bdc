// igtrace -- ray list in, contribution per ray out; the contract of the reference's src/frontend/trace/main.cpp:
//   input : one ray per line, `ox oy oz dx dy dz [tmin [tmax]]`; fewer than six numbers = line ignored; an empty line
//           or EOF ends the list; tmax <= tmin means FLT_MAX                                   (main.cpp:15-61)
//   run   : film = N x 1, samples per iteration fixed to 1, ceil(spp / spi) iterations, every iteration traces the
//           same rays                                                                          (main.cpp:79-81,124-137)
//   output: one line per ray, `r\tg\tb` in std::scientific, the film divided by the iteration count (main.cpp:63-67)
#include "Options.h"

#include <cmath>
#include <fstream>
#include <iterator>
#include <limits>
#include <sstream>

using namespace IG;

static std::vector<Ray> read_input(std::istream& is)
{
    std::vector<Ray> rays;
    std::string      line;
    while (std::getline(is, line) && !line.empty()) {
        std::vector<float> data;
        std::stringstream  stream(line);
        std::copy(std::istream_iterator<float>(stream), std::istream_iterator<float>(), std::back_inserter(data));
        if (data.size() < 6) continue;
        Ray ray;
        for (int k = 0; k < 3; ++k) {
            ray.Origin(k)    = data[k];
            ray.Direction(k) = data[3 + k];
        }
        ray.Range(0) = data.size() > 6 ? data[6] : 0;
        ray.Range(1) = data.size() > 7 ? data[7] : 0;
        if (ray.Range(1) <= ray.Range(0)) ray.Range(1) = std::numeric_limits<float>::max();
        rays.push_back(ray);
    }
    return rays;
}

static void write_output(std::ostream& os, const float* data, size_t count, size_t iterations)
{
    for (size_t i = 0; i < count; ++i)
        os << std::scientific << data[3 * i + 0] / iterations << "\t" << data[3 * i + 1] / iterations << "\t" << data[3 * i + 2] / iterations << "\n";
    os.flush();
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

    std::vector<Ray> rays;
    if (cmd.InputRay.empty()) {
        rays = read_input(std::cin);
    } else {
        std::ifstream stream(cmd.InputRay);
        rays = read_input(stream);
    }
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

    std::vector<float> accum_data, iter_data;
    for (size_t iter = 0; iter < desired_iter; ++iter) {
        runtime.trace(rays, iter_data); // the film accumulates on the device: the last read is the sum
        accum_data = iter_data;
    }
    if (accum_data.size() != rays.size() * 3) {
        std::cerr << "[FATAL] Got trace output size " << accum_data.size() << " but expected " << rays.size() * 3 << std::endl;
        return EXIT_FAILURE;
    }
    if (cmd.Output.empty()) {
        write_output(std::cout, accum_data.data(), rays.size(), runtime.currentIterationCount());
    } else {
        std::ofstream stream(cmd.Output);
        write_output(stream, accum_data.data(), rays.size(), runtime.currentIterationCount());
    }
    return EXIT_SUCCESS;
}
