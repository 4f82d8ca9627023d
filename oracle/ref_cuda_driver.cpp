// oracle/ref_cuda_driver.cpp — TEST/BASELINE INFRASTRUCTURE. Second same-box bar: the reference's OWN
// FP64 CUDA kernels (cuda/ForceKernels.cu, IntegrationKernels.cu, BenchmarkKernels.cu, CudaUtils.cu, Constants.cu,
// BHKernels2.cu), compiled unmodified for sm_100a by `make -C oracle refcuda`, driven through the reference's
// runBenchmarkOnGPU exactly as System::runGPUBenchmark does (core/System.cpp:197-207). Prints one JSON line.
//   usage: ref_cuda_bench <n> <steps> <force 0|2> <block_size (0 = reference heuristic)>
#include "System.hpp"
#include "Initializer.hpp"
#include "BenchmarkKernels.h"
#include <chrono>
#include <cstdio>
#include <cstdlib>

// the viewer entry point System.cpp references; there is no graphics stack here
extern "C" void runVisualizationOnGPU(Particles&, IntegrationMethod, ForceMethod, double, int, double, int) { std::abort(); }

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 65536;
    const int steps = argc > 2 ? std::atoi(argv[2]) : 5;
    const ForceMethod force = static_cast<ForceMethod>(argc > 3 ? std::atoi(argv[3]) : 0);
    const int block = argc > 4 ? std::atoi(argv[4]) : 256;
    try {
        InitResult init = Initializer::initSpiralGalaxy(n, 1e4, 2, 3.09398e3, 6.18795e2, 5.0);
        Particles p(init.particles);
        p.setExecutionMode(ExecutionMode::GPU);
        runBenchmarkOnGPU(p, IntegrationMethod::VELOCITY_VERLET, force, 0.04, 1, block);          // warm-up (context, clocks)
        auto t0 = std::chrono::high_resolution_clock::now();
        runBenchmarkOnGPU(p, IntegrationMethod::VELOCITY_VERLET, force, 0.04, steps, block);
        auto t1 = std::chrono::high_resolution_clock::now();
        const double sec = std::chrono::duration<double>(t1 - t0).count();
        const double evals = steps + 1.0;                                                            // initial force + one per step
        std::printf("{\"impl\": \"reference-cuda-fp64 (sm_100a build of the reference kernels)\", \"n\": %d, \"steps\": %d, \"force\": %d, "
                    "\"block_size\": %d, \"seconds\": %.6f, \"Gint_per_s\": %.3f}\n",
                    n, steps, (int)force, block, sec, (double)n * n * evals / sec / 1e9);
        return 0;
    } catch (const std::exception& ex) { std::fprintf(stderr, "ref_cuda_bench: %s\n", ex.what()); return 1; }
}
