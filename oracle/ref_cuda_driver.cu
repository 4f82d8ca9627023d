// oracle/ref_cuda_driver.cpp — TEST/BASELINE INFRASTRUCTURE. Second same-box bar: the reference's OWN
// FP64 CUDA kernels (cuda/ForceKernels.cu, IntegrationKernels.cu, BenchmarkKernels.cu, CudaUtils.cu, Constants.cu,
// BHKernels2.cu), compiled unmodified for sm_100a by `make -C oracle refcuda`, driven through the reference's
// runBenchmarkOnGPU exactly as System::runGPUBenchmark does (core/System.cpp:197-207). Prints one JSON line.
//   usage: ref_cuda_bench <n> <steps> <force 0|1|2> <block_size (0 = reference heuristic)>
//          ref_cuda_bench force <in.nbin> <out.f64> <force 0|1|2>   one launchForceKernel (cuda/ForceKernels.cu:208-232) on the
//                         particles of a binary snapshot; writes n x (ax,ay,az,0) doubles — golden outputs of the reference's
//                         own GPU kernels for the parity tests.
#include "System.hpp"
#include "Initializer.hpp"
#include "BenchmarkKernels.h"
#include "ForceKernels.cuh"
#include <cstring>
#include <fstream>
#include <vector>
#include <chrono>
#include <cstdio>
#include <cstdlib>

// the viewer entry point System.cpp references; there is no graphics stack here
extern "C" void runVisualizationOnGPU(Particles&, IntegrationMethod, ForceMethod, double, int, double, int) { std::abort(); }

static int force_mode(int argc, char** argv) {
    if (argc < 5) { std::fprintf(stderr, "usage: ref_cuda_bench force <in.nbin> <out.f64> <method>\n"); return 2; }
    std::ifstream f(argv[2], std::ios::binary);
    char magic[8]; unsigned long long n = 0;
    f.read(magic, 8); f.read(reinterpret_cast<char*>(&n), 8);
    if (!f || std::memcmp(magic, "NBIN0001", 8) != 0) { std::fprintf(stderr, "not a NBIN0001 snapshot\n"); return 2; }
    std::vector<double4> pos(n);
    f.read(reinterpret_cast<char*>(pos.data()), (std::streamsize)(n * sizeof(double4)));
    const ForceMethod method = static_cast<ForceMethod>(std::atoi(argv[4]));
    double avg = 0.0; for (auto& p : pos) avg += p.w; avg /= (double)n;          // cuda/BenchmarkKernels.cu:59-65
    double4 *d_pos = nullptr, *d_acc = nullptr;
    gpuErrchk(cudaMalloc(&d_pos, n * sizeof(double4))); gpuErrchk(cudaMalloc(&d_acc, n * sizeof(double4)));
    gpuErrchk(cudaMemcpy(d_pos, pos.data(), n * sizeof(double4), cudaMemcpyHostToDevice));
    const int block = 256;                                                        // core/System.hpp:98
    launchForceKernel(method, calculateGrid((int)n, block), block, block * sizeof(double4), 0, d_pos, d_acc, avg, (int)n);
    gpuErrchk(cudaDeviceSynchronize());
    std::vector<double4> acc(n);
    gpuErrchk(cudaMemcpy(acc.data(), d_acc, n * sizeof(double4), cudaMemcpyDeviceToHost));
    std::ofstream o(argv[3], std::ios::binary);
    o.write(reinterpret_cast<const char*>(acc.data()), (std::streamsize)(n * sizeof(double4)));
    cudaFree(d_pos); cudaFree(d_acc);
    return o ? 0 : 1;
}

int main(int argc, char** argv) {
    if (argc > 1 && std::strcmp(argv[1], "force") == 0) return force_mode(argc, argv);
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
