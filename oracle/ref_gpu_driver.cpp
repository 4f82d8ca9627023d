// oracle/ref_gpu_driver.cpp — TEST INFRASTRUCTURE. Drop-in proof: the reference's UNMODIFIED System.cpp / Forces.cpp /
// Initializer.cpp linked against n-body-simulation_b200/host/ref_shims.cpp + libnbody_b200.so instead of the reference's CUDA drivers.
// Runs the same Particles through System(..., ExecutionMode::GPU).runBenchmark (-> runBenchmarkOnGPU -> the B200 library)
// and through System(..., ExecutionMode::CPU).runBenchmark (the reference's OpenMP path) and prints the differences.
//   usage: ref_system_on_b200 <n> <steps> <dt> <force 0|2> <integ 0|1> [sistema.dat path]
#include "System.hpp"
#include "Initializer.hpp"
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

int main(int argc, char** argv) {
    const int n = argc > 1 ? std::atoi(argv[1]) : 2048;
    const int steps = argc > 2 ? std::atoi(argv[2]) : 10;
    const double dt = argc > 3 ? std::atof(argv[3]) : 0.01;
    const ForceMethod force = static_cast<ForceMethod>(argc > 4 ? std::atoi(argv[4]) : 0);
    const IntegrationMethod integ = static_cast<IntegrationMethod>(argc > 5 ? std::atoi(argv[5]) : 1);
    try {
        InitResult init = argc > 6 ? Initializer::initFromSistemadat(argv[6], n)
                                   : Initializer::initRandomParticles(n, 1e-10, 1e-5, 100.0, 1.0, 1e-3);
        Particles cpuCopy(init.particles);
        Particles gpuCopy(init.particles);
        System cpu(std::move(cpuCopy), integ, force, ExecutionMode::CPU, OutputMode::BENCHMARK);
        System gpu(std::move(gpuCopy), integ, force, ExecutionMode::GPU, OutputMode::BENCHMARK);
        const double e0 = cpu.computeTotalEnergy();
        cpu.runBenchmark(dt, steps);
        gpu.runBenchmark(dt, steps);          // reference System::runGPUBenchmark -> runBenchmarkOnGPU (shim) -> nb_run_benchmark
        const Particles& a = cpu.getParticles(); const Particles& b = gpu.getParticles();
        double dpos = 0, dvel = 0;
        for (size_t i = 0; i < a.n; ++i) {
            dpos = std::fmax(dpos, std::fmax(std::fabs(a.posMass[i].x - b.posMass[i].x), std::fmax(std::fabs(a.posMass[i].y - b.posMass[i].y), std::fabs(a.posMass[i].z - b.posMass[i].z))));
            dvel = std::fmax(dvel, std::fmax(std::fabs(a.vel[i].x - b.vel[i].x), std::fmax(std::fabs(a.vel[i].y - b.vel[i].y), std::fabs(a.vel[i].z - b.vel[i].z))));
        }
        const double ec = cpu.computeTotalEnergy(), eg = gpu.computeTotalEnergy();
        if (std::getenv("NB_FRAME_DIR")) {     // the reference's VISUALIZATION dispatch (System.cpp:96-99,210-214,367-382) on the headless shim
            Particles vizCopy(init.particles);
            System viz(std::move(vizCopy), integ, force, ExecutionMode::GPU, OutputMode::VISUALIZATION);
            viz.runSimulation(dt, steps, 2.0, "", "");
        }
        std::printf("{\"n\": %zu, \"steps\": %d, \"max_abs_dpos\": %.3e, \"max_abs_dvel\": %.3e, \"E0\": %.15e, \"E_cpu\": %.15e, \"E_gpu\": %.15e}\n",
                    a.n, steps, dpos, dvel, e0, ec, eg);
        return 0;
    } catch (const std::exception& ex) {
        std::fprintf(stderr, "ref_system_on_b200: %s\n", ex.what());
        return 1;
    }
}
