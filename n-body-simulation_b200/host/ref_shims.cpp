// ref_shims.cpp — the three unmangled symbols the reference's System.cpp binds (core/System.cpp:25-38,164,205,375),
// implemented on libnbody_b200's C ABI. Compiled AGAINST THE REFERENCE'S OWN HEADERS (Particles.hpp, System.hpp,
// OutputData.hpp), so it is only built where the reference tree is present (oracle/Makefile target `refgpu`); it is the
// file a maintainer of the reference drops in place of cuda/BenchmarkKernels.cu + cuda/SimulationKernels.cu.
#include "System.hpp"
#include "nbody_b200.h"

#include <cstdio>
#include <cstdlib>
#include <stdexcept>

namespace {
void check(int rc) { if (rc != NB_OK) throw std::runtime_error(nb_last_error()); }   // core/main.cpp:597-615 catches it
int env_int(const char* name, int dflt) { const char* v = std::getenv(name); return v ? std::atoi(v) : dflt; }
}

// cuda/BenchmarkKernels.cu:15-126
extern "C" void runBenchmarkOnGPU(Particles& p, IntegrationMethod m, ForceMethod f, double dt, int steps, int BLOCK_SIZE) {
    check(nb_run_benchmark((long)p.n, &p.posMass[0].x, &p.vel[0].x, static_cast<int>(m), static_cast<int>(f), dt, steps,
                           BLOCK_SIZE, env_int("NB_PRECISION", NB_PREC_FP32), env_int("NB_SUBTRACT_COM", 1), -1));
}
// cuda/SimulationKernels.cu:22-286
extern "C" void runSimulationOnGPU(Particles& p, IntegrationMethod m, ForceMethod f, double dt, int steps, int stepFreq, OutputData& out) {
    check(nb_run_simulation((long)p.n, &p.posMass[0].x, &p.vel[0].x, static_cast<int>(m), static_cast<int>(f), dt, steps,
                            stepFreq, out.h_particleData, out.h_systemData, out.h_staticData, (long)out.numTimeSteps,
                            env_int("NB_PRECISION", NB_PREC_FP64), env_int("NB_SUBTRACT_COM", 1), -1));
}
// visualization/cuda/VisualizationKernels.cu:52-303 — the reference opens an OpenGL window: unavailable on this box. With
// NB_FRAME_DIR set, the same loop runs headless and every stepFreq-th frame lands there as a PPM (nb_run_visualization);
// without it the call fails the way the reference's main already handles (core/main.cpp:597-615).
extern "C" void runVisualizationOnGPU(Particles& p, IntegrationMethod m, ForceMethod f, double dt, int steps, double stepFreq,
                                      int physicsStepsPerFrame) {
    const char* dir = std::getenv("NB_FRAME_DIR");
    if (!dir) throw std::runtime_error("VISUALIZATION is reference-only (no graphics stack on the target machine); "
                                       "set NB_FRAME_DIR for headless PPM frames");
    long frames[2] = {0, 0};
    check(nb_run_visualization((long)p.n, &p.posMass[0].x, &p.vel[0].x, static_cast<int>(m), static_cast<int>(f), dt, steps,
                               physicsStepsPerFrame, env_int("NB_FRAME_WIDTH", 1280), env_int("NB_FRAME_HEIGHT", 720), 0.0f, dir,
                               stepFreq >= 1.0 ? (int)stepFreq : 1, nullptr, frames, env_int("NB_PRECISION", NB_PREC_FP32),
                               env_int("NB_SUBTRACT_COM", 1), -1));
    std::printf("Headless visualization: %ld frames rendered, %ld written to %s\n", frames[0], frames[1], dir);
}
