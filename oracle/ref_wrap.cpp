// oracle/ref_wrap.cpp — TEST INFRASTRUCTURE ONLY.
// Thin extern "C" wrapper around the UNMODIFIED reference CPU sources, which are compiled where they lie
// under /root/reference by oracle/Makefile into oracle/_ref/libnbref.so (git-ignored). Nothing in the
// product path may link or call this; only tests/, __graft_entry__.smoke() and bench.py's CPU baseline /
// `--impl reference` arm do.
//
// Wrapped reference entry points:
//   calculateForces                 simulator/core/Forces.cpp:67-83  (pairwiseAcceleration :98-179,
//                                                                     pairwiseAccelerationAdaptiveMutual :301-387)
//   computeAverageMass              simulator/core/Forces.cpp:85-95
//   System::runBenchmark (CPU)      simulator/core/System.cpp:168-194 -> velocityVerlet :432-489 / eulerIntegration :385-429
//   System::computeTotalEnergy      simulator/core/System.cpp:60-85
//   Initializer::init*              simulator/core/Initializer.cpp:17-53,71-154,168-239,360-469
#include "System.hpp"
#include "Initializer.hpp"
#include "OutputData.hpp"
#include "OutputUtils.hpp"
#include <omp.h>
#include <cstring>
#include <cstdio>
#include <string>

// The reference's System.cpp references the three GPU drivers; the oracle is CPU-only.
extern "C" void runBenchmarkOnGPU(Particles&, IntegrationMethod, ForceMethod, double, int, int) {
    std::fprintf(stderr, "nbref: GPU path is not part of the CPU oracle\n"); std::abort();
}
extern "C" void runSimulationOnGPU(Particles&, IntegrationMethod, ForceMethod, double, int, int, OutputData&) {
    std::fprintf(stderr, "nbref: GPU path is not part of the CPU oracle\n"); std::abort();
}
extern "C" void runVisualizationOnGPU(Particles&, IntegrationMethod, ForceMethod, double, int, double, int) {
    std::fprintf(stderr, "nbref: viewer unavailable on this box\n"); std::abort();
}

namespace {
Particles make_particles(long n, const double* posMass4, const double* vel4) {
    Particles p((size_t)n);
    std::memcpy(p.posMass, posMass4, (size_t)n * sizeof(double4));
    if (vel4) std::memcpy(p.vel, vel4, (size_t)n * sizeof(double4));
    return p;
}
void export_particles(const Particles& p, long cap, long* n_out, double* posMass4, double* vel4) {
    long n = (long)p.n; if (n > cap) n = cap;
    *n_out = (long)p.n;
    std::memcpy(posMass4, p.posMass, (size_t)n * sizeof(double4));
    std::memcpy(vel4, p.vel, (size_t)n * sizeof(double4));
}
}

extern "C" {

int nbref_max_threads() { return omp_get_max_threads(); }
void nbref_set_threads(int t) { if (t > 0) omp_set_num_threads(t); }

// method: ForceMethod enum value (0 PAIRWISE, 2 ADAPTIVE_MUTUAL). out: n x 3 doubles.
int nbref_forces(long n, const double* posMass4, int method, double* accel3) {
    try {
        Particles p = make_particles(n, posMass4, nullptr);
        Accelerations a = calculateForces(p, static_cast<ForceMethod>(method), 0.5);
        std::memcpy(accel3, a.accel, (size_t)n * sizeof(double3));
        return 0;
    } catch (...) { return 1; }
}

double nbref_average_mass(long n, const double* posMass4) {
    Particles p = make_particles(n, posMass4, nullptr);
    return computeAverageMass(p);
}

// In-place CPU run: System::runBenchmark in CPU mode (what config executionMode:"CPU", BENCHMARK does).
int nbref_run(long n, double* posMass4, double* vel4, int integ, int force, double dt, int steps) {
    try {
        Particles p = make_particles(n, posMass4, vel4);
        System sys(std::move(p), static_cast<IntegrationMethod>(integ), static_cast<ForceMethod>(force),
                   ExecutionMode::CPU, OutputMode::BENCHMARK);
        sys.runBenchmark(dt, steps);
        const Particles& q = sys.getParticles();
        std::memcpy(posMass4, q.posMass, (size_t)n * sizeof(double4));
        std::memcpy(vel4, q.vel, (size_t)n * sizeof(double4));
        return 0;
    } catch (...) { return 1; }
}

double nbref_energy(long n, const double* posMass4, const double* vel4) {
    Particles p = make_particles(n, posMass4, vel4);
    System sys(std::move(p), IntegrationMethod::VELOCITY_VERLET, ForceMethod::PAIRWISE,
               ExecutionMode::CPU, OutputMode::BENCHMARK);
    return sys.computeTotalEnergy();
}

static std::string g_last_metadata;
// Metadata text of the last nbref_init_* call (what the reference prepends to its CSV files).
long nbref_last_metadata(char* buf, long cap) {
    if (buf && cap > 0) { std::strncpy(buf, g_last_metadata.c_str(), (size_t)cap - 1); buf[cap - 1] = 0; }
    return (long)g_last_metadata.size();
}
// Reference generators (seeded from std::random_device inside the reference: NOT reproducible; used for
// statistical comparison with the seeded generators and as CPU-baseline workloads).
int nbref_init_random(int n, double minMass, double maxMass, double L, double maxVel, double minDist,
                      long cap, long* n_out, double* posMass4, double* vel4) {
    try { InitResult r = Initializer::initRandomParticles(n, minMass, maxMass, L, maxVel, minDist);
          g_last_metadata = r.metadata; export_particles(r.particles, cap, n_out, posMass4, vel4); return 0; } catch (...) { return 1; }
}
int nbref_init_spiral(int nStars, double totalMass, int nArms, double radius, double thickness, double perturbation,
                      long cap, long* n_out, double* posMass4, double* vel4) {
    try { InitResult r = Initializer::initSpiralGalaxy(nStars, totalMass, nArms, radius, thickness, perturbation);
          g_last_metadata = r.metadata; export_particles(r.particles, cap, n_out, posMass4, vel4); return 0; } catch (...) { return 1; }
}
int nbref_init_galaxy(int nStars, double bhMass, double mMin, double mMax, double maxPos, double maxVel,
                      int addPlanets, int minPl, int maxPl, double pmMin, double pmMax, double pDist, double pVel,
                      long cap, long* n_out, double* posMass4, double* vel4) {
    try { InitResult r = Initializer::initGalaxy(nStars, bhMass, mMin, mMax, maxPos, maxVel, addPlanets != 0,
                                                 minPl, maxPl, pmMin, pmMax, pDist, pVel);
          g_last_metadata = r.metadata; export_particles(r.particles, cap, n_out, posMass4, vel4); return 0; } catch (...) { return 1; }
}
int nbref_init_stellar(double starMass, int nPlanets, double pmMin, double pmMax, double pdMin, double pdMax, int addMoons,
                       int minMoons, int maxMoons, double mmMin, double mmMax, double mdMin, double mdMax,
                       long cap, long* n_out, double* posMass4, double* vel4) {
    try { InitResult r = Initializer::initStellarSystem(starMass, nPlanets, pmMin, pmMax, pdMin, pdMax, addMoons != 0, minMoons,
                                                        maxMoons, mmMin, mmMax, mdMin, mdMax);
          g_last_metadata = r.metadata; export_particles(r.particles, cap, n_out, posMass4, vel4); return 0; } catch (...) { return 1; }
}
int nbref_init_file(const char* path, int count, long cap, long* n_out, double* posMass4, double* vel4) {
    try { InitResult r = Initializer::initFromSistemadat(std::string(path), count);
          g_last_metadata = r.metadata; export_particles(r.particles, cap, n_out, posMass4, vel4); return 0; } catch (...) { return 1; }
}

// The reference's FILE_CSV sink (output/OutputUtils.cpp:8-59) on caller-provided OutputData contents: fixture for the
// byte-for-byte comparison with the host's writer.
int nbref_write_csv(const char* filename, const char* metadata, long n, long T, const double* particleData6,
                    const double* systemData2, const double* staticData2) {
    OutputData out((size_t)n, (size_t)T, ExecutionMode::CPU);
    std::memcpy(out.h_particleData, particleData6, (size_t)n * T * 6 * sizeof(double));
    std::memcpy(out.h_systemData, systemData2, (size_t)T * 2 * sizeof(double));
    std::memcpy(out.h_staticData, staticData2, (size_t)n * 2 * sizeof(double));
    flushCSVOutput(out, filename, metadata);
    return 0;
}
} // extern "C"
