// nbh_host.hpp — C++17 host side above the C ABI: the reference's config.json surface, initial-condition generators
// and output sinks (core/ConfigParser.hpp, core/Initializer.cpp, output/OutputUtils.cpp), re-provided without the
// reference's network (nlohmann/json) and OpenGL dependencies. Same keys, enum names, generator parameters and file
// formats; optional additions (all have defaults, so existing config files stay valid):
//     "seed": integer   — reproducible initial conditions (the reference always seeds from std::random_device)
//     "precision": "FP32" | "FP64"   (default FP32 for BENCHMARK, FP64 for FILE_CSV)
//     "subtractCom": true | false    (default true: the reference CPU integrators' behaviour)
//     "device": integer
//     "frames": {"dir": "frames", "width": 1280, "height": 720, "physicsStepsPerFrame": 1, "scale": 0, "every": 0}
//                                    — with outputMode "VISUALIZATION": headless PPM frames (the viewer's shader arithmetic
//                                      as CUDA kernels) instead of the reference's OpenGL window; "every" = write every
//                                      k-th rendered frame (0: the reference's print frequency max(1, steps/500))
//     "snapshotOut": "path.nbin"     — write the final state as a binary snapshot (checkpoint); a "FROM_FILE" filePath
//                                      ending in ".nbin" reads one back (resume; partialCount <= 0 means all particles)
#pragma once
#include <cstddef>
#include <string>
#include <vector>

#include "nbh_json.hpp"

namespace nbh {

// Same numeric values as the reference enums (core/System.hpp:15-18, core/Forces.hpp:40-46, core/ExecutionModes.hpp,
// output/OutputModes.hpp:4-8) == the NB_* constants of include/nbody_b200.h.
enum class IntegrationMethod { EULER = 0, VELOCITY_VERLET = 1 };
enum class ForceMethod { PAIRWISE = 0, PAIRWISE_AVX2_FP32 = 1, ADAPTIVE_MUTUAL = 2, BARNES_HUT = 3 };
enum class ExecutionMode { CPU = 0, GPU = 1 };
enum class OutputMode { BENCHMARK = 0, FILE_CSV = 1, VISUALIZATION = 2 };

// Particle container with the reference's memory layout (core/Particles.hpp:9-16): AoS (x,y,z,m) / (vx,vy,vz,0).
struct ParticleSet {
    std::vector<double> posMass;   // 4 doubles per particle
    std::vector<double> vel;       // 4 doubles per particle
    size_t n = 0;
    explicit ParticleSet(size_t cap = 0) : posMass(cap * 4, 0.0), vel(cap * 4, 0.0), n(cap) {}
    void set(size_t i, double m, double x, double y, double z, double vx, double vy, double vz) {
        posMass[4 * i] = x; posMass[4 * i + 1] = y; posMass[4 * i + 2] = z; posMass[4 * i + 3] = m;
        vel[4 * i] = vx; vel[4 * i + 1] = vy; vel[4 * i + 2] = vz; vel[4 * i + 3] = 0.0;
    }
    void shrink(size_t count) { n = count; posMass.resize(count * 4); vel.resize(count * 4); }
};

struct InitResult { ParticleSet particles; std::string metadata; };

struct Config {
    std::string initSelected; Json initParams;
    int threads = 1; double years = 0.0, dtYears = 0.0;
    std::string outputDir, outputFile;
    IntegrationMethod integrationMethod = IntegrationMethod::VELOCITY_VERLET;
    ForceMethod forceMethod = ForceMethod::PAIRWISE;
    ExecutionMode executionMode = ExecutionMode::GPU;
    OutputMode outputMode = OutputMode::BENCHMARK;
    // optional extensions
    bool hasSeed = false; unsigned long seed = 0;
    int precision = -1;            // -1: choose by output mode
    bool subtractCom = true;
    int device = -1;
    std::string snapshotOut;
    bool deviceInit = false;       // SPIRAL_GALAXY + BENCHMARK: initial conditions generated on the GPU (nb_init_spiral), needs "seed"
    bool hasFrames = false; std::string frameDir = "frames"; int frameWidth = 1280, frameHeight = 720;   // VisualizationUtils.cpp:15-16
    int physicsStepsPerFrame = 0, frameEvery = 0; double frameScale = 0.0;
};

Config parse_config_text(const std::string& text);
Config parse_config(const std::string& path);
InitResult create_initial_conditions(const Config& cfg);

// Generators (core/Initializer.cpp). seed == nullptr -> std::random_device, as the reference.
InitResult init_from_file(const std::string& path, int count);
InitResult init_random(int n, double minMass, double maxMass, double L, double maxVelocity, double minDistance, const unsigned long* seed);
InitResult init_galaxy(int nStars, double blackHoleMass, double starMassMin, double starMassMax, double maxPosition,
                       double maxVelocity, bool addPlanets, int minPlanets, int maxPlanets, double planetMassMin,
                       double planetMassMax, double planetDistanceMax, double planetVelocityMax, const unsigned long* seed);
InitResult init_stellar_system(double starMass, int nPlanets, double planetMassMin, double planetMassMax,
                               double minPlanetDistance, double maxPlanetDistance, bool addMoons, int minMoons, int maxMoons,
                               double moonMassMin, double moonMassMax, double moonDistanceMin, double moonDistanceMax,
                               const unsigned long* seed);
InitResult init_spiral_galaxy(int nStars, double totalMass, int nArms, double galaxyRadius, double thickness,
                              double perturbation, const unsigned long* seed);

// Binary snapshot (no counterpart in the reference, which can only read sistema.dat and never reads its own output):
// "NBIN0001" | uint64 n | n x (x,y,z,m) doubles | n x (vx,vy,vz,0) doubles, little endian.
void save_snapshot(const std::string& path, const ParticleSet& p);
InitResult load_snapshot(const std::string& path, long count);

// FILE_CSV sink (output/OutputUtils.cpp:8-59): metadata, header, one row per particle per snapshot.
bool write_csv(const std::string& filename, const std::string& metadata, size_t n, size_t numTimeSteps,
               const double* particleData6, const double* systemData2, const double* staticData2);

// Run a parsed configuration through the C ABI. Returns 0 on success; throws std::runtime_error on library errors.
struct RunReport { double init_seconds = 0, run_seconds = 0; long steps = 0; size_t n = 0; std::string output_path; long frames_rendered = 0, frames_written = 0; };
RunReport run_config(const Config& cfg, bool quiet);

// --auto-benchmark harness (core/main.cpp:13-452): size sweep x repeated runs on random initial conditions -> statistics
// and a CSV with the reference's columns. Implemented in nbh_sweep.cpp.
struct SampleStats { double mean = 0, std = 0, min = 0, max = 0, cv = 0; };
SampleStats sample_stats(const std::vector<double>& values);
double sweep_years_for(int numParticles);
struct SweepCase {
    int numParticles = 0; double years = 0, dtYears = 0.01;
    ForceMethod forceMethod = ForceMethod::PAIRWISE; IntegrationMethod integrationMethod = IntegrationMethod::VELOCITY_VERLET;
    int numThreads = 1; int blockSize = -1; std::string initType = "random";
    int precision = 0; int device = -1;
};
struct SweepResult {
    SweepCase params; int numRuns = 0; long totalSteps = 0;
    SampleStats execTime, stepsPerSecond, particleStepsPerSecond;
    bool success = false; std::string errorMessage;
};
struct SweepSpec {
    std::vector<int> particleCounts{64, 128, 256, 512, 1024, 2048, 4096, 8192, 16384, 32768, 65536, 131072};   // main.cpp:295
    std::vector<ForceMethod> forceMethods{ForceMethod::PAIRWISE};
    std::vector<IntegrationMethod> integrationMethods{IntegrationMethod::VELOCITY_VERLET};
    double years = 0;          // <= 0: sweep_years_for(N)
    double dtYears = 0.01;     // main.cpp:306
    int runs = 5;              // main.cpp:308
    bool warm = true;          // one untimed run per configuration first
    bool hasSeed = false; unsigned long seed = 0;
    int precision = 0; int device = -1;
    std::string outputFile;
};
SweepResult sweep_one(const SweepCase& c, int runs, const unsigned long* seed, bool warm);
std::string sweep_csv(const std::vector<SweepResult>& results);
std::vector<SweepResult> run_sweep(const SweepSpec& spec, bool quiet);

} // namespace nbh
