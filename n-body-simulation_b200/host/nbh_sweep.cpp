// nbh_sweep.cpp — the `--auto-benchmark` harness of the reference executable (core/main.cpp:52-452), pointed at the
// B200 library: a sweep over particle counts, every configuration repeated `runs` times on freshly generated random
// initial conditions, run statistics (sample standard deviation, coefficient of variation) and a CSV with the
// reference's column schema (main.cpp:238-245) so the reference's plotting scripts read it unchanged.
//
// Differences, all deliberate: the execution mode is always GPU (the CPU arm is the reference's own code); the sweep
// lists are arguments instead of compile-time vectors; a seed makes the initial conditions reproducible; one untimed
// first run per configuration absorbs CUDA context creation, which the reference's wall clock folds into run 0.
#include <chrono>
#include <cmath>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <sstream>
#include <stdexcept>

#include "nbh_host.hpp"
#include "nbody_b200.h"

namespace nbh {

// main.cpp:52-80 — mean, unbiased (n-1) standard deviation, cv = std / mean; a single value has std = cv = 0.
SampleStats sample_stats(const std::vector<double>& v) {
    SampleStats s;
    if (v.empty()) return s;
    s.min = std::numeric_limits<double>::max();
    s.max = std::numeric_limits<double>::lowest();
    double acc = 0.0;
    for (double x : v) { acc += x; s.min = std::min(s.min, x); s.max = std::max(s.max, x); }
    s.mean = acc / (double)v.size();
    if (v.size() > 1) {
        double ss = 0.0;
        for (double x : v) ss += (x - s.mean) * (x - s.mean);
        s.std = std::sqrt(ss / (double)(v.size() - 1));
        s.cv = s.std / s.mean;
    }
    return s;
}

// main.cpp:300-304 — years per run so that small systems take many steps and large ones few (dt = 0.01).
double sweep_years_for(int numParticles) {
    if (numParticles < 1000) return 1.50;
    if (numParticles < 10000) return 0.25;
    return 0.10;
}

static const char* force_label(ForceMethod f) {        // main.cpp:250-256 (note the '/' in the AVX2 label)
    switch (f) {
        case ForceMethod::PAIRWISE: return "PAIRWISE";
        case ForceMethod::PAIRWISE_AVX2_FP32: return "PAIRWISE_AVX2/FP32";
        case ForceMethod::ADAPTIVE_MUTUAL: return "ADAPTIVE_MUTUAL";
        case ForceMethod::BARNES_HUT: return "BARNES_HUT";
    }
    return "?";
}

// main.cpp:150-234 (runMultipleBenchmarks): every run draws new random particles with the generator parameters the
// reference hard-codes (main.cpp:103-108), then times one BENCHMARK-mode run end to end on the host clock.
SweepResult sweep_one(const SweepCase& c, int runs, const unsigned long* seed, bool warm) {
    SweepResult r;
    r.params = c; r.numRuns = runs;
    try {
        if (c.initType != "random") throw std::runtime_error("auto-benchmark supports initType 'random' only");
        if (runs < 1) throw std::runtime_error("auto-benchmark needs at least one run");
        const long steps = (long)(c.years / c.dtYears);
        r.totalSteps = steps;
        std::vector<double> secs, sps, psps;
        for (int run = warm ? -1 : 0; run < runs; ++run) {
            unsigned long s = seed ? *seed + 7919UL * (unsigned long)(run + 1) + (unsigned long)c.numParticles : 0;
            InitResult init = init_random(c.numParticles, 1e-10, 1e-5, 1e2, 1.0, 1e-3, seed ? &s : nullptr);
            ParticleSet& p = init.particles;
            auto t0 = std::chrono::high_resolution_clock::now();
            const int rc = nb_run_benchmark((long)p.n, p.posMass.data(), p.vel.data(), (int)c.integrationMethod,
                                            (int)c.forceMethod, c.dtYears, (int)steps, c.blockSize, c.precision, 1, c.device);
            auto t1 = std::chrono::high_resolution_clock::now();
            if (rc != NB_OK) throw std::runtime_error(nb_last_error());
            if (run < 0) continue;
            const double t = std::chrono::duration<double>(t1 - t0).count();
            secs.push_back(t);
            sps.push_back((double)steps / t);
            psps.push_back((double)c.numParticles * (double)steps / t);
        }
        r.execTime = sample_stats(secs);
        r.stepsPerSecond = sample_stats(sps);
        r.particleStepsPerSecond = sample_stats(psps);
        r.success = true;
    } catch (const std::exception& ex) {
        r.errorMessage = ex.what();
        r.success = false;
    }
    return r;
}

// main.cpp:236-288 — same header, same column order, same number formats.
std::string sweep_csv(const std::vector<SweepResult>& results) {
    std::ostringstream f;
    f << "NumParticles,Years,DtYears,ExecutionMode,ForceMethod,IntegrationMethod,NumThreads,BlockSize,InitType,"
         "NumRuns,TotalSteps,MeanExecutionTime,StdExecutionTime,"
         "MeanStepsPerSecond,StdStepsPerSecond,CVStepsPerSecond,MinStepsPerSecond,MaxStepsPerSecond,"
         "MeanParticleStepsPerSecond,StdParticleStepsPerSecond,Success,ErrorMessage\n";
    for (const SweepResult& r : results) {
        const SweepCase& p = r.params;
        f << std::defaultfloat << std::setprecision(6);
        f << p.numParticles << ',' << p.years << ',' << p.dtYears << ",GPU," << force_label(p.forceMethod) << ','
          << (p.integrationMethod == IntegrationMethod::EULER ? "EULER" : "VELOCITY_VERLET") << ',' << p.numThreads << ','
          << p.blockSize << ',' << p.initType << ',' << r.numRuns << ',' << r.totalSteps << ',';
        f << std::fixed << std::setprecision(6) << r.execTime.mean << ',' << r.execTime.std << ',';
        f << std::setprecision(2) << r.stepsPerSecond.mean << ',' << r.stepsPerSecond.std << ',';
        f << std::setprecision(4) << r.stepsPerSecond.cv << ',';
        f << std::setprecision(2) << r.stepsPerSecond.min << ',' << r.stepsPerSecond.max << ',';
        f << std::scientific << std::setprecision(3) << r.particleStepsPerSecond.mean << ',' << r.particleStepsPerSecond.std << ',';
        f << (r.success ? "true" : "false") << ",\"" << r.errorMessage << "\"\n";
    }
    return f.str();
}

// main.cpp:291-452 (runAutomaticBenchmark), with the reference's default size list and run count.
std::vector<SweepResult> run_sweep(const SweepSpec& spec, bool quiet) {
    std::vector<SweepCase> cases;
    for (int n : spec.particleCounts)
        for (ForceMethod fm : spec.forceMethods)
            for (IntegrationMethod im : spec.integrationMethods) {
                SweepCase c;
                c.numParticles = n; c.years = spec.years > 0 ? spec.years : sweep_years_for(n); c.dtYears = spec.dtYears;
                c.forceMethod = fm; c.integrationMethod = im; c.precision = spec.precision; c.device = spec.device;
                cases.push_back(c);
            }
    // The reference writes into presentation/benchmarks/ without creating it, so its CSV is lost when the directory is
    // missing (main.cpp:432-436); here the directory is created and the file opened before any GPU time is spent.
    std::ofstream csv;
    if (!spec.outputFile.empty()) {
        const std::filesystem::path parent = std::filesystem::path(spec.outputFile).parent_path();
        if (!parent.empty()) std::filesystem::create_directories(parent);
        csv.open(spec.outputFile);
        if (!csv) throw std::runtime_error("cannot write " + spec.outputFile);
    }
    std::vector<SweepResult> out;
    auto w0 = std::chrono::high_resolution_clock::now();
    for (size_t i = 0; i < cases.size(); ++i) {
        const SweepCase& c = cases[i];
        if (!quiet)
            std::cout << "[" << std::fixed << std::setprecision(1) << 100.0 * (double)i / (double)cases.size() << "%] test "
                      << i + 1 << "/" << cases.size() << "  N=" << c.numParticles << " " << force_label(c.forceMethod) << " ("
                      << (long)(c.years / c.dtYears) << " steps x " << spec.runs << " runs)" << std::flush;
        auto t0 = std::chrono::high_resolution_clock::now();
        SweepResult r = sweep_one(c, spec.runs, spec.hasSeed ? &spec.seed : nullptr, spec.warm);
        auto t1 = std::chrono::high_resolution_clock::now();
        if (!quiet) {
            if (r.success)
                std::cout << "  ok " << std::setprecision(2) << std::chrono::duration<double>(t1 - t0).count() << " s, "
                          << r.stepsPerSecond.mean << " steps/s, "
                          << std::setprecision(1) << (double)c.numParticles * r.particleStepsPerSecond.mean / 1e9
                          << " G-interactions/s (CV " << 100.0 * r.stepsPerSecond.cv << "%)\n";
            else
                std::cout << "  FAILED: " << r.errorMessage << "\n";
        }
        out.push_back(std::move(r));
    }
    if (!quiet) {
        size_t ok = 0;
        for (const SweepResult& r : out) ok += r.success ? 1 : 0;
        std::cout << "Benchmark finished in " << std::setprecision(2)
                  << std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - w0).count() << " s, " << ok
                  << "/" << out.size() << " configurations succeeded\n";
    }
    if (csv.is_open()) {
        csv << sweep_csv(out);
        if (!quiet) std::cout << "Results saved to " << spec.outputFile << "\n";
    }
    return out;
}

} // namespace nbh
