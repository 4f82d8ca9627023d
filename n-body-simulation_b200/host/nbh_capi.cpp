// nbh_capi.cpp — plain-C exports of the host layer (libnbody_host.so) so that tests and bench.py can call the C++
// generators / config parser / run driver through ctypes. All functions return 0 on success; nbh_last_error() has the text.
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "nbh_host.hpp"

namespace { thread_local std::string g_err; 
int fill(const nbh::InitResult& r, long cap, long* n_out, double* posMass4, double* vel4, char* meta, long meta_cap) {
    *n_out = (long)r.particles.n;
    const long n = (long)r.particles.n < cap ? (long)r.particles.n : cap;
    std::memcpy(posMass4, r.particles.posMass.data(), (size_t)n * 32);
    std::memcpy(vel4, r.particles.vel.data(), (size_t)n * 32);
    if (meta && meta_cap > 0) { std::strncpy(meta, r.metadata.c_str(), (size_t)meta_cap - 1); meta[meta_cap - 1] = 0; }
    return 0;
}
template <typename F> int guard(F f) { try { return f(); } catch (const std::exception& e) { g_err = e.what(); return 1; } } }

extern "C" {
const char* nbh_last_error(void) { return g_err.c_str(); }

// Initial conditions from a config.json text (the "init" block + optional "seed"); cap = rows available in the outputs.
int nbh_init_from_config(const char* config_text, long cap, long* n_out, double* posMass4, double* vel4, char* meta, long meta_cap) {
    return guard([&] { nbh::Config c = nbh::parse_config_text(config_text);
                       return fill(nbh::create_initial_conditions(c), cap, n_out, posMass4, vel4, meta, meta_cap); });
}
int nbh_init_spiral(int nStars, double totalMass, int nArms, double radius, double thickness, double perturbation,
                    unsigned long seed, long cap, long* n_out, double* posMass4, double* vel4) {
    return guard([&] { return fill(nbh::init_spiral_galaxy(nStars, totalMass, nArms, radius, thickness, perturbation, &seed), cap, n_out, posMass4, vel4, nullptr, 0); });
}
int nbh_init_random(int n, double minMass, double maxMass, double L, double maxVel, double minDist, unsigned long seed,
                    long cap, long* n_out, double* posMass4, double* vel4) {
    return guard([&] { return fill(nbh::init_random(n, minMass, maxMass, L, maxVel, minDist, &seed), cap, n_out, posMass4, vel4, nullptr, 0); });
}
int nbh_save_snapshot(const char* path, long n, const double* posMass4, const double* vel4) {
    return guard([&] { nbh::ParticleSet p((size_t)n); std::memcpy(p.posMass.data(), posMass4, (size_t)n * 32);
                       std::memcpy(p.vel.data(), vel4, (size_t)n * 32); nbh::save_snapshot(path, p); return 0; });
}
// Parse + validate only: fills steps, stepFreq and the enum values. out[0..7] = steps, stepFreq, integ, force, exec, output, threads, hasSeed
int nbh_parse_config(const char* config_text, long out[8]) {
    return guard([&] { nbh::Config c = nbh::parse_config_text(config_text);
                       const long steps = (long)(c.years / c.dtYears);
                       out[0] = steps; out[1] = steps / 500 > 1 ? steps / 500 : 1; out[2] = (long)c.integrationMethod; out[3] = (long)c.forceMethod;
                       out[4] = (long)c.executionMode; out[5] = (long)c.outputMode; out[6] = c.threads; out[7] = c.hasSeed ? 1 : 0; return 0; });
}
// Whole run of a config text through the C ABI of libnbody_b200 (BENCHMARK or FILE_CSV).
int nbh_run_config(const char* config_text, double* run_seconds, long* n_out, long* steps_out) {
    return guard([&] { nbh::Config c = nbh::parse_config_text(config_text); nbh::RunReport r = nbh::run_config(c, true);
                       if (run_seconds) *run_seconds = r.run_seconds; if (n_out) *n_out = (long)r.n; if (steps_out) *steps_out = r.steps; return 0; });
}
// --auto-benchmark pieces. out5 = mean, std, min, max, cv (core/main.cpp:52-80).
int nbh_sample_stats(const double* values, long count, double out5[5]) {
    return guard([&] { nbh::SampleStats s = nbh::sample_stats(std::vector<double>(values, values + count));
                       out5[0] = s.mean; out5[1] = s.std; out5[2] = s.min; out5[3] = s.max; out5[4] = s.cv; return 0; });
}
double nbh_sweep_years_for(int n) { return nbh::sweep_years_for(n); }
// CSV text for one synthetic result whose per-run execution times are given (no GPU involved): checks schema + formats.
int nbh_sweep_csv_from_times(int n, double years, double dt, int force, int integ, const double* secs, long runs, char* out, long cap) {
    return guard([&] { nbh::SweepResult r; r.params.numParticles = n; r.params.years = years; r.params.dtYears = dt;
                       r.params.forceMethod = (nbh::ForceMethod)force; r.params.integrationMethod = (nbh::IntegrationMethod)integ;
                       r.numRuns = (int)runs; r.totalSteps = (long)(years / dt); r.success = true;
                       std::vector<double> t(secs, secs + runs), sps, psps;
                       for (double x : t) { sps.push_back((double)r.totalSteps / x); psps.push_back((double)n * (double)r.totalSteps / x); }
                       r.execTime = nbh::sample_stats(t); r.stepsPerSecond = nbh::sample_stats(sps); r.particleStepsPerSecond = nbh::sample_stats(psps);
                       const std::string csv = nbh::sweep_csv({r});
                       if ((long)csv.size() + 1 > cap) throw std::runtime_error("csv buffer too small");
                       std::memcpy(out, csv.c_str(), csv.size() + 1); return 0; });
}
// A real sweep on the GPU; sizes[count]; CSV text into out. Returns 0 when every configuration succeeded.
int nbh_run_sweep(const int* sizes, long count, int force, int integ, int runs, double years, double dt, unsigned long seed,
                  int precision, char* out, long cap) {
    return guard([&] { nbh::SweepSpec spec; spec.particleCounts.assign(sizes, sizes + count);
                       spec.forceMethods = {(nbh::ForceMethod)force}; spec.integrationMethods = {(nbh::IntegrationMethod)integ};
                       spec.runs = runs; spec.years = years; spec.dtYears = dt; spec.hasSeed = true; spec.seed = seed; spec.precision = precision;
                       const std::vector<nbh::SweepResult> res = nbh::run_sweep(spec, true);
                       const std::string csv = nbh::sweep_csv(res);
                       if ((long)csv.size() + 1 > cap) throw std::runtime_error("csv buffer too small");
                       std::memcpy(out, csv.c_str(), csv.size() + 1);
                       for (const auto& r : res) if (!r.success) throw std::runtime_error(r.errorMessage);
                       return 0; });
}
// FILE_CSV sink on caller-provided arrays (output/OutputUtils.cpp:8-59).
int nbh_write_csv(const char* filename, const char* metadata, long n, long T, const double* particleData6,
                  const double* systemData2, const double* staticData2) {
    return guard([&] { return nbh::write_csv(filename, metadata, (size_t)n, (size_t)T, particleData6, systemData2, staticData2) ? 0 : 1; });
}
}
