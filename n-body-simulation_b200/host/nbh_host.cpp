// nbh_host.cpp — config parsing, initial conditions, CSV sink and the run driver of the C++17 host drop-in.
#include "nbh_host.hpp"

#include <array>
#include <charconv>
#include <cstring>
#include <vector>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <filesystem>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <random>
#include <sstream>
#include <unordered_map>

#include "nbody_b200.h"

namespace nbh {
namespace {
const double kG = 39.47841760435743;      // AU^3 / (Msun yr^2), core/Initializer.cpp:11

std::mt19937 make_rng(const unsigned long* seed) {
    if (seed) return std::mt19937((std::mt19937::result_type)*seed);
    std::random_device rd;
    return std::mt19937(rd());
}
template <typename T> T pick(const std::string& s, std::initializer_list<std::pair<const char*, T>> table, const char* what) {
    for (auto& kv : table) if (s == kv.first) return kv.second;
    throw std::runtime_error(std::string("Unknown ") + what + ": " + s);
}
} // namespace

// core/ConfigParser.hpp:29-95 — same keys, same error texts for unknown enum strings.
Config parse_config_text(const std::string& text) {
    const Json j = Json::parse(text);
    Config c;
    c.initSelected = j["init"]["selected"].as_string();
    c.initParams = j["init"][c.initSelected];
    c.threads = j["threads"].as_int();
    c.years = j["years"].as_double();
    c.dtYears = j["dtYears"].as_double();
    c.outputDir = j["output"]["dir"].as_string();
    c.outputFile = j["output"]["file"].as_string();
    c.integrationMethod = pick<IntegrationMethod>(j["integrationMethod"].as_string(),
        {{"EULER", IntegrationMethod::EULER}, {"VELOCITY_VERLET", IntegrationMethod::VELOCITY_VERLET}}, "integration method");
    c.forceMethod = pick<ForceMethod>(j["forceMethod"].as_string(),
        {{"PAIRWISE", ForceMethod::PAIRWISE}, {"PAIRWISE_AVX2_FP32", ForceMethod::PAIRWISE_AVX2_FP32},
         {"ADAPTIVE_MUTUAL", ForceMethod::ADAPTIVE_MUTUAL}, {"BARNES_HUT", ForceMethod::BARNES_HUT}}, "force method");
    c.executionMode = pick<ExecutionMode>(j["executionMode"].as_string(),
        {{"CPU", ExecutionMode::CPU}, {"GPU", ExecutionMode::GPU}}, "execution mode");
    c.outputMode = pick<OutputMode>(j["outputMode"].as_string(),
        {{"BENCHMARK", OutputMode::BENCHMARK}, {"FILE_CSV", OutputMode::FILE_CSV}, {"VISUALIZATION", OutputMode::VISUALIZATION}}, "output mode");
    if (j.contains("seed")) { c.hasSeed = true; c.seed = (unsigned long)j["seed"].as_double(); }
    if (j.contains("precision")) c.precision = pick<int>(j["precision"].as_string(), {{"FP32", NB_PREC_FP32}, {"FP64", NB_PREC_FP64}}, "precision");
    if (j.contains("subtractCom")) c.subtractCom = j["subtractCom"].as_bool();
    if (j.contains("device")) c.device = j["device"].as_int();
    if (j.contains("snapshotOut")) c.snapshotOut = j["snapshotOut"].as_string();
    if (j.contains("deviceInit")) c.deviceInit = j["deviceInit"].as_bool();
    if (j.contains("frames")) {
        const Json& f = j["frames"];
        c.hasFrames = true;
        if (f.contains("dir")) c.frameDir = f["dir"].as_string();
        if (f.contains("width")) c.frameWidth = f["width"].as_int();
        if (f.contains("height")) c.frameHeight = f["height"].as_int();
        if (f.contains("physicsStepsPerFrame")) c.physicsStepsPerFrame = f["physicsStepsPerFrame"].as_int();
        if (f.contains("every")) c.frameEvery = f["every"].as_int();
        if (f.contains("scale")) c.frameScale = f["scale"].as_double();
    }
    return c;
}

Config parse_config(const std::string& path) {
    std::ifstream in(path);
    if (!in.is_open()) throw std::runtime_error("Failed to open config file: " + path);
    std::stringstream ss; ss << in.rdbuf();
    return parse_config_text(ss.str());
}

// core/ConfigParser.hpp:98-168
InitResult create_initial_conditions(const Config& cfg) {
    const Json& p = cfg.initParams;
    const unsigned long* seed = cfg.hasSeed ? &cfg.seed : nullptr;
    if (cfg.initSelected == "FROM_FILE")
        return init_from_file(p["filePath"].as_string(), p["partialCount"].as_int());
    if (cfg.initSelected == "RANDOM")
        return init_random(p["nParticles"].as_int(), p["minMass"].as_double(), p["maxMass"].as_double(), p["L"].as_double(),
                           p["maxVelocity"].as_double(), p["minDistance"].as_double(), seed);
    if (cfg.initSelected == "GALAXY")
        return init_galaxy(p["nStars"].as_int(), p["blackHoleMass"].as_double(), p["starMassMin"].as_double(),
                           p["starMassMax"].as_double(), p["maxPosition"].as_double(), p["maxVelocity"].as_double(),
                           p["addPlanets"].as_bool(), p["minPlanets"].as_int(), p["maxPlanets"].as_int(),
                           p["planetMassMin"].as_double(), p["planetMassMax"].as_double(), p["planetDistanceMax"].as_double(),
                           p["planetVelocityMax"].as_double(), seed);
    if (cfg.initSelected == "STELLAR_SYSTEM")
        // The reference passes moonDistanceMax / moonVelocityMax into the (moonDistanceMin, moonDistanceMax) slots
        // (core/ConfigParser.hpp:146-152 vs core/Initializer.hpp:59-61); kept so the same file gives the same system.
        return init_stellar_system(p["starMass"].as_double(), p["nPlanets"].as_int(), p["planetMassMin"].as_double(),
                                   p["planetMassMax"].as_double(), p["minPlanetDistance"].as_double(), p["maxPlanetDistance"].as_double(),
                                   p["addMoons"].as_bool(), p["minMoons"].as_int(), p["maxMoons"].as_int(), p["moonMassMin"].as_double(),
                                   p["moonMassMax"].as_double(), p["moonDistanceMax"].as_double(), p["moonVelocityMax"].as_double(), seed);
    if (cfg.initSelected == "SPIRAL_GALAXY")
        return init_spiral_galaxy(p["nStars"].as_int(), p["totalMass"].as_double(), p["nArms"].as_int(), p["galaxyRadius"].as_double(),
                                  p["thickness"].as_double(), p["perturbation"].as_double(), seed);
    throw std::runtime_error("Unknown initialization method: " + cfg.initSelected);
}

// core/Initializer.cpp:17-53 — first token: body count; per body mass[kg] x y z[km] vx vy vz[km/s].
void save_snapshot(const std::string& path, const ParticleSet& p) {
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Cannot open snapshot for writing: " + path);
    const uint64_t n = p.n;
    f.write("NBIN0001", 8);
    f.write(reinterpret_cast<const char*>(&n), 8);
    f.write(reinterpret_cast<const char*>(p.posMass.data()), (std::streamsize)(n * 32));
    f.write(reinterpret_cast<const char*>(p.vel.data()), (std::streamsize)(n * 32));
    if (!f) throw std::runtime_error("Short write to snapshot: " + path);
}

InitResult load_snapshot(const std::string& path, long count) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("Cannot open file: " + path);
    char magic[8]; uint64_t n = 0;
    f.read(magic, 8); f.read(reinterpret_cast<char*>(&n), 8);
    if (!f || std::string(magic, 8) != "NBIN0001") throw std::runtime_error("Not a NBIN0001 snapshot: " + path);
    const uint64_t use = (count > 0 && (uint64_t)count < n) ? (uint64_t)count : n;
    ParticleSet p((size_t)use);
    f.read(reinterpret_cast<char*>(p.posMass.data()), (std::streamsize)(use * 32));
    f.seekg((std::streamoff)(16 + n * 32));
    f.read(reinterpret_cast<char*>(p.vel.data()), (std::streamsize)(use * 32));
    if (!f) throw std::runtime_error("Truncated snapshot: " + path);
    std::ostringstream md;
    md << "# Initializer: binary snapshot\n# Total in file: " << n << "\n# Using first: " << use << "\n# Total particle count: "
       << use << "\n# filePath: " << path << "\n";
    return {std::move(p), md.str()};
}

InitResult init_from_file(const std::string& path, int count) {
    if (path.size() > 5 && path.compare(path.size() - 5, 5, ".nbin") == 0) return load_snapshot(path, count);
    std::ifstream f(path);
    if (!f) throw std::runtime_error("Cannot open file: " + path);
    int total = 0; f >> total;
    if (!f.good()) throw std::runtime_error("Failed reading first line from sistema.dat.");
    ParticleSet p((size_t)count);
    for (int i = 0; i < count; ++i) {
        double m, x, y, z, vx, vy, vz;
        f >> m >> x >> y >> z >> vx >> vy >> vz;
        if (!f.good()) break;          // as the reference (core/Initializer.cpp:36): a last body whose final token ends the file
                                       // without a trailing newline sets eofbit and is dropped there too
        p.set((size_t)i, m / 1.989e30, x / 1.496e8, y / 1.496e8, z / 1.496e8, vx * 0.2108, vy * 0.2108, vz * 0.2108);
    }
    std::ostringstream md;
    md << "# Initializer: From sistema.dat\n# Total in file: " << total << "\n# Using first: " << p.n
       << "\n# Total particle count: " << p.n << "\n# filePath: " << path << "\n";
    return {std::move(p), md.str()};
}

// core/Initializer.cpp:168-239 — uniform box, rejection of candidates closer than minDistance to a placed particle
// (<= 1000 retries). Draw order per attempt as in the reference: mass, x, y, z, vx, vy, vz. The neighbour test uses
// a cell list (cells of minDistance) instead of the reference's O(N^2) scan; the accepted set is identical.
InitResult init_random(int n, double minMass, double maxMass, double L, double maxVelocity, double minDistance, const unsigned long* seed) {
    ParticleSet p((size_t)n);
    std::mt19937 gen = make_rng(seed);
    std::uniform_real_distribution<double> xD(-L / 2, L / 2), yD(-L / 2, L / 2), zD(-L / 2, L / 2), vD(-maxVelocity, maxVelocity), mD(minMass, maxMass);
    std::unordered_map<uint64_t, std::vector<int>> cells;
    const bool useCells = minDistance > 0.0 && (L / minDistance) < 2.0e6;
    auto cellOf = [&](double x) { return (int64_t)std::floor((x + L / 2) / minDistance); };
    auto key = [](int64_t a, int64_t b, int64_t c) { return ((uint64_t)(a & 0x1fffff) << 42) | ((uint64_t)(b & 0x1fffff) << 21) | (uint64_t)(c & 0x1fffff); };
    for (int i = 0; i < n; ++i) {
        double m = 0, x = 0, y = 0, z = 0, vx = 0, vy = 0, vz = 0;
        bool valid = false; int retries = 0;
        while (!valid) {
            m = mD(gen); x = xD(gen); y = yD(gen); z = zD(gen); vx = vD(gen); vy = vD(gen); vz = vD(gen);
            valid = true;
            if (minDistance > 0.0) {
                auto tooClose = [&](int j) {
                    const double dx = x - p.posMass[4 * j], dy = y - p.posMass[4 * j + 1], dz = z - p.posMass[4 * j + 2];
                    return std::sqrt(dx * dx + dy * dy + dz * dz) < minDistance;
                };
                if (useCells) {
                    const int64_t cx = cellOf(x), cy = cellOf(y), cz = cellOf(z);
                    for (int64_t a = cx - 1; a <= cx + 1 && valid; ++a)
                        for (int64_t b = cy - 1; b <= cy + 1 && valid; ++b)
                            for (int64_t c = cz - 1; c <= cz + 1 && valid; ++c) {
                                auto it = cells.find(key(a, b, c));
                                if (it == cells.end()) continue;
                                for (int j : it->second) if (tooClose(j)) { valid = false; break; }
                            }
                } else {
                    for (int j = 0; j < i; ++j) if (tooClose(j)) { valid = false; break; }
                }
            }
            if (!valid && ++retries > 1000) throw std::runtime_error("Failed to place particle " + std::to_string(i));
        }
        p.set((size_t)i, m, x, y, z, vx, vy, vz);
        if (useCells) cells[key(cellOf(x), cellOf(y), cellOf(z))].push_back(i);
    }
    std::ostringstream md;
    md << "# Random particles\n# Number of particles: " << n << "\n# Box size (AU): " << L << "\n# Velocity range (AU/year): [-"
       << maxVelocity << ", " << maxVelocity << "]\n# Mass range (solar masses): [" << minMass << ", " << maxMass << "]\n";
    if (minDistance > 0.0) md << "# Minimum distance between particles: " << minDistance << " AU\n";
    md << "# Total particle count: " << p.n << "\n# Format: time,x,y,z,vx,vy,vz,energy\n# Units: years, AU, AU/year, energy/(solar mass)\n";
    return {std::move(p), md.str()};
}

// core/Initializer.cpp:71-154 — black hole at the origin, stars uniform in a cube, optional planets offset from their star.
InitResult init_galaxy(int nStars, double blackHoleMass, double starMassMin, double starMassMax, double maxPosition,
                       double maxVelocity, bool addPlanets, int minPlanets, int maxPlanets, double planetMassMin,
                       double planetMassMax, double planetDistanceMax, double planetVelocityMax, const unsigned long* seed) {
    const size_t cap = 1 + (size_t)nStars * (1 + (addPlanets ? (size_t)maxPlanets : 0));
    ParticleSet p(cap);
    size_t k = 0;
    p.set(k++, blackHoleMass, 0, 0, 0, 0, 0, 0);
    std::mt19937 gen = make_rng(seed);
    std::uniform_real_distribution<double> mS(starMassMin, starMassMax), pos(-maxPosition, maxPosition), vel(-maxVelocity, maxVelocity);
    std::uniform_real_distribution<double> mP(planetMassMin, planetMassMax), dP(-planetDistanceMax, planetDistanceMax), vP(-planetVelocityMax, planetVelocityMax);
    std::uniform_int_distribution<int> cnt(minPlanets, maxPlanets);
    for (int i = 0; i < nStars; ++i) {
        const double m = mS(gen);
        const double x = pos(gen), y = pos(gen), z = pos(gen);
        const double vx = vel(gen), vy = vel(gen), vz = vel(gen);
        p.set(k++, m, x, y, z, vx, vy, vz);
        if (addPlanets) {
            const int np = cnt(gen);
            for (int q = 0; q < np; ++q) {
                const double pm = mP(gen);
                const double ox = dP(gen), oy = dP(gen), oz = dP(gen);
                const double ovx = vP(gen), ovy = vP(gen), ovz = vP(gen);
                p.set(k++, pm, x + ox, y + oy, z + oz, vx + ovx, vy + ovy, vz + ovz);
            }
        }
    }
    p.shrink(k);
    std::ostringstream md;
    md << "# Initializer: Galaxy\n# Units: AU (distance), Solar Masses (mass), AU/yr (velocity)\n# nStars: " << nStars
       << "\n# blackHoleMass: " << blackHoleMass << " Msun\n# starMass range: [" << starMassMin << ", " << starMassMax
       << "] Msun\n# maxPosition: " << maxPosition << " AU\n# maxVelocity: " << maxVelocity << " AU/yr\n# addPlanets: "
       << std::boolalpha << addPlanets << "\n";
    if (addPlanets)
        md << "# Planets per star: [" << minPlanets << ", " << maxPlanets << "]\n# planetMass range: [" << planetMassMin << ", "
           << planetMassMax << "] Msun\n# planetDistanceMax: " << planetDistanceMax << " AU\n# planetVelocityMax: " << planetVelocityMax << " AU/yr\n";
    md << "# Total particle count: " << p.n << "\n";
    return {std::move(p), md.str()};
}

// core/Initializer.cpp:257-347 — star at the origin, planets on circular orbits in the XY plane, optional moons.
InitResult init_stellar_system(double starMass, int nPlanets, double planetMassMin, double planetMassMax,
                               double minPlanetDistance, double maxPlanetDistance, bool addMoons, int minMoons, int maxMoons,
                               double moonMassMin, double moonMassMax, double moonDistanceMin, double moonDistanceMax,
                               const unsigned long* seed) {
    const size_t cap = 1 + (size_t)nPlanets * (1 + (addMoons ? (size_t)maxMoons : 0));
    ParticleSet p(cap);
    size_t k = 0;
    std::mt19937 gen = make_rng(seed);
    std::uniform_real_distribution<double> mP(planetMassMin, planetMassMax), rP(minPlanetDistance, maxPlanetDistance), ang(0.0, 2 * M_PI);
    std::uniform_int_distribution<int> cnt(minMoons, maxMoons);
    std::uniform_real_distribution<double> mM(moonMassMin, moonMassMax), rM(moonDistanceMin, moonDistanceMax), angM(0.0, 2 * M_PI);
    p.set(k++, starMass, 0, 0, 0, 0, 0, 0);
    for (int i = 0; i < nPlanets; ++i) {
        const double pm = mP(gen), r = rP(gen), th = ang(gen);
        const double x = r * std::cos(th), y = r * std::sin(th);
        const double sp = std::sqrt(kG * starMass / r);
        const double vx = -sp * std::sin(th), vy = sp * std::cos(th);
        p.set(k++, pm, x, y, 0.0, vx, vy, 0.0);
        if (addMoons) {
            const int nm = cnt(gen);
            for (int q = 0; q < nm; ++q) {
                const double mm = mM(gen), d = rM(gen), ph = angM(gen);
                const double ms = std::sqrt(kG * pm / d);
                p.set(k++, mm, x + d * std::cos(ph), y + d * std::sin(ph), 0.0, vx - ms * std::sin(ph), vy + ms * std::cos(ph), 0.0);
            }
        }
    }
    p.shrink(k);
    std::ostringstream md;
    md << "# Initializer: Stellar System \n# starMass: " << starMass << " Msol\n# nPlanets: " << nPlanets << "\n# planetMass range: ["
       << planetMassMin << ", " << planetMassMax << "] Msol\n# planetDistance range: [" << minPlanetDistance << ", " << maxPlanetDistance
       << "] AU\n# addMoons: " << std::boolalpha << addMoons << "\n";
    if (addMoons)
        md << "# moons per planet: [" << minMoons << ", " << maxMoons << "]\n# moonMass range: [" << moonMassMin << ", " << moonMassMax
           << "] Msol\n# moonDistance range: [" << moonDistanceMin << ", " << moonDistanceMax << "] AU\n";
    md << "# Total particle count: " << p.n << "\n";
    return {std::move(p), md.str()};
}

// core/Initializer.cpp:360-469 — equal-mass stars, uniform-area disc r = R sqrt(u), arms by index, random disc orientation,
// rotation curve v = v0 sqrt(r/R) plus a uniform velocity perturbation.
InitResult init_spiral_galaxy(int nStars, double totalMass, int nArms, double galaxyRadius, double thickness,
                              double perturbation, const unsigned long* seed) {
    ParticleSet p((size_t)nStars);
    std::mt19937 gen = make_rng(seed);
    std::uniform_real_distribution<double> uD(0.0, 1.0), armD(-M_PI / 8, M_PI / 8), zD(-thickness / 2, thickness / 2),
        vD(-perturbation, perturbation), thD(0, M_PI), phD(0, 2 * M_PI);
    const double v0 = std::sqrt(kG * totalMass / galaxyRadius);
    const double thN = thD(gen), phN = phD(gen);
    const std::array<double, 3> nrm = {std::sin(thN) * std::cos(phN), std::sin(thN) * std::sin(phN), std::cos(thN)};
    std::array<double, 3> e1;
    if (std::fabs(nrm[0]) < 1e-6 && std::fabs(nrm[1]) < 1e-6) e1 = {1, 0, 0};
    else {
        e1 = {-nrm[1], nrm[0], 0};
        const double len = std::sqrt(e1[0] * e1[0] + e1[1] * e1[1] + e1[2] * e1[2]);
        for (double& c : e1) c /= len;
    }
    const std::array<double, 3> e2 = {nrm[1] * e1[2] - nrm[2] * e1[1], nrm[2] * e1[0] - nrm[0] * e1[2], nrm[0] * e1[1] - nrm[1] * e1[0]};
    const double mass = totalMass / nStars;
    for (int i = 0; i < nStars; ++i) {
        const double u = uD(gen);
        const double r = galaxyRadius * std::sqrt(u);
        const double phi = (2 * M_PI * (i % nArms)) / nArms + (r / galaxyRadius) * 2 * M_PI + armD(gen);
        const double xd = r * std::cos(phi), yd = r * std::sin(phi), zd = zD(gen);
        const double sp = v0 * std::sqrt(r / galaxyRadius);
        const double tx = -std::sin(phi), ty = std::cos(phi);
        double pos[3], vel[3];
        for (int c = 0; c < 3; ++c) pos[c] = xd * e1[c] + yd * e2[c] + zd * nrm[c];
        for (int c = 0; c < 3; ++c) vel[c] = sp * (tx * e1[c] + ty * e2[c]) + vD(gen);
        p.set((size_t)i, mass, pos[0], pos[1], pos[2], vel[0], vel[1], vel[2]);
    }
    std::ostringstream md;
    md << "# Initializer: Spiral Galaxy\n# nStars: " << nStars << "\n# totalMass: " << totalMass << " Msol\n# nArms: " << nArms
       << "\n# galaxyRadius: " << galaxyRadius << " AU\n# thickness: " << thickness << " AU\n# base circular velocity at galaxyRadius: "
       << v0 << " AU/yr\n# random disk orientation (unit normal): (" << nrm[0] << ", " << nrm[1] << ", " << nrm[2]
       << ")\n# Total particle count: " << p.n << "\n";
    return {std::move(p), md.str()};
}

// output/OutputUtils.cpp:8-59 — metadata, blank line, header, then per snapshot and particle one row with 8-digit fixed
// floats and the 15-digit energy.
// Same text as the reference's flushCSVOutput (output/OutputUtils.cpp:8-59: std::fixed, 8 digits, energy 15 digits), but
// formatted with std::to_chars into a 1 MB buffer instead of ten operator<< calls per row: the file of a large run has
// tens of millions of rows. tests/test_host.py compares the bytes with the reference writer's.
namespace {
inline char* put_fixed(char* p, char* end, double v, int digits) {
    if (!std::isfinite(v)) {                         // iostream spelling of inf / nan, as the reference would print them
        std::ostringstream o; o << std::fixed << std::setprecision(digits) << v;
        const std::string t = o.str();
        std::memcpy(p, t.data(), t.size());
        return p + t.size();
    }
    const auto r = std::to_chars(p, end, v, std::chars_format::fixed, digits);
    return r.ptr;
}
} // namespace

bool write_csv(const std::string& filename, const std::string& metadata, size_t n, size_t T,
               const double* pd, const double* sd, const double* st) {
    std::ofstream f(filename, std::ios::binary);
    if (!f) { std::cerr << "Could not open CSV file " << filename << " for writing.\n"; return false; }
    f << metadata << "\n" << "particle_id,time,x,y,z,vx,vy,vz,mass,energy\n";
    constexpr size_t kBuf = 1 << 20, kRowMax = 10 * 340;     // a double in fixed notation has at most ~310 + 17 characters
    std::vector<char> buf(kBuf + kRowMax);
    char* p = buf.data();
    char* const flush_at = buf.data() + kBuf;
    char* const end = buf.data() + buf.size();
    for (size_t t = 0; t < T; ++t) {
        char tail[2][360];                                   // time and energy repeat for every particle of the snapshot
        const size_t time_len = (size_t)(put_fixed(tail[0], tail[0] + 360, sd[2 * t], 8) - tail[0]);
        const size_t energy_len = (size_t)(put_fixed(tail[1], tail[1] + 360, sd[2 * t + 1], 15) - tail[1]);
        for (size_t i = 0; i < n; ++i) {
            const double* q = pd + (t * n + i) * 6;
            p = put_fixed(p, end, st[2 * i], 8); *p++ = ',';
            std::memcpy(p, tail[0], time_len); p += time_len; *p++ = ',';
            for (int c = 0; c < 6; ++c) { p = put_fixed(p, end, q[c], 8); *p++ = ','; }
            p = put_fixed(p, end, st[2 * i + 1], 8); *p++ = ',';
            std::memcpy(p, tail[1], energy_len); p += energy_len; *p++ = '\n';
            if (p >= flush_at) { f.write(buf.data(), p - buf.data()); p = buf.data(); }
        }
    }
    f.write(buf.data(), p - buf.data());
    f.close();
    std::cout << "CSV output written to " << filename << "\n";
    return (bool)f;
}

namespace {
void check(int rc) { if (rc != NB_OK) throw std::runtime_error(nb_last_error()); }
const char* name_of(IntegrationMethod m) { return m == IntegrationMethod::EULER ? "Euler" : "Velocity Verlet"; }
const char* name_of(ForceMethod f) {
    switch (f) { case ForceMethod::PAIRWISE: return "Pairwise"; case ForceMethod::PAIRWISE_AVX2_FP32: return "Pairwise AVX2/FP32";
                 case ForceMethod::ADAPTIVE_MUTUAL: return "Adaptive Mutual"; default: return "Barnes-Hut"; }
}
const char* name_of(OutputMode o) {
    return o == OutputMode::BENCHMARK ? "None (Benchmark)" : o == OutputMode::FILE_CSV ? "CSV File" : "Real-time Visualization";
}
} // namespace

// "deviceInit": true — SPIRAL_GALAXY initial conditions made on the GPU (nb_init_spiral: the reference generator's formulas with
// counter-based draws), BENCHMARK mode: nothing but the final state ever crosses PCIe. What a 16M-particle run wants: the host
// generator takes seconds and 1 GB of upload there.
static RunReport run_device_init(const Config& cfg, long steps, bool quiet) {
    if (cfg.initSelected != "SPIRAL_GALAXY" || cfg.outputMode != OutputMode::BENCHMARK || !cfg.hasSeed)
        throw std::runtime_error("\"deviceInit\": true needs initializationMethod SPIRAL_GALAXY, outputMode BENCHMARK and a \"seed\"");
    RunReport rep;
    const Json& q = cfg.initParams;
    const long n = q["nStars"].as_int();
    ParticleSet p((size_t)n);
    const int prec = cfg.precision >= 0 ? cfg.precision : NB_PREC_FP32;
    nb_ctx* c = nullptr;
    auto t0 = std::chrono::high_resolution_clock::now();
    check(nb_create(&c, n, cfg.device, prec));
    double nrm[3] = {0, 0, 0};
    int rc = nb_set_methods(c, (int)cfg.integrationMethod, (int)cfg.forceMethod, cfg.subtractCom ? 1 : 0);
    if (!rc) rc = nb_init_spiral(c, (unsigned long long)cfg.seed, q["totalMass"].as_double(), q["nArms"].as_int(), q["galaxyRadius"].as_double(),
                                 q["thickness"].as_double(), q["perturbation"].as_double(), nrm);
    auto t1 = std::chrono::high_resolution_clock::now();
    if (!rc) rc = nb_compute_accel(c);
    if (!rc) rc = nb_step(c, cfg.dtYears, (int)steps);
    if (!rc) rc = nb_download(c, p.posMass.data(), p.vel.data());
    auto t2 = std::chrono::high_resolution_clock::now();
    nb_destroy(c);
    check(rc);
    rep.init_seconds = std::chrono::duration<double>(t1 - t0).count();
    rep.run_seconds = std::chrono::duration<double>(t2 - t1).count();
    rep.steps = steps; rep.n = (size_t)n;
    if (!quiet)
        std::cout << "Initialization time: " << rep.init_seconds << " seconds (on the device; disc normal " << nrm[0] << ", " << nrm[1] << ", "
                  << nrm[2] << ")\nInitialization mode: SPIRAL_GALAXY\nForce method: " << name_of(cfg.forceMethod)
                  << "\nExecution mode: GPU\nOutput mode: BENCHMARK\n";
    if (!cfg.snapshotOut.empty()) save_snapshot(cfg.snapshotOut, p);
    return rep;
}

// The flow of the reference's main (core/main.cpp:454-636) + System::runSimulation (core/System.cpp:87-157) for
// executionMode "GPU": BENCHMARK -> nb_run_benchmark, FILE_CSV -> nb_run_simulation + write_csv.
RunReport run_config(const Config& cfg, bool quiet) {
    RunReport rep;
    if (cfg.executionMode != ExecutionMode::GPU)
        throw std::runtime_error("executionMode \"CPU\" is the reference's OpenMP path; this host drives the B200 library only (no CPU fallback)");
    if (cfg.outputMode == OutputMode::VISUALIZATION && !cfg.hasFrames)
        throw std::runtime_error("outputMode \"VISUALIZATION\": the interactive OpenGL viewer is reference-only (no graphics stack on "
                                 "this machine); add a \"frames\" object to the config for headless PPM frames of the same shader arithmetic");
    const long steps = (long)(cfg.years / cfg.dtYears);                 // main.cpp:496
    const long stepFreq = std::max(1L, steps / 500);                    // main.cpp:499,525
    if (cfg.deviceInit) return run_device_init(cfg, steps, quiet);
    auto t0 = std::chrono::high_resolution_clock::now();
    InitResult init = create_initial_conditions(cfg);
    auto t1 = std::chrono::high_resolution_clock::now();
    rep.init_seconds = std::chrono::duration<double>(t1 - t0).count();
    rep.steps = steps; rep.n = init.particles.n;
    if (!quiet) std::cout << "Initialization time: " << rep.init_seconds << " seconds\n";
    std::string md = "# years: " + std::to_string(cfg.years) + "\n" + init.metadata;
    md += "# automatic steps: " + std::to_string(steps) + "\n# print frequency (steps): " + std::to_string(stepFreq) + "\n";
    md += std::string("# integration method: ") + name_of(cfg.integrationMethod) + "\n# force method: " + name_of(cfg.forceMethod) + "\n";
    md += "# execution mode: GPU\n# output mode: " + std::string(name_of(cfg.outputMode)) + "\n";
    md += "# time basis: years\n# dt (years): " + std::to_string(cfg.dtYears) + "\n# total years: " + std::to_string((long double)cfg.years) + "\n";
    if (!quiet)
        std::cout << "Initialization mode: " << cfg.initSelected << "\nForce method: " << name_of(cfg.forceMethod)
                  << "\nExecution mode: GPU\nOutput mode: " << name_of(cfg.outputMode) << "\n";
    ParticleSet& p = init.particles;
    const int prec = cfg.precision >= 0 ? cfg.precision : (cfg.outputMode == OutputMode::BENCHMARK ? NB_PREC_FP32 : NB_PREC_FP64);
    auto r0 = std::chrono::high_resolution_clock::now();
    if (cfg.outputMode == OutputMode::BENCHMARK) {
        check(nb_run_benchmark((long)p.n, p.posMass.data(), p.vel.data(), (int)cfg.integrationMethod, (int)cfg.forceMethod,
                               cfg.dtYears, (int)steps, 256, prec, cfg.subtractCom ? 1 : 0, cfg.device));
    } else if (cfg.outputMode == OutputMode::VISUALIZATION) {
        // runVisualizationOnGPU (visualization/cuda/VisualizationKernels.cu:52-300) without the window
        const std::string dir = cfg.outputDir + cfg.frameDir;
        std::filesystem::create_directories(dir);
        long frames[2] = {0, 0};
        if (!quiet) std::cout << "Starting headless visualization with " << p.n << " particles, frames -> " << dir << std::endl;
        check(nb_run_visualization((long)p.n, p.posMass.data(), p.vel.data(), (int)cfg.integrationMethod, (int)cfg.forceMethod,
                                   cfg.dtYears, (int)steps, cfg.physicsStepsPerFrame, cfg.frameWidth, cfg.frameHeight,
                                   (float)cfg.frameScale, dir.c_str(), cfg.frameEvery > 0 ? cfg.frameEvery : (int)stepFreq, nullptr,
                                   frames, prec, cfg.subtractCom ? 1 : 0, cfg.device));
        rep.frames_rendered = frames[0]; rep.frames_written = frames[1]; rep.output_path = dir;
    } else {
        const size_t T = (size_t)(steps / stepFreq + 1);                // core/System.cpp:107
        std::vector<double> pd(T * p.n * 6, 0.0), sd(T * 2, 0.0), st(p.n * 2, 0.0);
        if (!quiet) std::cout << "Running simulation on GPU..." << std::endl;
        check(nb_run_simulation((long)p.n, p.posMass.data(), p.vel.data(), (int)cfg.integrationMethod, (int)cfg.forceMethod,
                                cfg.dtYears, (int)steps, (int)stepFreq, pd.data(), sd.data(), st.data(), (long)T, prec,
                                cfg.subtractCom ? 1 : 0, cfg.device));
        rep.output_path = cfg.outputDir + cfg.outputFile;
        write_csv(rep.output_path, md, p.n, T, pd.data(), sd.data(), st.data());
    }
    auto r1 = std::chrono::high_resolution_clock::now();
    rep.run_seconds = std::chrono::duration<double>(r1 - r0).count();
    if (!cfg.snapshotOut.empty()) save_snapshot(cfg.snapshotOut, p);
    return rep;
}

} // namespace nbh
