// nbody_host — C++17 drop-in for the reference executable (core/main.cpp:454-636) on top of libnbody_b200.so.
//   nbody_host [config.json]                     run one configuration (default: ./config.json, like the reference)
//   nbody_host --auto-benchmark [options]        the reference's sweep harness (core/main.cpp:291-452) on the GPU
//       --sizes 1024,4096,...   particle counts            (default: the reference's 64 ... 131072 list)
//       --forces PAIRWISE,ADAPTIVE_MUTUAL,PAIRWISE_AVX2_FP32 (default PAIRWISE)
//       --integrators VELOCITY_VERLET,EULER                (default VELOCITY_VERLET)
//       --runs K (default 5)  --years Y (default: by size)  --dt D (default 0.01)  --seed S  --precision FP32|FP64
//       --device I  --no-warm  --out file.csv (default presentation/benchmarks/benchmark_results_B200_<unix time>.csv)
#include <chrono>
#include <iostream>
#include <sstream>

#include "nbh_host.hpp"

namespace {
std::vector<std::string> split_list(const std::string& s) {
    std::vector<std::string> out; std::stringstream ss(s); std::string tok;
    while (std::getline(ss, tok, ',')) if (!tok.empty()) out.push_back(tok);
    return out;
}
nbh::ForceMethod force_of(const std::string& s) {
    if (s == "PAIRWISE") return nbh::ForceMethod::PAIRWISE;
    if (s == "PAIRWISE_AVX2_FP32" || s == "PAIRWISE_AVX2/FP32") return nbh::ForceMethod::PAIRWISE_AVX2_FP32;
    if (s == "ADAPTIVE_MUTUAL") return nbh::ForceMethod::ADAPTIVE_MUTUAL;
    if (s == "BARNES_HUT") return nbh::ForceMethod::BARNES_HUT;
    throw std::runtime_error("unknown force method '" + s + "'");
}
int auto_benchmark(int argc, char** argv) {
    nbh::SweepSpec spec;
    spec.outputFile = "presentation/benchmarks/benchmark_results_B200_" +
        std::to_string(std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count()) + ".csv";
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&]() -> std::string { if (i + 1 >= argc) throw std::runtime_error("missing value after " + a); return argv[++i]; };
        if (a == "--sizes") { spec.particleCounts.clear(); for (auto& t : split_list(val())) spec.particleCounts.push_back(std::stoi(t)); }
        else if (a == "--forces") { spec.forceMethods.clear(); for (auto& t : split_list(val())) spec.forceMethods.push_back(force_of(t)); }
        else if (a == "--integrators") {
            spec.integrationMethods.clear();
            for (auto& t : split_list(val())) {
                if (t == "EULER") spec.integrationMethods.push_back(nbh::IntegrationMethod::EULER);
                else if (t == "VELOCITY_VERLET") spec.integrationMethods.push_back(nbh::IntegrationMethod::VELOCITY_VERLET);
                else throw std::runtime_error("unknown integration method '" + t + "'");
            }
        }
        else if (a == "--runs") spec.runs = std::stoi(val());
        else if (a == "--years") spec.years = std::stod(val());
        else if (a == "--dt") spec.dtYears = std::stod(val());
        else if (a == "--seed") { spec.hasSeed = true; spec.seed = std::stoul(val()); }
        else if (a == "--precision") { const std::string p = val(); if (p == "FP64") spec.precision = 1; else if (p == "FP32") spec.precision = 0; else throw std::runtime_error("precision must be FP32 or FP64"); }
        else if (a == "--device") spec.device = std::stoi(val());
        else if (a == "--no-warm") spec.warm = false;
        else if (a == "--out") spec.outputFile = val();
        else throw std::runtime_error("unknown option " + a);
    }
    const std::vector<nbh::SweepResult> res = nbh::run_sweep(spec, false);
    for (const nbh::SweepResult& r : res) if (!r.success) return 2;
    return 0;
}
} // namespace

int main(int argc, char** argv) {
    try {
        if (argc > 1 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) {
            std::cout << "usage: nbody_host [config.json]\n"
                         "       nbody_host --auto-benchmark [--sizes a,b,..] [--forces PAIRWISE,ADAPTIVE_MUTUAL,PAIRWISE_AVX2_FP32]\n"
                         "                  [--integrators VELOCITY_VERLET,EULER] [--runs K] [--years Y] [--dt D] [--seed S]\n"
                         "                  [--precision FP32|FP64] [--device I] [--no-warm] [--out file.csv]\n"
                         "config.json: the reference's keys, plus optional \"seed\", \"precision\", \"subtractCom\", \"device\",\n"
                         "             \"snapshotOut\", \"frames\" (see n-body-simulation_b200/host/nbh_host.hpp)\n";
            return 0;
        }
        if (argc > 1 && std::string(argv[1]) == "--auto-benchmark") return auto_benchmark(argc, argv);   // main.cpp:457-460
        const std::string path = argc > 1 ? argv[1] : "config.json";
        const nbh::Config cfg = nbh::parse_config(path);
        const nbh::RunReport rep = nbh::run_config(cfg, false);
        const double n2 = (double)rep.n * (double)rep.n;
        std::cout << "\nSimulation time: " << rep.run_seconds << " seconds (using GPU)" << std::endl;   // main.cpp:617-627
        if (rep.frames_rendered > 0)
            std::cout << "Frames: " << rep.frames_rendered << " rendered, " << rep.frames_written << " written to " << rep.output_path
                      << " (" << rep.frames_rendered / rep.run_seconds << " FPS)" << std::endl;    // VisualizationKernels.cu:287-291
        if (rep.run_seconds > 0 && rep.steps > 0)
            std::cout << "Throughput: " << rep.steps / rep.run_seconds << " steps/s, "
                      << n2 * (double)(rep.steps + 1) / rep.run_seconds / 1e9 << " G-interactions/s (N=" << rep.n << ", "
                      << rep.steps << " steps + initial force, copies included)" << std::endl;
        return 0;
    } catch (const std::exception& ex) {
        std::cerr << "Error: " << ex.what() << "\n";
        return 1;
    }
}
