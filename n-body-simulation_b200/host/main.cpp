// nbody_host — C++17 drop-in for the reference executable's config.json flow (core/main.cpp:454-636) on top of
// libnbody_b200.so. Usage: nbody_host [config.json]   (default: ./config.json, like the reference)
#include <iostream>

#include "nbh_host.hpp"

int main(int argc, char** argv) {
    try {
        const std::string path = argc > 1 ? argv[1] : "config.json";
        const nbh::Config cfg = nbh::parse_config(path);
        const nbh::RunReport rep = nbh::run_config(cfg, false);
        const double n2 = (double)rep.n * (double)rep.n;
        std::cout << "\nSimulation time: " << rep.run_seconds << " seconds (using GPU)" << std::endl;   // main.cpp:617-627
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
