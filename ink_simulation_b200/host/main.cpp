// main.cpp — Qt-free equivalent of the reference's src/main.cpp (Qt there only parses --help): create the output
// directory and run InkSim::simulate(.35, 2000).  Optional arguments: [output dir] [render timestep] [total timesteps].
#include <cstdlib>
#include <filesystem>
#include <iostream>

#include "inkSim.h"

#define path "output"  // Directory name to write to

int main(int argc, char* argv[]) {
    if (argc > 1 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) {
        std::cout << "usage: " << argv[0] << " [output dir = output] [render timestep = .35] [total timesteps = 2000]" << std::endl;
        return 0;
    }
    const std::string dir = argc > 1 ? argv[1] : path;
    const float render = argc > 2 ? std::atof(argv[2]) : .35f;
    const int steps = argc > 3 ? std::atoi(argv[3]) : 2000;
    std::filesystem::create_directory(dir);
    try {
        InkSim sim(dir);
        sim.simulate(render, steps);
    } catch (const std::exception& e) {
        std::cerr << "inksim_b200: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
