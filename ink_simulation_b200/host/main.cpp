// main.cpp — Qt-free equivalent of the reference's src/main.cpp (Qt there only parses --help): create the output
// directory and run InkSim::simulate(.35, 2000).  Optional arguments: [output dir] [render timestep] [total timesteps],
// then any of `--resume` (continue from <dir>/checkpoint if there is one) and `--checkpoint` (write one at the end).
#include <cstdlib>
#include <filesystem>
#include <iostream>

#include "inkSim.h"

#define path "output"  // Directory name to write to

int main(int argc, char* argv[]) {
    if (argc > 1 && (std::string(argv[1]) == "--help" || std::string(argv[1]) == "-h")) {
        std::cout << "usage: " << argv[0] << " [output dir = output] [render timestep = .35] [total timesteps = 2000] [--resume] [--checkpoint]" << std::endl;
        return 0;
    }
    bool resume = false, checkpoint = false;
    std::vector<std::string> pos;
    for (int a = 1; a < argc; ++a) {
        const std::string s = argv[a];
        if (s == "--resume") resume = true;
        else if (s == "--checkpoint") checkpoint = true;
        else pos.push_back(s);
    }
    const std::string dir = pos.size() > 0 ? pos[0] : path;
    const float render = pos.size() > 1 ? std::atof(pos[1].c_str()) : .35f;
    const int steps = pos.size() > 2 ? std::atoi(pos[2].c_str()) : 2000;
    std::filesystem::create_directory(dir);
    try {
        InkSim sim(dir);
        if (resume && !sim.resume()) std::cerr << "inksim_b200: no checkpoint in " << dir << ", starting from the initial scene" << std::endl;
        sim.simulate(render, steps);
        if (checkpoint) sim.writeCheckpoint();
    } catch (const std::exception& e) {
        std::cerr << "inksim_b200: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
