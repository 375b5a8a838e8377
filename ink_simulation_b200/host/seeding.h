// seeding.h — host-side particle seeding of System::init (reference src/system.cpp:62-148, src/inkSolver.cpp:11-42),
// kept free of the GPU context so it can be checked on a CPU-only machine against the compiled reference.
#pragma once
#include <string>
#include <vector>

#include "system.h"

namespace inkseed {
void box(int X, int Y, int Z, const std::vector<int>& per_drop, std::vector<std::vector<Particle>>& drops);
std::vector<Particle> from_file(const std::string& path);
void hemisphere(int X, int Y, int Z, float radius, std::vector<Particle>& out);
void emitter_scene(int X, int Y, int Z, std::vector<Particle>& out);   // the 120-radius loop of System::init
// loop-variable tables of that scene for the device seeder (ink_seed_emitter)
void emitter_tables(std::vector<float>& radius, std::vector<float>& sin_i, std::vector<float>& cos_i, std::vector<float>& sin_j, std::vector<float>& cos_j);
}  // namespace inkseed
