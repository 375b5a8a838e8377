// seeding.cpp — see seeding.h.
#include "seeding.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

using Eigen::Vector3f;

namespace inkseed {

static double randBetween(double lo, double hi) { return lo + ((double)rand() / RAND_MAX) * (hi - lo); }

// initParticles (src/system.cpp:93-115): per_drop[i] particles per drop, uniform (glibc rand(), never seeded) in the
// box x in X/2 -+ X/5, y in [Y-1.5, Y-1], z in Z/2 -+ Z/3; draw order x, y, z.
void box(int X, int Y, int Z, const std::vector<int>& per_drop, std::vector<std::vector<Particle>>& drops) {
    for (size_t d = 0; d < per_drop.size(); ++d) {
        std::vector<Particle> ink;
        ink.reserve(per_drop[d]);
        for (int n = 0; n < per_drop[d]; ++n) {
            Particle p;
            const float x = randBetween(X / 2.f - X / 5.f, X / 2.f + X / 5.f);
            const float y = randBetween(Y - 1.5, Y - 1.);
            const float z = randBetween(Z / 2.f - Z / 3.f, Z / 2.f + Z / 3.f);
            p.position = Vector3f(x, y, z);
            p.velocity = Vector3f(0, 0, 0);
            p.opacity = 1.f;
            p.lifeTime = 5.f;
            ink.push_back(p);
        }
        drops.push_back(ink);
    }
}

// initFromFile (src/system.cpp:117-148): one particle per line, "x y z vx vy vz" separated by single spaces.
std::vector<Particle> from_file(const std::string& path) {
    std::ifstream pf(path);
    if (!pf.is_open()) throw std::runtime_error("PART_FILE cannot be opened: " + path);
    std::vector<Particle> ink;
    std::string line;
    while (std::getline(pf, line)) {
        float vals[6] = {0, 0, 0, 0, 0, 0};
        std::stringstream s(line);
        std::string tok;
        int idx = 0;
        while (idx < 6 && std::getline(s, tok, ' ')) vals[idx++] = std::stof(tok);
        Particle p;
        p.position = Vector3f(vals[0], vals[1], vals[2]);
        p.velocity = Vector3f(vals[3], vals[4], vals[5]);
        p.opacity = 1.f;
        p.lifeTime = 5.f;
        ink.push_back(p);
    }
    return ink;
}

// emitParticleHemisphere (src/inkSolver.cpp:11-42).  The reference rotates by AngleAxisf(pi, (n + up).normalized()) with
// n = -up, i.e. about the ZERO axis: Eigen leaves a zero vector un-normalised and the rotation matrix degenerates to
// cos(pi) * I = -I, so velocity = -(x, y, z) and position = ceiling centre + radius * velocity.  The float
// loop counters (73 x 37 iterations per radius) are kept as float accumulations.
void hemisphere(int X, int Y, int Z, float radius, std::vector<Particle>& out) {
    // double-precision M_PI expressions rounded to float once, exactly as the reference writes them
    const float fraction = M_PI / 36.f;
    const float twoPi = M_PI * 2.f;
    const float c = std::cos((float)M_PI);   // -1: the only surviving term of the degenerate rotation
    for (float i = 0; i < twoPi; i += fraction) {
        for (float j = 0; j < M_PI; j += fraction) {
            const float x = radius * std::sin(j) * std::sin(i);
            const float y = radius * std::cos(j);
            const float z = radius * std::sin(j) * std::cos(i);
            Particle p;
            // rows of the degenerate rotation matrix c*I applied the way a 3x3 product is: (c*x + 0*y) + 0*z — this keeps
            // the sign of zero identical to the reference (-0 + 0 = +0), so the .ply bytes agree
            const float zf = 0.f;
            p.velocity = Vector3f((c * x + zf * y) + zf * z, (zf * x + c * y) + zf * z, (zf * x + zf * y) + c * z);
            p.position = Vector3f(X / 2.f, Y - 0.1f, Z / 2.f) + radius * p.velocity;
            p.opacity = 1.f;
            p.lifeTime = 5.f;
            out.push_back(p);
        }
    }
}

// The loop-variable tables of the emitter scene for ink_seed_emitter (device seeding): the float loop counters of System::init
// (src/system.cpp:18-23) and emitParticleHemisphere (src/inkSolver.cpp:13-14,19-20) and glibc's sinf / cosf of them — the only
// parts of the seeder that are sequential or libm-dependent; the per-particle products run on the GPU.
void emitter_tables(std::vector<float>& radius, std::vector<float>& sin_i, std::vector<float>& cos_i, std::vector<float>& sin_j, std::vector<float>& cos_j) {
    const float fraction = M_PI / 36.f;
    const float twoPi = M_PI * 2.f;
    for (float i = 0; i < 2 * M_PI; i += (M_PI / 60.f)) radius.push_back(i);
    for (float i = 0; i < twoPi; i += fraction) { sin_i.push_back(std::sin(i)); cos_i.push_back(std::cos(i)); }
    for (float j = 0; j < M_PI; j += fraction) { sin_j.push_back(std::sin(j)); cos_j.push_back(std::cos(j)); }
}

void emitter_scene(int X, int Y, int Z, std::vector<Particle>& out) {
    for (float i = 0; i < 2 * M_PI; i += (M_PI / 60.f)) hemisphere(X, Y, Z, i, out);
}

}  // namespace inkseed

// C hooks for the CPU-side tests (ctypes): the emitter scene as flat arrays
extern "C" long inkhost_emitter_scene(int X, int Y, int Z, float* pos, float* vel, long cap) {
    std::vector<Particle> out;
    inkseed::emitter_scene(X, Y, Z, out);
    const long n = (long)out.size();
    for (long i = 0; i < n && i < cap; ++i)
        for (int a = 0; a < 3; ++a) { pos[3 * i + a] = out[i].position[a]; vel[3 * i + a] = out[i].velocity[a]; }
    return n;
}

// one drop of n particles in the initParticles box, drawn the way a fresh process draws them: glibc rand() is never seeded
// by the reference, i.e. the sequence of srand(1)
// SIDE EFFECT: re-seeds the process-global rand() (inkseed::box draws from rand() like the reference's initParticles does); a
// caller that uses rand() itself should take inkhost_rand_stream + ink_seed_box instead, which leave the global generator alone.
extern "C" long inkhost_box_scene(int X, int Y, int Z, int n, float* pos, long cap) {
    std::srand(1);
    std::vector<std::vector<Particle>> drops;
    inkseed::box(X, Y, Z, {n}, drops);
    const long m = (long)drops[0].size();
    for (long i = 0; i < m && i < cap; ++i)
        for (int a = 0; a < 3; ++a) pos[3 * i + a] = drops[0][i].position[a];
    return m;
}

// initFromFile: returns the particle count (call with cap = 0 to size the arrays), -1 if the file cannot be read
extern "C" long inkhost_from_file(const char* path, float* pos, float* vel, long cap) {
    try {
        const std::vector<Particle> ink = inkseed::from_file(path);
        const long n = (long)ink.size();
        for (long i = 0; i < n && i < cap; ++i)
            for (int a = 0; a < 3; ++a) { pos[3 * i + a] = ink[i].position[a]; vel[3 * i + a] = ink[i].velocity[a]; }
        return n;
    } catch (const std::exception&) { return -1; }
}

// device seeding (ink_seed_emitter / ink_seed_box): the tables and the rand() stream the Python host hands to the C ABI
extern "C" void inkhost_emitter_tables(float* radius, int* nr, float* sin_i, float* cos_i, int* ni, float* sin_j, float* cos_j, int* nj, int cap) {
    std::vector<float> r, si, ci, sj, cj;
    inkseed::emitter_tables(r, si, ci, sj, cj);
    *nr = (int)r.size(); *ni = (int)si.size(); *nj = (int)sj.size();
    for (int a = 0; a < *nr && a < cap; ++a) radius[a] = r[a];
    for (int a = 0; a < *ni && a < cap; ++a) { sin_i[a] = si[a]; cos_i[a] = ci[a]; }
    for (int a = 0; a < *nj && a < cap; ++a) { sin_j[a] = sj[a]; cos_j[a] = cj[a]; }
}
// a fresh process's rand() sequence (the reference never seeds it).  glibc's rand() is random() on the default 128-byte TYPE_3
// state; the same generator on a PRIVATE state (initstate_r with seed 1) gives the same numbers without re-seeding the
// process-global one under the caller's feet (tests/test_host_cpp.py compares both).
extern "C" void inkhost_rand_stream(int32_t* out, long n) {
    char statebuf[128];
    struct random_data rd;
    memset(&rd, 0, sizeof rd);
    initstate_r(1u, statebuf, sizeof statebuf, &rd);
    for (long a = 0; a < n; ++a) { int32_t v; random_r(&rd, &v); out[a] = v; }
}
