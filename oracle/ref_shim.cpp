// TEST INFRASTRUCTURE — not product code.
//
// C-ABI shim over the *compiled, unmodified-in-arithmetic* reference (zhang-tianran/Ink-Simulation).
// Built by oracle/build_ref.sh from a scratch copy of /root/reference/src (nothing is copied into
// this repository) into oracle/_ref/libinkref.so.  The only edit the recipe applies to the copy is
// turning the three grid-size constants of src/system.h:15-17 from `const int` into `inline int`
// so ONE shared object serves every grid size (the constants are only ever used in run-time
// expressions).  Private members of `System` are reached with the usual `#define private public`
// include hack (std / Eigen headers are included first so only the reference's own classes are
// affected).
//
// Array convention shared by ref / oracle / GPU C-ABI: dense (X, Y, Z) C-order, k (z) fastest,
// i.e. idx = (i*Y + j)*Z + k.  Particle arrays are AoS xyz (N x 3) floats.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.

#include <iostream>
#include <vector>
#include <unordered_map>
#include <unordered_set>
#include <string>
#include <fstream>
#include <sstream>
#include <chrono>
#include <cmath>
#include <cstring>
#include <Eigen/Dense>
#include <Eigen/Sparse>

#define private public
#include "system.h"
#include "inkSim.h"
#undef private

using Eigen::Vector3f;
using Eigen::Vector3i;

namespace {
struct RefCtx {
    System sys;
    int X, Y, Z;
};
inline void copy_old_from_curr(System& s) {
    // src/waterSolver.cpp:56-59 (and the four identical sweeps after each phase)
    for (auto& kv : s.m_waterGrid) kv.second.oldVelocity = kv.second.currVelocity;
}
}  // namespace

extern "C" {

// Sets the (process-global) grid size, builds the grid + pressure factorisation
// (System::initWaterGrid, src/system.cpp:40-60 -> initPressureA src/waterSolver.cpp:240-261).
// No particles are created; use ref_set_particles.
void* ref_create(int X, int Y, int Z) {
    WATERGRID_X = X;
    WATERGRID_Y = Y;
    WATERGRID_Z = Z;
    RefCtx* c = new RefCtx();
    c->X = X; c->Y = Y; c->Z = Z;
    c->sys.initWaterGrid();
    return c;
}

void ref_destroy(void* h) { delete static_cast<RefCtx*>(h); }

long ref_grid_entries(void* h) { return (long)static_cast<RefCtx*>(h)->sys.m_waterGrid.size(); }

int ref_num_drops(void* h) { return (int)static_cast<RefCtx*>(h)->sys.m_ink.size(); }

long ref_num_particles(void* h, int drop) {
    RefCtx* c = static_cast<RefCtx*>(h);
    if (drop < 0 || drop >= (int)c->sys.m_ink.size()) return 0;
    return (long)c->sys.m_ink[drop].size();
}

// pos/vel: N x 3 AoS floats (vel may be null => zero).  Replaces drop `drop` (appending drops as needed).
void ref_set_particles(void* h, int drop, const float* pos, const float* vel, long n) {
    RefCtx* c = static_cast<RefCtx*>(h);
    while ((int)c->sys.m_ink.size() <= drop) c->sys.m_ink.emplace_back();
    std::vector<Particle>& ink = c->sys.m_ink[drop];
    ink.clear();
    ink.reserve(n);
    for (long p = 0; p < n; ++p) {
        Particle q;
        q.position = Vector3f(pos[3 * p], pos[3 * p + 1], pos[3 * p + 2]);
        q.velocity = vel ? Vector3f(vel[3 * p], vel[3 * p + 1], vel[3 * p + 2]) : Vector3f(0, 0, 0);
        q.opacity = 1.f;
        q.lifeTime = 5.f;
        ink.push_back(q);
    }
}

void ref_get_particles(void* h, int drop, float* pos, float* vel) {
    RefCtx* c = static_cast<RefCtx*>(h);
    const std::vector<Particle>& ink = c->sys.m_ink[drop];
    for (size_t p = 0; p < ink.size(); ++p) {
        if (pos) { pos[3 * p] = ink[p].position[0]; pos[3 * p + 1] = ink[p].position[1]; pos[3 * p + 2] = ink[p].position[2]; }
        if (vel) { vel[3 * p] = ink[p].velocity[0]; vel[3 * p + 1] = ink[p].velocity[1]; vel[3 * p + 2] = ink[p].velocity[2]; }
    }
}

// The reference's own seeding routines (src/system.cpp:93-115, src/inkSolver.cpp:11-42).
// NUM_PARTICLES is the header constant {10000, 10000}; initParticles uses glibc rand().
void ref_init_particles_box(void* h, unsigned seed) {
    RefCtx* c = static_cast<RefCtx*>(h);
    srand(seed);
    c->sys.m_ink.clear();
    c->sys.initParticles();
}
void ref_init_emitter(void* h) {  // System::init()'s default branch, src/system.cpp:17-23
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.m_ink.clear();
    std::vector<Particle> ink;
    ink.reserve(400000);
    c->sys.m_ink.push_back(ink);
    for (float i = 0; i < 2 * M_PI; i += (M_PI / 60.f)) c->sys.emitParticleHemisphere(i);
}

// Sets BOTH oldVelocity and currVelocity (the step-entry invariant, src/waterSolver.cpp:43-46).
void ref_set_velocity(void* h, const float* u, const float* v, const float* w) {
    RefCtx* c = static_cast<RefCtx*>(h);
    for (int i = 0; i < c->X; i++)
        for (int j = 0; j < c->Y; j++)
            for (int k = 0; k < c->Z; k++) {
                size_t idx = ((size_t)i * c->Y + j) * c->Z + k;
                Cell& cell = c->sys.m_waterGrid.at(Vector3i(i, j, k));
                cell.oldVelocity = Vector3f(u[idx], v[idx], w[idx]);
                cell.currVelocity = cell.oldVelocity;
            }
}

// which: 0 = oldVelocity, 1 = currVelocity, 2 = curl
void ref_get_field(void* h, int which, float* u, float* v, float* w) {
    RefCtx* c = static_cast<RefCtx*>(h);
    for (int i = 0; i < c->X; i++)
        for (int j = 0; j < c->Y; j++)
            for (int k = 0; k < c->Z; k++) {
                size_t idx = ((size_t)i * c->Y + j) * c->Z + k;
                const Cell& cell = c->sys.m_waterGrid.at(Vector3i(i, j, k));
                const Vector3f& s = which == 0 ? cell.oldVelocity : (which == 1 ? cell.currVelocity : cell.curl);
                u[idx] = s[0]; v[idx] = s[1]; w[idx] = s[2];
            }
}

// max over ALL map entries (phantoms included) of |oldVelocity| — what calcTimeStep sees.
float ref_max_speed_all_entries(void* h) {
    RefCtx* c = static_cast<RefCtx*>(h);
    float m = 0;
    for (auto& kv : c->sys.m_waterGrid) m = std::max(m, kv.second.oldVelocity.norm());
    return m;
}

double ref_calc_dt(void* h, double timeToNextRender) {
    return static_cast<RefCtx*>(h)->sys.calcTimeStep(timeToNextRender);
}

// Each phase hook = the phase followed by the reference's `old = curr` sweep, exactly as
// System::updateVelocityField sequences them (src/waterSolver.cpp:41-96).
void ref_advect(void* h, float dt) {
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.applyConvection(dt);
    copy_old_from_curr(c->sys);
}
void ref_forces(void* h, float dt) {
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.applyExternalForces(dt);
    copy_old_from_curr(c->sys);
}
void ref_viscosity(void* h, float dt) {
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.applyViscosity(dt);
    copy_old_from_curr(c->sys);
}
// RHS-free view of the pressure the reference solves for (calculatePressure, src/waterSolver.cpp:264-281);
// p_out in (X,Y,Z) C-order (un-scrambling grid2mat, src/system.h:110-112).  Does not modify the grid.
void ref_calc_pressure(void* h, float dt, float* p_out) {
    RefCtx* c = static_cast<RefCtx*>(h);
    Eigen::VectorXf p = c->sys.calculatePressure(dt);
    for (int i = 0; i < c->X; i++)
        for (int j = 0; j < c->Y; j++)
            for (int k = 0; k < c->Z; k++)
                p_out[((size_t)i * c->Y + j) * c->Z + k] = p[c->sys.grid2mat(i, j, k)];
}
void ref_pressure(void* h, float dt) {
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.applyPressure(dt);
    copy_old_from_curr(c->sys);
}
void ref_vorticity(void* h, float dt) {
    RefCtx* c = static_cast<RefCtx*>(h);
    c->sys.applyVorticity(dt);
    copy_old_from_curr(c->sys);
}
void ref_divergence(void* h, float* div_out) {  // getDivergence on oldVelocity, src/systemHelpers.cpp:30-43
    RefCtx* c = static_cast<RefCtx*>(h);
    for (int i = 0; i < c->X; i++)
        for (int j = 0; j < c->Y; j++)
            for (int k = 0; k < c->Z; k++)
                div_out[((size_t)i * c->Y + j) * c->Z + k] = c->sys.getDivergence(i, j, k);
}
void ref_sample_velocity(void* h, const float* pos, long n, float* out) {  // getVelocity, src/systemHelpers.cpp:110-115
    RefCtx* c = static_cast<RefCtx*>(h);
    for (long p = 0; p < n; ++p) {
        Vector3f v = c->sys.getVelocity(Vector3f(pos[3 * p], pos[3 * p + 1], pos[3 * p + 2]));
        out[3 * p] = v[0]; out[3 * p + 1] = v[1]; out[3 * p + 2] = v[2];
    }
}
void ref_velocity_field_update(void* h, float dt) { static_cast<RefCtx*>(h)->sys.updateVelocityField(dt); }
void ref_particles(void* h, float dt) { static_cast<RefCtx*>(h)->sys.updateParticles(dt); }
double ref_solve(void* h, double timeToNextRender) { return static_cast<RefCtx*>(h)->sys.solve(timeToNextRender); }

// Per-phase wall-clock of one System::solve (for the CPU baseline): t[0..6] =
// dt, convection, forces, viscosity, pressure, vorticity, particles (seconds). Returns dt.
double ref_solve_timed(void* h, double timeToNextRender, double* t) {
    RefCtx* c = static_cast<RefCtx*>(h);
    using clk = std::chrono::steady_clock;
    auto s = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
    auto t0 = clk::now();
    double dt = c->sys.calcTimeStep(timeToNextRender);
    float fdt = (float)dt;
    auto t1 = clk::now();
    c->sys.applyConvection(fdt); copy_old_from_curr(c->sys);
    auto t2 = clk::now();
    c->sys.applyExternalForces(fdt); copy_old_from_curr(c->sys);
    auto t3 = clk::now();
    c->sys.applyViscosity(fdt); copy_old_from_curr(c->sys);
    auto t4 = clk::now();
    c->sys.applyPressure(fdt); copy_old_from_curr(c->sys);
    auto t5 = clk::now();
    c->sys.applyVorticity(fdt); copy_old_from_curr(c->sys);
    auto t6 = clk::now();
    c->sys.updateParticles(fdt);
    auto t7 = clk::now();
    t[0] = s(t0, t1); t[1] = s(t1, t2); t[2] = s(t2, t3); t[3] = s(t3, t4);
    t[4] = s(t4, t5); t[5] = s(t5, t6); t[6] = s(t6, t7);
    return dt;
}

// InkSim::writeToFile (src/inkSim.cpp:43-96) on caller-provided particles: writes
// <dir>/drop<i>/<frame>.ply with the reference's own code (dir/drop<i> must exist).
// The InkSim constructor runs System::init() (emitter scene) first, so keep the grid small.
void ref_write_ply(const char* dir, int frame, int ndrops, const float* const* pos, const long* n) {
    InkSim sim{std::string(dir)};
    sim.ink_system.m_ink.clear();
    for (int d = 0; d < ndrops; ++d) {
        std::vector<Particle> ink;
        for (long p = 0; p < n[d]; ++p) {
            Particle q;
            q.position = Vector3f(pos[d][3 * p], pos[d][3 * p + 1], pos[d][3 * p + 2]);
            q.velocity = Vector3f(0, 0, 0);
            q.opacity = 1.f; q.lifeTime = 5.f;
            ink.push_back(q);
        }
        sim.ink_system.m_ink.push_back(ink);
    }
    sim.writeToFile(frame);
}

}  // extern "C"
