// system.h — drop-in for the reference's src/system.h (zhang-tianran/Ink-Simulation): same global parameters
// (src/system.h:13-38), same public types (Cell, Particle, hash_func :41-68) and the same public `System` interface
// (:70-84).  What changed is underneath: `System` owns an opaque ink_ctx* (include/ink_capi.h) — grid and particles
// live in B200 HBM, System::solve is ONE call into libinkb200.so, and the host containers the getters hand out are
// mirrors refreshed lazily.  A caller written against the reference (src/inkSim.cpp, src/main.cpp) compiles unchanged.
//
// Every `const` below keeps the reference's name, type and committed value; each is overridable at compile time
// (-DINK_WATERGRID_X=8 ...) so one source tree builds every scene of the README without editing the header.
#pragma once
#include <cassert>
#include <iostream>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "ink_eigen_lite.h"

struct ink_ctx;

// ============== Global Constants (reference src/system.h:13-38) ==============
#ifndef INK_PART_FILE
#define INK_PART_FILE ""
#endif
#ifndef INK_WATERGRID_X
#define INK_WATERGRID_X 20
#endif
#ifndef INK_WATERGRID_Y
#define INK_WATERGRID_Y 45
#endif
#ifndef INK_WATERGRID_Z
#define INK_WATERGRID_Z 20
#endif
#ifndef INK_NUM_PARTICLES
#define INK_NUM_PARTICLES {10000, 10000}
#endif

const std::string PART_FILE = INK_PART_FILE;

const int WATERGRID_X = INK_WATERGRID_X;  /// Water grid length
const int WATERGRID_Y = INK_WATERGRID_Y;  /// Water grid height
const int WATERGRID_Z = INK_WATERGRID_Z;  /// Water grid width

const std::vector<int> NUM_PARTICLES = INK_NUM_PARTICLES;  // particles per drop

const bool USE_LIFETIME = false;
const float CELL_DIM = 1;     /// Cell dimension (is a cube, so length == width == height)
const int BUFFER_SIZE = 3;    /// Dictates the number/levels of neighbors

const float DENSITY = .95;    /// Fluid density

const float K_VORT = 1;       /// strength of vorticity

const float VISCOSITY = 1.0016;          /// Fluid viscosity
const float ATMOSPHERIC_PRESSURE = 1;

const Eigen::Vector3f gravity = Eigen::Vector3f(0, -0.1, 0);

const float K_CFL = 0.2f;
const float MIN_TIMESTEP = 0.005f;
const float MAX_TIMESTEP = 1.f;
// ==============================================================================

// How System::init seeds particles when PART_FILE is empty.  The committed reference calls the hemisphere emitter
// (src/system.cpp:18-23) and has the random box (initParticles, :93-115) commented out; both are kept.
#ifndef INK_SEED_BOX
#define INK_SEED_BOX 0
#endif

typedef struct Cell {
    Eigen::Vector3f oldVelocity;
    Eigen::Vector3f currVelocity;

    Eigen::Vector3f curl;
    std::vector<Eigen::Vector3i> neighbors;

    friend std::ostream& operator<<(std::ostream& strm, const Cell& obj);
} Cell;

typedef struct Particle {
    Eigen::Vector3f position;
    Eigen::Vector3f velocity;
    float opacity;   /// In [0, 1]
    float lifeTime;

    friend std::ostream& operator<<(std::ostream& strm, const Particle& obj);
} Particle;

struct hash_func {
    size_t operator()(const Eigen::Vector3i& v) const {
        assert(v.x() >= 0 && v.y() >= 0 && v.z() >= 0);
        return 541 * v.x() + 79 * v.y() + 31 * v.z();
    }
};

class System {
public:
    System();
    ~System();
    System(const System&) = delete;
    System& operator=(const System&) = delete;

    void init();
    double solve(double timeToNextRender);
    const std::vector<std::vector<Particle>>& getInkParticles();
    const std::unordered_map<Eigen::Vector3i, Cell, hash_func> getWaterGrid();

    // check if particles or watergrid values for pos/vel have inf or nans
    void checkNanAndInf();

    friend std::ostream& operator<<(std::ostream& strm, const System& obj);

    // ---- additions (not in the reference): zero-copy access for the run loop, and the raw handle -----------------
    // Positions of one drop straight into caller memory (N x 3 floats, particle index order) — what the per-frame
    // .ply payload is (src/inkSim.cpp:85-87) — without materialising std::vector<Particle>.
    int numDrops() const;
    long long numParticles(int drop) const;
    void copyPositions(int drop, float* xyz);
    // same, asynchronously (ink_fetch_positions_async): returns at once, `xyz` (pinned host memory) is complete after
    // syncIO(); the device snapshot is taken in stream order, so later solve() calls do not disturb it
    void fetchPositionsAsync(int drop, float* xyz);
    void syncIO();
    // dense currVelocity, (X, Y, Z) C-order, three arrays of X*Y*Z floats
    void copyVelocityField(float* u, float* v, float* w);
    ink_ctx* handle() { return m_ctx; }
    // checkpoint / resume (InkSim::writeCheckpoint / resume; SURVEY 8f-3 "matching load" of the velocity dump)
    void copyParticles(int drop, float* xyz, float* vel);        // positions AND stored velocities (Q4)
    void copyWarmStart(float* q);                                // the Poisson solver's start vector, X*Y*Z floats
    void setVelocityField(const float* u, const float* v, const float* w);
    void setWarmStart(const float* q);
    void setInk(int drop, const float* xyz, const float* vel, long long n);   // replace (or append) one drop
    bool getPhantom(double v[3]);                                // phantom-cell state for checkpoints
    void setPhantom(bool have, const double v[3]);

private:
    ink_ctx* m_ctx = nullptr;
    void fail(const char* what) const;

    // init particles (src/system.cpp:93-148, src/inkSolver.cpp:11-42): host-side seeding, uploaded once
    void initFromFile();
    void initParticles();
    void emitParticleHemisphere(float radius);
    void uploadInk();

    /// host mirrors, refreshed by the getters
    std::unordered_map<Eigen::Vector3i, Cell, hash_func> m_waterGrid;
    std::vector<std::vector<Particle>> m_ink;
    bool m_inkStale = false;
};
