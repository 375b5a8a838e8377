// system.cpp — the reference's `System` (src/system.{h,cpp}, src/inkSolver.cpp:5-9) as a thin owner of an ink_ctx.
// All per-timestep arithmetic runs in libinkb200.so on the GPU; this file only seeds particles on the host the way
// System::init does, forwards solve(), and refreshes the host mirrors that the reference's getters expose.
#include "system.h"

#include <cmath>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "../../include/ink_capi.h"
#include "seeding.h"

using Eigen::Vector3f;
using Eigen::Vector3i;

System::System() {}

System::~System() {
    if (m_ctx) ink_destroy(m_ctx);
}

void System::fail(const char* what) const {
    throw std::runtime_error(std::string(what) + ": " + ink_last_error(m_ctx));
}

// System::init (src/system.cpp:10-26): grid first (zero velocity; no LU factorisation to set up — the pressure solve
// is matrix-free), then the particles.
void System::init() {
    ink_params p;
    ink_default_params(&p);
    p.X = WATERGRID_X; p.Y = WATERGRID_Y; p.Z = WATERGRID_Z;
    p.cell_dim = CELL_DIM; p.density = DENSITY; p.k_vort = K_VORT; p.viscosity = VISCOSITY; p.p_atm = ATMOSPHERIC_PRESSURE;
    p.gravity[0] = gravity[0]; p.gravity[1] = gravity[1]; p.gravity[2] = gravity[2];
    p.k_cfl = K_CFL; p.dt_min = MIN_TIMESTEP; p.dt_max = MAX_TIMESTEP; p.buffer_size = BUFFER_SIZE;
    if (const char* e = std::getenv("INK_PRECISION")) p.precision = std::atoi(e);
    if (const char* e = std::getenv("INK_ARITH")) p.arith = std::string(e) == "fast" ? INK_ARITH_FAST : INK_ARITH_STRICT;
    if (const char* e = std::getenv("INK_DEVICE")) p.device = std::atoi(e);
    if (ink_create(&p, &m_ctx) != 0) throw std::runtime_error(std::string("ink_create: ") + ink_last_error(nullptr));

    // Particle seeding.  The emitter scene and the initParticles box are produced ON THE DEVICE (ink_seed_emitter /
    // ink_seed_box): the host only supplies the float loop counters with their sinf / cosf, or the rand() stream.
    // INK_SEED_ON_HOST=1 keeps the host seeders (seeding.cpp, the code the CPU tests pin to the compiled reference) + one upload.
    const bool on_host = std::getenv("INK_SEED_ON_HOST") && std::atoi(std::getenv("INK_SEED_ON_HOST")) != 0;
    Particle blank;
    blank.position = blank.velocity = Vector3f(0, 0, 0);
    blank.opacity = 1.f;
    blank.lifeTime = 5.f;
    if (PART_FILE != "") {
        initFromFile();
        uploadInk();
    } else if (INK_SEED_BOX) {
        if (on_host) { initParticles(); uploadInk(); return; }
        for (size_t d = 0; d < NUM_PARTICLES.size(); ++d) {   // src/system.cpp:62-64: rand() is drawn x, y, z per particle, drop after drop
            std::vector<int32_t> rnd(3 * (size_t)NUM_PARTICLES[d]);
            for (auto& v : rnd) v = rand();
            if (ink_seed_box(m_ctx, (int)d, rnd.data(), (int64_t)NUM_PARTICLES[d]) != 0) fail("ink_seed_box");
            m_ink.emplace_back((size_t)NUM_PARTICLES[d], blank);
        }
        m_inkStale = true;   // the mirror is filled from the device by the first getInkParticles()
    } else {
        if (on_host) {
            std::vector<Particle> ink;
            ink.reserve(400000);
            m_ink.push_back(ink);
            for (float i = 0; i < 2 * M_PI; i += (M_PI / 60.f)) emitParticleHemisphere(i);
            uploadInk();
            return;
        }
        std::vector<float> radius, si, ci, sj, cj;
        inkseed::emitter_tables(radius, si, ci, sj, cj);
        if (ink_seed_emitter(m_ctx, 0, radius.data(), (int)radius.size(), si.data(), ci.data(), (int)si.size(), sj.data(), cj.data(), (int)sj.size()) != 0)
            fail("ink_seed_emitter");
        m_ink.emplace_back(radius.size() * si.size() * sj.size(), blank);
        m_inkStale = true;
    }
}

void System::uploadInk() {
    std::vector<float> pos, vel;
    for (size_t d = 0; d < m_ink.size(); ++d) {
        const std::vector<Particle>& ink = m_ink[d];
        pos.resize(ink.size() * 3);
        vel.resize(ink.size() * 3);
        for (size_t i = 0; i < ink.size(); ++i)
            for (int a = 0; a < 3; ++a) { pos[3 * i + a] = ink[i].position[a]; vel[3 * i + a] = ink[i].velocity[a]; }
        if (ink_set_particles(m_ctx, (int)d, pos.data(), vel.data(), (int64_t)ink.size(), 32) != 0) fail("ink_set_particles");
    }
    m_inkStale = false;
}

// ---- particle seeding (seeding.cpp holds the arithmetic; these are the reference's member names) ---------------------
void System::initParticles() { inkseed::box(WATERGRID_X, WATERGRID_Y, WATERGRID_Z, NUM_PARTICLES, m_ink); }
void System::initFromFile() { m_ink.push_back(inkseed::from_file(PART_FILE)); }
void System::emitParticleHemisphere(float radius) { inkseed::hemisphere(WATERGRID_X, WATERGRID_Y, WATERGRID_Z, radius, m_ink[0]); }

// ---- the hot path ----------------------------------------------------------------------------------------------------
double System::solve(double timeToNextRender) {
    double dt = 0;
    if (ink_step(m_ctx, timeToNextRender, &dt) != 0) fail("ink_step");
    m_inkStale = true;
    return dt;
}

// ---- getters -----------------------------------------------------------------------------------------------------------
int System::numDrops() const { return ink_num_drops(m_ctx); }
long long System::numParticles(int drop) const { return ink_num_particles(m_ctx, drop); }
void System::copyPositions(int drop, float* xyz) {
    if (ink_get_particles(m_ctx, drop, xyz, nullptr, 32) != 0) fail("ink_get_particles");
}
void System::fetchPositionsAsync(int drop, float* xyz) {
    if (ink_fetch_positions_async(m_ctx, drop, xyz, 32) != 0) fail("ink_fetch_positions_async");
}
void System::syncIO() {
    if (ink_sync_io(m_ctx) != 0) fail("ink_sync_io");
}
void System::copyVelocityField(float* u, float* v, float* w) {
    if (ink_get_field(m_ctx, INK_FIELD_CURR, u, v, w, 32) != 0) fail("ink_get_field");
}

// ---- checkpoint / resume -------------------------------------------------------------------------------------------------
void System::copyParticles(int drop, float* xyz, float* vel) {
    if (ink_get_particles(m_ctx, drop, xyz, vel, 32) != 0) fail("ink_get_particles");
}
void System::copyWarmStart(float* q) {
    if (ink_get_warm_start(m_ctx, q, 32) != 0) fail("ink_get_warm_start");
}
void System::setVelocityField(const float* u, const float* v, const float* w) {
    if (ink_set_velocity(m_ctx, u, v, w, 32) != 0) fail("ink_set_velocity");
}
void System::setWarmStart(const float* q) {
    if (ink_set_warm_start(m_ctx, q, 32) != 0) fail("ink_set_warm_start");
}
bool System::getPhantom(double v[3]) {
    int32_t have = 0;
    if (ink_get_phantom(m_ctx, &have, v) != 0) fail("ink_get_phantom");
    return have != 0;
}
void System::setPhantom(bool have, const double v[3]) {
    if (ink_set_phantom(m_ctx, have ? 1 : 0, v) != 0) fail("ink_set_phantom");
}
void System::setInk(int drop, const float* xyz, const float* vel, long long n) {
    if (drop < 0 || drop > (int)m_ink.size()) throw std::runtime_error("setInk: drop index out of range");
    if (drop == (int)m_ink.size()) m_ink.emplace_back();
    Particle blank;
    blank.position = blank.velocity = Vector3f(0, 0, 0);
    blank.opacity = 1.f;
    blank.lifeTime = 5.f;
    m_ink[drop].assign((size_t)n, blank);
    if (ink_set_particles(m_ctx, drop, xyz, vel, (int64_t)n, 32) != 0) fail("ink_set_particles");
    m_inkStale = true;   // the mirror is refilled from the device by the next getInkParticles()
}

const std::vector<std::vector<Particle>>& System::getInkParticles() {
    if (!m_inkStale) return m_ink;
    std::vector<float> pos, vel;
    for (size_t d = 0; d < m_ink.size(); ++d) {
        std::vector<Particle>& ink = m_ink[d];
        pos.resize(ink.size() * 3);
        vel.resize(ink.size() * 3);
        if (ink.empty()) continue;
        if (ink_get_particles(m_ctx, (int)d, pos.data(), vel.data(), 32) != 0) fail("ink_get_particles");
        for (size_t i = 0; i < ink.size(); ++i) {
            ink[i].position = Vector3f(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
            ink[i].velocity = Vector3f(vel[3 * i], vel[3 * i + 1], vel[3 * i + 2]);
        }
    }
    m_inkStale = false;
    return m_ink;
}

// The reference returns its whole hash map BY VALUE (src/system.cpp:155-157); so does this, rebuilt from the dense
// device arrays.  old == curr between steps (src/waterSolver.cpp:92-95); curl is the last confinement pass's.
const std::unordered_map<Vector3i, Cell, hash_func> System::getWaterGrid() {
    const size_t n = (size_t)WATERGRID_X * WATERGRID_Y * WATERGRID_Z;
    std::vector<float> u(n), v(n), w(n), cx(n), cy(n), cz(n);
    if (ink_get_field(m_ctx, INK_FIELD_CURR, u.data(), v.data(), w.data(), 32) != 0) fail("ink_get_field");
    if (ink_get_field(m_ctx, INK_FIELD_CURL, cx.data(), cy.data(), cz.data(), 32) != 0) fail("ink_get_field");
    m_waterGrid.clear();
    m_waterGrid.reserve(n);
    for (int i = 0; i < WATERGRID_X; ++i)
        for (int j = 0; j < WATERGRID_Y; ++j)
            for (int k = 0; k < WATERGRID_Z; ++k) {
                const size_t c = ((size_t)i * WATERGRID_Y + j) * WATERGRID_Z + k;
                Cell cell;
                cell.oldVelocity = cell.currVelocity = Vector3f(u[c], v[c], w[c]);
                cell.curl = Vector3f(cx[c], cy[c], cz[c]);
                const int nb[6][3] = {{i + 1, j, k}, {i - 1, j, k}, {i, j + 1, k}, {i, j - 1, k}, {i, j, k + 1}, {i, j, k - 1}};
                for (const auto& q : nb)
                    if (q[0] >= 0 && q[0] < WATERGRID_X && q[1] >= 0 && q[1] < WATERGRID_Y && q[2] >= 0 && q[2] < WATERGRID_Z)
                        cell.neighbors.push_back(Vector3i(q[0], q[1], q[2]));
                m_waterGrid.insert({Vector3i(i, j, k), cell});
            }
    return m_waterGrid;
}

// checkNanAndInf (src/systemHelpers.cpp:212-241): assert-only in the reference; here it throws.  The tests themselves (no NaN / Inf
// component, norm() < 10000 on every cell velocity and particle velocity) run on the device as one reduction
// (ink_check_finite): two counters come back, not the grid and every Particle.
void System::checkNanAndInf() {
    int64_t bad[2] = {0, 0};
    if (ink_check_finite(m_ctx, bad) != 0) fail("ink_check_finite");
    if (bad[0] != 0) throw std::runtime_error("checkNanAndInf: " + std::to_string(bad[0]) + " grid cells / particles hold a NaN or Inf");
    if (bad[1] != 0) throw std::runtime_error("checkNanAndInf: " + std::to_string(bad[1]) + " velocities with norm() >= 10000");
}

std::ostream& operator<<(std::ostream& strm, const Cell& obj) {
    return strm << "Cell(old " << obj.oldVelocity << ", curr " << obj.currVelocity << ", curl " << obj.curl << ")";
}
std::ostream& operator<<(std::ostream& strm, const Particle& obj) {
    return strm << "Particle(pos " << obj.position << ", vel " << obj.velocity << ", opacity " << obj.opacity << ", life " << obj.lifeTime << ")";
}
std::ostream& operator<<(std::ostream& strm, const System& obj) {
    return strm << "System(" << WATERGRID_X << "x" << WATERGRID_Y << "x" << WATERGRID_Z << ", " << obj.m_ink.size() << " drops)";
}
