// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  CPU oracle for the Ink-Simulation hot path.
//
// A flat-array restatement of the reference's per-timestep algorithm (System::solve), following
// the reference operation-for-operation so that, instantiated with Real = float and compiled
// without FMA contraction, every phase except the pressure *solve* is bit-identical to the
// compiled reference (oracle/_ref/libinkref.so; checked by tests/test_oracle_vs_ref.py and pinned
// by the fixtures under tests/golden/).  Real = double gives the fp64 oracle the GPU's fp64 mode is
// compared with.  The pressure system A p = b (src/waterSolver.cpp:240-281) has a unique solution;
// the reference solves it with Eigen::SparseLU<float> (vendored Eigen 3.4.0), the oracle solves it
// exactly (fp64 separable DST-I — A is diagonalised by sine transforms) and rounds p to Real.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  Parity pin: compiled reference outputs (no tests / golden vectors exist in
// the reference itself — SURVEY.md §4, §8c).
//
// Array convention: dense (X, Y, Z) C-order, z fastest: idx = (i*Y + j)*Z + k.  Particles AoS xyz.
//
// Reference citations (file:line under /root/reference) are given per function.

#include <algorithm>
#include <array>
#include <type_traits>
#include <chrono>
#include <omp.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <tuple>
#include <vector>

namespace {

struct Params {          // src/system.h:13-38
    int X, Y, Z;
    float cell_dim;      // CELL_DIM
    float density;       // DENSITY
    float k_vort;        // K_VORT
    float viscosity;     // VISCOSITY
    float p_atm;         // ATMOSPHERIC_PRESSURE
    float gravity[3];    // gravity
    float k_cfl;         // K_CFL
    float dt_min;        // MIN_TIMESTEP
    float dt_max;        // MAX_TIMESTEP
    int buffer_size;     // BUFFER_SIZE
};

template <typename Real>
struct Oracle {
    Params P;
    size_t n;
    // oldVelocity / currVelocity / curl of struct Cell (src/system.h:41-50), SoA
    std::vector<Real> ou, ov, ow, cu, cv, cw, curlx, curly, curlz;
    std::vector<Real> pressure;                   // last solved p, (X,Y,Z) order
    std::vector<std::vector<Real>> pos, vel;      // per drop, AoS xyz (struct Particle, src/system.h:52-60)
    // phantom map cells created by out-of-bounds particles (Q8): key -> (old, curr) velocity
    std::map<std::tuple<int, int, int>, std::array<Real, 6>> phantom;
    // DST-I sine tables (fp64) for the exact Poisson solve
    std::vector<double> Sx, Sy, Sz;
    int last_solver_iters = 1;

    explicit Oracle(const Params& p) : P(p) {
        n = (size_t)P.X * P.Y * P.Z;
        for (auto* v : {&ou, &ov, &ow, &cu, &cv, &cw, &curlx, &curly, &curlz, &pressure}) v->assign(n, Real(0));
    }

    inline size_t id(int i, int j, int k) const { return ((size_t)i * P.Y + j) * P.Z + k; }

    // src/systemHelpers.cpp:141-147 (float position test; also applied to integer indices)
    inline bool inBounds(Real x, Real y, Real z) const {
        bool xIs = (x >= 0) && (x < P.X * P.cell_dim);
        bool yIs = (y >= 0) && (y < P.Y * P.cell_dim);
        bool zIs = (z >= 0) && (z < P.Z * P.cell_dim);
        return xIs && yIs && zIs;
    }
    // src/systemHelpers.cpp:150-155
    inline bool inBoundsIdx(int i, int j, int k) const {
        return (i >= 0) && (i < P.X) && (j >= 0) && (j < P.Y) && (k >= 0) && (k < P.Z);
    }
    inline const std::vector<Real>& oldc(int c) const { return c == 0 ? ou : (c == 1 ? ov : ow); }

    // src/systemHelpers.cpp:158-209 — staggered trilinear sample with zero extension; out-of-grid
    // corners contribute 0 to the total but their weight still enters weightAccum (Q13).
    Real interp(Real x, Real y, Real z, int c) const {
        int i = (int)std::floor(x);
        int j = (int)std::floor(y);
        int k = (int)std::floor(z);
        const std::vector<Real>& f = oldc(c);
        Real weightAccum = 0, totalAccum = 0;
        if (inBounds(i, j, k))             totalAccum += (i + 1 - x) * (j + 1 - y) * (k + 1 - z) * f[id(i, j, k)];
        weightAccum += (i + 1 - x) * (j + 1 - y) * (k + 1 - z);
        if (inBounds(i + 1, j, k))         totalAccum += (x - i) * (j + 1 - y) * (k + 1 - z) * f[id(i + 1, j, k)];
        weightAccum += (x - i) * (j + 1 - y) * (k + 1 - z);
        if (inBounds(i, j + 1, k))         totalAccum += (i + 1 - x) * (y - j) * (k + 1 - z) * f[id(i, j + 1, k)];
        weightAccum += (i + 1 - x) * (y - j) * (k + 1 - z);
        if (inBounds(i + 1, j + 1, k))     totalAccum += (x - i) * (y - j) * (k + 1 - z) * f[id(i + 1, j + 1, k)];
        weightAccum += (x - i) * (y - j) * (k + 1 - z);
        if (inBounds(i, j, k + 1))         totalAccum += (i + 1 - x) * (j + 1 - y) * (z - k) * f[id(i, j, k + 1)];
        weightAccum += (i + 1 - x) * (j + 1 - y) * (z - k);
        if (inBounds(i + 1, j, k + 1))     totalAccum += (x - i) * (j + 1 - y) * (z - k) * f[id(i + 1, j, k + 1)];
        weightAccum += (x - i) * (j + 1 - y) * (z - k);
        if (inBounds(i, j + 1, k + 1))     totalAccum += (i + 1 - x) * (y - j) * (z - k) * f[id(i, j + 1, k + 1)];
        weightAccum += (i + 1 - x) * (y - j) * (z - k);
        if (inBounds(i + 1, j + 1, k + 1)) totalAccum += (x - i) * (y - j) * (z - k) * f[id(i + 1, j + 1, k + 1)];
        weightAccum += (x - i) * (y - j) * (z - k);
        if (weightAccum == 0) return 0;
        return totalAccum / weightAccum;
    }

    // src/systemHelpers.cpp:110-115
    inline void getVelocity(const Real p[3], Real out[3]) const {
        const Real h = P.cell_dim;
        out[0] = interp(p[0] / h, (p[1] / h) - Real(0.5f), (p[2] / h) - Real(0.5f), 0);
        out[1] = interp((p[0] / h) - Real(0.5f), p[1] / h, (p[2] / h) - Real(0.5f), 1);
        out[2] = interp((p[0] / h) - Real(0.5f), (p[1] / h) - Real(0.5f), p[2] / h, 2);
    }

    // Eigen 3.4 Vector3f::norm(): sqrt of the unrolled tree reduction x*x + (y*y + z*z)
    // (redux_novec_unroller splits 3 coefficients as 1 + 2); verified bit-for-bit against the compiled reference.
    static inline Real norm3(Real x, Real y, Real z) { return std::sqrt(x * x + (y * y + z * z)); }

    void commit() {  // the `old = curr` sweeps, src/waterSolver.cpp:56-59,65-68,74-77,83-86,92-95
        ou = cu; ov = cv; ow = cw;
        for (auto& kv : phantom) { kv.second[0] = kv.second[3]; kv.second[1] = kv.second[4]; kv.second[2] = kv.second[5]; }
    }

    // src/waterSolver.cpp:13-38 (Q12).  Max over every map entry incl. phantom cells (Q8).
    double calcTimeStep(double timeToNextRender) const {
        float maxVelocity = 0;
        float m = 0;
#pragma omp parallel for reduction(max : m)
        for (size_t c = 0; c < n; ++c) {
            float s = (float)norm3(ou[c], ov[c], ow[c]);
            if (m < s) m = s;
        }
        maxVelocity = m;
        for (auto& kv : phantom) {
            float s = (float)norm3(kv.second[0], kv.second[1], kv.second[2]);
            if (maxVelocity < s) maxVelocity = s;
        }
        if (std::isinf(maxVelocity)) return P.dt_min;
        if (maxVelocity == 0) return P.dt_min;
        float timeStep = P.k_cfl * (P.cell_dim / maxVelocity);
        timeStep = std::max(std::min(timeStep, P.dt_max), P.dt_min);
        // `timeStep = std::min(timeToNextRender, (double)timeStep)` assigns to the FLOAT local (src/waterSolver.cpp:35):
        // the returned step is narrowed to float even when timeToNextRender wins
        timeStep = (float)std::min(timeToNextRender, (double)timeStep);
        return timeStep;
    }

    // src/waterSolver.cpp:98-128 (Q6, Q7): RK2 back-trace of the cell centre, nearest-cell copy of
    // the whole old velocity vector; skipped when the traced point leaves the grid.
    void applyConvection(Real timeStep) {
        const Real t = -timeStep;
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    const Real x = i + (P.cell_dim / 2.f), y = j + (P.cell_dim / 2.f), z = k + (P.cell_dim / 2.f);
                    Real p0[3] = {x, y, z}, v[3];
                    getVelocity(p0, v);
                    // midpoint computed in double then rounded to Real (src/waterSolver.cpp:100)
                    Real pm[3] = {(Real)(x + 0.5 * t * v[0]), (Real)(y + 0.5 * t * v[1]), (Real)(z + 0.5 * t * v[2])};
                    getVelocity(pm, v);
                    const Real px = x + t * v[0], py = y + t * v[1], pz = z + t * v[2];
                    if (!inBounds(px, py, pz)) continue;
                    const size_t src = id((int)std::floor(px), (int)std::floor(py), (int)std::floor(pz));
                    const size_t dst = id(i, j, k);
                    cu[dst] = ou[src]; cv[dst] = ov[src]; cw[dst] = ow[src];
                }
    }

    // src/waterSolver.cpp:158-191 (Q8): cells within an in-bounds neighbour chain of length
    // <= BUFFER_SIZE (== Manhattan ball clipped to the box) of any particle's cell get += dt*g once.
    // `mark` (optional) receives the 0/1 flag grid.
    void applyExternalForces(Real timeStep, uint8_t* mark_out = nullptr) {
        std::vector<uint8_t> seed(n, 0), mark(n, 0);
        std::vector<std::tuple<int, int, int>> oob;
        for (auto& drop : pos) {
            const size_t np = drop.size() / 3;
            for (size_t p = 0; p < np; ++p) {
                int ci = (int)std::floor(drop[3 * p]), cj = (int)std::floor(drop[3 * p + 1]), ck = (int)std::floor(drop[3 * p + 2]);
                if (inBoundsIdx(ci, cj, ck)) seed[id(ci, cj, ck)] = 1;
                else oob.emplace_back(ci, cj, ck);
            }
        }
        const int R = P.buffer_size;
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    uint8_t m = 0;
                    for (int di = -R; di <= R && !m; ++di)
                        for (int dj = -(R - std::abs(di)); dj <= R - std::abs(di) && !m; ++dj) {
                            int rk = R - std::abs(di) - std::abs(dj);
                            for (int dk = -rk; dk <= rk; ++dk)
                                if (inBoundsIdx(i + di, j + dj, k + dk) && seed[id(i + di, j + dj, k + dk)]) { m = 1; break; }
                        }
                    mark[id(i, j, k)] = m;
                }
        // updateForce: curr += (float)dt * gravity  (Q15: the double dt is narrowed to the vector's scalar)
        const Real gx = timeStep * (Real)P.gravity[0], gy = timeStep * (Real)P.gravity[1], gz = timeStep * (Real)P.gravity[2];
#pragma omp parallel for
        for (size_t c = 0; c < n; ++c)
            if (mark[c]) { cu[c] += gx; cv[c] += gy; cw[c] += gz; }
        // phantom cells: the map entry default-constructs to zero velocity with no neighbours
        std::sort(oob.begin(), oob.end());
        oob.erase(std::unique(oob.begin(), oob.end()), oob.end());
        for (auto& key : oob) {
            auto it = phantom.find(key);
            if (it == phantom.end()) it = phantom.emplace(key, std::array<Real, 6>{0, 0, 0, 0, 0, 0}).first;
            it->second[3] += gx; it->second[4] += gy; it->second[5] += gz;
        }
        if (mark_out) std::memcpy(mark_out, mark.data(), n);
    }

    // src/systemHelpers.cpp:88-107
    inline Real laplacian(const std::vector<Real>& f, int i, int j, int k) const {
        Real l = 0;
        l += (i + 1 < P.X) ? f[id(i + 1, j, k)] : 0;
        l += (i - 1 >= 0) ? f[id(i - 1, j, k)] : 0;
        l += (j + 1 < P.Y) ? f[id(i, j + 1, k)] : 0;
        l += (j - 1 >= 0) ? f[id(i, j - 1, k)] : 0;
        l += (k + 1 < P.Z) ? f[id(i, j, k + 1)] : 0;
        l += (k - 1 >= 0) ? f[id(i, j, k - 1)] : 0;
        l -= 6 * f[id(i, j, k)];
        return l;
    }
    // src/waterSolver.cpp:225-238 (Q9)
    void applyViscosity(Real timeStep) {
        const Real s = timeStep * (Real)P.viscosity;
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    const size_t c = id(i, j, k);
                    Real lx = laplacian(ou, i, j, k), ly = laplacian(ov, i, j, k), lz = laplacian(ow, i, j, k);
                    cu[c] += s * lx; cv[c] += s * ly; cw[c] += s * lz;
                }
    }

    // src/systemHelpers.cpp:30-43
    inline Real divergence(int i, int j, int k) const {
        const size_t c = id(i, j, k);
        Real d = -ou[c] - ov[c] - ow[c];
        if (inBoundsIdx(i + 1, j, k)) d += ou[id(i + 1, j, k)];
        if (inBoundsIdx(i, j + 1, k)) d += ov[id(i, j + 1, k)];
        if (inBoundsIdx(i, j, k + 1)) d += ow[id(i, j, k + 1)];
        return d;
    }
    inline int numNeighbors(int i, int j, int k) const {  // getGridNeighbors().size(), src/systemHelpers.cpp:117-138
        return (i + 1 < P.X) + (i - 1 >= 0) + (j + 1 < P.Y) + (j - 1 >= 0) + (k + 1 < P.Z) + (k - 1 >= 0);
    }
    // RHS of src/waterSolver.cpp:264-278 in (X,Y,Z) order
    void pressureRhs(Real timeStep, std::vector<Real>& b) const {
        b.resize(n);
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    Real dv = divergence(i, j, k);
                    int ki = 6 - numNeighbors(i, j, k);
                    b[id(i, j, k)] = (((Real)P.density * (Real)P.cell_dim) / timeStep) * dv - ki * (Real)P.p_atm;
                }
    }

    void buildSine() {
        auto mk = [](int m, std::vector<double>& S) {
            S.resize((size_t)m * m);
            for (int a = 0; a < m; ++a)
                for (int b = 0; b < m; ++b) S[(size_t)a * m + b] = std::sin(M_PI * (a + 1.0) * (b + 1.0) / (m + 1.0));
        };
        if (Sx.empty()) { mk(P.X, Sx); mk(P.Y, Sy); mk(P.Z, Sz); }
    }
    // Unnormalised DST-I along all three axes, fp64, in place (w = scratch of n doubles).
    void dst3(std::vector<double>& f, std::vector<double>& w) const {
        const int X = P.X, Y = P.Y, Z = P.Z;
        // z axis
#pragma omp parallel for
        for (long ij = 0; ij < (long)X * Y; ++ij) {
            const double* in = &f[(size_t)ij * Z];
            double* out = &w[(size_t)ij * Z];
            for (int c = 0; c < Z; ++c) {
                const double* s = &Sz[(size_t)c * Z];
                double acc = 0;
                for (int k = 0; k < Z; ++k) acc += s[k] * in[k];
                out[c] = acc;
            }
        }
        // y axis: f[i][b][k] = sum_j Sy[b][j] w[i][j][k]
#pragma omp parallel for collapse(2)
        for (int i = 0; i < X; ++i)
            for (int b = 0; b < Y; ++b) {
                double* out = &f[((size_t)i * Y + b) * Z];
                for (int k = 0; k < Z; ++k) out[k] = 0;
                for (int j = 0; j < Y; ++j) {
                    const double s = Sy[(size_t)b * Y + j];
                    const double* in = &w[((size_t)i * Y + j) * Z];
                    for (int k = 0; k < Z; ++k) out[k] += s * in[k];
                }
            }
        // x axis: w[a][j][k] = sum_i Sx[a][i] f[i][j][k]
        const size_t plane = (size_t)Y * Z;
#pragma omp parallel for
        for (int a = 0; a < X; ++a) {
            double* out = &w[(size_t)a * plane];
            for (size_t q = 0; q < plane; ++q) out[q] = 0;
            for (int i = 0; i < X; ++i) {
                const double s = Sx[(size_t)a * X + i];
                const double* in = &f[(size_t)i * plane];
                for (size_t q = 0; q < plane; ++q) out[q] += s * in[q];
            }
        }
        f.swap(w);
    }
    // Exact solve of A p = b, A = 7-point operator with diag -6 and +1 per in-bounds neighbour
    // (src/waterSolver.cpp:240-261).  Eigenpairs: sin modes, lambda = -6 + 2 sum cos(pi a/(n+1)).
    void solvePoisson(const std::vector<Real>& b, std::vector<Real>& p) {
        buildSine();
        std::vector<double> f(n), w(n);
        for (size_t c = 0; c < n; ++c) f[c] = (double)b[c];
        dst3(f, w);
        const int X = P.X, Y = P.Y, Z = P.Z;
        const double scale = 8.0 / ((X + 1.0) * (Y + 1.0) * (Z + 1.0));
#pragma omp parallel for collapse(2)
        for (int a = 0; a < X; ++a)
            for (int bb = 0; bb < Y; ++bb)
                for (int c = 0; c < Z; ++c) {
                    double lam = -6.0 + 2.0 * std::cos(M_PI * (a + 1.0) / (X + 1.0)) + 2.0 * std::cos(M_PI * (bb + 1.0) / (Y + 1.0)) +
                                 2.0 * std::cos(M_PI * (c + 1.0) / (Z + 1.0));
                    f[id(a, bb, c)] *= scale / lam;
                }
        dst3(f, w);
        p.resize(n);
        for (size_t c = 0; c < n; ++c) p[c] = (Real)f[c];
    }
    // max |A p - b| / max|b| in fp64 — used by tests to certify the solve.
    double poissonResidual(const std::vector<Real>& p, const std::vector<Real>& b) const {
        double worst = 0, bmax = 0;
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    double s = -6.0 * (double)p[id(i, j, k)];
                    if (i + 1 < P.X) s += p[id(i + 1, j, k)];
                    if (i - 1 >= 0) s += p[id(i - 1, j, k)];
                    if (j + 1 < P.Y) s += p[id(i, j + 1, k)];
                    if (j - 1 >= 0) s += p[id(i, j - 1, k)];
                    if (k + 1 < P.Z) s += p[id(i, j, k + 1)];
                    if (k - 1 >= 0) s += p[id(i, j, k - 1)];
                    worst = std::max(worst, std::fabs(s - (double)b[id(i, j, k)]));
                    bmax = std::max(bmax, std::fabs((double)b[id(i, j, k)]));
                }
        return worst / std::max(bmax, 1e-300);
    }

    // src/systemHelpers.cpp:5-28 + src/waterSolver.cpp:284-296: curr -= dt/(rho h) * backward-difference
    // gradient with low-side ghost = 1 (the literal 1 in the reference, == ATMOSPHERIC_PRESSURE).
    void subtractGradient(Real timeStep, const std::vector<Real>& p) {
        const Real s = timeStep / ((Real)P.density * (Real)P.cell_dim);
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    const size_t c = id(i, j, k);
                    const Real val = p[c];
                    Real gx = val, gy = val, gz = val;
                    gx -= (i - 1 >= 0) ? p[id(i - 1, j, k)] : Real(1);
                    gy -= (j - 1 >= 0) ? p[id(i, j - 1, k)] : Real(1);
                    gz -= (k - 1 >= 0) ? p[id(i, j, k - 1)] : Real(1);
                    cu[c] -= s * gx; cv[c] -= s * gy; cw[c] -= s * gz;
                }
    }
    void applyPressure(Real timeStep) {
        std::vector<Real> b;
        pressureRhs(timeStep, b);
        solvePoisson(b, pressure);
        subtractGradient(timeStep, pressure);
    }

    // src/systemHelpers.cpp:45-69 (Q11): central differences without 1/2, zero outside.
    inline void curlAt(int i, int j, int k, Real c[3]) const {
        c[0] = c[1] = c[2] = 0;
        c[0] += (j + 1 < P.Y) ? ow[id(i, j + 1, k)] : 0;
        c[0] -= (j - 1 >= 0) ? ow[id(i, j - 1, k)] : 0;
        c[0] -= (k + 1 < P.Z) ? ov[id(i, j, k + 1)] : 0;
        c[0] += (k - 1 >= 0) ? ov[id(i, j, k - 1)] : 0;
        c[1] += (k + 1 < P.Z) ? ou[id(i, j, k + 1)] : 0;
        c[1] -= (k - 1 >= 0) ? ou[id(i, j, k - 1)] : 0;
        c[1] -= (i + 1 < P.X) ? ow[id(i + 1, j, k)] : 0;
        c[1] += (i - 1 >= 0) ? ow[id(i - 1, j, k)] : 0;
        c[2] += (i + 1 < P.X) ? ov[id(i + 1, j, k)] : 0;
        c[2] -= (i - 1 >= 0) ? ov[id(i - 1, j, k)] : 0;
        c[2] -= (j + 1 < P.Y) ? ou[id(i, j + 1, k)] : 0;
        c[2] += (j - 1 >= 0) ? ou[id(i, j - 1, k)] : 0;
    }
    inline Real curlNorm(int i, int j, int k) const {
        const size_t c = id(i, j, k);
        return norm3(curlx[c], curly[c], curlz[c]);
    }
    // src/waterSolver.cpp:193-222 + src/systemHelpers.cpp:71-80 (Q2: the k-1 term is always 0)
    void applyVorticity(Real timeStep) {
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    Real c[3];
                    curlAt(i, j, k, c);
                    const size_t q = id(i, j, k);
                    curlx[q] = c[0]; curly[q] = c[1]; curlz[q] = c[2];
                }
#pragma omp parallel for collapse(2)
        for (int i = 0; i < P.X; i++)
            for (int j = 0; j < P.Y; j++)
                for (int k = 0; k < P.Z; k++) {
                    const size_t q = id(i, j, k);
                    const Real wx = curlx[q], wy = curly[q], wz = curlz[q];
                    if (wx == 0 && wy == 0 && wz == 0) continue;
                    Real g[3] = {0, 0, 0};
                    g[0] += (i + 1 < P.X) ? curlNorm(i + 1, j, k) : 0;
                    g[0] -= (i - 1 >= 0) ? curlNorm(i - 1, j, k) : 0;
                    g[1] += (j + 1 < P.Y) ? curlNorm(i, j + 1, k) : 0;
                    g[1] -= (j - 1 >= 0) ? curlNorm(i, j - 1, k) : 0;
                    g[2] += (k + 1 < P.Z) ? curlNorm(i, j, k + 1) : 0;
                    // (k-1 < 0) ? phantom-cell curl norm (== 0) : 0   — src/systemHelpers.cpp:78
                    g[2] -= 0;
                    const Real nrm = norm3(wx, wy, wz);
                    const Real Nx = g[0] / nrm, Ny = g[1] / nrm, Nz = g[2] / nrm;
                    const Real kv = (Real)P.k_vort;
                    const Real Fx = kv * (Ny * wz - Nz * wy);
                    const Real Fy = kv * (Nz * wx - Nx * wz);
                    const Real Fz = kv * (Nx * wy - Ny * wx);
                    cu[q] += timeStep * Fx; cv[q] += timeStep * Fy; cw[q] += timeStep * Fz;
                }
    }

    // src/waterSolver.cpp:41-96
    void updateVelocityField(Real timeStep) {
        applyConvection(timeStep); commit();
        applyExternalForces(timeStep); commit();
        applyViscosity(timeStep); commit();
        applyPressure(timeStep); commit();
        applyVorticity(timeStep); commit();
    }

    // src/inkSolver.cpp:44-84 (Q3, Q4, Q5)
    void updateParticles(Real timeStep) {
        const Real rho = (Real)P.density;
        const Real hiX = P.X * P.cell_dim, hiY = P.Y * P.cell_dim, hiZ = P.Z * P.cell_dim;
        const Real cX = P.X * P.cell_dim - 0.01f, cY = P.Y * P.cell_dim - 0.01f, cZ = P.Z * P.cell_dim - 0.01f;
        for (size_t d = 0; d < pos.size(); ++d) {
            Real* ps = pos[d].data();
            Real* vs = vel[d].data();
            const long np = (long)(pos[d].size() / 3);
#pragma omp parallel for
            for (long q = 0; q < np; ++q) {
                Real* p = ps + 3 * q;
                Real* v = vs + 3 * q;
                Real s[3];
                getVelocity(p, s);
                v[0] = rho * s[0]; v[1] = rho * s[1]; v[2] = rho * s[2];
                Real mid[3] = {p[0] + (v[0] * timeStep * Real(.5)), p[1] + (v[1] * timeStep * Real(.5)), p[2] + (v[2] * timeStep * Real(.5))};
                getVelocity(mid, s);
                const Real mv[3] = {rho * s[0], rho * s[1], rho * s[2]};
                p[0] = mid[0] + (timeStep * mv[0]);
                p[1] = mid[1] + (timeStep * mv[1]);
                p[2] = mid[2] + (timeStep * mv[2]);
                if (p[0] < 0 || p[0] > hiX) { v[0] *= -1.f; p[0] = std::min(std::max(p[0], Real(0.01f)), cX); }
                if (p[1] < 0 || p[1] > hiY) { v[1] *= -1.f; p[1] = std::min(std::max(p[1], Real(0.01f)), cY); }
                if (p[2] < 0 || p[2] > hiZ) { v[2] *= -1.f; p[2] = std::min(std::max(p[2], Real(0.01f)), cZ); }
            }
        }
    }

    // src/inkSolver.cpp:5-9 + src/waterSolver.cpp:7-11
    double solve(double timeToNextRender) {
        double dt = calcTimeStep(timeToNextRender);
        updateVelocityField((Real)(float)dt);   // `float timeStep` parameter narrows the double (Q15)
        updateParticles((Real)(float)dt);
        return dt;
    }
};

struct Handle {
    int dtype;  // 32 or 64
    Oracle<float>* f;
    Oracle<double>* d;
};

template <typename Real, typename T>
void copy_in(std::vector<Real>& dst, const T* src, size_t n) {
    dst.resize(n);
    for (size_t i = 0; i < n; ++i) dst[i] = (Real)src[i];
}
template <typename Real, typename T>
void copy_out(T* dst, const std::vector<Real>& src) {
    for (size_t i = 0; i < src.size(); ++i) dst[i] = (T)src[i];
}

}  // namespace

#define DISPATCH(h, ...)                         \
    do {                                         \
        Handle* H_ = static_cast<Handle*>(h);    \
        if (H_->dtype == 32) { auto& o = *H_->f; __VA_ARGS__; } \
        else { auto& o = *H_->d; __VA_ARGS__; }         \
    } while (0)

extern "C" {

// params: 16 floats-or-ints packed by the Python side; see oracle/__init__.py (field order below).
void* orc_create(int dtype, int X, int Y, int Z, float cell_dim, float density, float k_vort, float viscosity, float p_atm,
                 float gx, float gy, float gz, float k_cfl, float dt_min, float dt_max, int buffer_size) {
    Params P{X, Y, Z, cell_dim, density, k_vort, viscosity, p_atm, {gx, gy, gz}, k_cfl, dt_min, dt_max, buffer_size};
    Handle* h = new Handle{dtype, nullptr, nullptr};
    if (dtype == 32) h->f = new Oracle<float>(P);
    else h->d = new Oracle<double>(P);
    return h;
}
void orc_destroy(void* h) {
    Handle* H = static_cast<Handle*>(h);
    delete H->f; delete H->d; delete H;
}

// All array I/O through the C API is in the oracle's own dtype (float for 32, double for 64).
void orc_set_particles(void* h, int drop, const void* pos, const void* vel, long np) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        while ((int)o.pos.size() <= drop) { o.pos.emplace_back(); o.vel.emplace_back(); }
        copy_in(o.pos[drop], (const R*)pos, (size_t)np * 3);
        if (vel) copy_in(o.vel[drop], (const R*)vel, (size_t)np * 3);
        else o.vel[drop].assign((size_t)np * 3, R(0));
    });
}
long orc_num_particles(void* h, int drop) {
    long r = 0;
    DISPATCH(h, { r = drop < (int)o.pos.size() ? (long)(o.pos[drop].size() / 3) : 0; });
    return r;
}
void orc_get_particles(void* h, int drop, void* pos, void* vel) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        if (pos) copy_out((R*)pos, o.pos[drop]);
        if (vel) copy_out((R*)vel, o.vel[drop]);
    });
}
void orc_set_velocity(void* h, const void* u, const void* v, const void* w) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        copy_in(o.ou, (const R*)u, o.n); copy_in(o.ov, (const R*)v, o.n); copy_in(o.ow, (const R*)w, o.n);
        o.cu = o.ou; o.cv = o.ov; o.cw = o.ow;
    });
}
// which: 0 old, 1 curr, 2 curl
void orc_get_field(void* h, int which, void* u, void* v, void* w) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        if (which == 0) { copy_out((R*)u, o.ou); copy_out((R*)v, o.ov); copy_out((R*)w, o.ow); }
        else if (which == 1) { copy_out((R*)u, o.cu); copy_out((R*)v, o.cv); copy_out((R*)w, o.cw); }
        else { copy_out((R*)u, o.curlx); copy_out((R*)v, o.curly); copy_out((R*)w, o.curlz); }
    });
}
void orc_get_pressure(void* h, void* p) {
    DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; copy_out((R*)p, o.pressure); });
}
double orc_calc_dt(void* h, double ttnr) { double r = 0; DISPATCH(h, { r = o.calcTimeStep(ttnr); }); return r; }
void orc_advect(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.applyConvection((R)(float)dt); o.commit(); }); }
void orc_forces(void* h, double dt, uint8_t* mark) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.applyExternalForces((R)(float)dt, mark); o.commit(); }); }
void orc_viscosity(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.applyViscosity((R)(float)dt); o.commit(); }); }
void orc_pressure_rhs(void* h, double dt, void* b_out) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        std::vector<R> b; o.pressureRhs((R)(float)dt, b); copy_out((R*)b_out, b);
    });
}
void orc_pressure(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.applyPressure((R)(float)dt); o.commit(); }); }
// applyPressure with a caller-supplied pressure field (e.g. the compiled reference's LU solution),
// to check the gradient-subtract arithmetic bit-for-bit independently of the solver.
void orc_pressure_given_p(void* h, double dt, const void* p) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        copy_in(o.pressure, (const R*)p, o.n); o.subtractGradient((R)(float)dt, o.pressure); o.commit();
    });
}
// solve A p = b for arbitrary b; returns fp64 max-norm relative residual
double orc_poisson_solve(void* h, const void* b, void* p) {
    double r = 0;
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        std::vector<R> bb, pp; copy_in(bb, (const R*)b, o.n); o.solvePoisson(bb, pp); copy_out((R*)p, pp);
        r = o.poissonResidual(pp, bb);
    });
    return r;
}
void orc_vorticity(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.applyVorticity((R)(float)dt); o.commit(); }); }
void orc_divergence(void* h, void* out) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        R* d = (R*)out;
        for (int i = 0; i < o.P.X; i++) for (int j = 0; j < o.P.Y; j++) for (int k = 0; k < o.P.Z; k++) d[o.id(i, j, k)] = o.divergence(i, j, k);
    });
}
void orc_sample_velocity(void* h, const void* pos, long np, void* out) {
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        const R* p = (const R*)pos; R* q = (R*)out;
        for (long a = 0; a < np; ++a) o.getVelocity(p + 3 * a, q + 3 * a);
    });
}
void orc_velocity_field_update(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.updateVelocityField((R)(float)dt); }); }
void orc_particles(void* h, double dt) { DISPATCH(h, { using R = typename std::remove_reference<decltype(o.ou[0])>::type; o.updateParticles((R)(float)dt); }); }
double orc_solve(void* h, double ttnr) { double r = 0; DISPATCH(h, { r = o.solve(ttnr); }); return r; }

// per-phase wall clock, same slots as ref_solve_timed
double orc_solve_timed(void* h, double ttnr, double* t) {
    double dt = 0;
    DISPATCH(h, {
        using R = typename std::remove_reference<decltype(o.ou[0])>::type;
        using clk = std::chrono::steady_clock;
        auto s = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
        auto t0 = clk::now();
        dt = o.calcTimeStep(ttnr);
        R fdt = (R)(float)dt;
        auto t1 = clk::now(); o.applyConvection(fdt); o.commit();
        auto t2 = clk::now(); o.applyExternalForces(fdt); o.commit();
        auto t3 = clk::now(); o.applyViscosity(fdt); o.commit();
        auto t4 = clk::now(); o.applyPressure(fdt); o.commit();
        auto t5 = clk::now(); o.applyVorticity(fdt); o.commit();
        auto t6 = clk::now(); o.updateParticles(fdt);
        auto t7 = clk::now();
        t[0] = s(t0, t1); t[1] = s(t1, t2); t[2] = s(t2, t3); t[3] = s(t3, t4); t[4] = s(t4, t5); t[5] = s(t5, t6); t[6] = s(t6, t7);
    });
    return dt;
}
// OpenMP team size of this library: torchrun exports OMP_NUM_THREADS=1 to every rank, so bench.py sets the count explicitly
// (and reports what omp_get_max_threads() then says, not os.cpu_count()).
void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_max_threads() { return omp_get_max_threads(); }
int orc_num_phantom(void* h) { int r = 0; DISPATCH(h, { r = (int)o.phantom.size(); }); return r; }

}  // extern "C"
