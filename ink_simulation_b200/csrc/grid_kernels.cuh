// grid_kernels.cuh — MAC water-grid phases of System::updateVelocityField (reference src/waterSolver.cpp:41-96)
// as sm_100a CUDA kernels.  Every kernel cites the reference loop it replaces.  All are HBM/L2 streaming
// kernels: one thread per cell, z fastest => a warp reads/writes whole 128-byte lines; no tensor cores.
//
// Phase protocol (faithfulness ledger Q14): the reference runs phase(old -> curr) then `old = curr`.
// Point-wise phases (forces, pressure-gradient subtract, vorticity force) update the committed field in
// place; gather/stencil-on-velocity phases (convection, viscosity) write the second buffer and the host
// swaps pointers.  No copy sweeps exist.
#pragma once
#include "common.cuh"

namespace ink {

// ------------------------------------------------------------------------------------------------
// Staggered trilinear sampling — System::getInterpolatedValue / getVelocity (src/systemHelpers.cpp:110-115,158-209)
// ------------------------------------------------------------------------------------------------
template <class Real>
__device__ __forceinline__ bool axis_ok(Real f, Real bound, int n) {
    // isInBounds() applied to an integer index converted to float (Q13) AND array safety for h != 1
    return (f >= Real(0)) && (f < bound) && (f < Real(n));
}

// STRICT: the reference's operation order — weight = (wx*wy)*wz, total += weight*f over the 8 corners in the
// order 000,100,010,110,001,101,011,111; out-of-grid corners add nothing to total but their weight still
// enters weightAccum; result = total / weightAccum.
template <class Real>
__device__ __forceinline__ Real interp_strict(const Real* __restrict__ f, const Grid& g, const Consts<Real>& C, Real x, Real y, Real z) {
    const Real fi = rfloor(x), fj = rfloor(y), fk = rfloor(z);
    const Real fi1 = fi + Real(1), fj1 = fj + Real(1), fk1 = fk + Real(1);
    const Real wx0 = fi1 - x, wx1 = x - fi;
    const Real wy0 = fj1 - y, wy1 = y - fj;
    const Real wz0 = fk1 - z, wz1 = z - fk;
    const bool xi0 = axis_ok(fi, C.bx, g.X), xi1 = axis_ok(fi1, C.bx, g.X);
    const bool yj0 = axis_ok(fj, C.by, g.Y), yj1 = axis_ok(fj1, C.by, g.Y);
    const bool zk0 = axis_ok(fk, C.bz, g.Z), zk1 = axis_ok(fk1, C.bz, g.Z);
    const long long base = (long long)(int)fj * g.sy + (long long)(int)fi * g.sx + (int)fk;
    const Real* __restrict__ p = f + base;
    const long long sx = g.sx, sy = g.sy;
    // issue all in-bounds loads first (MLP), then accumulate in the reference order
    const Real f000 = (xi0 && yj0 && zk0) ? __ldg(p) : Real(0);
    const Real f100 = (xi1 && yj0 && zk0) ? __ldg(p + sx) : Real(0);
    const Real f010 = (xi0 && yj1 && zk0) ? __ldg(p + sy) : Real(0);
    const Real f110 = (xi1 && yj1 && zk0) ? __ldg(p + sx + sy) : Real(0);
    const Real f001 = (xi0 && yj0 && zk1) ? __ldg(p + 1) : Real(0);
    const Real f101 = (xi1 && yj0 && zk1) ? __ldg(p + sx + 1) : Real(0);
    const Real f011 = (xi0 && yj1 && zk1) ? __ldg(p + sy + 1) : Real(0);
    const Real f111 = (xi1 && yj1 && zk1) ? __ldg(p + sx + sy + 1) : Real(0);
    const Real w00 = wx0 * wy0, w10 = wx1 * wy0, w01 = wx0 * wy1, w11 = wx1 * wy1;
    Real tot = Real(0), wacc = Real(0), w;
    w = w00 * wz0; if (xi0 && yj0 && zk0) tot += w * f000; wacc += w;
    w = w10 * wz0; if (xi1 && yj0 && zk0) tot += w * f100; wacc += w;
    w = w01 * wz0; if (xi0 && yj1 && zk0) tot += w * f010; wacc += w;
    w = w11 * wz0; if (xi1 && yj1 && zk0) tot += w * f110; wacc += w;
    w = w00 * wz1; if (xi0 && yj0 && zk1) tot += w * f001; wacc += w;
    w = w10 * wz1; if (xi1 && yj0 && zk1) tot += w * f101; wacc += w;
    w = w01 * wz1; if (xi0 && yj1 && zk1) tot += w * f011; wacc += w;
    w = w11 * wz1; if (xi1 && yj1 && zk1) tot += w * f111; wacc += w;
    if (wacc == Real(0)) return Real(0);
    return tot / wacc;
}

// FAST: same stencil, zero extension, nested FMA lerps; the (== 1 up to rounding) weight sum is not divided out.
// Differs from STRICT by a few ulp of the sampled value.
template <class Real>
__device__ __forceinline__ Real interp_fast(const Real* __restrict__ f, const Grid& g, const Consts<Real>& C, Real x, Real y, Real z) {
    const Real fi = rfloor(x), fj = rfloor(y), fk = rfloor(z);
    const Real tx = x - fi, ty = y - fj, tz = z - fk;
    const int i = (int)fi, j = (int)fj, k = (int)fk;
    const long long sx = g.sx, sy = g.sy;
    const Real* __restrict__ p = f + ((long long)j * sy + (long long)i * sx + k);
    Real f000, f100, f010, f110, f001, f101, f011, f111;
    if (i >= 0 && i + 1 < g.X && j >= 0 && j + 1 < g.Y && k >= 0 && k + 1 < g.Z) {   // interior: no per-corner predicates
        f000 = __ldg(p); f001 = __ldg(p + 1);
        f100 = __ldg(p + sx); f101 = __ldg(p + sx + 1);
        f010 = __ldg(p + sy); f011 = __ldg(p + sy + 1);
        f110 = __ldg(p + sx + sy); f111 = __ldg(p + sx + sy + 1);
    } else {
        const bool xi0 = axis_ok(fi, C.bx, g.X), xi1 = axis_ok(fi + Real(1), C.bx, g.X);
        const bool yj0 = axis_ok(fj, C.by, g.Y), yj1 = axis_ok(fj + Real(1), C.by, g.Y);
        const bool zk0 = axis_ok(fk, C.bz, g.Z), zk1 = axis_ok(fk + Real(1), C.bz, g.Z);
        f000 = (xi0 && yj0 && zk0) ? __ldg(p) : Real(0);
        f100 = (xi1 && yj0 && zk0) ? __ldg(p + sx) : Real(0);
        f010 = (xi0 && yj1 && zk0) ? __ldg(p + sy) : Real(0);
        f110 = (xi1 && yj1 && zk0) ? __ldg(p + sx + sy) : Real(0);
        f001 = (xi0 && yj0 && zk1) ? __ldg(p + 1) : Real(0);
        f101 = (xi1 && yj0 && zk1) ? __ldg(p + sx + 1) : Real(0);
        f011 = (xi0 && yj1 && zk1) ? __ldg(p + sy + 1) : Real(0);
        f111 = (xi1 && yj1 && zk1) ? __ldg(p + sx + sy + 1) : Real(0);
    }
    const Real a00 = rfma(tz, f001 - f000, f000);
    const Real a10 = rfma(tz, f101 - f100, f100);
    const Real a01 = rfma(tz, f011 - f010, f010);
    const Real a11 = rfma(tz, f111 - f110, f110);
    const Real b0 = rfma(tx, a10 - a00, a00);
    const Real b1 = rfma(tx, a11 - a01, a01);
    return rfma(ty, b1 - b0, b0);
}

template <class Real, bool FAST>
__device__ __forceinline__ void get_velocity(const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                             const Grid& g, const Consts<Real>& C, Real px, Real py, Real pz, Real& ox, Real& oy, Real& oz) {
    const Real xs = C.unit_h ? px : px / C.h;
    const Real ys = C.unit_h ? py : py / C.h;
    const Real zs = C.unit_h ? pz : pz / C.h;
    const Real xh = xs - Real(0.5f), yh = ys - Real(0.5f), zh = zs - Real(0.5f);
    if (FAST) {
        ox = interp_fast(u, g, C, xs, yh, zh);
        oy = interp_fast(v, g, C, xh, ys, zh);
        oz = interp_fast(w, g, C, xh, yh, zs);
    } else {
        ox = interp_strict(u, g, C, xs, yh, zh);
        oy = interp_strict(v, g, C, xh, ys, zh);
        oz = interp_strict(w, g, C, xh, yh, zs);
    }
}

// thread -> cell of this rank's slab; returns false past the end
__device__ __forceinline__ bool cell_of_thread(const Grid& g, long long t, int& i, int& j, int& k, long long& c) {
    const long long rows = (long long)(g.j1 - g.j0) * g.X;
    if (t >= rows * g.Z) return false;
    k = (int)(t % g.Z);
    const long long row = t / g.Z;
    i = (int)(row % g.X);
    j = g.j0 + (int)(row / g.X);
    c = (long long)g.j0 * g.sy + t;
    return true;
}

// ------------------------------------------------------------------------------------------------
// K0  max |oldVelocity| — System::calcTimeStep (src/waterSolver.cpp:13-21).  Non-negative floats order like
// their bit patterns, so one atomicMax(uint) per block gives the exact, order-independent maximum.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void block_max_to_global(float m, unsigned int* out) {
    __shared__ float smax[32];
    m = warp_max(m);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) smax[wid] = m;
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        m = lane < nw ? smax[lane] : 0.f;
        m = warp_max(m);
        if (lane == 0) atomicMax(out, __float_as_uint(m));
    }
}

template <class Real>
__global__ void __launch_bounds__(256) k_max_speed(Grid g, const Real* __restrict__ u, const Real* __restrict__ v,
                                                   const Real* __restrict__ w, unsigned int* out) {
    INK_PDL();
    float m = 0.f;
    const long long cells = (long long)(g.j1 - g.j0) * g.sy;
    const long long base = (long long)g.j0 * g.sy;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < cells; t += (long long)gridDim.x * blockDim.x) {
        const float s = (float)norm3(u[base + t], v[base + t], w[base + t]);
        if (m < s) m = s;
    }
    block_max_to_global(m, out);
}

// System::checkNanAndInf (src/systemHelpers.cpp:212-241) as ONE device reduction over n (x, y, z) triples: the reference
// asserts, per grid cell and per particle, no NaN / Inf component and norm() < 10000.  counts[0] += triples with a
// non-finite component, counts[1] += triples whose norm is not < 10000 (a NaN norm fails that test, as in the reference).
// check_norm = 0: positions (the reference only tests them for NaN / Inf).  8 bytes come back instead of the whole state.
template <class Real>
__global__ void __launch_bounds__(256) k_check_finite(long long n, const Real* __restrict__ x, const Real* __restrict__ y, const Real* __restrict__ z,
                                                      int check_norm, unsigned long long* __restrict__ counts) {
    INK_PDL();
    unsigned int bad = 0, big = 0;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const Real a = x[t], b = y[t], c = z[t];
        if (!(isfinite(a) && isfinite(b) && isfinite(c))) ++bad;
        if (check_norm && !(norm3(a, b, c) < Real(10000))) ++big;
    }
    bad = __reduce_add_sync(0xffffffffu, bad);
    big = __reduce_add_sync(0xffffffffu, big);
    if ((threadIdx.x & 31) == 0) {
        if (bad) atomicAdd(counts, (unsigned long long)bad);
        if (big) atomicAdd(counts + 1, (unsigned long long)big);
    }
}

// ------------------------------------------------------------------------------------------------
// K1  nearest-cell semi-Lagrangian convection — System::applyConvection + traceParticle
// (src/waterSolver.cpp:98-128; Q6, Q7).  old (u,v,w) -> curr (nu,nv,nw).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float trace_mid(float x, double t, float v) { return (float)((double)x + 0.5 * t * (double)v); }
__device__ __forceinline__ double trace_mid(double x, double t, double v) { return x + 0.5 * t * v; }

template <class Real>
__global__ void __launch_bounds__(256) k_advect(Grid g, Consts<Real> C, Real dt, const Real* __restrict__ u, const Real* __restrict__ v,
                                                const Real* __restrict__ w, Real* __restrict__ nu, Real* __restrict__ nv, Real* __restrict__ nw) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const Real t = -dt;
    // cell centre i + CELL_DIM/2.f is a float expression in the reference (traceParticle takes floats)
    const float halff = (float)C.h / 2.f;
    const Real x = (Real)(i + halff), y = (Real)(j + halff), z = (Real)(k + halff);
    Real vx, vy, vz;
    get_velocity<Real, false>(u, v, w, g, C, x, y, z, vx, vy, vz);
    // midpoint: 0.5 * t * vel evaluated in double, rounded to Real (src/waterSolver.cpp:100)
    const Real mx = trace_mid(x, (double)t, vx), my = trace_mid(y, (double)t, vy), mz = trace_mid(z, (double)t, vz);
    get_velocity<Real, false>(u, v, w, g, C, mx, my, mz, vx, vy, vz);
    const Real px = x + t * vx, py = y + t * vy, pz = z + t * vz;
    long long src = c;
    const bool inb = (px >= 0) && (px < C.bx) && (py >= 0) && (py < C.by) && (pz >= 0) && (pz < C.bz);
    if (inb) {
        const int si = (int)rfloor(px), sj = (int)rfloor(py), sk = (int)rfloor(pz);
        if (in_idx(g, si, sj, sk)) src = cidx(g, si, sj, sk);
    }
    nu[c] = u[src]; nv[c] = v[src]; nw[c] = w[src];
}

// ------------------------------------------------------------------------------------------------
// K2  particle-localised body force — System::applyExternalForces / updateForce (src/waterSolver.cpp:158-191; Q8).
// seed[] (1 byte per cell) is written by the particle kernel / k_mark_particle_cells.  K2b dilates it by the
// box-clipped Manhattan ball of radius BUFFER_SIZE and adds dt*g once per marked cell, in place.
// ------------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_dilate_add_gravity(Grid g, int R, const uint8_t* __restrict__ seed, uint8_t* __restrict__ mark,
                                                            Real gx, Real gy, Real gz, Real* __restrict__ u, Real* __restrict__ v, Real* __restrict__ w) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    uint8_t m = 0;
    for (int dj = -R; dj <= R && !m; ++dj) {
        const int jj = j + dj;
        if (jj < 0 || jj >= g.Y) continue;
        const int rj = R - abs(dj);
        for (int di = -rj; di <= rj && !m; ++di) {
            const int ii = i + di;
            if (ii < 0 || ii >= g.X) continue;
            const int rk = rj - abs(di);
            const int k0 = max(k - rk, 0), k1 = min(k + rk, g.Z - 1);
            const uint8_t* __restrict__ row = seed + ((long long)jj * g.sy + (long long)ii * g.sx);
            for (int kk = k0; kk <= k1; ++kk)
                if (row[kk]) { m = 1; break; }
        }
    }
    mark[c] = m;
    if (m) {
        if (gx != Real(0)) u[c] += gx;   // adding an exact zero cannot change a value the reference would keep
        if (gy != Real(0)) v[c] += gy;
        if (gz != Real(0)) w[c] += gz;
    }
}

// ------------------------------------------------------------------------------------------------
// K3  explicit viscosity — System::applyViscosity + laplacianOperatorOnVelocity
// (src/waterSolver.cpp:225-238, src/systemHelpers.cpp:88-107; Q9).  old -> curr.
// ------------------------------------------------------------------------------------------------
template <class Real>
__device__ __forceinline__ Real lap7(const Real* __restrict__ f, const Grid& g, int i, int j, int k, long long c) {
    Real l = Real(0);
    l += (i + 1 < g.X) ? f[c + g.sx] : Real(0);
    l += (i - 1 >= 0) ? f[c - g.sx] : Real(0);
    l += (j + 1 < g.Y) ? f[c + g.sy] : Real(0);
    l += (j - 1 >= 0) ? f[c - g.sy] : Real(0);
    l += (k + 1 < g.Z) ? f[c + 1] : Real(0);
    l += (k - 1 >= 0) ? f[c - 1] : Real(0);
    l -= Real(6) * f[c];
    return l;
}
template <class Real>
__global__ void __launch_bounds__(256) k_viscosity(Grid g, Real s, const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                   Real* __restrict__ nu, Real* __restrict__ nv, Real* __restrict__ nw) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const Real lx = lap7(u, g, i, j, k, c), ly = lap7(v, g, i, j, k, c), lz = lap7(w, g, i, j, k, c);
    nu[c] = u[c] + s * lx;
    nv[c] = v[c] + s * ly;
    nw[c] = w[c] + s * lz;
}

// ------------------------------------------------------------------------------------------------
// K4  divergence / pressure right-hand side — System::getDivergence (src/systemHelpers.cpp:30-43) and the RHS of
// calculatePressure (src/waterSolver.cpp:264-278; Q10).  mode 0: out = div; mode 1: out = (rho*h/dt)*div - k_i*p_atm
// (the reference's b); mode 2: out = -(rho*h/dt)*div  (RHS of the SPD system (-A) q = -b', q = p - p_atm).
// ------------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_divergence(Grid g, int mode, Real scale, Real patm, const Real* __restrict__ u, const Real* __restrict__ v,
                                                    const Real* __restrict__ w, Real* __restrict__ out) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    Real d = -u[c] - v[c] - w[c];
    if (i + 1 < g.X) d += u[c + g.sx];
    if (j + 1 < g.Y) d += v[c + g.sy];
    if (k + 1 < g.Z) d += w[c + 1];
    if (mode == 0) { out[c] = d; return; }
    if (mode == 1) {
        const int nn = (i + 1 < g.X) + (i - 1 >= 0) + (j + 1 < g.Y) + (j - 1 >= 0) + (k + 1 < g.Z) + (k - 1 >= 0);
        const int ki = 6 - nn;
        out[c] = scale * d - ki * patm;
        return;
    }
    out[c] = -(scale * d);
}

// ------------------------------------------------------------------------------------------------
// K6  pressure-gradient subtract — System::applyPressure + getGradient (src/waterSolver.cpp:284-296,
// src/systemHelpers.cpp:5-28; Q10): backward differences, low-side ghost = the literal 1.  In place.
// ------------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_grad_subtract(Grid g, Real s, const Real* __restrict__ p, Real* __restrict__ u, Real* __restrict__ v, Real* __restrict__ w) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const Real val = p[c];
    Real gx = val, gy = val, gz = val;
    gx -= (i - 1 >= 0) ? p[c - g.sx] : Real(1);
    gy -= (j - 1 >= 0) ? p[c - g.sy] : Real(1);
    gz -= (k - 1 >= 0) ? p[c - 1] : Real(1);
    u[c] -= s * gx; v[c] -= s * gy; w[c] -= s * gz;
}

template <class Tin, class Tout>
__global__ void __launch_bounds__(256) k_convert(long long n, const Tin* __restrict__ in, Tout* __restrict__ out) {
    INK_PDL();
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) out[t] = (Tout)in[t];
}
template <class Real>
__global__ void __launch_bounds__(256) k_scale(long long n, Real s, Real* __restrict__ x) {
    INK_PDL();
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) x[t] = s * x[t];
}
// q -> p = q + p_atm (and back) for the shifted Poisson unknown
template <class Real>
__global__ void __launch_bounds__(256) k_add_scalar(long long n, Real s, const Real* __restrict__ in, Real* __restrict__ out) {
    INK_PDL();
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) out[t] = in[t] + s;
}

// ------------------------------------------------------------------------------------------------
// K7a curl — System::getCurl (src/systemHelpers.cpp:45-69; Q11): central differences without 1/2, zero outside.
// ------------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_curl(Grid g, const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                              Real* __restrict__ cx, Real* __restrict__ cy, Real* __restrict__ cz) {
    INK_PDL();
    int i, j, k; long long c;
    if (!cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const bool ip = i + 1 < g.X, im = i - 1 >= 0, jp = j + 1 < g.Y, jm = j - 1 >= 0, kp = k + 1 < g.Z, km = k - 1 >= 0;
    Real a = Real(0), b = Real(0), d = Real(0);
    a += jp ? w[c + g.sy] : Real(0);
    a -= jm ? w[c - g.sy] : Real(0);
    a -= kp ? v[c + 1] : Real(0);
    a += km ? v[c - 1] : Real(0);
    b += kp ? u[c + 1] : Real(0);
    b -= km ? u[c - 1] : Real(0);
    b -= ip ? w[c + g.sx] : Real(0);
    b += im ? w[c - g.sx] : Real(0);
    d += ip ? v[c + g.sx] : Real(0);
    d -= im ? v[c - g.sx] : Real(0);
    d -= jp ? u[c + g.sy] : Real(0);
    d += jm ? u[c - g.sy] : Real(0);
    cx[c] = a; cy[c] = b; cz[c] = d;
}

// ------------------------------------------------------------------------------------------------
// K7b vorticity confinement — System::applyVorticity / getVort / getCurlGradient (src/waterSolver.cpp:193-222,
// src/systemHelpers.cpp:71-80; Q2: the k-1 term of the z gradient is always 0; Q11).  In place on the velocity.
// Fused K0: the block maximum of the NEW |velocity| feeds the next step's calcTimeStep.
// ------------------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_vorticity(Grid g, Real dt, Real kvort, const Real* __restrict__ cx, const Real* __restrict__ cy,
                                                   const Real* __restrict__ cz, Real* __restrict__ u, Real* __restrict__ v, Real* __restrict__ w,
                                                   unsigned int* maxspeed) {
    INK_PDL();
    int i, j, k; long long c;
    float speed = 0.f;
    if (cell_of_thread(g, (long long)blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) {
        const Real wx = cx[c], wy = cy[c], wz = cz[c];
        Real nu_ = u[c], nv_ = v[c], nw_ = w[c];
        if (!(wx == Real(0) && wy == Real(0) && wz == Real(0))) {
            Real g0 = Real(0), g1 = Real(0), g2 = Real(0);
            long long q;
            q = c + g.sx; g0 += (i + 1 < g.X) ? norm3(cx[q], cy[q], cz[q]) : Real(0);
            q = c - g.sx; g0 -= (i - 1 >= 0) ? norm3(cx[q], cy[q], cz[q]) : Real(0);
            q = c + g.sy; g1 += (j + 1 < g.Y) ? norm3(cx[q], cy[q], cz[q]) : Real(0);
            q = c - g.sy; g1 -= (j - 1 >= 0) ? norm3(cx[q], cy[q], cz[q]) : Real(0);
            q = c + 1;    g2 += (k + 1 < g.Z) ? norm3(cx[q], cy[q], cz[q]) : Real(0);
            g2 -= Real(0);
            const Real nrm = norm3(wx, wy, wz);
            const Real Nx = g0 / nrm, Ny = g1 / nrm, Nz = g2 / nrm;
            const Real Fx = kvort * (Ny * wz - Nz * wy);
            const Real Fy = kvort * (Nz * wx - Nx * wz);
            const Real Fz = kvort * (Nx * wy - Ny * wx);
            nu_ += dt * Fx; nv_ += dt * Fy; nw_ += dt * Fz;
            u[c] = nu_; v[c] = nv_; w[c] = nw_;
        }
        speed = (float)norm3(nu_, nv_, nw_);
    }
    block_max_to_global(speed, maxspeed);
}

// ------------------------------------------------------------------------------------------------
// Layout conversion between the public (X,Y,Z) C-order (i*Y + j)*Z + k and the internal (j*X + i)*Z + k,
// with dtype conversion.  Not on the hot path (state injection / readback only).
// ------------------------------------------------------------------------------------------------
template <class Tin, class Tout>
__global__ void __launch_bounds__(256) k_xyz_to_internal(Grid g, const Tin* __restrict__ in, Tout* __restrict__ out) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.n) return;
    const int k = (int)(t % g.Z);
    const long long row = t / g.Z;
    const int i = (int)(row % g.X), j = (int)(row / g.X);
    out[t] = (Tout)in[((long long)i * g.Y + j) * g.Z + k];
}
template <class Tin, class Tout>
__global__ void __launch_bounds__(256) k_internal_to_xyz(Grid g, const Tin* __restrict__ in, Tout* __restrict__ out) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.n) return;
    const int k = (int)(t % g.Z);
    const long long row = t / g.Z;
    const int i = (int)(row % g.X), j = (int)(row / g.X);
    out[((long long)i * g.Y + j) * g.Z + k] = (Tout)in[t];
}
__global__ void __launch_bounds__(256) k_mark_to_xyz(Grid g, const uint8_t* __restrict__ in, uint8_t* __restrict__ out) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.n) return;
    const int k = (int)(t % g.Z);
    const long long row = t / g.Z;
    const int i = (int)(row % g.X), j = (int)(row / g.X);
    out[((long long)i * g.Y + j) * g.Z + k] = in[t];
}

}  // namespace ink
