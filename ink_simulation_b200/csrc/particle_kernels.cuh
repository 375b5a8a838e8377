// particle_kernels.cuh — RK2 ink-particle update (reference src/inkSolver.cpp:44-84) and the particle
// storage around it, as sm_100a CUDA kernels.
//
// Storage: SoA px,py,pz,vx,vy,vz in the compute precision, kept BINNED BY GRID CELL (counting sort on the
// internal cell index, y slowest / z fastest) so that the 32 particles of a warp gather from the same few
// 128-byte lines of the velocity field (L1 hits) and consecutive CTAs walk the field once (L2 -> HBM read
// once per step).  orig[s] = index of slot s in the caller's order; getters un-permute, so .ply output keeps
// the reference's stable particle order (render.py keys by vertex index).
#pragma once
#include "common.cuh"
#include "grid_kernels.cuh"
#include "sample3.cuh"

namespace ink {

template <class Real>
struct Particles {
    Real* p[3];
    Real* v[3];
    uint32_t* orig;
};

__device__ __forceinline__ float rmin(float a, float b) { return (b < a) ? b : a; }   // std::min
__device__ __forceinline__ float rmax(float a, float b) { return (a < b) ? b : a; }   // std::max
__device__ __forceinline__ double rmin(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double rmax(double a, double b) { return (a < b) ? b : a; }

// cell containing a point, floor(pos) per axis with NO division by h (System::getCellIndexFromPoint,
// src/waterSolver.cpp:104-106, as used by applyExternalForces :163)
template <class Real>
__device__ __forceinline__ bool cell_of_point(const Grid& g, Real x, Real y, Real z, long long& c) {
    const Real fx = rfloor(x), fy = rfloor(y), fz = rfloor(z);
    if (!(fx >= Real(0) && fx < Real(g.X) && fy >= Real(0) && fy < Real(g.Y) && fz >= Real(0) && fz < Real(g.Z))) return false;
    c = cidx(g, (int)fx, (int)fy, (int)fz);
    return true;
}

// ------------------------------------------------------------------------------------------------
// K8  System::updateParticles (src/inkSolver.cpp:44-84; Q3, Q4, Q5) fused with K2a (seed the force mask with the
// cell of the NEW position: it is what the next step's applyExternalForces sees, src/waterSolver.cpp:161-163).
// Algorithmic traffic: read pos 3R, write pos 3R, write vel 3R (R = sizeof(Real)) + the velocity field once.
// ------------------------------------------------------------------------------------------------
template <class Real, bool FAST, bool UNIT_H>
__global__ void __launch_bounds__(256) k_particles_rk2(Grid g, Consts<Real> C, Real dt, long long n, Particles<Real> P,
                                                       const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                       uint8_t* __restrict__ seed) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    Real x = ld_stream(P.p[0] + s), y = ld_stream(P.p[1] + s), z = ld_stream(P.p[2] + s);
    Real sx, sy, sz;
    get_velocity3<Real, FAST, UNIT_H>(u, v, w, g, C, x, y, z, sx, sy, sz);
    Real vx = C.rho * sx, vy = C.rho * sy, vz = C.rho * sz;                       // equation 9
    const Real mx = x + (vx * dt * Real(.5)), my = y + (vy * dt * Real(.5)), mz = z + (vz * dt * Real(.5));  // equation 7
    get_velocity3<Real, FAST, UNIT_H>(u, v, w, g, C, mx, my, mz, sx, sy, sz);
    const Real wx = C.rho * sx, wy = C.rho * sy, wz = C.rho * sz;
    x = mx + (dt * wx); y = my + (dt * wy); z = mz + (dt * wz);                  // equation 8 (from the MIDPOINT, Q3)
    if (x < Real(0) || x > C.bx) { vx *= Real(-1.f); x = rmin(rmax(x, C.lo), C.cx); }
    if (y < Real(0) || y > C.by) { vy *= Real(-1.f); y = rmin(rmax(y, C.lo), C.cy); }
    if (z < Real(0) || z > C.bz) { vz *= Real(-1.f); z = rmin(rmax(z, C.lo), C.cz); }
    st_stream(P.p[0] + s, x); st_stream(P.p[1] + s, y); st_stream(P.p[2] + s, z);
    st_stream(P.v[0] + s, vx); st_stream(P.v[1] + s, vy); st_stream(P.v[2] + s, vz);
    if (UNIT_H) {   // launcher guarantees X*Y*Z < 2^31 here: 32-bit cell index; same test as cell_of_point
        const int ci = (int)rfloor(x), cj = (int)rfloor(y), ck = (int)rfloor(z);
        if ((unsigned)ci < (unsigned)g.X && (unsigned)cj < (unsigned)g.Y && (unsigned)ck < (unsigned)g.Z)
            seed[(cj * g.X + ci) * g.Z + ck] = 1;
    } else {
        long long c;
        if (cell_of_point(g, x, y, z, c)) seed[c] = 1;
    }
}

// K2a stand-alone: seed mask from the current positions (after ink_set_particles); counts out-of-grid particles.
template <class Real>
__global__ void __launch_bounds__(256) k_mark_particle_cells(Grid g, long long n, Particles<Real> P, uint8_t* __restrict__ seed, unsigned int* oob_count) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    long long c;
    if (cell_of_point(g, P.p[0][s], P.p[1][s], P.p[2][s], c)) seed[c] = 1;
    else atomicAdd(oob_count, 1u);
}

// getVelocity at arbitrary points (test hook; System::getVelocity src/systemHelpers.cpp:110-115)
template <class Real, bool FAST>
__global__ void __launch_bounds__(256) k_sample(Grid g, Consts<Real> C, long long n, const Real* __restrict__ pos, Real* __restrict__ out,
                                                const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    Real a, b, c;
    get_velocity<Real, FAST>(u, v, w, g, C, pos[3 * s], pos[3 * s + 1], pos[3 * s + 2], a, b, c);
    out[3 * s] = a; out[3 * s + 1] = b; out[3 * s + 2] = c;
}

// ------------------------------------------------------------------------------------------------
// Import / export (AoS xyz in the caller's dtype <-> internal SoA) and cell binning.
// ------------------------------------------------------------------------------------------------
template <class Tin, class Real>
__global__ void __launch_bounds__(256) k_import_aos(long long n, long long dst_off, const Tin* __restrict__ pos, const Tin* __restrict__ vel, Particles<Real> P) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const long long d = dst_off + s;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        P.p[a][d] = (Real)pos[3 * s + a];
        P.v[a][d] = vel ? (Real)vel[3 * s + a] : Real(0);
    }
    P.orig[d] = (uint32_t)d;
}
// Streaming input (ink_commit_positions): new positions of one drop, caller order -> the binned slots the particles
// already occupy (orig keeps the mapping).  No re-binning, velocities untouched (K8 overwrites them).
template <class Tin, class Real>
__global__ void __launch_bounds__(256) k_import_pos_perm(long long n, long long off, long long cnt, Particles<Real> P, const Tin* __restrict__ pos) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const long long o = (long long)P.orig[s] - off;
    if (o < 0 || o >= cnt) return;
#pragma unroll
    for (int a = 0; a < 3; ++a) P.p[a][s] = (Real)pos[3 * o + a];
}
// out[(orig - off)*3 + a] for the particles of one drop (orig in [off, off+cnt))
template <class Real, class Tout>
__global__ void __launch_bounds__(256) k_export_aos(long long n, long long off, long long cnt, Particles<Real> P, Tout* __restrict__ pos, Tout* __restrict__ vel) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const long long o = (long long)P.orig[s] - off;
    if (o < 0 || o >= cnt) return;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (pos) pos[3 * o + a] = (Tout)P.p[a][s];
        if (vel) vel[3 * o + a] = (Tout)P.v[a][s];
    }
}

// bin key: particles are binned by QUARTER COLUMN — the half-cell the position falls in along x and y, then the cell
// along z:  key = ((2j + hy) * 2X + (2i + hx)) * Z + k,  hx = frac(x) >= 0.5.  The staggered sampler reads rows
// floor(x), floor(x - 0.5), floor(y), floor(y - 0.5): inside one quarter column all four are the same for every
// particle, so each of a warp's 48 corner gathers falls in ONE z-row of the field (1-3 128-byte lines) instead of the
// 2 (u, v) or 4 (w) rows a whole-cell bin mixes (ncu r1_b: 3.0 L1 wavefronts per gather, the kernel's bottleneck).
constexpr int kBinsPerCell = 4;
template <class Real>
__device__ __forceinline__ uint32_t bin_key(const Grid& g, Real x, Real y, Real z) {
    const Real fx = rfloor(x), fy = rfloor(y);
    int i = (int)fx, j = (int)fy, k = (int)rfloor(z);
    const int hx = (x - fx >= Real(0.5)) ? 1 : 0, hy = (y - fy >= Real(0.5)) ? 1 : 0;
    i = min(max(2 * i + hx, 0), 2 * g.X - 1); j = min(max(2 * j + hy, 0), 2 * g.Y - 1); k = min(max(k, 0), g.Z - 1);
    return (uint32_t)(((long long)j * (2 * g.X) + i) * g.Z + k);
}
template <class Real>
__global__ void __launch_bounds__(256) k_bin_count(Grid g, long long n, Particles<Real> P, uint32_t* __restrict__ count, uint32_t* __restrict__ key, uint32_t* __restrict__ rank) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t c = bin_key(g, P.p[0][s], P.p[1][s], P.p[2][s]);
    key[s] = c;
    rank[s] = atomicAdd(count + c, 1u);
}
template <class Real>
__global__ void __launch_bounds__(256) k_bin_scatter(long long n, Particles<Real> src, Particles<Real> dst, const uint32_t* __restrict__ offset,
                                                     const uint32_t* __restrict__ key, const uint32_t* __restrict__ rank) {
    INK_PDL();
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const long long d = (long long)offset[key[s]] + rank[s];
#pragma unroll
    for (int a = 0; a < 3; ++a) { dst.p[a][d] = src.p[a][s]; dst.v[a][d] = src.v[a][s]; }
    dst.orig[d] = src.orig[s];
}

// ---- exclusive scan of uint32 (tile = 256 threads x 8) ------------------------------------------
constexpr int kScanTile = 2048;
__global__ void __launch_bounds__(256) k_scan_tiles(long long n, const uint32_t* in, uint32_t* out, uint32_t* __restrict__ tilesum) {
    INK_PDL();  // in may alias out
    __shared__ uint32_t wsum[8];
    const long long base = (long long)blockIdx.x * kScanTile + (long long)threadIdx.x * 8;
    uint32_t v[8], t = 0;
#pragma unroll
    for (int a = 0; a < 8; ++a) { v[a] = (base + a < n) ? in[base + a] : 0u; t += v[a]; }
    // warp inclusive scan of per-thread totals
    uint32_t inc = t;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    uint32_t woff = 0;
    for (int q = 0; q < wid; ++q) woff += wsum[q];
    uint32_t run = woff + inc - t;
#pragma unroll
    for (int a = 0; a < 8; ++a) { if (base + a < n) out[base + a] = run; run += v[a]; }
    if (threadIdx.x == 255) tilesum[blockIdx.x] = woff + inc;
}
__global__ void __launch_bounds__(256) k_scan_add(long long n, uint32_t* __restrict__ out, const uint32_t* __restrict__ tileoff) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) out[t] += tileoff[t / kScanTile];
}

}  // namespace ink
