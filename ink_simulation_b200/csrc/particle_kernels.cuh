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
#include "tma.cuh"

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

// ------------------------------------------------------------------------------------------------
// K8, tile version: the velocity rows a chunk of binned particles can touch are staged in SHARED MEMORY by the TMA unit
// (cp.async.bulk + mbarrier, tma.cuh) and the 48 corner gathers per particle read shared memory instead of L1/L2.
//
// Why: ncu r1_e — the global-gather kernel moves 7.2 sectors per request through L1 (a warp's 32 lanes hit 2.6 distinct
// 128-byte lines per gather, 48 gathers per particle) and is L1-wavefront bound at 0.27 of the HBM roofline while DRAM idles.
// Particles are binned by quarter column (bin_key): one HALF-ROW (j, hy) of the grid is a contiguous slot range ordered by x
// then z.  A block takes a chunk of <= 256 * kTileKP consecutive slots of one half-row (k_tile_worklist), finds the half-cell
// bounding box of its particles (so drift since the last re-binning only widens the tile), and stages
//     u rows  x in [ux0, ux1], y in [uy0, uy1]     v rows  x in [vx0, vx1], y in [vy0, vy1]     w rows  x as v, y as u
// — each (component, y) an x-run of whole z-rows, i.e. ONE contiguous bulk copy — with one half-cell of margin for the RK2
// midpoint (CFL: it moves < 0.1 cell).  A sample whose 24 corners are not all inside the tile (wild dt, walls: the tile is
// clipped to the grid, so "inside the tile" implies "inside the grid") takes the global-memory sampler: results never depend
// on the tile.  Arithmetic per sample is the same eval_strict / eval_fast on the same eight values: strict stays bit-exact.
// ------------------------------------------------------------------------------------------------
constexpr int kTileThreads = 256;
constexpr int kTileKP = 4;                       // particles per thread at most (chunk <= 1024 slots)
constexpr int kTileSmemBytes = 56 * 1024;        // 3 blocks per SM: 3 x (56 KB + 1 KB reserved + the static mbarrier / reduction arrays) fits the 228 KB carve-out, 4 x does not

// one block per half-row r = 2j + hy: split its slot range into equal chunks of at most `chunk` slots
__global__ void __launch_bounds__(128) k_tile_worklist(int X, int Y, int Z, long long np, const uint32_t* __restrict__ bin_off, int chunk,
                                                       uint2* __restrict__ work, unsigned int* __restrict__ nwork) {
    INK_PDL();
    __shared__ unsigned int base;
    const int r = blockIdx.x;
    const long long row_keys = 2ll * X * Z;
    const long long begin = bin_off[(long long)r * row_keys];
    const long long end = (r + 1 == 2 * Y) ? np : (long long)bin_off[(long long)(r + 1) * row_keys];
    const long long cnt = end - begin;
    if (cnt <= 0) return;
    const long long nch = (cnt + chunk - 1) / chunk, per = (cnt + nch - 1) / nch;
    if (threadIdx.x == 0) base = atomicAdd(nwork, (unsigned int)nch);
    __syncthreads();
    for (long long c = threadIdx.x; c < nch; c += blockDim.x) {
        const long long b0 = begin + c * per, b1 = (b0 + per < end) ? b0 + per : end;
        work[base + c] = (b0 < b1) ? make_uint2((uint32_t)b0, (uint32_t)b1) : make_uint2(0u, 0u);
    }
}

template <class Real>
struct TileGeom {
    int ux0, uy0, vx0, vy0;          // first staged row of the u / v tiles (w: x as v, y as u)
    unsigned int unx1, uny1, vnx1, vny1;   // extents minus one: (unsigned)(ix - ux0) < unx1  <=>  ix and ix + 1 are staged
    int Z, usy, vsy;                 // row stride (elements) and y strides of the u and v / w tiles
    int bv, bw;                      // element offsets of the v and w tiles
};

template <class Real>
__device__ __forceinline__ void load8s(const Real* p, int sx, int sy, Real (&f)[8]) {   // shared-memory corners, order of load8
    f[0] = p[0];       f[4] = p[1];
    f[1] = p[sx];      f[5] = p[sx + 1];
    f[2] = p[sy];      f[6] = p[sy + 1];
    f[3] = p[sy + sx]; f[7] = p[sy + sx + 1];
}

// get_velocity3's interior path on the staged tile; false = not all corners staged (caller takes the global sampler)
template <class Real, bool FAST>
__device__ __forceinline__ bool get_velocity3_tile(const Real* tile, const TileGeom<Real>& T, Real px, Real py, Real pz, Real& ox, Real& oy, Real& oz) {
    const Real xh = px - Real(0.5f), yh = py - Real(0.5f), zh = pz - Real(0.5f);
    const Real fx = rfloor(px), fy = rfloor(py), fz = rfloor(pz);
    const Real fxh = rfloor(xh), fyh = rfloor(yh), fzh = rfloor(zh);
    // NaN / huge coordinates: the float -> int conversion saturates (cvt.rzi), the range tests below then fail
    const int ix = (int)fx, jy = (int)fy, kz = (int)fz, ixh = (int)fxh, jyh = (int)fyh, kzh = (int)fzh;
    const unsigned int ax = (unsigned int)(ix - T.ux0), ay = (unsigned int)(jyh - T.uy0), bx = (unsigned int)(ixh - T.vx0), by = (unsigned int)(jy - T.vy0);
    if (!(ax < T.unx1 && ay < T.uny1 && bx < T.vnx1 && by < T.vny1 && kzh >= 0 && kz <= T.Z - 2)) return false;
    const int Z = T.Z;
    const Real* pu = tile + ((int)ay * (int)(T.unx1 + 1) + (int)ax) * Z + kzh;
    const Real* pv = tile + T.bv + ((int)by * (int)(T.vnx1 + 1) + (int)bx) * Z + kzh;
    const Real* pw = tile + T.bw + ((int)ay * (int)(T.vnx1 + 1) + (int)bx) * Z + kz;
    Real f[8];
    if (FAST) {
        const Real tx = px - fx, ty = py - fy, tz = pz - fz, txh = xh - fxh, tyh = yh - fyh, tzh = zh - fzh;
        load8s(pu, Z, T.usy, f);  ox = eval_fast(f, tx, tyh, tzh);
        load8s(pv, Z, T.vsy, f);  oy = eval_fast(f, txh, ty, tzh);
        load8s(pw, Z, T.vsy, f);  oz = eval_fast(f, txh, tyh, tz);
    } else {
        load8s(pu, Z, T.usy, f);  ox = eval_strict(f, px, fx, yh, fyh, zh, fzh);
        load8s(pv, Z, T.vsy, f);  oy = eval_strict(f, xh, fxh, py, fy, zh, fzh);
        load8s(pw, Z, T.vsy, f);  oz = eval_strict(f, xh, fxh, yh, fyh, pz, fz);
    }
    return true;
}

template <class Real, bool FAST>
__global__ void __launch_bounds__(kTileThreads, 3) k_particles_rk2_tile(Grid g, Consts<Real> C, Real dt, Particles<Real> P,
                                                                         const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                                         uint8_t* __restrict__ seed, const uint2* __restrict__ work,
                                                                         const unsigned int* __restrict__ nwork) {
    INK_PDL();
    if (blockIdx.x >= *nwork) return;
    const uint2 wk = work[blockIdx.x];
    if (wk.x >= wk.y) return;
    extern __shared__ __align__(16) unsigned char tile_raw[];
    Real* tile = reinterpret_cast<Real*>(tile_raw);
    __shared__ uint64_t bar;
    __shared__ int red[4][kTileThreads / 32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) mbar_init(&bar, 1);
    // this thread's particles + the half-cell bounding box of the block's particles
    Real x[kTileKP], y[kTileKP], z[kTileKP];
    int gx0 = 0x7fffffff, gx1 = -1, hy0 = 0x7fffffff, hy1 = -1;
    const Real gxm = Real(2 * g.X - 1), hym = Real(2 * g.Y - 1);
#pragma unroll
    for (int q = 0; q < kTileKP; ++q) {
        const unsigned int s = wk.x + tid + q * kTileThreads;
        if (s < wk.y) {
            x[q] = ld_stream(P.p[0] + s); y[q] = ld_stream(P.p[1] + s); z[q] = ld_stream(P.p[2] + s);
            const int gx = (int)fmin(fmax(rfloor(x[q] + x[q]), Real(0)), gxm), hy = (int)fmin(fmax(rfloor(y[q] + y[q]), Real(0)), hym);
            gx0 = min(gx0, gx); gx1 = max(gx1, gx); hy0 = min(hy0, hy); hy1 = max(hy1, hy);
        }
    }
    gx0 = __reduce_min_sync(0xffffffffu, gx0); gx1 = __reduce_max_sync(0xffffffffu, gx1);
    hy0 = __reduce_min_sync(0xffffffffu, hy0); hy1 = __reduce_max_sync(0xffffffffu, hy1);
    if (lane == 0) { red[0][wid] = gx0; red[1][wid] = gx1; red[2][wid] = hy0; red[3][wid] = hy1; }
    __syncthreads();   // also: the initialised mbarrier is visible to every thread
#pragma unroll
    for (int q = 0; q < kTileThreads / 32; ++q) { gx0 = min(gx0, red[0][q]); gx1 = max(gx1, red[1][q]); hy0 = min(hy0, red[2][q]); hy1 = max(hy1, red[3][q]); }
    // tile geometry: one half-cell of margin each way for the midpoint, clipped to the grid
    TileGeom<Real> T;
    const int ux1 = min(((gx1 + 1) >> 1) + 1, g.X - 1), vx1 = min((gx1 >> 1) + 1, g.X - 1);
    const int uy1 = min((hy1 >> 1) + 1, g.Y - 1), vy1 = min(((hy1 + 1) >> 1) + 1, g.Y - 1);
    T.ux0 = max((gx0 - 1) >> 1, 0); T.vx0 = max((gx0 - 2) >> 1, 0);
    T.uy0 = max((hy0 - 2) >> 1, 0); T.vy0 = max((hy0 - 1) >> 1, 0);
    const int unx = ux1 - T.ux0 + 1, vnx = vx1 - T.vx0 + 1, uny = uy1 - T.uy0 + 1, vny = vy1 - T.vy0 + 1;
    T.Z = g.Z; T.usy = unx * g.Z; T.vsy = vnx * g.Z;
    T.bv = unx * uny * g.Z; T.bw = T.bv + vnx * vny * g.Z;
    const long long elems = (long long)T.bw + (long long)vnx * uny * g.Z;
    const bool tile_ok = unx >= 2 && vnx >= 2 && uny >= 2 && vny >= 2 && elems * (long long)sizeof(Real) <= (long long)kTileSmemBytes;
    T.unx1 = tile_ok ? (unsigned int)(unx - 1) : 0u; T.uny1 = (unsigned int)(uny - 1); T.vnx1 = (unsigned int)(vnx - 1); T.vny1 = (unsigned int)(vny - 1);
    if (tile_ok) {
        if (wid == 0) {
            if (lane == 0) mbar_arrive_expect_tx(&bar, (unsigned int)(elems * sizeof(Real)));
            __syncwarp();
            // one bulk copy per (component, y): an x-run of whole z-rows is contiguous in memory
            for (int c = lane; c < uny + vny + uny; c += 32) {
                const Real* src; Real* dst; int n;
                if (c < uny) { src = u + ((long long)(T.uy0 + c) * g.X + T.ux0) * g.Z; dst = tile + c * T.usy; n = T.usy; }
                else if (c < uny + vny) { const int a = c - uny; src = v + ((long long)(T.vy0 + a) * g.X + T.vx0) * g.Z; dst = tile + T.bv + a * T.vsy; n = T.vsy; }
                else { const int a = c - uny - vny; src = w + ((long long)(T.uy0 + a) * g.X + T.vx0) * g.Z; dst = tile + T.bw + a * T.vsy; n = T.vsy; }
                tma_load_1d(dst, src, (unsigned int)(n * sizeof(Real)), &bar);
            }
        }
        mbar_wait(&bar, 0);
    }
#pragma unroll
    for (int q = 0; q < kTileKP; ++q) {
        const unsigned int s = wk.x + tid + q * kTileThreads;
        if (s >= wk.y) continue;
        Real px = x[q], py = y[q], pz = z[q];
        Real sx, sy, sz;
        if (!get_velocity3_tile<Real, FAST>(tile, T, px, py, pz, sx, sy, sz)) get_velocity3<Real, FAST, true>(u, v, w, g, C, px, py, pz, sx, sy, sz);
        Real vx = C.rho * sx, vy = C.rho * sy, vz = C.rho * sz;                       // equation 9
        const Real mx = px + (vx * dt * Real(.5)), my = py + (vy * dt * Real(.5)), mz = pz + (vz * dt * Real(.5));  // equation 7
        if (!get_velocity3_tile<Real, FAST>(tile, T, mx, my, mz, sx, sy, sz)) get_velocity3<Real, FAST, true>(u, v, w, g, C, mx, my, mz, sx, sy, sz);
        const Real wx = C.rho * sx, wy = C.rho * sy, wz = C.rho * sz;
        px = mx + (dt * wx); py = my + (dt * wy); pz = mz + (dt * wz);              // equation 8 (from the MIDPOINT, Q3)
        if (px < Real(0) || px > C.bx) { vx *= Real(-1.f); px = rmin(rmax(px, C.lo), C.cx); }
        if (py < Real(0) || py > C.by) { vy *= Real(-1.f); py = rmin(rmax(py, C.lo), C.cy); }
        if (pz < Real(0) || pz > C.bz) { vz *= Real(-1.f); pz = rmin(rmax(pz, C.lo), C.cz); }
        st_stream(P.p[0] + s, px); st_stream(P.p[1] + s, py); st_stream(P.p[2] + s, pz);
        st_stream(P.v[0] + s, vx); st_stream(P.v[1] + s, vy); st_stream(P.v[2] + s, vz);
        const int ci = (int)rfloor(px), cj = (int)rfloor(py), ck = (int)rfloor(pz);   // launcher guarantees X*Y*Z < 2^31
        if ((unsigned)ci < (unsigned)g.X && (unsigned)cj < (unsigned)g.Y && (unsigned)ck < (unsigned)g.Z) seed[(cj * g.X + ci) * g.Z + ck] = 1;
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
// Particle seeding ON THE DEVICE.  The reference seeds on the host with FLOAT loop counters and glibc's
// sinf / cosf / rand(); what is sequential or libm-dependent (the accumulated loop values, their sines and cosines, the raw
// rand() stream) comes in as small tables, every per-particle operation happens here, in the reference's float / double
// expression order (no FMA contraction in this translation unit), so the arrays equal the compiled reference's bit for bit.
// ------------------------------------------------------------------------------------------------
// emitParticleHemisphere for all radii (src/inkSolver.cpp:11-42 called from src/system.cpp:18-23): particle
// n = (r * ni + i) * nj + j.  The rotation AngleAxisf(pi, (n + up).normalized()) is about the ZERO axis and degenerates to
// cos(pi) * I: velocity = c * (x, y, z) with the zero terms of the 3x3 product kept (sign of zero), position =
// ceiling centre + radius * velocity.  Writes AoS xyz of particles [first, first + count).
template <class Real>
__global__ void __launch_bounds__(256) k_seed_emitter(long long first, long long count, int ni, int nj, const float* __restrict__ radius,
                                                      const float* __restrict__ sin_i, const float* __restrict__ cos_i, const float* __restrict__ sin_j,
                                                      const float* __restrict__ cos_j, float c, float cx, float cy, float cz, Real* __restrict__ pos,
                                                      Real* __restrict__ vel) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const long long n = first + t;
    const int j = (int)(n % nj);
    const long long q = n / nj;
    const int i = (int)(q % ni), r = (int)(q / ni);
    const float rad = radius[r];
    const float x = rad * sin_j[j] * sin_i[i];
    const float y = rad * cos_j[j];
    const float z = rad * sin_j[j] * cos_i[i];
    const float zf = 0.f;
    const float vx = (c * x + zf * y) + zf * z, vy = (zf * x + c * y) + zf * z, vz = (zf * x + zf * y) + c * z;
    pos[3 * t] = (Real)(cx + rad * vx); pos[3 * t + 1] = (Real)(cy + rad * vy); pos[3 * t + 2] = (Real)(cz + rad * vz);
    vel[3 * t] = (Real)vx; vel[3 * t + 1] = (Real)vy; vel[3 * t + 2] = (Real)vz;
}
// initParticles (src/system.cpp:62-64, 93-115): position = float(lo + (double(rand()) / RAND_MAX) * (hi - lo)) per axis, draw
// order x, y, z; rnd holds the caller's rand() stream, 3 draws per particle.
template <class Real>
__global__ void __launch_bounds__(256) k_seed_box(long long count, const int32_t* __restrict__ rnd, double lox, double hix, double loy, double hiy,
                                                  double loz, double hiz, Real* __restrict__ pos, Real* __restrict__ vel) {
    INK_PDL();
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const double rmax = 2147483647.0;   // RAND_MAX of glibc
    const float x = (float)(lox + ((double)rnd[3 * t] / rmax) * (hix - lox));
    const float y = (float)(loy + ((double)rnd[3 * t + 1] / rmax) * (hiy - loy));
    const float z = (float)(loz + ((double)rnd[3 * t + 2] / rmax) * (hiz - loz));
    pos[3 * t] = (Real)x; pos[3 * t + 1] = (Real)y; pos[3 * t + 2] = (Real)z;
    vel[3 * t] = Real(0); vel[3 * t + 1] = Real(0); vel[3 * t + 2] = Real(0);
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
