// mg_tri.cuh — second-generation multigrid preconditioner for the matrix-free CG pressure solve (cg.cuh):
// V(1,1) cycle with cell-centred TRILINEAR prolongation P and full-weighting restriction R = P^T.
//
// Why (scripts/mg_proto.py, numpy): the piecewise-constant transfer of mg.cuh needs V(2,2) to reach ~13-16 CG
// iterations; trilinear transfer reaches 10 with V(1,1) (8 with V(2,2)) at 64x128x64, size-independent, and V(1,1)
// needs far fewer passes over a level:
//      down : res = r - s M (c r) = r - (omega/6) M r                      (the sweep from zero is never stored)
//             r_{l+1} = P^T res
//      up   : x' = c r + P e ;  z = x' + c (r - s M x') = c r + (1-omega) x' + (omega/6) N x'
//                                                                         (x' is never stored: ONE pass r, e -> z)
// with M x = 6x - N x, N = sum of the in-grid (in-slab) face neighbours, c = omega / (6 s_l), s_l = 2^l.
// Fine-level traffic per CG iteration: 8 (res) + 4.5 (restrict) + 8.5 (up) = 21 B/cell instead of 46 B/cell.
//
// 1-D transfer (tensor product in 3-D): fine cell f has parent c = f >> 1 and second donor o = c + (f odd ? +1 : -1);
// P[f, c] = 3/4, P[f, o] = 1/4, donors outside the coarse grid (or slab) are dropped (zero Dirichlet ghost, like
// the ghost pressure of the reference's matrix, src/waterSolver.cpp:240-261).  The cycle is a fixed symmetric
// positive-definite operator: R = P^T exactly (same masks), Jacobi is self-adjoint, coarsest level = a fixed
// polynomial in the level operator (kMgCoarsestSweeps damped-Jacobi sweeps from zero).
//
// The level operators are (thread id, thread count)-strided INK_HD functions: the global-memory kernels, the
// shared-memory tail and the HOST harness (tests/mg_host_harness.cu, checked against the numpy prototype without a
// GPU) all execute the same code.
#pragma once
#include "mg.cuh"
#include "mg_vec.cuh"

namespace ink {

constexpr int kMg3CoarsestCells = 2048;   // stop coarsening at <= this many cells (prototype: same iteration count as 16); 2048 so that
                                          // a 1024-cell level widened by the multi-GPU overlap rows (1088 cells) is still the coarsest

// res = r - (omega/6) M r
template <class Real>
INK_HD void mg3_residual(const MgDims& d, const Real* __restrict__ r, Real* __restrict__ res, long long tid, long long nth) {
    const Real w6 = Real(kMgOmega) / Real(6);
    INK_MG_CELL_LOOP(d, tid, nth) {
        INK_MG_DECODE(d)
        const Real rc = r[c];
        res[c] = rc - w6 * (Real(6) * rc - mg_nbr_sum(d, r, i, j, k, c));
    }
}

// rc = P^T res : coarse cell (I,J,K) gathers the 4x4x4 fine cells 2I-1..2I+2 (weights 1/4, 3/4, 3/4, 1/4 per axis)
template <class Real>
INK_HD void mg3_restrict(const MgDims& df, const MgDims& dc, const Real* __restrict__ res, Real* __restrict__ rc, long long tid, long long nth) {
    INK_MG_CELL_LOOP(dc, tid, nth) {
        INK_MG_DECODE(dc)
        Real acc = Real(0);
#pragma unroll
        for (int n = 0; n < 4; ++n) {
            const int fj = 2 * j - 1 + n;
            if (fj < df.j0 || fj >= df.j1) continue;
            const Real wy = (n == 1 || n == 2) ? Real(0.75) : Real(0.25);
#pragma unroll
            for (int l = 0; l < 4; ++l) {
                const int fi = 2 * i - 1 + l;
                if (fi < 0 || fi >= df.X) continue;
                const Real wxy = wy * ((l == 1 || l == 2) ? Real(0.75) : Real(0.25));
                const long long row = mg_idx(df, fi, fj, 0);
                Real a = Real(0);
#pragma unroll
                for (int m = 0; m < 4; ++m) {
                    const int fk = 2 * k - 1 + m;
                    if (fk < 0 || fk >= df.Z) continue;
                    a += ((m == 1 || m == 2) ? Real(0.75) : Real(0.25)) * res[row + fk];
                }
                acc += wxy * a;
            }
        }
        rc[c] = acc;
    }
}

// ---- down pass in ONE operator: rc = P^T (r - (omega/6) M r) without storing the residual ------------------------------
// With res = (1-omega) r + (omega/6) N r (N = sum of the in-grid face neighbours) and the separable restriction weights
// a = (0, 1/4, 3/4, 3/4, 1/4, 0) on the fine positions 2I-2 .. 2I+3 of each axis (zero where the fine cell does not exist),
//     rc[I,J,K] = sum_n r_n [ (1-omega) ax ay az + (omega/6) (bx ay az + ax by az + ax ay bz) ],   b[p] = a[p-1] + a[p+1]
// (b zero where the fine cell n does not exist: N only sums existing neighbours).  Per (x, y) column of the 6x6 window two
// z-contractions (with az over 4 cells, with bz over 6) — 160 loads per coarse cell instead of a stored residual and a second
// kernel.  Used on the levels below the finest, where a kernel costs its launch latency, not its traffic.
template <class Real>
INK_HD void mg3_axis_weights(int f0, int lo, int hi, Real (&a)[6], Real (&b)[6]) {   // f0 = fine index of position 0 (= 2 I - 2)
    const Real w[6] = {Real(0), Real(0.25), Real(0.75), Real(0.75), Real(0.25), Real(0)};
    Real am[8];
    am[0] = Real(0); am[7] = Real(0);
#pragma unroll
    for (int p = 0; p < 6; ++p) { const int f = f0 + p; a[p] = (f >= lo && f < hi) ? w[p] : Real(0); am[p + 1] = a[p]; }
#pragma unroll
    for (int p = 0; p < 6; ++p) { const int f = f0 + p; b[p] = (f >= lo && f < hi) ? am[p] + am[p + 2] : Real(0); }
}
template <class Real>
INK_HD void mg3_down_fused(const MgDims& df, const MgDims& dc, const Real* __restrict__ r, Real* __restrict__ rc, long long tid, long long nth) {
    const Real c0 = Real(1) - Real(kMgOmega), c1 = Real(kMgOmega) / Real(6);
    INK_MG_CELL_LOOP(dc, tid, nth) {
        INK_MG_DECODE(dc)
        Real ax[6], bx[6], ay[6], by[6], az[6], bz[6];
        mg3_axis_weights<Real>(2 * i - 2, 0, df.X, ax, bx);
        mg3_axis_weights<Real>(2 * j - 2, df.j0, df.j1, ay, by);
        mg3_axis_weights<Real>(2 * k - 2, 0, df.Z, az, bz);
        const int fk0 = 2 * k - 2;
        Real acc = Real(0);
#pragma unroll
        for (int py = 0; py < 6; ++py)
#pragma unroll
            for (int px = 0; px < 6; ++px) {
                const bool ox = (px == 0 || px == 5), oy = (py == 0 || py == 5);
                if (ox && oy) continue;                                   // both weights vanish identically
                const Real wa = c0 * (ax[px] * ay[py]) + c1 * (bx[px] * ay[py] + ax[px] * by[py]);
                const Real wb = c1 * (ax[px] * ay[py]);
                if (wa == Real(0) && wb == Real(0)) continue;             // column outside the grid / slab
                const Real* __restrict__ row = r + mg_idx(df, 2 * i - 2 + px, 2 * j - 2 + py, 0);
                Real v[6];
#pragma unroll
                for (int pz = 0; pz < 6; ++pz) {
                    const int fk = fk0 + pz;
                    const bool need = (pz >= 1 && pz <= 4) || !(ox || oy);   // the outer z positions only enter through bz
                    v[pz] = (need && fk >= 0 && fk < df.Z) ? row[fk] : Real(0);
                }
                const Real sa = (az[1] * v[1] + az[2] * v[2]) + (az[3] * v[3] + az[4] * v[4]);
                acc += wa * sa;
                if (!(ox || oy)) {
                    const Real sb = ((bz[0] * v[0] + bz[1] * v[1]) + (bz[2] * v[2] + bz[3] * v[3])) + (bz[4] * v[4] + bz[5] * v[5]);
                    acc += wb * sb;
                }
            }
        rc[c] = acc;
    }
}

// the four fine positions 2c-1, 2c, 2c+1, 2c+2 interpolated from coarse c-1, c, c+1
template <class Real>
INK_HD void mg3_interp4(Real v0, Real v1, Real v2, Real (&o)[4]) {
    o[0] = Real(0.75) * v0 + Real(0.25) * v1;
    o[1] = Real(0.75) * v1 + Real(0.25) * v0;
    o[2] = Real(0.75) * v1 + Real(0.25) * v2;
    o[3] = Real(0.75) * v2 + Real(0.25) * v1;
}

INK_HD float hd_fma(float a, float b, float c) { return fmaf(a, b, c); }
INK_HD double hd_fma(double a, double b, double c) { return fma(a, b, c); }
// same, unscaled (x4): 3 * main donor + second donor, one FMA each
template <class Real>
INK_HD void mg3_interp4u(Real v0, Real v1, Real v2, Real (&o)[4]) {
    o[0] = hd_fma(Real(3), v0, v1);
    o[1] = hd_fma(Real(3), v1, v0);
    o[2] = hd_fma(Real(3), v1, v2);
    o[3] = hd_fma(Real(3), v2, v1);
}

// mg3_up for ONE coarse cell whose 3x3 coarse columns and 4x4 fine columns all exist (x, y interior; warp-uniform when a
// warp walks z) on a level with even Z and 2*sizeof(Real)-aligned arrays: no x/y predicates, the z ends are handled by
// two flags, the fine pairs (2K, 2K+1) move as one 64/128-bit access, interpolation is 1 FMA per value (x64 scaling
// folded into the x' FMA).  Same operator as the generic path (different rounding order).
template <class Real>
INK_HD double mg3_up_cell_fast(const MgDims& df, const MgDims& dc, Real cj, const Real* __restrict__ r, const Real* __restrict__ e, Real* __restrict__ z,
                               int I, int J, int K, int dj0, int dj1) {
    const Real a0 = Real(1) - Real(kMgOmega), a1 = Real(kMgOmega) / Real(6), sc = Real(1) / Real(64);
    const bool zlo = K > 0, zhi = K + 1 < dc.Z;
    // 32-bit element offsets (a level has < 2^31 cells): one IMAD.WIDE per access instead of 64-bit index arithmetic,
    // which was 40 % of this kernel's instructions (profiles/r1_c_mg3_up.txt)
    const int sxc = dc.Z, syc = dc.X * dc.Z, sxf = df.Z, syf = df.X * df.Z;
    const Real* __restrict__ eb = e + (((J - 1) * dc.X + (I - 1)) * dc.Z + K);
    Real tz[3][3][4];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int b = 0; b < 3; ++b) {
            const Real* __restrict__ p = eb + a * sxc + b * syc;
            const Real e1 = p[0], e0 = zlo ? p[-1] : Real(0), e2 = zhi ? p[1] : Real(0);
            mg3_interp4u(e0, e1, e2, tz[a][b]);
        }
    Real ty[3][4][4];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int m = 0; m < 4; ++m) {
            Real o[4];
            mg3_interp4u(tz[a][0][m], tz[a][1][m], tz[a][2][m], o);
#pragma unroll
            for (int n = 0; n < 4; ++n) ty[a][n][m] = o[n];
        }
    Real xp[4][4][4], rin[2][2][2];
    const int fb = ((2 * J - 1) * df.X + (2 * I - 1)) * df.Z + 2 * K;
#pragma unroll
    for (int n = 0; n < 4; ++n)
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const bool lo = (l == 0 || l == 3), no = (n == 0 || n == 3);
            if (lo && no) continue;
            const Real* __restrict__ p = r + fb + l * sxf + n * syf;
            const Pack<Real, 2> q = ldp<Real, 2>(p);
            Real pe[4];
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                Real o[4];
                mg3_interp4u(ty[0][n][m], ty[1][n][m], ty[2][n][m], o);
                pe[m] = o[l];
            }
            xp[l][n][1] = hd_fma(sc, pe[1], cj * q.a[0]);
            xp[l][n][2] = hd_fma(sc, pe[2], cj * q.a[1]);
            if (!lo && !no) {
                rin[l - 1][n - 1][0] = q.a[0]; rin[l - 1][n - 1][1] = q.a[1];
                xp[l][n][0] = zlo ? hd_fma(sc, pe[0], cj * p[-1]) : Real(0);
                xp[l][n][3] = zhi ? hd_fma(sc, pe[3], cj * p[2]) : Real(0);
            }
        }
    double acc = 0;
#pragma unroll
    for (int n = 1; n < 3; ++n) {
        const int fj = 2 * J - 1 + n;
        const bool in_dot = (fj >= dj0) && (fj < dj1);
        Real dsum = Real(0);   // the 4 products of one fine row in working precision, then one fp64 add
#pragma unroll
        for (int l = 1; l < 3; ++l) {
            Pack<Real, 2> o;
#pragma unroll
            for (int m = 1; m < 3; ++m) {
                const Real nb = (xp[l - 1][n][m] + xp[l + 1][n][m]) + (xp[l][n - 1][m] + xp[l][n + 1][m]) + (xp[l][n][m - 1] + xp[l][n][m + 1]);
                const Real rr = rin[l - 1][n - 1][m - 1];
                const Real v = hd_fma(a1, nb, hd_fma(a0, xp[l][n][m], cj * rr));
                o.a[m - 1] = v;
                dsum = hd_fma(rr, v, dsum);
            }
            stp<Real, 2>(z + fb + l * sxf + n * syf, o);
        }
        if (in_dot) acc += (double)dsum;
    }
    return acc;
}

// z = c r + (1-omega) x' + (omega/6) N x',  x' = c r + P e.  One thread per COARSE cell (I,J,K) = the 2x2x2 fine block
// (2I+a, 2J+b, 2K+q): the 27 coarse values around it give P e on the 4x4x4 fine cells 2I-1..2I+2 (separable
// interpolation in registers); x' is formed there (0 outside the grid / slab, which is exactly "N sums in-grid neighbours
// only") and the 8 outputs follow.  Per fine cell: 4 loads of r + 3.4 loads of e instead of 7 x (1 + 8).
// Returns this thread's partial of <r, z> over rows [dj0, dj1).
template <class Real>
INK_HD double mg3_up(const MgDims& df, const MgDims& dc, Real s, const Real* __restrict__ r, const Real* __restrict__ e, Real* __restrict__ z,
                     long long tid, long long nth, int dj0 = 0, int dj1 = 0x7fffffff) {
    const Real cj = Real(kMgOmega) / (Real(6) * s), a0 = Real(1) - Real(kMgOmega), a1 = Real(kMgOmega) / Real(6);
    double acc = 0;
    const bool vec_ok = ((df.Z & 1) == 0) && ((((unsigned long long)r | (unsigned long long)z) & (2 * sizeof(Real) - 1)) == 0);
    INK_MG_CELL_LOOP(dc, tid, nth) {
        const int I = it_.i, K = it_.k, J = dc.j0 + it_.jl;
        if (vec_ok && I >= 1 && I + 1 < dc.X && 2 * I + 2 < df.X && J - 1 >= dc.j0 && J + 1 < dc.j1 && 2 * J - 1 >= df.j0 && 2 * J + 2 < df.j1) {
            acc += mg3_up_cell_fast<Real>(df, dc, cj, r, e, z, I, J, K, dj0, dj1);
            continue;
        }
        // P e on the 4x4x4 block: z, then y, then x
        Real tz[3][3][4];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int ii = I + a - 1;
            const bool vx = (ii >= 0) && (ii < dc.X);
#pragma unroll
            for (int b = 0; b < 3; ++b) {
                const int jj = J + b - 1;
                const bool vxy = vx && (jj >= dc.j0) && (jj < dc.j1);
                Real ev[3];
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const int kk = K + q - 1;
                    ev[q] = (vxy && kk >= 0 && kk < dc.Z) ? e[mg_idx(dc, ii, jj, kk)] : Real(0);
                }
                mg3_interp4(ev[0], ev[1], ev[2], tz[a][b]);
            }
        }
        Real ty[3][4][4];   // [a][n][m]
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                Real o[4];
                mg3_interp4(tz[a][0][m], tz[a][1][m], tz[a][2][m], o);
#pragma unroll
                for (int n = 0; n < 4; ++n) ty[a][n][m] = o[n];
            }
        Real xp[4][4][4];   // [l][n][m] <-> fine (2I-1+l, 2J-1+n, 2K-1+m)
        Real rin[2][2][2];
#pragma unroll
        for (int n = 0; n < 4; ++n)
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                Real o[4];
                mg3_interp4(ty[0][n][m], ty[1][n][m], ty[2][n][m], o);
#pragma unroll
                for (int l = 0; l < 4; ++l) xp[l][n][m] = o[l];
            }
        // x' = c r + P e where the fine cell exists, else 0 (only the 8 block cells and their 24 face neighbours are used)
#pragma unroll
        for (int l = 0; l < 4; ++l) {
            const int fi = 2 * I - 1 + l;
            const bool vx = (fi >= 0) && (fi < df.X);
#pragma unroll
            for (int n = 0; n < 4; ++n) {
                const int fj = 2 * J - 1 + n;
                const bool vxy = vx && (fj >= df.j0) && (fj < df.j1);
#pragma unroll
                for (int m = 0; m < 4; ++m) {
                    const int outer = (l == 0 || l == 3) + (n == 0 || n == 3) + (m == 0 || m == 3);
                    if (outer >= 2) continue;   // edges / corners of the 4x4x4 block are not face neighbours of the 2x2x2 block
                    const int fk = 2 * K - 1 + m;
                    Real v = Real(0);
                    if (vxy && fk >= 0 && fk < df.Z) {
                        const Real rr = r[mg_idx(df, fi, fj, fk)];
                        v = cj * rr + xp[l][n][m];
                        if (outer == 0) rin[l - 1][n - 1][m - 1] = rr;
                    }
                    xp[l][n][m] = v;
                }
            }
        }
#pragma unroll
        for (int l = 1; l < 3; ++l) {
            const int fi = 2 * I - 1 + l;
            if (fi >= df.X) continue;
#pragma unroll
            for (int n = 1; n < 3; ++n) {
                const int fj = 2 * J - 1 + n;
                if (fj < df.j0 || fj >= df.j1) continue;
#pragma unroll
                for (int m = 1; m < 3; ++m) {
                    const int fk = 2 * K - 1 + m;
                    if (fk >= df.Z) continue;
                    const Real nb = (xp[l - 1][n][m] + xp[l + 1][n][m]) + (xp[l][n - 1][m] + xp[l][n + 1][m]) + (xp[l][n][m - 1] + xp[l][n][m + 1]);
                    const Real rr = rin[l - 1][n - 1][m - 1];
                    const Real v = cj * rr + a0 * xp[l][n][m] + a1 * nb;
                    z[mg_idx(df, fi, fj, fk)] = v;
                    if (fj >= dj0 && fj < dj1) acc += (double)rr * (double)v;
                }
            }
        }
    }
    return acc;
}

#ifdef __CUDACC__
// ---- global-memory kernels ---------------------------------------------------------------------------------------
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_residual(const CgScal* S, MgDims d, const Real* __restrict__ r, Real* __restrict__ res) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    mg3_residual<Real>(d, r, res, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}
// Z % 4 == 0: one thread per 4 z-consecutive cells, 128-bit accesses (same arithmetic per cell)
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_residual_v4(const CgScal* S, MgDims d, const Real* __restrict__ r, Real* __restrict__ res) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    const Real w6 = Real(kMgOmega) / Real(6);
    int i, j, k; long long c;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; mg_vec_cell(d, t, i, j, k, c); t += gridDim.x * blockDim.x) {
        const P4<Real> rc = ldp<Real, 4>(r + c);
        Real L[4];
        mg_lap4<Real>(d, r, rc, i, j, k, c, L);
        P4<Real> o;
#pragma unroll
        for (int a = 0; a < 4; ++a) o.a[a] = rc.a[a] - w6 * L[a];
        stp<Real, 4>(res + c, o);
    }
}
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_restrict(const CgScal* S, MgDims df, MgDims dc, const Real* __restrict__ res, Real* __restrict__ rc) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    mg3_restrict<Real>(df, dc, res, rc, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}
// FINAL: level 0 — also produces <r, z> for CG (rz_old <- rz, rz <- <r,z>)
template <class Real, bool FINAL>
__global__ void __launch_bounds__(256, 3) k_mg3_up(CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r, const Real* __restrict__ e,
                                                Real* __restrict__ z, int dj0, int dj1) {
    INK_PDL();
    if (S != nullptr && S->done) return;   // S may be nullptr for the non-FINAL levels (INK_MG_NODONE)
    const double acc = mg3_up<Real>(df, dc, s, r, e, z, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x, dj0, dj1);
    if (FINAL) {
        double tot;
        if (block_sum_last(acc, S, 0, tot)) {
            if (threadIdx.x == 0) { if (S->dist) S->loc[2] = tot; else { S->rz_old = S->rz; S->rz = tot; } }
        }
    }
}

// ---- up pass through a shared-memory tile (the finest levels) ----------------------------------------------------------------
// k_mg3_up gives one thread a coarse cell and lets it form x' = c r + P e on the whole 4x4x4 fine neighbourhood of its 2x2x2
// block: every x' is computed 4 times, ~680 warp instructions per coarse-cell warp, 80 registers, 39 % issue utilisation — the
// largest single kernel of a PCG iteration (33 us of ~110 at 128x256x128, 92 us of 340 at 256^3: 1.2-1.5 TB/s for 8.5 B/cell).
// Here a block owns a kUpTX x kUpTY x kUpTZ fine tile:
//   stage  the (TX/2+2)(TY/2+2)(TZ/2+2) coarse values around it in shared memory (zero outside the coarse grid / slab),
//   A      x' = c r + P e for the tile plus a one-cell halo, four z-consecutive cells per thread (separable: 28 FMAs per
//          quad from 16 shared coarse values, one LDG.128 of r), into shared memory — each x' once (1.6x with the halo),
//   B      z = c r + (1-omega) x' + (omega/6) N x' from shared memory (5 LDS.128 + 2 LDS.32 per quad), one STG.128, <r,z>.
// (Measured and rejected, profiles/r2_f_*: a thread layout that does the row arithmetic once per two quads — 8 z-lanes x 32 row-lanes —
// is slower, 30.6 / 90 us against 26.7 / 75.8 us: four rows per warp instruction instead of two.)
// ~38 instructions per fine cell instead of ~85.  Same operator as mg3_up (cells outside the grid / slab hold x' = 0, which is
// "N sums existing neighbours only"); different rounding order.  Z % 4 == 0 (vector accesses); any X, Y, slab start.
constexpr int kUpTX = 8, kUpTY = 8, kUpTZ = 64;   // measured: 8x8x32 tiles (8 blocks per SM, no wave tail) are SLOWER — 32.8 / 104.6 us against 26.7 / 75.8 us
                                                 // at 128x256x128 / 256^3 (profiles/r2_f_*): the halo and the staging per tile cost more than the tail
constexpr int kUpRows = (kUpTX + 2) * (kUpTY + 2);       // fine rows of the tile + halo
constexpr int kUpRow = kUpTZ + 8;                        // [3 pad][k = -1][TZ cells][k = TZ][3 pad]: cell k at 4 + k, 16-byte aligned quads
constexpr int kUpCX = kUpTX / 2 + 2, kUpCY = kUpTY / 2 + 2, kUpCZ = kUpTZ / 2 + 2;
constexpr int kUpCRow = kUpCZ + 2;                       // even row length: 8-byte aligned pairs
template <class Real>
constexpr size_t mg3_up_tile_smem() { return (size_t)(kUpRows * kUpRow + kUpCX * kUpCY * kUpCRow) * sizeof(Real); }

// (3 * main + other) donors of fine index f on one axis, as local indices into the staged coarse tile whose origin is c0
__device__ __forceinline__ void mg3_donors(int f, int c0, int& pm, int& po) {
    const int c = f >> 1;
    pm = c - c0;
    po = pm + ((f & 1) ? 1 : -1);
}

template <class Real, bool FINAL>
__global__ void __launch_bounds__(256) k_mg3_up_tile(CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r, const Real* __restrict__ e,
                                                     Real* __restrict__ z, int dj0, int dj1, int ntx, int nty, int ntz) {
    INK_PDL();
    if (S != nullptr && S->done) return;   // S may be nullptr for the non-FINAL levels (INK_MG_NODONE)
    extern __shared__ __align__(16) unsigned char up_smem_raw[];
    Real* xs = reinterpret_cast<Real*>(up_smem_raw);            // x' rows
    Real* es = xs + kUpRows * kUpRow;                           // coarse values
    const Real cj = Real(kMgOmega) / (Real(6) * s), a0 = Real(1) - Real(kMgOmega), a1 = Real(kMgOmega) / Real(6), sc = Real(1) / Real(64);
    const int tid = threadIdx.x;
    const int jbase = df.j0 & ~1;                               // tiles start on even rows so that a tile's coarse origin is fy0 / 2 - 1
    const int sxf = df.Z, syf = df.X * df.Z;                    // 32-bit element strides (a level has < 2^31 cells)
    double acc = 0;
    for (int tile = blockIdx.x; tile < ntx * nty * ntz; tile += gridDim.x) {
        const int tz_ = tile % ntz, txy = tile / ntz, tx_ = txy % ntx, ty_ = txy / ntx;
        const int fx0 = tx_ * kUpTX, fy0 = jbase + ty_ * kUpTY, fz0 = tz_ * kUpTZ;
        const int cx0 = (fx0 >> 1) - 1, cy0 = (fy0 >> 1) - 1, cz0 = (fz0 >> 1) - 1;
        // ---- stage the coarse values (zero outside the coarse grid / slab: dropped donors)
        for (int col = tid >> 5; col < kUpCX * kUpCY; col += 8) {   // one warp per coarse column: validity and base address once per column
            const int ci = cx0 + col % kUpCX, cjj = cy0 + col / kUpCX;
            const bool okc = ci >= 0 && ci < dc.X && cjj >= dc.j0 && cjj < dc.j1;
            const Real* src = e + ((long long)cjj * dc.X + ci) * dc.Z + cz0;
            for (int kk = tid & 31; kk < kUpCZ; kk += 32) {
                const int ck = cz0 + kk;
                es[col * kUpCRow + kk] = (okc && ck >= 0 && ck < dc.Z) ? src[kk] : Real(0);
            }
        }
        __syncthreads();
        // ---- A: x' on the tile + halo rows, one z-quad per thread and pass
        for (int q = tid; q < kUpRows * (kUpTZ / 4); q += 256) {
            const int zq = q % (kUpTZ / 4), row = q / (kUpTZ / 4), lx = row % (kUpTX + 2), ly = row / (kUpTX + 2);
            if ((lx == 0 || lx == kUpTX + 1) && (ly == 0 || ly == kUpTY + 1)) continue;   // corner rows are nobody's face neighbours
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = fz0 + 4 * zq;
            P4<Real> xp = p4_zero<Real>();
            if (fi >= 0 && fi < df.X && fj >= df.j0 && fj < df.j1 && fk < df.Z) {
                const P4<Real> rq = ldp<Real, 4>(r + (fj * syf + fi * sxf + fk));
                int xm, xo, ym, yo;
                mg3_donors(fi, cx0, xm, xo);
                mg3_donors(fj, cy0, ym, yo);
                const int lk = 2 * zq;                           // local coarse index of K - 1, K = fk / 2
                // x-y combination FIRST, on the four coarse z values (9 mm + 3 mo + 3 om + oo per value: 12 FMAs), then the z
                // interpolation of the four fine cells (4 FMAs): 16 FMAs per quad instead of 28 for z-first
                Real g[4];
                {
                    const Real* cmm = es + (ym * kUpCX + xm) * kUpCRow + lk;
                    const Real* cmo = es + (yo * kUpCX + xm) * kUpCRow + lk;   // main x, other y
                    const Real* com = es + (ym * kUpCX + xo) * kUpCRow + lk;
                    const Real* coo = es + (yo * kUpCX + xo) * kUpCRow + lk;
                    const Pack<Real, 2> a01 = ldp<Real, 2>(cmm), a23 = ldp<Real, 2>(cmm + 2), b01 = ldp<Real, 2>(cmo), b23 = ldp<Real, 2>(cmo + 2);
                    const Pack<Real, 2> c01 = ldp<Real, 2>(com), c23 = ldp<Real, 2>(com + 2), d01 = ldp<Real, 2>(coo), d23 = ldp<Real, 2>(coo + 2);
                    g[0] = hd_fma(Real(3), hd_fma(Real(3), a01.a[0], b01.a[0]), hd_fma(Real(3), c01.a[0], d01.a[0]));
                    g[1] = hd_fma(Real(3), hd_fma(Real(3), a01.a[1], b01.a[1]), hd_fma(Real(3), c01.a[1], d01.a[1]));
                    g[2] = hd_fma(Real(3), hd_fma(Real(3), a23.a[0], b23.a[0]), hd_fma(Real(3), c23.a[0], d23.a[0]));
                    g[3] = hd_fma(Real(3), hd_fma(Real(3), a23.a[1], b23.a[1]), hd_fma(Real(3), c23.a[1], d23.a[1]));
                }
                xp.a[0] = hd_fma(sc, hd_fma(Real(3), g[1], g[0]), cj * rq.a[0]);
                xp.a[1] = hd_fma(sc, hd_fma(Real(3), g[1], g[2]), cj * rq.a[1]);
                xp.a[2] = hd_fma(sc, hd_fma(Real(3), g[2], g[1]), cj * rq.a[2]);
                xp.a[3] = hd_fma(sc, hd_fma(Real(3), g[2], g[3]), cj * rq.a[3]);
            }
            stp<Real, 4>(xs + row * kUpRow + 4 + 4 * zq, xp);
        }
        // the two z-halo cells (k = fz0 - 1, fz0 + TZ) of the INTERIOR rows
        for (int q = tid; q < kUpTX * kUpTY * 2; q += 256) {
            const int side = q & 1, col = q >> 1, lx = 1 + col % kUpTX, ly = 1 + col / kUpTX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = side ? fz0 + kUpTZ : fz0 - 1;
            Real v = Real(0);
            if (fi < df.X && fj >= df.j0 && fj < df.j1 && fk >= 0 && fk < df.Z) {
                int xm, xo, ym, yo, zm, zo;
                mg3_donors(fi, cx0, xm, xo);
                mg3_donors(fj, cy0, ym, yo);
                mg3_donors(fk, cz0, zm, zo);
                Real u2[2];
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    const Real* c0 = es + (ym * kUpCX + (a ? xo : xm)) * kUpCRow;
                    const Real* c1 = es + (yo * kUpCX + (a ? xo : xm)) * kUpCRow;
                    u2[a] = hd_fma(Real(3), hd_fma(Real(3), c0[zm], c0[zo]), hd_fma(Real(3), c1[zm], c1[zo]));
                }
                v = hd_fma(sc, hd_fma(Real(3), u2[0], u2[1]), cj * r[fj * syf + fi * sxf + fk]);
            }
            xs[(ly * (kUpTX + 2) + lx) * kUpRow + (side ? 4 + kUpTZ : 3)] = v;
        }
        __syncthreads();
        // ---- B: z = c r + (1 - omega) x' + (omega / 6) N x' on the tile
        for (int q = tid; q < kUpTX * kUpTY * (kUpTZ / 4); q += 256) {
            const int zq = q % (kUpTZ / 4), col = q / (kUpTZ / 4), lx = 1 + col % kUpTX, ly = 1 + col / kUpTX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = fz0 + 4 * zq;
            if (fi >= df.X || fj < df.j0 || fj >= df.j1 || fk >= df.Z) continue;
            const Real* c = xs + (ly * (kUpTX + 2) + lx) * kUpRow + 4 + 4 * zq;
            const P4<Real> xc = ldp<Real, 4>(c), xl = ldp<Real, 4>(c - kUpRow), xr = ldp<Real, 4>(c + kUpRow);
            const P4<Real> yl = ldp<Real, 4>(c - (kUpTX + 2) * kUpRow), yr = ldp<Real, 4>(c + (kUpTX + 2) * kUpRow);
            const Real zl = c[-1], zr = c[4];
            const int g = fj * syf + fi * sxf + fk;
            const P4<Real> rq = ldp<Real, 4>(r + g);
            P4<Real> o;
            Real dsum = Real(0);
#pragma unroll
            for (int m = 0; m < 4; ++m) {
                const Real lo = (m == 0) ? zl : xc.a[m > 0 ? m - 1 : 0], hi = (m == 3) ? zr : xc.a[m < 3 ? m + 1 : 3];
                const Real nb = (xl.a[m] + xr.a[m]) + (yl.a[m] + yr.a[m]) + (lo + hi);
                const Real v = hd_fma(a1, nb, hd_fma(a0, xc.a[m], cj * rq.a[m]));
                o.a[m] = v;
                dsum = hd_fma(rq.a[m], v, dsum);
            }
            stp<Real, 4>(z + g, o);
            if (FINAL && fj >= dj0 && fj < dj1) acc += (double)dsum;
        }
        __syncthreads();   // the next tile of this block reuses the shared arrays
    }
    if (FINAL) {
        double tot;
        if (block_sum_last(acc, S, 0, tot)) {
            if (threadIdx.x == 0) { if (S->dist) S->loc[2] = tot; else { S->rz_old = S->rz; S->rz = tot; } }
        }
    }
}

// ---- down pass through a shared-memory tile: residual and full-weighting restriction in ONE launch ----------------------------
// The kernel pair below (k_mg3_residual_rz_v4 + k_mg3_restrict_xy_v4) stores a z-restricted residual (2 B/cell) and reads it back;
// on the levels below the finest each of the two launches costs its latency (~4.5 us for < 1 us of work).  Here a block owns a
// kDnCX x kDnCY x kDnCZ COARSE tile: A) res = r - (omega/6) M r on the 2x fine tile plus its one-cell halo (zero where the fine
// cell does not exist: exactly the cells the restriction skips), four z-consecutive cells per thread with the vector stencil of
// mg_lap4, into shared memory; B) one thread per coarse cell gathers its 4x4x4 residuals with the separable weights (1/4, 3/4,
// 3/4, 1/4).  Same operator as the pair (and as mg3_residual + mg3_restrict), different rounding order.  df.Z % 4 == 0.
constexpr int kDnCX = 4, kDnCY = 4, kDnCZ = 16;
constexpr int kDnFX = 2 * kDnCX + 2, kDnFY = 2 * kDnCY + 2, kDnFZ = 2 * kDnCZ;   // fine rows (with halo) and fine z cells (without)
constexpr int kDnRow = kDnFZ + 8;                        // [3 pad][k = -1][FZ cells][k = FZ][3 pad]: cell k at 4 + k
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_down_tile(const CgScal* S, MgDims df, MgDims dc, const Real* __restrict__ r, Real* __restrict__ rc,
                                                       int ntx, int nty, int ntz) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    __shared__ __align__(16) Real rs[kDnFX * kDnFY * kDnRow];
    const Real w6 = Real(kMgOmega) / Real(6);
    const int tid = threadIdx.x;
    const int sxf = df.Z, syf = df.X * df.Z;
    for (int tile = blockIdx.x; tile < ntx * nty * ntz; tile += gridDim.x) {
        const int tz_ = tile % ntz, txy = tile / ntz, tx_ = txy % ntx, ty_ = txy / ntx;
        const int I0 = tx_ * kDnCX, J0 = dc.j0 + ty_ * kDnCY, K0 = tz_ * kDnCZ;
        const int fx0 = 2 * I0, fy0 = 2 * J0, fz0 = 2 * K0;
        // ---- A: residual on the fine tile + halo rows, z-quads
        for (int q = tid; q < kDnFX * kDnFY * (kDnFZ / 4); q += 256) {
            const int zq = q % (kDnFZ / 4), row = q / (kDnFZ / 4), lx = row % kDnFX, ly = row / kDnFX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = fz0 + 4 * zq;
            P4<Real> o = p4_zero<Real>();
            if (fi >= 0 && fi < df.X && fj >= df.j0 && fj < df.j1 && fk < df.Z) {
                const long long c = (long long)(fj * syf + fi * sxf + fk);
                const P4<Real> c4 = ldp<Real, 4>(r + c);
                Real L[4];
                mg_lap4<Real>(df, r, c4, fi, fj, fk, c, L);
#pragma unroll
                for (int a = 0; a < 4; ++a) o.a[a] = c4.a[a] - w6 * L[a];
            }
            stp<Real, 4>(rs + row * kDnRow + 4 + 4 * zq, o);
        }
        for (int q = tid; q < kDnFX * kDnFY * 2; q += 256) {   // the two z-halo cells of every row
            const int side = q & 1, row = q >> 1, lx = row % kDnFX, ly = row / kDnFX;
            const int fi = fx0 - 1 + lx, fj = fy0 - 1 + ly, fk = side ? fz0 + kDnFZ : fz0 - 1;
            Real v = Real(0);
            if (fi >= 0 && fi < df.X && fj >= df.j0 && fj < df.j1 && fk >= 0 && fk < df.Z) {
                const long long c = mg_idx(df, fi, fj, fk);
                const Real x0 = r[c];
                v = x0 - w6 * (Real(6) * x0 - mg_nbr_sum(df, r, fi, fj, fk, c));
            }
            rs[row * kDnRow + (side ? 4 + kDnFZ : 3)] = v;
        }
        __syncthreads();
        // ---- B: one coarse cell per thread
        {
            const int kc = tid % kDnCZ, ic = (tid / kDnCZ) % kDnCX, jc = tid / (kDnCZ * kDnCX);
            const int I = I0 + ic, J = J0 + jc, K = K0 + kc;
            if (I < dc.X && J < dc.j1 && K < dc.Z) {
                Real acc = Real(0);
#pragma unroll
                for (int n = 0; n < 4; ++n) {
                    const Real wy = (n == 1 || n == 2) ? Real(0.75) : Real(0.25);
                    Real a = Real(0);
#pragma unroll
                    for (int l = 0; l < 4; ++l) {
                        // fine rows 2I-1+l, 2J-1+n are tile rows (2 ic + l, 2 jc + n); fine z 2K-1 .. 2K+2 start at local index 3 + 2 kc
                        const Real* p = rs + ((2 * jc + n) * kDnFX + (2 * ic + l)) * kDnRow + 3 + 2 * kc;
                        const Real zsum = Real(0.25) * (p[0] + p[3]) + Real(0.75) * (p[1] + p[2]);
                        a += ((l == 1 || l == 2) ? Real(0.75) : Real(0.25)) * zsum;
                    }
                    acc += wy * a;
                }
                rc[((long long)J * dc.X + I) * dc.Z + K] = acc;
            }
        }
        __syncthreads();
    }
}

// ---- fused down-pass for levels with Z % 4 == 0 --------------------------------------------------------------------
// k_mg3_residual_rz_v4: res = r - (omega/6) M r for 4 z-consecutive cells per thread, restricted along z on the fly:
//   resz[i, j, K] = 1/4 res[2K-1] + 3/4 res[2K] + 3/4 res[2K+1] + 1/4 res[2K+2]      (layout (Y, X, Z/2), z fastest)
// The two values a thread lacks (res[k-1], res[k+4]) come from the neighbouring lanes by warp shuffle; a lane at a warp
// edge whose neighbour cell is in another warp recomputes that one residual itself.  The full-size residual is never
// written: 2 B/cell out instead of 4, and the x-y part (k_mg3_restrict_xy) reads 2 B/cell, unit stride along z.
template <class Real>
__device__ __forceinline__ Real mg3_res_at(const MgDims& d, const Real* __restrict__ r, int i, int j, int k) {
    const long long c = mg_idx(d, i, j, k);
    const Real rc = r[c];
    return rc - (Real(kMgOmega) / Real(6)) * (Real(6) * rc - mg_nbr_sum(d, r, i, j, k, c));
}
// gtid / gnth: global thread id and thread count of the (possibly multi-block) team; gnth must be a multiple of 32
template <class Real>
__device__ __forceinline__ void mg3_residual_rz_v4_dev(const MgDims& d, const Real* __restrict__ r, Real* __restrict__ resz, unsigned int gtid, unsigned int gnth) {
    const Real w6 = Real(kMgOmega) / Real(6);
    const unsigned int ZV = (unsigned int)d.Z >> 2;
    const unsigned int tot = (unsigned int)(d.j1 - d.j0) * (unsigned int)d.X * ZV;
    const unsigned int lane = threadIdx.x & 31u;
    for (unsigned int base = gtid - lane; base < tot; base += gnth) {   // warp-uniform trip count
        int i = 0, j = 0, k = 0; long long c = 0;
        Real res[4] = {Real(0), Real(0), Real(0), Real(0)};
        const bool act = mg_vec_cell(d, base + lane, i, j, k, c);
        if (act) {
            const P4<Real> rc = ldp<Real, 4>(r + c);
            Real L[4];
            mg_lap4<Real>(d, r, rc, i, j, k, c, L);
#pragma unroll
            for (int a = 0; a < 4; ++a) res[a] = rc.a[a] - w6 * L[a];
        }
        Real left = __shfl_up_sync(0xffffffffu, res[3], 1);
        Real right = __shfl_down_sync(0xffffffffu, res[0], 1);
        if (act) {
            if (k == 0) left = Real(0);
            else if (lane == 0) left = mg3_res_at<Real>(d, r, i, j, k - 1);
            if (k + 4 >= d.Z) right = Real(0);
            else if (lane == 31) right = mg3_res_at<Real>(d, r, i, j, k + 4);
            Pack<Real, 2> o;
            o.a[0] = Real(0.25) * (left + res[2]) + Real(0.75) * (res[0] + res[1]);
            o.a[1] = Real(0.25) * (res[1] + right) + Real(0.75) * (res[2] + res[3]);
            stp<Real, 2>(resz + (c >> 1), o);   // c = row * Z + k, Z and k even: (row * Z/2 + k/2)
        }
    }
}
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_residual_rz_v4(const CgScal* S, MgDims d, const Real* __restrict__ r, Real* __restrict__ resz) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    mg3_residual_rz_v4_dev<Real>(d, r, resz, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}
// Rows longer than one warp of quads (Z > 128): the two residuals a thread lacks come from its neighbours THROUGH SHARED MEMORY (a block covers 256 consecutive
// quads, i.e. whole rows whenever Z/4 divides 256): no warp-edge recomputation.  ncu at 256^3 (profiles/r2_l_c4_level0_kernels.txt)
// showed the shuffle version issue-bound — 74 % issue slots, 233 warp instructions per warp of quads — because with Z/4 = 64 every
// warp's lane 0 or lane 31 recomputed a 7-point residual in a divergent branch.  Only a block-edge thread in the middle of a row
// (Z > 1024) still recomputes.  Measured: down L0 57 -> 47 us at 256^3; on rows of <= 32 quads the shuffle kernel above never
// recomputes and is faster (no block barriers), so the launcher picks by row length.
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_residual_rz_v4_blk(const CgScal* S, MgDims d, const Real* __restrict__ r, Real* __restrict__ resz) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    __shared__ Real edge[2][256];
    const Real w6 = Real(kMgOmega) / Real(6);
    const unsigned int ZV = (unsigned int)d.Z >> 2;
    const unsigned int tot = (unsigned int)(d.j1 - d.j0) * (unsigned int)d.X * ZV;
    const unsigned int tid = threadIdx.x;
    for (unsigned int base = blockIdx.x * 256u; base < tot; base += gridDim.x * 256u) {   // block-uniform trip count
        int i = 0, j = 0, k = 0; long long c = 0;
        Real res[4] = {Real(0), Real(0), Real(0), Real(0)};
        const bool act = mg_vec_cell(d, base + tid, i, j, k, c);
        if (act) {
            const P4<Real> rc = ldp<Real, 4>(r + c);
            Real L[4];
            mg_lap4<Real>(d, r, rc, i, j, k, c, L);
#pragma unroll
            for (int a = 0; a < 4; ++a) res[a] = rc.a[a] - w6 * L[a];
        }
        edge[0][tid] = res[0]; edge[1][tid] = res[3];
        __syncthreads();
        if (act) {
            Real left = Real(0), right = Real(0);
            if (k > 0) left = (tid > 0) ? edge[1][tid - 1] : mg3_res_at<Real>(d, r, i, j, k - 1);          // quad t - 1 is the same row
            if (k + 4 < d.Z) right = (tid < 255u) ? edge[0][tid + 1] : mg3_res_at<Real>(d, r, i, j, k + 4);
            Pack<Real, 2> o;
            o.a[0] = Real(0.25) * (left + res[2]) + Real(0.75) * (res[0] + res[1]);
            o.a[1] = Real(0.25) * (res[1] + right) + Real(0.75) * (res[2] + res[3]);
            stp<Real, 2>(resz + (c >> 1), o);   // c = row * Z + k, Z and k even: (row * Z/2 + k/2)
        }
        __syncthreads();
    }
}
// rc[I,J,K] = sum over the 4x4 fine columns (2I-1..2I+2, 2J-1..2J+2) of wx wy resz[.,.,K]
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_restrict_xy(const CgScal* S, MgDims df, MgDims dc, const Real* __restrict__ resz, Real* __restrict__ rc) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    INK_MG_CELL_LOOP(dc, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x) {
        INK_MG_DECODE(dc)
        Real acc = Real(0);
#pragma unroll
        for (int n = 0; n < 4; ++n) {
            const int fj = 2 * j - 1 + n;
            if (fj < df.j0 || fj >= df.j1) continue;
            const Real wy = (n == 1 || n == 2) ? Real(0.75) : Real(0.25);
            Real a = Real(0);
#pragma unroll
            for (int l = 0; l < 4; ++l) {
                const int fi = 2 * i - 1 + l;
                if (fi < 0 || fi >= df.X) continue;
                a += ((l == 1 || l == 2) ? Real(0.75) : Real(0.25)) * resz[((long long)fj * df.X + fi) * dc.Z + k];
            }
            acc += wy * a;
        }
        rc[c] = acc;
    }
}

// same, coarse Z % 4 == 0: one thread per 4 z-consecutive coarse cells, 16 LDG.128 + 1 STG.128 (the scalar kernel is
// issue-bound on index arithmetic and validity tests: 62 % issue slots at 7 % of DRAM peak, profiles/r1_c_mg3_down.txt)
template <class Real>
__device__ __forceinline__ void mg3_restrict_xy_v4_dev(const MgDims& df, const MgDims& dc, const Real* __restrict__ resz, Real* __restrict__ rc, unsigned int gtid,
                                                       unsigned int gnth) {
    int i, j, k; long long c;
    for (unsigned int t = gtid; mg_vec_cell(dc, t, i, j, k, c); t += gnth) {
        Real acc[4] = {Real(0), Real(0), Real(0), Real(0)};
#pragma unroll
        for (int n = 0; n < 4; ++n) {
            const int fj = 2 * j - 1 + n;
            if (fj < df.j0 || fj >= df.j1) continue;
            const Real wy = (n == 1 || n == 2) ? Real(0.75) : Real(0.25);
#pragma unroll
            for (int l = 0; l < 4; ++l) {
                const int fi = 2 * i - 1 + l;
                if (fi < 0 || fi >= df.X) continue;
                const Real w = wy * ((l == 1 || l == 2) ? Real(0.75) : Real(0.25));
                const P4<Real> q = ldp<Real, 4>(resz + ((long long)fj * df.X + fi) * dc.Z + k);
#pragma unroll
                for (int a = 0; a < 4; ++a) acc[a] = hd_fma(w, q.a[a], acc[a]);
            }
        }
        P4<Real> o;
#pragma unroll
        for (int a = 0; a < 4; ++a) o.a[a] = acc[a];
        stp<Real, 4>(rc + c, o);
    }
}
template <class Real>
__global__ void __launch_bounds__(256) k_mg3_restrict_xy_v4(const CgScal* S, MgDims df, MgDims dc, const Real* __restrict__ resz, Real* __restrict__ rc) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    mg3_restrict_xy_v4_dev<Real>(df, dc, resz, rc, blockIdx.x * blockDim.x + threadIdx.x, gridDim.x * blockDim.x);
}

template <class Real>
__global__ void __launch_bounds__(256) k_mg3_down_fused(const CgScal* S, MgDims df, MgDims dc, const Real* __restrict__ r, Real* __restrict__ rc) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    mg3_down_fused<Real>(df, dc, r, rc, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}

// ---- the coarsest level alone (<= kMg3CoarsestCells cells): kMgCoarsestSweeps damped-Jacobi sweeps from zero ----------------
// One block, up to 2 cells per thread, state in shared memory, the six neighbour slots resolved ONCE (a missing neighbour points
// at a zero slot): a sweep is 7 shared loads, 8 flops and a barrier.  Same expressions in the same order as mg_smooth (the
// generic tail, k_mg3_tail, took 11.5 us of every ~115 us PCG iteration for this: strided iterators, predicated neighbour sums
// and 308 bytes of spills under its 64-register cap).
constexpr int kMg3CoarsestThreads = 1024;
template <class Real>
__global__ void __launch_bounds__(kMg3CoarsestThreads) k_mg3_coarsest(const CgScal* S, MgDims d, Real s, const Real* __restrict__ rin, Real* __restrict__ xout) {
    INK_PDL();
    if (S != nullptr && S->done) return;
    __shared__ Real xs[2][kMg3CoarsestCells + 1];
    const int X = d.X, Z = d.Z, rows = d.j1 - d.j0, n = X * rows * Z;   // rin / xout already point at the slab's first cell
    const Real cj = Real(kMgOmega) / (Real(6) * s);
    int nb[2][6];
    Real rr[2], xc[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int c = threadIdx.x + q * kMg3CoarsestThreads;
        rr[q] = Real(0); xc[q] = Real(0);
#pragma unroll
        for (int a = 0; a < 6; ++a) nb[q][a] = n;
        if (c < n) {
            const int k = c % Z, row = c / Z, i = row % X, jl = row / X;
            if (i + 1 < X) nb[q][0] = c + Z;
            if (i > 0) nb[q][1] = c - Z;
            if (jl + 1 < rows) nb[q][2] = c + X * Z;
            if (jl > 0) nb[q][3] = c - X * Z;
            if (k + 1 < Z) nb[q][4] = c + 1;
            if (k > 0) nb[q][5] = c - 1;
            rr[q] = rin[c];
            xc[q] = cj * rr[q];
            xs[0][c] = xc[q];
        }
    }
    if (threadIdx.x == 0) { xs[0][n] = Real(0); xs[1][n] = Real(0); }
    __syncthreads();
    int cur = 0;
    for (int t = 1; t < kMgCoarsestSweeps; ++t) {
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int c = threadIdx.x + q * kMg3CoarsestThreads;
            if (c < n) {
                const Real* x = xs[cur];
                Real a = Real(0);
#pragma unroll
                for (int e = 0; e < 6; ++e) a += x[nb[q][e]];
                const Real Lx = Real(6) * xc[q] - a;
                xc[q] = xc[q] + cj * (rr[q] - s * Lx);
                xs[cur ^ 1][c] = xc[q];
            }
        }
        __syncthreads();
        cur ^= 1;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        const int c = threadIdx.x + q * kMg3CoarsestThreads;
        if (c < n) xout[c] = xc[q];
    }
}

// ---- shared-memory tail: all levels from `lt` down in ONE block (MgTailArgs of mg.cuh; 3 arrays per level) ----------
template <class Real>
__device__ __forceinline__ void mg3_tail_dev(CgScal* S, const MgTailArgs& A, const Real* __restrict__ rin, Real* __restrict__ xout) {
    extern __shared__ __align__(16) unsigned char mg_smem_raw[];
    Real* sm = reinterpret_cast<Real*>(mg_smem_raw);
    const long long tid = threadIdx.x, nth = blockDim.x;
    Real *xx[kMgMaxLevels], *tmp[kMgMaxLevels], *rr[kMgMaxLevels];   // xx: level result e; tmp: residual / Jacobi ping-pong
    {
        Real* p = sm;
        for (int q = 0; q < A.nlev; ++q) {
            const long long n = (long long)A.d[q].X * A.d[q].Y * A.d[q].Z;
            xx[q] = p; tmp[q] = p + n; rr[q] = p + 2 * n;
            p += 3 * n;
        }
    }
    const long long n0 = (long long)A.d[0].X * A.d[0].Y * A.d[0].Z;
    for (long long c = tid; c < n0; c += nth) rr[0][c] = rin[c];
    __syncthreads();
    Real s = (Real)A.s0;
    for (int q = 0; q + 1 < A.nlev; ++q) {   // down
        mg3_residual<Real>(A.d[q], rr[q], tmp[q], tid, nth);
        __syncthreads();
        mg3_restrict<Real>(A.d[q], A.d[q + 1], tmp[q], rr[q + 1], tid, nth);
        __syncthreads();
        s *= Real(2);
    }
    const Real* ecoarse;
    {   // coarsest: kMgCoarsestSweeps damped-Jacobi sweeps from zero
        const int q = A.nlev - 1;
        const long long n = (long long)A.d[q].X * A.d[q].Y * A.d[q].Z;
        const Real cj = Real(kMgOmega) / (Real(6) * s);
        Real *a = xx[q], *b = tmp[q];
        for (long long c = tid; c < n; c += nth) a[c] = cj * rr[q][c];
        __syncthreads();
        for (int t = 1; t < kMgCoarsestSweeps; ++t) {
            mg_smooth<Real>(A.d[q], s, rr[q], a, b, tid, nth);
            __syncthreads();
            Real* w = a; a = b; b = w;
        }
        ecoarse = a;
    }
    for (int q = A.nlev - 2; q >= 1; --q) {   // up
        s *= Real(0.5);
        mg3_up<Real>(A.d[q], A.d[q + 1], s, rr[q], ecoarse, xx[q], tid, nth);
        __syncthreads();
        ecoarse = xx[q];
    }
    double acc = 0;
    if (A.nlev > 1) {
        s *= Real(0.5);
        acc = mg3_up<Real>(A.d[0], A.d[1], s, rr[0], ecoarse, xout, tid, nth, A.dj0, A.dj1);
    } else
        for (long long c = tid; c < n0; c += nth) {
            const Real v = ecoarse[c];
            xout[c] = v;
            const int jrow = (int)(c / ((long long)A.d[0].X * A.d[0].Z));
            if (jrow >= A.dj0 && jrow < A.dj1) acc += (double)rr[0][c] * (double)v;
        }
    if (A.final_dot) {
        __shared__ double red[kMgTailThreads / 32];
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
            if (S->dist) S->loc[2] = t; else { S->rz_old = S->rz; S->rz = t; }
        }
    }
}
template <class Real>
__global__ void __launch_bounds__(kMgTailThreads) k_mg3_tail(CgScal* S, MgTailArgs A, const Real* __restrict__ rin, Real* __restrict__ xout) {
    INK_PDL();
    if (S->done) return;
    mg3_tail_dev<Real>(S, A, rin, xout);
}

// ---- persistent coarse-level kernel: every level below kMg3CoarseCells cells in ONE launch ----------------------------------
// Levels with <= ~100 k cells are launch-latency bound: each of their 3 kernels costs 4-5 us for < 1 us of work, ~55 us of
// a ~120 us PCG iteration at 128x256x128 (INK_TRACE_ITER timeline, profiles/r1_e_trace.txt).  Here a small co-resident team
// of blocks walks down and up those levels with a device-side grid barrier between passes (arrays stay in global memory =
// L2; block 0 runs the shared-memory tail).  Safe with plain/non-coherent loads: inside one launch every array element is
// written in exactly one pass and only read in later passes, so no SM can hold a stale L1 line for it.
constexpr int kMg3CoarseCells = 131072;
constexpr int kMg3CoarseBlocks = 32;
constexpr int kMg3ClusterCtas = 16;     // non-portable cluster size (opt-in attribute); 8 is the portable maximum
constexpr int kMg3CoarseThreads = 512;
constexpr int kMg3CoarseMaxLevels = 6;

__device__ __forceinline__ void mg_grid_barrier(unsigned int* bar) {   // bar[0] = arrivals, bar[1] = generation
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int* gen = bar + 1;
        __threadfence();
        const unsigned int g = *gen;
        if (atomicAdd(bar, 1u) == gridDim.x - 1) {
            bar[0] = 0;
            __threadfence();
            *gen = g + 1;
        } else {
            while (*gen == g) { }
        }
        __threadfence();
    }
    __syncthreads();
}

template <class Real>
struct Mg3CoarseArgs {
    int nlev;                                  // levels handled here; the last one is the top of the shared-memory tail
    MgLevelPtrs<Real> lv[kMg3CoarseMaxLevels];
    MgTailArgs tail;
    long long tail_off;                        // element offset of the tail's slab in its level arrays
    unsigned int* bar;
};

// hardware barrier across the CTAs of a thread-block cluster (sm_90+): release on arrive, acquire on wait, cluster scope —
// global-memory writes of every CTA before the barrier are visible to every CTA after it
__device__ __forceinline__ void mg_cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// CLUSTER = false: a co-resident grid of kMg3CoarseBlocks blocks with a global-memory barrier (round 1: 63 us for levels 2-4 at
// 128x256x128, slower than the launch train — each barrier is an atomic + a spin through L2).
// CLUSTER = true : ONE thread-block cluster (kMg3ClusterCtas CTAs, launched with cudaLaunchAttributeClusterDimension); the
// barrier between passes is the hardware cluster barrier (sub-microsecond).  Level arrays stay in global memory = L2 (written
// in exactly one pass, read only in later passes, so no SM can hold a stale L1 line); CTA 0 runs the shared-memory tail.
template <class Real, bool CLUSTER>
__global__ void __launch_bounds__(kMg3CoarseThreads) k_mg3_coarse(CgScal* S, Mg3CoarseArgs<Real> A) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (S->done) return;                       // uniform: nobody writes `done` during this launch
    const unsigned int gtid = blockIdx.x * blockDim.x + threadIdx.x, gnth = gridDim.x * blockDim.x;
    auto barrier = [&]() { if (CLUSTER) mg_cluster_barrier(); else mg_grid_barrier(A.bar); };
    for (int q = 0; q + 1 < A.nlev; ++q) {     // down
        const MgDims &df = A.lv[q].d, &dc = A.lv[q + 1].d;
        if (df.Z % 4 == 0) mg3_residual_rz_v4_dev<Real>(df, A.lv[q].r, A.lv[q].xb, gtid, gnth);
        else mg3_residual<Real>(df, A.lv[q].r, A.lv[q].xb, gtid, gnth);
        barrier();
        if (df.Z % 4 == 0 && dc.Z % 4 == 0) mg3_restrict_xy_v4_dev<Real>(df, dc, A.lv[q].xb, A.lv[q + 1].r, gtid, gnth);
        else if (df.Z % 4 == 0) {              // z-restricted residual, scalar x-y gather
            INK_MG_CELL_LOOP(dc, gtid, gnth) {
                INK_MG_DECODE(dc)
                Real acc = Real(0);
                for (int n = 0; n < 4; ++n) {
                    const int fj = 2 * j - 1 + n;
                    if (fj < df.j0 || fj >= df.j1) continue;
                    const Real wy = (n == 1 || n == 2) ? Real(0.75) : Real(0.25);
                    for (int l = 0; l < 4; ++l) {
                        const int fi = 2 * i - 1 + l;
                        if (fi < 0 || fi >= df.X) continue;
                        acc += wy * ((l == 1 || l == 2) ? Real(0.75) : Real(0.25)) * A.lv[q].xb[((long long)fj * df.X + fi) * dc.Z + k];
                    }
                }
                A.lv[q + 1].r[c] = acc;
            }
        } else mg3_restrict<Real>(df, dc, A.lv[q].xb, A.lv[q + 1].r, gtid, gnth);
        barrier();
    }
    if (blockIdx.x == 0) mg3_tail_dev<Real>(S, A.tail, A.lv[A.nlev - 1].r + A.tail_off, A.lv[A.nlev - 1].xa + A.tail_off);
    barrier();
    for (int q = A.nlev - 2; q >= 0; --q) {    // up
        mg3_up<Real>(A.lv[q].d, A.lv[q + 1].d, A.lv[q].s, A.lv[q].r, A.lv[q + 1].xa, A.lv[q].xa, gtid, gnth);
        if (q > 0) barrier();
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
#endif  // __CUDACC__

}  // namespace ink
