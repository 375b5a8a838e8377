// mg.cuh — geometric-multigrid V-cycle used as the PRECONDITIONER of the matrix-free CG pressure solve (cg.cuh).
//
// The reference pre-factors its Poisson matrix with Eigen::SparseLU (src/waterSolver.cpp:240-261) and back-solves
// every step (:279).  The B200 path is matrix-free CG; un-preconditioned it needs ~450 iterations at 128x256x128
// (kappa ~ 1e4) — 97 % of a step.  One V-cycle per iteration brings that to ~15-20 iterations, independent of size.
//
// System on level l (cell-centred 2x coarsening, ceil for odd extents; level 0 = the pressure grid):
//      s_l * (6 e - sum_{in-grid nbrs} e) = r_l ,   s_l = 2^l      (M of cg.cuh is level 0)
//   restriction  r_{l+1}[I] = sum over the (<= 8) children of (r_l - s_l L e_l)        (= P^T)
//   prolongation e_l[c]    += e_{l+1}[parent(c)]                                        (piecewise constant P)
//   smoother     damped Jacobi, omega = 0.8, two sweeps before and two after (the first pair starts from e = 0 and
//                is fused into ONE stencil pass over r:  e = c (2 r - (omega/6) L r),  c = omega / (6 s_l))
// The cycle is a fixed symmetric positive-definite linear operator (R = P^T, same smoother both sides), which is
// what CG requires of a preconditioner.  Levels small enough for shared memory (<= kMgTailCells cells) are
// processed by ONE thread block in ONE launch (k_mg_tail) to avoid a train of launch-latency-bound tiny kernels.
//
// All loops are (thread id, thread count)-strided so the same device functions serve the global-memory kernels
// and the shared-memory tail.  y-slab ranges [j0, j1) restrict the rows a kernel updates (multi-GPU: every rank
// preconditions its own slab, treating slab interfaces like walls => block-diagonal SPD preconditioner, no traffic).
#pragma once
#include "cg.cuh"
#include "common.cuh"

namespace ink {

constexpr int kMgMaxLevels = 12;
constexpr int kMgTailThreads = 1024;
constexpr float kMgOmega = 0.8f;
constexpr int kMgCoarsestSweeps = 8;

struct MgDims {
    int X, Y, Z;
    int j0, j1;   // rows [j0, j1) updated here; neighbours outside [j0, j1) read as 0 (slab-local operator)
};

template <class Real>
struct MgLevelPtrs {
    MgDims d;
    Real *xa, *xb, *r;   // xa/xb: Jacobi ping-pong; r: right-hand side of this level
    Real s;              // operator scale 2^l
};

INK_HD long long mg_idx(const MgDims& d, int i, int j, int k) { return ((long long)j * d.X + i) * d.Z + k; }

// sum of the in-grid (and in-slab) face neighbours of cell (i,j,k)
template <class Real>
INK_HD Real mg_nbr_sum(const MgDims& d, const Real* __restrict__ x, int i, int j, int k, long long c) {
    const long long sx = d.Z, sy = (long long)d.X * d.Z;
    Real a = Real(0);
    if (i + 1 < d.X) a += x[c + sx];
    if (i > 0) a += x[c - sx];
    if (j + 1 < d.j1) a += x[c + sy];
    if (j > d.j0) a += x[c - sy];
    if (k + 1 < d.Z) a += x[c + 1];
    if (k > 0) a += x[c - 1];
    return a;
}

// Strided cell iterator with INCREMENTAL (i, j, k) decode: two 32-bit divisions per thread at start, then adds and
// compares per cell.  (Per-cell div/mod made the single-block tail issue-bound: 8 cells/thread x ~50 instructions.)
struct MgIter {
    unsigned int t, tot, nth;
    int i, jl, k, di, dj, dk, X, Z;
    INK_HD MgIter(const MgDims& d, unsigned int tid, unsigned int nthreads) {
        X = d.X; Z = d.Z; nth = nthreads; t = tid;
        tot = (unsigned int)(d.j1 - d.j0) * (unsigned int)d.X * (unsigned int)d.Z;
        unsigned int row = tid / (unsigned int)Z;
        k = (int)(tid - row * (unsigned int)Z);
        jl = (int)(row / (unsigned int)X);
        i = (int)(row - (unsigned int)jl * (unsigned int)X);
        row = nthreads / (unsigned int)Z;
        dk = (int)(nthreads - row * (unsigned int)Z);
        dj = (int)(row / (unsigned int)X);
        di = (int)(row - (unsigned int)dj * (unsigned int)X);
    }
    INK_HD bool ok() const { return t < tot; }
    INK_HD void next() {
        t += nth;
        k += dk; if (k >= Z) { k -= Z; ++i; }
        i += di; if (i >= X) { i -= X; ++jl; }
        jl += dj;
    }
};
#define INK_MG_CELL_LOOP(d, tid, nth) for (MgIter it_((d), (unsigned int)(tid), (unsigned int)(nth)); it_.ok(); it_.next())

#define INK_MG_DECODE(d)                                   \
    const int i = it_.i, k = it_.k, j = (d).j0 + it_.jl;   \
    const long long c = mg_idx((d), i, j, k);

// two Jacobi sweeps from a zero guess, fused: x = c (2 r - (omega/6) L r)
template <class Real>
__device__ __forceinline__ void mg_presmooth(const MgDims& d, Real s, const Real* __restrict__ r, Real* __restrict__ x, long long tid, long long nth) {
    const Real cj = Real(kMgOmega) / (Real(6) * s), w6 = Real(kMgOmega) / Real(6);
    INK_MG_CELL_LOOP(d, tid, nth) {
        INK_MG_DECODE(d)
        const Real rc = r[c];
        const Real Lr = Real(6) * rc - mg_nbr_sum(d, r, i, j, k, c);
        x[c] = cj * (Real(2) * rc - w6 * Lr);
    }
}

// one Jacobi sweep xo = xi + c (r - s L xi); returns this thread's partial of <r, xo> (used by the final sweep)
template <class Real>
INK_HD double mg_smooth(const MgDims& d, Real s, const Real* __restrict__ r, const Real* __restrict__ xi, Real* __restrict__ xo,
                                            long long tid, long long nth, int dj0 = 0, int dj1 = 0x7fffffff) {
    const Real cj = Real(kMgOmega) / (Real(6) * s);
    double acc = 0;
    INK_MG_CELL_LOOP(d, tid, nth) {
        INK_MG_DECODE(d)
        const Real xc = xi[c], rc = r[c];
        const Real Lx = Real(6) * xc - mg_nbr_sum(d, xi, i, j, k, c);
        const Real v = xc + cj * (rc - s * Lx);
        xo[c] = v;
        if (j >= dj0 && j < dj1) acc += (double)rc * (double)v;
    }
    return acc;
}

// r_c[I] = sum over the children of (r - s L x).  df = fine dims, dc = coarse dims (dc.j0 = df.j0/2, dc.j1 = ceil(df.j1/2)).
template <class Real>
__device__ __forceinline__ void mg_residual_restrict(const MgDims& df, const MgDims& dc, Real s, const Real* __restrict__ r, const Real* __restrict__ x,
                                                     Real* __restrict__ rc, long long tid, long long nth) {
    INK_MG_CELL_LOOP(dc, tid, nth) {
        INK_MG_DECODE(dc)
        Real acc = Real(0);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int fj = 2 * j + b;
            if (fj >= df.j1 || fj < df.j0) continue;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int fi = 2 * i + a;
                if (fi >= df.X) continue;
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int fk = 2 * k + e;
                    if (fk >= df.Z) continue;
                    const long long f = mg_idx(df, fi, fj, fk);
                    const Real Lx = Real(6) * x[f] - mg_nbr_sum(df, x, fi, fj, fk, f);
                    acc += r[f] - s * Lx;
                }
            }
        }
        rc[c] = acc;
    }
}

// x' = xi + P e ; xo = x' + c (r - s L x')   (prolongation fused with the first post-smoothing sweep)
template <class Real>
__device__ __forceinline__ void mg_prolong_smooth(const MgDims& df, const MgDims& dc, Real s, const Real* __restrict__ r, const Real* __restrict__ xi,
                                                  const Real* __restrict__ e, Real* __restrict__ xo, long long tid, long long nth) {
    const Real cj = Real(kMgOmega) / (Real(6) * s);
    const long long sx = df.Z, sy = (long long)df.X * df.Z;
    INK_MG_CELL_LOOP(df, tid, nth) {
        INK_MG_DECODE(df)
#define INK_XP(ii, jj, kk, off) (xi[c + (off)] + e[mg_idx(dc, (ii) >> 1, (jj) >> 1, (kk) >> 1)])
        const Real xc = INK_XP(i, j, k, 0);
        Real nb = Real(0);
        if (i + 1 < df.X) nb += INK_XP(i + 1, j, k, sx);
        if (i > 0) nb += INK_XP(i - 1, j, k, -sx);
        if (j + 1 < df.j1) nb += INK_XP(i, j + 1, k, sy);
        if (j > df.j0) nb += INK_XP(i, j - 1, k, -sy);
        if (k + 1 < df.Z) nb += INK_XP(i, j, k + 1, 1);
        if (k > 0) nb += INK_XP(i, j, k - 1, -1);
#undef INK_XP
        const Real Lx = Real(6) * xc - nb;
        xo[c] = xc + cj * (r[c] - s * Lx);
    }
}

// ---- global-memory kernels (levels too large for shared memory) --------------------------------------------------
constexpr int kMgBlocks = 148 * 8;

template <class Real>
__global__ void __launch_bounds__(256) k_mg_presmooth(const CgScal* S, MgDims d, Real s, const Real* __restrict__ r, Real* __restrict__ x) {
    INK_PDL();
    if (S->done) return;
    mg_presmooth<Real>(d, s, r, x, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}
template <class Real>
__global__ void __launch_bounds__(256) k_mg_residual_restrict(const CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r,
                                                              const Real* __restrict__ x, Real* __restrict__ rc) {
    INK_PDL();
    if (S->done) return;
    mg_residual_restrict<Real>(df, dc, s, r, x, rc, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}
template <class Real>
__global__ void __launch_bounds__(256) k_mg_prolong_smooth(const CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r,
                                                           const Real* __restrict__ xi, const Real* __restrict__ e, Real* __restrict__ xo) {
    INK_PDL();
    if (S->done) return;
    mg_prolong_smooth<Real>(df, dc, s, r, xi, e, xo, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}
// FINAL: the last sweep of level 0 also produces <r, z> for CG (rz_old <- rz, rz <- <r,z>)
template <class Real, bool FINAL>
__global__ void __launch_bounds__(256) k_mg_smooth(CgScal* S, MgDims d, Real s, const Real* __restrict__ r, const Real* __restrict__ xi, Real* __restrict__ xo,
                                                   int dj0, int dj1) {
    INK_PDL();
    if (S->done) return;
    const double acc = mg_smooth<Real>(d, s, r, xi, xo, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x, dj0, dj1);
    if (FINAL) {
        double tot;
        if (block_sum_last(acc, S, 0, tot)) {
            if (threadIdx.x == 0) { if (S->dist) S->loc[2] = tot; else { S->rz_old = S->rz; S->rz = tot; } }
        }
    }
}

// ---- shared-memory tail: all levels from `lt` down in ONE block ---------------------------------------------------
struct MgTailArgs {
    int nlev;                      // levels in the tail (>= 1)
    MgDims d[kMgMaxLevels];        // slab-local dims of each tail level (j0 = 0)
    float s0;                      // operator scale of the first tail level
    int final_dot;                 // tail starts at level 0: also produce <r, z> for CG ...
    int dj0, dj1;                  // ... over the slab-local rows [dj0, dj1)
};

template <class Real>
__global__ void __launch_bounds__(kMgTailThreads) k_mg_tail(CgScal* S, MgTailArgs A, const Real* __restrict__ rin, Real* __restrict__ xout) {
    INK_PDL();
    if (S->done) return;
    extern __shared__ __align__(16) unsigned char mg_smem_raw[];
    Real* sm = reinterpret_cast<Real*>(mg_smem_raw);
    const long long tid = threadIdx.x, nth = blockDim.x;
    Real *xa[kMgMaxLevels], *xb[kMgMaxLevels], *rr[kMgMaxLevels];
    {
        Real* p = sm;
        for (int q = 0; q < A.nlev; ++q) {
            const long long n = (long long)A.d[q].X * A.d[q].Y * A.d[q].Z;
            xa[q] = p; xb[q] = p + n; rr[q] = p + 2 * n;
            p += 3 * n;
        }
    }
    const long long n0 = (long long)A.d[0].X * A.d[0].Y * A.d[0].Z;
    for (long long c = tid; c < n0; c += nth) rr[0][c] = rin[c];
    __syncthreads();
    Real s = (Real)A.s0;
    // down
    for (int q = 0; q + 1 < A.nlev; ++q) {
        mg_presmooth<Real>(A.d[q], s, rr[q], xa[q], tid, nth);
        __syncthreads();
        mg_residual_restrict<Real>(A.d[q], A.d[q + 1], s, rr[q], xa[q], rr[q + 1], tid, nth);
        __syncthreads();
        s *= Real(2);
    }
    // coarsest: kMgCoarsestSweeps damped-Jacobi sweeps from zero (fixed polynomial in the operator => symmetric)
    {
        const int q = A.nlev - 1;
        const long long n = (long long)A.d[q].X * A.d[q].Y * A.d[q].Z;
        const Real cj = Real(kMgOmega) / (Real(6) * s);
        for (long long c = tid; c < n; c += nth) xa[q][c] = cj * rr[q][c];
        __syncthreads();
        for (int t = 1; t < kMgCoarsestSweeps; ++t) {   // kMgCoarsestSweeps is even: the result ends in xb, copy back below
            mg_smooth<Real>(A.d[q], s, rr[q], (t & 1) ? xa[q] : xb[q], (t & 1) ? xb[q] : xa[q], tid, nth);
            __syncthreads();
        }
        if ((kMgCoarsestSweeps & 1) == 0) {
            for (long long c = tid; c < n; c += nth) xa[q][c] = xb[q][c];
            __syncthreads();
        }
    }
    // up
    for (int q = A.nlev - 2; q >= 0; --q) {
        s *= Real(0.5);
        mg_prolong_smooth<Real>(A.d[q], A.d[q + 1], s, rr[q], xa[q], xa[q + 1], xb[q], tid, nth);
        __syncthreads();
        if (q > 0) {
            mg_smooth<Real>(A.d[q], s, rr[q], xb[q], xa[q], tid, nth);
            __syncthreads();
        }
    }
    // result of the first tail level goes to global memory (with the last smoothing sweep when there is one)
    double acc = 0;
    if (A.nlev > 1) acc = mg_smooth<Real>(A.d[0], s, rr[0], xb[0], xout, tid, nth, A.dj0, A.dj1);
    else
        for (long long c = tid; c < n0; c += nth) {
            const Real v = xa[0][c];
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

}  // namespace ink
