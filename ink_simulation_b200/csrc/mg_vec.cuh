// mg_vec.cuh — 128-bit vectorised versions of the global-memory multigrid kernels of mg.cuh for levels whose z extent
// is a multiple of 4 (every BASELINE grid on its fine levels).  One thread owns 4 consecutive z cells of one (i,j)
// row: the (i,j,k) decode is amortised over 4 cells and every row access is one LDG.128 / STG.128.  Arithmetic is
// identical to the scalar device functions (same expressions per cell), so scalar and vector levels can be mixed.
#pragma once
#include "mg.cuh"

namespace ink {

template <class Real>
using P4 = Pack<Real, 4>;

template <class Real>
__device__ __forceinline__ P4<Real> p4_zero() { P4<Real> z; z.a[0] = z.a[1] = z.a[2] = z.a[3] = Real(0); return z; }

// thread -> (i, j, k = 4*kv, c) of a level; returns false past the end
__device__ __forceinline__ bool mg_vec_cell(const MgDims& d, unsigned int t, int& i, int& j, int& k, long long& c) {
    const unsigned int ZV = (unsigned int)d.Z >> 2;
    const unsigned int tot = (unsigned int)(d.j1 - d.j0) * (unsigned int)d.X * ZV;
    if (t >= tot) return false;
    const unsigned int row = t / ZV;
    k = (int)(t - row * ZV) << 2;
    const unsigned int jl = row / (unsigned int)d.X;
    i = (int)(row - jl * (unsigned int)d.X);
    j = d.j0 + (int)jl;
    c = mg_idx(d, i, j, k);
    return true;
}

// 6 x - (in-grid, in-slab neighbours) for the 4 cells at c; xc = the centre pack (already loaded)
template <class Real>
__device__ __forceinline__ void mg_lap4(const MgDims& d, const Real* __restrict__ x, const P4<Real>& xc, int i, int j, int k, long long c, Real (&L)[4]) {
    const long long sx = d.Z, sy = (long long)d.X * d.Z;
    Real nb[4] = {Real(0), Real(0), Real(0), Real(0)};
#define INK_ROW(cond, off)                                         \
    if (cond) {                                                    \
        const P4<Real> q = ldp<Real, 4>(x + c + (off));            \
        nb[0] += q.a[0]; nb[1] += q.a[1]; nb[2] += q.a[2]; nb[3] += q.a[3]; \
    }
    INK_ROW(i + 1 < d.X, sx)
    INK_ROW(i > 0, -sx)
    INK_ROW(j + 1 < d.j1, sy)
    INK_ROW(j > d.j0, -sy)
#undef INK_ROW
    const Real zl = (k > 0) ? x[c - 1] : Real(0);
    const Real zr = (k + 4 < d.Z) ? x[c + 4] : Real(0);
    L[0] = Real(6) * xc.a[0] - (nb[0] + zl + xc.a[1]);
    L[1] = Real(6) * xc.a[1] - (nb[1] + xc.a[0] + xc.a[2]);
    L[2] = Real(6) * xc.a[2] - (nb[2] + xc.a[1] + xc.a[3]);
    L[3] = Real(6) * xc.a[3] - (nb[3] + xc.a[2] + zr);
}

template <class Real>
__global__ void __launch_bounds__(256) k_mg_presmooth_v4(const CgScal* S, MgDims d, Real s, const Real* __restrict__ r, Real* __restrict__ x) {
    INK_PDL();
    if (S->done) return;
    const Real cj = Real(kMgOmega) / (Real(6) * s), w6 = Real(kMgOmega) / Real(6);
    int i, j, k; long long c;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; mg_vec_cell(d, t, i, j, k, c); t += gridDim.x * blockDim.x) {
        const P4<Real> rc = ldp<Real, 4>(r + c);
        Real L[4];
        mg_lap4<Real>(d, r, rc, i, j, k, c, L);
        P4<Real> o;
#pragma unroll
        for (int a = 0; a < 4; ++a) o.a[a] = cj * (Real(2) * rc.a[a] - w6 * L[a]);
        stp<Real, 4>(x + c, o);
    }
}

template <class Real, bool FINAL>
__global__ void __launch_bounds__(256) k_mg_smooth_v4(CgScal* S, MgDims d, Real s, const Real* __restrict__ r, const Real* __restrict__ xi, Real* __restrict__ xo,
                                                      int dj0, int dj1) {
    INK_PDL();
    if (S->done) return;
    const Real cj = Real(kMgOmega) / (Real(6) * s);
    int i, j, k; long long c;
    double acc = 0;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; mg_vec_cell(d, t, i, j, k, c); t += gridDim.x * blockDim.x) {
        const P4<Real> xc = ldp<Real, 4>(xi + c), rc = ldp<Real, 4>(r + c);
        Real L[4];
        mg_lap4<Real>(d, xi, xc, i, j, k, c, L);
        P4<Real> o;
        const bool in_dot = (j >= dj0) && (j < dj1);
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            o.a[a] = xc.a[a] + cj * (rc.a[a] - s * L[a]);
            if (in_dot) acc += (double)rc.a[a] * (double)o.a[a];
        }
        stp<Real, 4>(xo + c, o);
    }
    if (FINAL) {
        double tot;
        if (block_sum_last(acc, S, 0, tot)) {
            if (threadIdx.x == 0) { if (S->dist) S->loc[2] = tot; else { S->rz_old = S->rz; S->rz = tot; } }
        }
    }
}

// x' = xi + P e ; xo = x' + c (r - s L x').  The 4 cells k..k+3 (k % 4 == 0) have parents K = k/2, k/2 + 1.
template <class Real>
__global__ void __launch_bounds__(256) k_mg_prolong_smooth_v4(const CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r,
                                                              const Real* __restrict__ xi, const Real* __restrict__ e, Real* __restrict__ xo) {
    INK_PDL();
    if (S->done) return;
    const Real cj = Real(kMgOmega) / (Real(6) * s);
    const long long sx = df.Z, sy = (long long)df.X * df.Z;
    int i, j, k; long long c;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; mg_vec_cell(df, t, i, j, k, c); t += gridDim.x * blockDim.x) {
        const int K = k >> 1;
        // x' of one fine row (ii, jj): xi row pack + the two parent values of coarse row (ii>>1, jj>>1)
        auto row = [&](int ii, int jj, long long off) -> P4<Real> {
            P4<Real> q = ldp<Real, 4>(xi + c + off);
            const Real* er = e + mg_idx(dc, ii >> 1, jj >> 1, K);
            const Real e0 = er[0], e1 = (K + 1 < dc.Z) ? er[1] : Real(0);   // Z % 4 == 0 => K + 1 < dc.Z always; guard is free
            q.a[0] += e0; q.a[1] += e0; q.a[2] += e1; q.a[3] += e1;
            return q;
        };
        const P4<Real> xc = row(i, j, 0);
        Real nb[4] = {Real(0), Real(0), Real(0), Real(0)};
#define INK_ROW(cond, ii, jj, off)                                                   \
        if (cond) {                                                                  \
            const P4<Real> q = row(ii, jj, off);                                     \
            nb[0] += q.a[0]; nb[1] += q.a[1]; nb[2] += q.a[2]; nb[3] += q.a[3];      \
        }
        INK_ROW(i + 1 < df.X, i + 1, j, sx)
        INK_ROW(i > 0, i - 1, j, -sx)
        INK_ROW(j + 1 < df.j1, i, j + 1, sy)
        INK_ROW(j > df.j0, i, j - 1, -sy)
#undef INK_ROW
        const long long ec = mg_idx(dc, i >> 1, j >> 1, K);
        const Real zl = (k > 0) ? xi[c - 1] + e[ec - 1] : Real(0);
        const Real zr = (k + 4 < df.Z) ? xi[c + 4] + e[ec + 2] : Real(0);
        const P4<Real> rc = ldp<Real, 4>(r + c);
        P4<Real> o;
        o.a[0] = xc.a[0] + cj * (rc.a[0] - s * (Real(6) * xc.a[0] - (nb[0] + zl + xc.a[1])));
        o.a[1] = xc.a[1] + cj * (rc.a[1] - s * (Real(6) * xc.a[1] - (nb[1] + xc.a[0] + xc.a[2])));
        o.a[2] = xc.a[2] + cj * (rc.a[2] - s * (Real(6) * xc.a[2] - (nb[2] + xc.a[1] + xc.a[3])));
        o.a[3] = xc.a[3] + cj * (rc.a[3] - s * (Real(6) * xc.a[3] - (nb[3] + xc.a[2] + zr)));
        stp<Real, 4>(xo + c, o);
    }
}

// r_c = P^T (r - s L x): one thread produces the two coarse cells (I, J, K..K+1), K even, from the 4 fine rows
// (2I+a, 2J+b) x cells 2K..2K+3.
template <class Real>
__global__ void __launch_bounds__(256) k_mg_residual_restrict_v4(const CgScal* S, MgDims df, MgDims dc, Real s, const Real* __restrict__ r,
                                                                 const Real* __restrict__ x, Real* __restrict__ rc) {
    INK_PDL();
    if (S->done) return;
    const unsigned int ZV = (unsigned int)dc.Z >> 1;   // df.Z % 4 == 0  =>  dc.Z even
    const unsigned int tot = (unsigned int)(dc.j1 - dc.j0) * (unsigned int)dc.X * ZV;
    for (unsigned int t = blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += gridDim.x * blockDim.x) {
        const unsigned int rowc = t / ZV;
        const int K = (int)(t - rowc * ZV) << 1;
        const unsigned int jl = rowc / (unsigned int)dc.X;
        const int I = (int)(rowc - jl * (unsigned int)dc.X), J = dc.j0 + (int)jl;
        Real acc0 = Real(0), acc1 = Real(0);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const int fj = 2 * J + b;
            if (fj >= df.j1 || fj < df.j0) continue;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int fi = 2 * I + a;
                if (fi >= df.X) continue;
                const int fk = 2 * K;
                const long long f = mg_idx(df, fi, fj, fk);
                const P4<Real> xc = ldp<Real, 4>(x + f), rr = ldp<Real, 4>(r + f);
                Real L[4];
                mg_lap4<Real>(df, x, xc, fi, fj, fk, f, L);
                acc0 += (rr.a[0] - s * L[0]) + (rr.a[1] - s * L[1]);
                acc1 += (rr.a[2] - s * L[2]) + (rr.a[3] - s * L[3]);
            }
        }
        Pack<Real, 2> o; o.a[0] = acc0; o.a[1] = acc1;
        stp<Real, 2>(rc + mg_idx(dc, I, J, K), o);
    }
}

// ---- multi-GPU: symmetric weighted overlapping Schwarz around the slab-local V-cycle -------------------------------
// Every rank preconditions its slab EXTENDED by `ovl` rows into each neighbour:  z = sum_i sqrt(W_i) B_i sqrt(W_i) r,
// W_i = partition of unity ramping linearly across the 2*ovl-row band around an interface (sw[j] = sqrt(W_i(j)) for
// this rank).  SPD and symmetric; at ovl = 8 it converges like the global V-cycle for two boundary-row
// exchanges per iteration instead of ~5 per level.  All three kernels run over one band of rows [ja, jb).
// VEC = elements per access (16 bytes when a row is a multiple of 16 bytes — always on the peer-memory path — else 1).
// t runs over the band in units of VEC elements; a band is < 2^31 cells.
// One launch covers BOTH bands of an interior rank (two neighbours): blocks [0, nb) work on side 0, [nb, 2 nb) on side 1 — three
// launches and three serialised peer waits fewer per PCG iteration than one launch per side.
template <class Real>
struct BandSide {
    int ja, jb;                              // band rows [ja, jb)
    Real* save;                              // unscaled own rows of the band (restored after the cycle)
    const Real* halo_src; int halo_j0;       // peer-memory path: where the neighbour pushed its r rows (nullptr: they are in r already)
    const unsigned long long* wait_r;        // flag the neighbour publishes after pushing r   (k_band_scale)
    Real* push_dst; unsigned long long* push_flag; unsigned int* ticket;   // weighted z band -> neighbour's arena (k_band_restore_weight)
    const Real* recv; const unsigned long long* wait_z;                    // neighbour's weighted band (k_band_combine)
};
template <class Real>
struct BandArgs {
    int nsides, nb;                          // blocks per side
    BandSide<Real> s[2];
    unsigned long long r_seq, z_seq;
};

template <class Real, int VEC>
__global__ void __launch_bounds__(256) k_band_scale(const CgScal* S, long long sy, BandArgs<Real> A, int own0, int own1, const Real* __restrict__ sw,
                                                    Real* __restrict__ r) {
    INK_PDL();
    if (S->done) return;
    const int side = (int)blockIdx.x >= A.nb ? 1 : 0;
    const BandSide<Real>& B = A.s[side];
    const unsigned int bid = blockIdx.x - side * A.nb;
    // peer-memory path: the neighbour's rows were pushed into this rank's arena by the kernel that produced them (P2pPush);
    // wait for its sequence number, then take them from there (L2) instead of from r
    if (B.wait_r) {
        if (threadIdx.x == 0) p2p_spin(reinterpret_cast<const volatile unsigned long long*>(B.wait_r), A.r_seq);
        __syncthreads();
    }
    const unsigned int syv = (unsigned int)(sy / VEC), n = (unsigned int)(B.jb - B.ja) * syv;
    for (unsigned int t = bid * blockDim.x + threadIdx.x; t < n; t += A.nb * blockDim.x) {
        const unsigned int jr = t / syv;
        const int j = B.ja + (int)jr;
        const long long e = (long long)t * VEC, c = (long long)B.ja * sy + e;
        Pack<Real, VEC> v;
        if (j >= own0 && j < own1) { v = ldp<Real, VEC>(r + c); stp<Real, VEC>(B.save + e, v); }   // own rows are restored after the cycle
        else if (B.halo_src) {
            const Real* hp = B.halo_src + ((long long)(j - B.halo_j0) * sy + (e - (long long)jr * sy));
#pragma unroll
            for (int a = 0; a < VEC; ++a) v.a[a] = __ldcg(hp + a);
        } else v = ldp<Real, VEC>(r + c);
        const Real w = sw[j];
#pragma unroll
        for (int a = 0; a < VEC; ++a) v.a[a] *= w;
        stp<Real, VEC>(r + c, v);
    }
}
template <class Real, int VEC>
__global__ void __launch_bounds__(256) k_band_restore_weight(const CgScal* S, long long sy, BandArgs<Real> A, int own0, int own1, const Real* __restrict__ sw,
                                                             Real* __restrict__ r, Real* __restrict__ z) {
    INK_PDL();
    if (S->done) return;
    const int side = (int)blockIdx.x >= A.nb ? 1 : 0;
    const BandSide<Real>& B = A.s[side];
    const unsigned int bid = blockIdx.x - side * A.nb;
    const unsigned int syv = (unsigned int)(sy / VEC), n = (unsigned int)(B.jb - B.ja) * syv;
    for (unsigned int t = bid * blockDim.x + threadIdx.x; t < n; t += A.nb * blockDim.x) {
        const int j = B.ja + (int)(t / syv);
        const long long e = (long long)t * VEC, c = (long long)B.ja * sy + e;
        if (j >= own0 && j < own1) stp<Real, VEC>(r + c, ldp<Real, VEC>(B.save + e));
        Pack<Real, VEC> zz = ldp<Real, VEC>(z + c);
        const Real w = sw[j];
#pragma unroll
        for (int a = 0; a < VEC; ++a) zz.a[a] *= w;
        stp<Real, VEC>(z + c, zz);
        if (B.push_dst) stp<Real, VEC>(B.push_dst + e, zz);   // peer-memory path: the weighted band goes straight into the neighbour's arena
    }
    if (B.push_dst) {   // the side's last block to finish publishes the sequence number (every block fenced its remote stores first)
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) {
            if (atomicAdd(B.ticket, 1u) == (unsigned int)A.nb - 1) {
                *B.ticket = 0;
                __threadfence_system();
                *reinterpret_cast<volatile unsigned long long*>(B.push_flag) = A.z_seq;
            }
        }
    }
}
// z += the neighbour's contribution on the band(s); adds <r, z> over this rank's own band rows to S->loc[2]; the last block of the
// launch — the iteration's last reducing kernel — also carries the folded all-reduce of <r,r>, <r,z> (ArFold)
template <class Real, int VEC>
__global__ void __launch_bounds__(256) k_band_combine(CgScal* S, long long sy, BandArgs<Real> A, int own0, int own1, const Real* __restrict__ r,
                                                      Real* __restrict__ z, ArFold fold) {
    INK_PDL();
    if (S->done) return;
    const int side = (int)blockIdx.x >= A.nb ? 1 : 0;
    const BandSide<Real>& B = A.s[side];
    const unsigned int bid = blockIdx.x - side * A.nb;
    if (B.wait_z) {   // peer-memory path: recv is this rank's arena, filled by the neighbour's k_band_restore_weight
        if (threadIdx.x == 0) p2p_spin(reinterpret_cast<const volatile unsigned long long*>(B.wait_z), A.z_seq);
        __syncthreads();
    }
    const unsigned int syv = (unsigned int)(sy / VEC), n = (unsigned int)(B.jb - B.ja) * syv;
    double acc = 0;
    for (unsigned int t = bid * blockDim.x + threadIdx.x; t < n; t += A.nb * blockDim.x) {
        const int j = B.ja + (int)(t / syv);
        const long long e = (long long)t * VEC, c = (long long)B.ja * sy + e;
        Pack<Real, VEC> v = ldp<Real, VEC>(z + c);
        const bool own = (j >= own0 && j < own1);
        Pack<Real, VEC> rc;
        if (own) rc = ldp<Real, VEC>(r + c);
#pragma unroll
        for (int a = 0; a < VEC; ++a) {
            v.a[a] += __ldcg(B.recv + e + a);
            if (own) acc += (double)rc.a[a] * (double)v.a[a];
        }
        stp<Real, VEC>(z + c, v);
    }
    double tot;
    if (block_sum_last(acc, S, 1, tot)) {
        if (threadIdx.x == 0) S->loc[2] += tot;
        if (fold.on && threadIdx.x < 32) p2p_fold(S, fold);
    }
}

}  // namespace ink
