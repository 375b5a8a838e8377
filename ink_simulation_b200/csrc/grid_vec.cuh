// grid_vec.cuh — 128-bit vectorised versions of the MAC stencil kernels of grid_kernels.cuh (K3 viscosity, K4 divergence,
// K6 gradient subtract, K7a curl, K7b vorticity confinement) for grids with Z % 4 == 0 (every BASELINE grid).
//
// One thread owns 4 z-consecutive cells of one (i, j) row: the (i, j, k) decode is two 32-bit divisions per 4 cells
// instead of two 64-bit divisions per cell, every row access is one LDG.128 / STG.128, and the z neighbours come from
// registers (one scalar load at each end).  The scalar kernels ran at 33-56 % of the HBM roofline, issue-bound on the
// decode and on scalar loads (profiles/r1_d_launches_summary.txt).
//
// PER-CELL ARITHMETIC IS IDENTICAL to the scalar kernels — same operations in the same order (the translation unit is
// compiled -fmad=false) — so the fp32 results stay bit-exact against the compiled reference (tests/test_gpu_parity.py).
#pragma once
#include "cg.cuh"
#include "grid_kernels.cuh"

namespace ink {

template <class Real>
using V4 = Pack<Real, 4>;

// thread -> (i, j, k = 4*kv, c) inside the rows [g.j0, g.j1) of this launch
__device__ __forceinline__ bool vec_cell_of_thread(const Grid& g, unsigned int t, int& i, int& j, int& k, long long& c) {
    const unsigned int ZV = (unsigned int)g.Z >> 2;
    const unsigned int tot = (unsigned int)(g.j1 - g.j0) * (unsigned int)g.X * ZV;
    if (t >= tot) return false;
    const unsigned int row = t / ZV;
    k = (int)(t - row * ZV) << 2;
    const unsigned int jl = row / (unsigned int)g.X;
    i = (int)(row - jl * (unsigned int)g.X);
    j = g.j0 + (int)jl;
    c = (long long)j * g.sy + (long long)i * g.sx + k;
    return true;
}
template <class Real>
__device__ __forceinline__ V4<Real> v4_load_or_zero(bool cond, const Real* p) {
    if (cond) return ldp<Real, 4>(p);
    V4<Real> z; z.a[0] = z.a[1] = z.a[2] = z.a[3] = Real(0);
    return z;
}

// the 7-point neighbourhood of 4 cells of field f: rows x+-1, y+-1 as packs (zero outside), z ends as scalars
template <class Real>
struct Nbr4 {
    V4<Real> c, xp, xm, yp, ym;
    Real zl, zr;     // f[k-1], f[k+4] (zero outside)
    __device__ __forceinline__ Real zp(int a) const { return a < 3 ? c.a[a < 3 ? a + 1 : 0] : zr; }
    __device__ __forceinline__ Real zm(int a) const { return a > 0 ? c.a[a > 0 ? a - 1 : 0] : zl; }
};
template <class Real>
__device__ __forceinline__ Nbr4<Real> nbr4(const Real* __restrict__ f, const Grid& g, int i, int j, int k, long long c) {
    Nbr4<Real> n;
    n.c = ldp<Real, 4>(f + c);
    n.xp = v4_load_or_zero<Real>(i + 1 < g.X, f + c + g.sx);
    n.xm = v4_load_or_zero<Real>(i - 1 >= 0, f + c - g.sx);
    n.yp = v4_load_or_zero<Real>(j + 1 < g.Y, f + c + g.sy);
    n.ym = v4_load_or_zero<Real>(j - 1 >= 0, f + c - g.sy);
    n.zl = (k > 0) ? f[c - 1] : Real(0);
    n.zr = (k + 4 < g.Z) ? f[c + 4] : Real(0);
    return n;
}

// ------------------------------------------------------------------------------------------------
// K2b  k_dilate_add_gravity for 4 z-consecutive cells per thread (System::applyExternalForces, src/waterSolver.cpp:158-191; Q8).
// The scalar kernel walks the clipped Manhattan ball byte by byte (up to 63 one-byte loads per cell, two 64-bit divisions
// for the decode); here a seed row contributes through at most three aligned 32-bit words: each word becomes a 4-bit
// "non-zero byte" nibble, the three nibbles form a 12-bit window (cells k-4 .. k+7), the dilation along z by the row's
// remaining radius is shifts and ORs, and bits 4..7 are this thread's cells.  Rows are visited ring by ring from the
// centre, so inside a particle cloud the first row settles all four cells.  The mask is an OR over the same set of seed
// bytes as the scalar kernel's, and a marked cell receives the same single addition: results are bit-identical.
// Valid for R <= 4 (a window of +-4 cells around 4 cells is exactly the 12 bits).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned int nz_nibble(uint32_t x) {   // bit b = (byte b of x != 0)
    const uint32_t f = (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
    return ((f >> 7) | (f >> 14) | (f >> 21) | (f >> 28)) & 0xFu;
}
template <class Real>
__device__ __forceinline__ void add_marked4(Real* __restrict__ p, Real gv, unsigned int m) {
    V4<Real> q = ldp<Real, 4>(p);
#pragma unroll
    for (int a = 0; a < 4; ++a)
        if ((m >> a) & 1u) q.a[a] += gv;
    stp<Real, 4>(p, q);
}
template <class Real>
__global__ void __launch_bounds__(256) k_dilate_add_gravity_v4(Grid g, int R, const uint8_t* __restrict__ seed, uint8_t* __restrict__ mark,
                                                               Real gx, Real gy, Real gz, Real* __restrict__ u, Real* __restrict__ v, Real* __restrict__ w) {
    INK_PDL();
    int i, j, k; long long c;
    if (!vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const bool zl = k > 0, zr = k + 4 < g.Z;
    unsigned int acc = 0;
    for (int ring = 0; ring <= R && acc != 0xFu; ++ring) {       // all (di, dj) with |di| + |dj| == ring
        const int rk = R - ring;                                  // radius left for z
        for (int dj = -ring; dj <= ring && acc != 0xFu; ++dj) {
            const int jj = j + dj;
            if (jj < 0 || jj >= g.Y) continue;
            const int a = ring - abs(dj);
            for (int s = 0; s < (a ? 2 : 1); ++s) {               // di = -a, +a
                const int ii = i + (s ? a : -a);
                if (ii < 0 || ii >= g.X) continue;
                const uint32_t* __restrict__ row = reinterpret_cast<const uint32_t*>(seed + ((long long)jj * g.sy + (long long)ii * g.sx + k));
                unsigned int m12 = nz_nibble(__ldg(row)) << 4;
                if (rk > 0) {
                    if (zl) m12 |= nz_nibble(__ldg(row - 1));
                    if (zr) m12 |= nz_nibble(__ldg(row + 1)) << 8;
                }
                unsigned int d = m12;
                for (int r = 1; r <= rk; ++r) d |= (m12 << r) | (m12 >> r);
                acc |= (d >> 4) & 0xFu;
            }
        }
    }
    *reinterpret_cast<uint32_t*>(mark + c) = (acc & 1u) | ((acc & 2u) << 7) | ((acc & 4u) << 14) | ((acc & 8u) << 21);
    if (acc) {
        if (gx != Real(0)) add_marked4<Real>(u + c, gx, acc);   // adding an exact zero cannot change a value the reference would keep
        if (gy != Real(0)) add_marked4<Real>(v + c, gy, acc);
        if (gz != Real(0)) add_marked4<Real>(w + c, gz, acc);
    }
}

// K3 (k_viscosity): l = ((((((0 + xp) + xm) + yp) + ym) + zp) + zm) - 6 f;  out = f + s l
template <class Real>
__device__ __forceinline__ V4<Real> visc4(const Real* __restrict__ f, const Grid& g, Real s, int i, int j, int k, long long c) {
    const Nbr4<Real> n = nbr4<Real>(f, g, i, j, k, c);
    V4<Real> o;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        Real l = Real(0);
        l += n.xp.a[a]; l += n.xm.a[a]; l += n.yp.a[a]; l += n.ym.a[a]; l += n.zp(a); l += n.zm(a);
        l -= Real(6) * n.c.a[a];
        o.a[a] = n.c.a[a] + s * l;
    }
    return o;
}
template <class Real>
__global__ void __launch_bounds__(256) k_viscosity_v4(Grid g, Real s, const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                      Real* __restrict__ nu, Real* __restrict__ nv, Real* __restrict__ nw) {
    INK_PDL();
    int i, j, k; long long c;
    if (!vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    stp<Real, 4>(nu + c, visc4<Real>(u, g, s, i, j, k, c));
    stp<Real, 4>(nv + c, visc4<Real>(v, g, s, i, j, k, c));
    stp<Real, 4>(nw + c, visc4<Real>(w, g, s, i, j, k, c));
}

// K4 (k_divergence), same modes
template <class Real>
__global__ void __launch_bounds__(256) k_divergence_v4(Grid g, int mode, Real scale, Real patm, const Real* __restrict__ u, const Real* __restrict__ v,
                                                       const Real* __restrict__ w, Real* __restrict__ out) {
    INK_PDL();
    int i, j, k; long long c;
    if (!vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const V4<Real> uc = ldp<Real, 4>(u + c), vc = ldp<Real, 4>(v + c), wc = ldp<Real, 4>(w + c);
    const bool ip = i + 1 < g.X, jp = j + 1 < g.Y;
    const V4<Real> ux = v4_load_or_zero<Real>(ip, u + c + g.sx), vy = v4_load_or_zero<Real>(jp, v + c + g.sy);
    const bool zr_in = k + 4 < g.Z;
    const Real wzr = zr_in ? w[c + 4] : Real(0);
    const int nxy = (ip ? 1 : 0) + (i - 1 >= 0 ? 1 : 0) + (jp ? 1 : 0) + (j - 1 >= 0 ? 1 : 0);
    V4<Real> o;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        Real d = -uc.a[a] - vc.a[a] - wc.a[a];
        if (ip) d += ux.a[a];
        if (jp) d += vy.a[a];
        const bool kp = (a < 3) || zr_in;
        if (kp) d += (a < 3) ? wc.a[a < 3 ? a + 1 : 0] : wzr;
        if (mode == 0) o.a[a] = d;
        else if (mode == 1) {
            const bool km = (a > 0) || (k > 0);
            const int ki = 6 - (nxy + (kp ? 1 : 0) + (km ? 1 : 0));
            o.a[a] = scale * d - ki * patm;
        } else o.a[a] = -(scale * d);
    }
    stp<Real, 4>(out + c, o);
}

// K6 (k_grad_subtract): backward differences, low-side ghost = the literal 1
template <class Real>
__global__ void __launch_bounds__(256) k_grad_subtract_v4(Grid g, Real s, const Real* __restrict__ p, Real* __restrict__ u, Real* __restrict__ v, Real* __restrict__ w) {
    INK_PDL();
    int i, j, k; long long c;
    if (!vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const V4<Real> pc = ldp<Real, 4>(p + c);
    const bool im = i - 1 >= 0, jm = j - 1 >= 0;
    V4<Real> px, py;
    if (im) px = ldp<Real, 4>(p + c - g.sx);
    if (jm) py = ldp<Real, 4>(p + c - g.sy);
    const Real pzl = (k > 0) ? p[c - 1] : Real(1);
    V4<Real> uc = ldp<Real, 4>(u + c), vc = ldp<Real, 4>(v + c), wc = ldp<Real, 4>(w + c);
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const Real val = pc.a[a];
        Real gx = val, gy = val, gz = val;
        gx -= im ? px.a[a] : Real(1);
        gy -= jm ? py.a[a] : Real(1);
        gz -= (a > 0) ? pc.a[a > 0 ? a - 1 : 0] : pzl;
        uc.a[a] -= s * gx; vc.a[a] -= s * gy; wc.a[a] -= s * gz;
    }
    stp<Real, 4>(u + c, uc); stp<Real, 4>(v + c, vc); stp<Real, 4>(w + c, wc);
}

// K7a (k_curl) + |curl| : the norm every neighbour of K7b needs is computed ONCE per cell here (same norm3 on the same
// three values => bit-identical to what k_vorticity recomputes 6 times per cell)
template <class Real>
__global__ void __launch_bounds__(256) k_curl_v4(Grid g, const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                 Real* __restrict__ cx, Real* __restrict__ cy, Real* __restrict__ cz, Real* __restrict__ cn) {
    INK_PDL();
    int i, j, k; long long c;
    if (!vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) return;
    const bool ip = i + 1 < g.X, im = i - 1 >= 0, jp = j + 1 < g.Y, jm = j - 1 >= 0;
    const V4<Real> uc = ldp<Real, 4>(u + c), vc = ldp<Real, 4>(v + c);
    const V4<Real> w_yp = v4_load_or_zero<Real>(jp, w + c + g.sy), w_ym = v4_load_or_zero<Real>(jm, w + c - g.sy);
    const V4<Real> w_xp = v4_load_or_zero<Real>(ip, w + c + g.sx), w_xm = v4_load_or_zero<Real>(im, w + c - g.sx);
    const V4<Real> v_xp = v4_load_or_zero<Real>(ip, v + c + g.sx), v_xm = v4_load_or_zero<Real>(im, v + c - g.sx);
    const V4<Real> u_yp = v4_load_or_zero<Real>(jp, u + c + g.sy), u_ym = v4_load_or_zero<Real>(jm, u + c - g.sy);
    const bool zl_in = k > 0, zr_in = k + 4 < g.Z;
    const Real u_zl = zl_in ? u[c - 1] : Real(0), u_zr = zr_in ? u[c + 4] : Real(0);
    const Real v_zl = zl_in ? v[c - 1] : Real(0), v_zr = zr_in ? v[c + 4] : Real(0);
    V4<Real> ox, oy, oz, on;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const Real v_zp = (q < 3) ? vc.a[q < 3 ? q + 1 : 0] : v_zr, v_zm = (q > 0) ? vc.a[q > 0 ? q - 1 : 0] : v_zl;
        const Real u_zp = (q < 3) ? uc.a[q < 3 ? q + 1 : 0] : u_zr, u_zm = (q > 0) ? uc.a[q > 0 ? q - 1 : 0] : u_zl;
        Real a = Real(0), b = Real(0), d = Real(0);
        a += w_yp.a[q]; a -= w_ym.a[q]; a -= v_zp; a += v_zm;
        b += u_zp; b -= u_zm; b -= w_xp.a[q]; b += w_xm.a[q];
        d += v_xp.a[q]; d -= v_xm.a[q]; d -= u_yp.a[q]; d += u_ym.a[q];
        ox.a[q] = a; oy.a[q] = b; oz.a[q] = d; on.a[q] = norm3(a, b, d);
    }
    stp<Real, 4>(cx + c, ox); stp<Real, 4>(cy + c, oy); stp<Real, 4>(cz + c, oz); stp<Real, 4>(cn + c, on);
}

// K7b (k_vorticity) with the precomputed |curl| (+ fused K0: max of the new |velocity|)
template <class Real>
__global__ void __launch_bounds__(256) k_vorticity_v4(Grid g, Real dt, Real kvort, const Real* __restrict__ cx, const Real* __restrict__ cy,
                                                      const Real* __restrict__ cz, const Real* __restrict__ cn, Real* __restrict__ u, Real* __restrict__ v,
                                                      Real* __restrict__ w, unsigned int* maxspeed) {
    INK_PDL();
    int i, j, k; long long c;
    float speed = 0.f;
    if (vec_cell_of_thread(g, blockIdx.x * blockDim.x + threadIdx.x, i, j, k, c)) {
        const V4<Real> wx = ldp<Real, 4>(cx + c), wy = ldp<Real, 4>(cy + c), wz = ldp<Real, 4>(cz + c), nc = ldp<Real, 4>(cn + c);
        V4<Real> uc = ldp<Real, 4>(u + c), vc = ldp<Real, 4>(v + c), wc = ldp<Real, 4>(w + c);
        const V4<Real> n_xp = v4_load_or_zero<Real>(i + 1 < g.X, cn + c + g.sx), n_xm = v4_load_or_zero<Real>(i - 1 >= 0, cn + c - g.sx);
        const V4<Real> n_yp = v4_load_or_zero<Real>(j + 1 < g.Y, cn + c + g.sy), n_ym = v4_load_or_zero<Real>(j - 1 >= 0, cn + c - g.sy);
        const Real n_zr = (k + 4 < g.Z) ? cn[c + 4] : Real(0);
        bool any = false;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (!(wx.a[q] == Real(0) && wy.a[q] == Real(0) && wz.a[q] == Real(0))) {
                Real g0 = Real(0), g1 = Real(0), g2 = Real(0);
                g0 += n_xp.a[q]; g0 -= n_xm.a[q];
                g1 += n_yp.a[q]; g1 -= n_ym.a[q];
                g2 += (q < 3) ? nc.a[q < 3 ? q + 1 : 0] : n_zr;
                g2 -= Real(0);
                const Real nrm = nc.a[q];
                const Real Nx = g0 / nrm, Ny = g1 / nrm, Nz = g2 / nrm;
                const Real Fx = kvort * (Ny * wz.a[q] - Nz * wy.a[q]);
                const Real Fy = kvort * (Nz * wx.a[q] - Nx * wz.a[q]);
                const Real Fz = kvort * (Nx * wy.a[q] - Ny * wx.a[q]);
                uc.a[q] += dt * Fx; vc.a[q] += dt * Fy; wc.a[q] += dt * Fz;
                any = true;
            }
            speed = fmaxf(speed, (float)norm3(uc.a[q], vc.a[q], wc.a[q]));
        }
        if (any) { stp<Real, 4>(u + c, uc); stp<Real, 4>(v + c, vc); stp<Real, 4>(w + c, wc); }
    }
    block_max_to_global(speed, maxspeed);
}

}  // namespace ink
