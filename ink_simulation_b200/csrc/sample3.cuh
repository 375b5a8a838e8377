// sample3.cuh — fused three-component staggered trilinear sample: System::getVelocity
// (reference src/systemHelpers.cpp:110-115) = 3 x getInterpolatedValue (:158-209), for the RK2 particle kernel.
//
// Why a second implementation beside get_velocity<> (grid_kernels.cuh): the particle kernel is instruction-issue
// bound, not HBM bound, until the per-particle SASS count drops well below ~350 (ncu r1_a: 1094 warp instructions
// per particle-warp, 75 % issue-slot utilisation, DRAM 10 %).  This version
//   * shares the six floor()/fraction computations between the three components (u, v, w use 3 of them each),
//   * uses 32-bit element offsets (one IMAD.WIDE per row pointer, immediate +1 for the k+1 corner),
//   * takes a predicate-free path when all 24 corners are inside the grid (true for every particle that is not
//     within ~1.5 cells of a wall) and falls back to the generic zero-extension samplers otherwise.
// STRICT arithmetic is operation-for-operation the generic interp_strict (hence the reference); FAST is interp_fast.
#pragma once
#include "common.cuh"
#include "grid_kernels.cuh"

namespace ink {

template <class Real>
__device__ __forceinline__ void load8(const Real* __restrict__ p, int sx, int sy, Real (&f)[8]) {
    const Real* __restrict__ px = p + sx;
    const Real* __restrict__ py = p + sy;
    const Real* __restrict__ pxy = py + sx;
    f[0] = __ldg(p);   f[4] = __ldg(p + 1);      // 000 001
    f[1] = __ldg(px);  f[5] = __ldg(px + 1);     // 100 101
    f[2] = __ldg(py);  f[6] = __ldg(py + 1);     // 010 011
    f[3] = __ldg(pxy); f[7] = __ldg(pxy + 1);    // 110 111
}

// reference order 000,100,010,110,001,101,011,111; weight = (wx*wy)*wz; total += weight*f; result = total/weightAccum
template <class Real>
__device__ __forceinline__ Real eval_strict(const Real (&f)[8], Real x, Real fi, Real y, Real fj, Real z, Real fk) {
    const Real fi1 = fi + Real(1), fj1 = fj + Real(1), fk1 = fk + Real(1);
    const Real wx0 = fi1 - x, wx1 = x - fi, wy0 = fj1 - y, wy1 = y - fj, wz0 = fk1 - z, wz1 = z - fk;
    const Real w00 = wx0 * wy0, w10 = wx1 * wy0, w01 = wx0 * wy1, w11 = wx1 * wy1;
    Real tot = Real(0), wacc, w;
    w = w00 * wz0; tot += w * f[0]; wacc = w;
    w = w10 * wz0; tot += w * f[1]; wacc += w;
    w = w01 * wz0; tot += w * f[2]; wacc += w;
    w = w11 * wz0; tot += w * f[3]; wacc += w;
    w = w00 * wz1; tot += w * f[4]; wacc += w;
    w = w10 * wz1; tot += w * f[5]; wacc += w;
    w = w01 * wz1; tot += w * f[6]; wacc += w;
    w = w11 * wz1; tot += w * f[7]; wacc += w;
    if (wacc == Real(0)) return Real(0);
    return tot / wacc;
}

template <class Real>
__device__ __forceinline__ Real eval_fast(const Real (&f)[8], Real tx, Real ty, Real tz) {
    const Real a00 = rfma(tz, f[4] - f[0], f[0]);
    const Real a10 = rfma(tz, f[5] - f[1], f[1]);
    const Real a01 = rfma(tz, f[6] - f[2], f[2]);
    const Real a11 = rfma(tz, f[7] - f[3], f[3]);
    const Real b0 = rfma(tx, a10 - a00, a00);
    const Real b1 = rfma(tx, a11 - a01, a01);
    return rfma(ty, b1 - b0, b0);
}

template <class Real, bool FAST>
__device__ __forceinline__ void get_velocity_generic(const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                                  const Grid& g, const Consts<Real>& C, Real px, Real py, Real pz, Real* out) {
    Real a, b, c;
    get_velocity<Real, FAST>(u, v, w, g, C, px, py, pz, a, b, c);
    out[0] = a; out[1] = b; out[2] = c;
}

// UNIT_H: CELL_DIM == 1 (every README / BASELINE scene): pos / CELL_DIM is the identity and isInBounds() on indices
// equals the index test, which is what makes the interior shortcut exact.
template <class Real, bool FAST, bool UNIT_H>
__device__ __forceinline__ void get_velocity3(const Real* __restrict__ u, const Real* __restrict__ v, const Real* __restrict__ w,
                                              const Grid& g, const Consts<Real>& C, Real px, Real py, Real pz, Real& ox, Real& oy, Real& oz) {
    if (UNIT_H) {
        const Real xh = px - Real(0.5f), yh = py - Real(0.5f), zh = pz - Real(0.5f);
        const Real fx = rfloor(px), fy = rfloor(py), fz = rfloor(pz);
        const Real fxh = rfloor(xh), fyh = rfloor(yh), fzh = rfloor(zh);
        const int ix = (int)fx, jy = (int)fy, kz = (int)fz, ixh = (int)fxh, jyh = (int)fyh, kzh = (int)fzh;
        const bool interior = ((ixh | jyh | kzh) >= 0) && (ix <= g.X - 2) && (jy <= g.Y - 2) && (kz <= g.Z - 2);
        if (interior) {
            const int sx = g.sx, sy = (int)g.sy;
            const int rx = ix * sx, rxh = ixh * sx, ry = jy * sy, ryh = jyh * sy;
            Real f[8];
            if (FAST) {
                const Real tx = px - fx, ty = py - fy, tz = pz - fz, txh = xh - fxh, tyh = yh - fyh, tzh = zh - fzh;
                load8(u + (ryh + rx + kzh), sx, sy, f);  ox = eval_fast(f, tx, tyh, tzh);
                load8(v + (ry + rxh + kzh), sx, sy, f);  oy = eval_fast(f, txh, ty, tzh);
                load8(w + (ryh + rxh + kz), sx, sy, f);  oz = eval_fast(f, txh, tyh, tz);
            } else {
                load8(u + (ryh + rx + kzh), sx, sy, f);  ox = eval_strict(f, px, fx, yh, fyh, zh, fzh);
                load8(v + (ry + rxh + kzh), sx, sy, f);  oy = eval_strict(f, xh, fxh, py, fy, zh, fzh);
                load8(w + (ryh + rxh + kz), sx, sy, f);  oz = eval_strict(f, xh, fxh, yh, fyh, pz, fz);
            }
            return;
        }
    }
    Real o[3];
    get_velocity_generic<Real, FAST>(u, v, w, g, C, px, py, pz, o);
    ox = o[0]; oy = o[1]; oz = o[2];
}

}  // namespace ink
