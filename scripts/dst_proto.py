"""Prototype of the fast-Poisson preconditioner (git branch r2-dst, csrc/dst.cuh): the pressure operator M = -A is exactly
diagonalised by three orthonormal type-I sine transforms, applied as DENSE matrices (GEMMs) because the FFT sizes would be
2(n+1) = 258 / 514.  Two questions, answered on the CPU:

1. Is one fp32 application accurate enough?            python scripts/dst_proto.py accuracy
   (128, 256, 128): fp32 relative residual 6.8e-07 (CG tolerance 1e-6), max error vs the exact solution 1.4e-06
   (the reference's own fp32 SparseLU is 4e-06 away), 3.3e-08 after one refinement step.
2. How many PCG iterations with exact slab solves inside the weighted overlapping Schwarz preconditioner (multi-GPU)?
                                                        python scripts/dst_proto.py schwarz
   (64, 256, 64): 1 rank 1 iteration; 2 ranks 5; 4 ranks 6; 8 ranks 6   (trilinear V-cycle today: 11.5 / 12.1 / 12.5 / 12.4)
3. How wide should the overlap be at the weak-scaling shapes (256 rows per rank, 128 x 128 cross-section)?
                                                        python scripts/dst_proto.py overlap      (minutes)
   (128, 512, 128), 2 ranks: overlap 4 -> 8 iterations, 8 -> 6, 16 -> 4, 32 -> 2
   (128, 1024, 128), 4 ranks: overlap 4 -> 10, 8 -> 6, 16 -> 4, 32 -> 3
   With an exact block solve the count keeps falling with the overlap (it does not with the V-cycle): 32 rows it is.
4. Which tensor-core arithmetic would do for the six GEMMs (tcgen05 kind::tf32 / kind::f16, fp32 accumulation)?
                                                        python scripts/dst_proto.py precision
   (64, 128, 64), residual of one application | PCG iterations to 1e-6:
     fp32 4.7e-07 | 1      tf32 7.5e-04 | 2      3 x tf32 (split operands) 4.9e-07 | 1
     bf16 5.4e-03 | 3      3 x bf16 1.1e-05 | 2  6 x bf16 2.2e-07 | 1
   A single-pass TF32 transform is a good enough PRECONDITIONER: two CG iterations, each six tensor-core GEMMs that are then
   bound by their 48 B/cell of traffic, not by arithmetic.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import mg_proto as P  # noqa: E402


def S(n, dt=np.float64):
    a = np.arange(1, n + 1)
    return (np.sin(np.pi * np.outer(a, a) / (n + 1)) * np.sqrt(2.0 / (n + 1))).astype(dt)


def lam1(n):
    return 2 - 2 * np.cos(np.pi * np.arange(1, n + 1) / (n + 1))


def dst_solve(f, dt=np.float64):
    X, Y, Z = f.shape
    Sx, Sy, Sz = S(X, dt), S(Y, dt), S(Z, dt)
    g = f.astype(dt)
    g = np.einsum("ck,abk->abc", Sz, g, optimize=True).astype(dt)
    g = np.einsum("ai,ijk->ajk", Sx, g, optimize=True).astype(dt)
    g = np.einsum("bj,ajk->abk", Sy, g, optimize=True).astype(dt)
    g = (g / (lam1(X)[:, None, None] + lam1(Y)[None, :, None] + lam1(Z)[None, None, :]).astype(dt)).astype(dt)
    g = np.einsum("bj,ajk->abk", Sy, g, optimize=True).astype(dt)
    g = np.einsum("ai,ijk->ajk", Sx, g, optimize=True).astype(dt)
    g = np.einsum("ck,abk->abc", Sz, g, optimize=True).astype(dt)
    return g


def rhs(shape):
    u = np.random.default_rng(0).standard_normal((3,) + shape)
    return np.diff(u[0], axis=0, append=0) + np.diff(u[1], axis=1, append=0) + np.diff(u[2], axis=2, append=0)


def accuracy():
    for shape in [(32, 64, 32), (64, 128, 64), (128, 256, 128)]:
        f = rhs(shape)
        q64, q32 = dst_solve(f, np.float64), dst_solve(f, np.float32)
        res = lambda q: np.linalg.norm(f - P.M(q.astype(np.float64))) / np.linalg.norm(f)  # noqa: E731
        r = (f - P.M(q32.astype(np.float64))).astype(np.float32)
        q2 = q32 + dst_solve(r, np.float32)
        print(shape, "fp64 residual %.1e | fp32 residual %.1e, max error vs exact %.1e | after one refinement %.1e"
              % (res(q64), res(q32), np.abs(q32 - q64).max() / np.abs(q64).max(), res(q2)))


def schwarz():
    from test_mg_host import schwarz_weights

    def B(r, world, dt):
        z = np.zeros_like(r)
        for rank in range(world):
            je0, je1, w = schwarz_weights(r.shape[1], world, rank)
            sw = np.sqrt(w)[None, :, None]
            z[:, je0:je1] += sw[:, je0:je1] * dst_solve((sw * r)[:, je0:je1], dt)
        return z

    for shape in [(32, 128, 32), (64, 256, 64)]:
        f = rhs(shape)
        for world in (1, 2, 4, 8):
            x, it = P.pcg(f, lambda r: B(r, world, np.float32).astype(np.float64), rtol=1e-6)
            print(shape, "ranks", world, "PCG iterations", it, "true residual %.1e" % (np.linalg.norm(f - P.M(x)) / np.linalg.norm(f)))


def _round_mantissa(x, keep_bits):
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).copy()
    drop = 23 - keep_bits
    u = (u + np.uint32(1 << (drop - 1))) & np.uint32((0xFFFFFFFF >> drop) << drop)
    return u.view(np.float32)


def precision():
    def mm(spec, Sm, g, mode):
        Sm, g = Sm.astype(np.float32), g.astype(np.float32)
        if mode == "fp32":
            return np.einsum(spec, Sm, g, optimize=True).astype(np.float32)
        rnd = (lambda x: _round_mantissa(x, 10)) if "tf32" in mode else (lambda x: _round_mantissa(x, 7))
        Sh, gh = rnd(Sm), rnd(g)
        out = np.einsum(spec, Sh, gh, optimize=True)
        if "x" in mode:   # split operands: hi*hi + hi*lo + lo*hi (+ three more terms for "x6")
            Sl, gl = rnd(Sm - Sh), rnd(g - gh)
            out = out + np.einsum(spec, Sh, gl, optimize=True) + np.einsum(spec, Sl, gh, optimize=True)
            if mode.endswith("x6"):
                Sl2, gl2 = rnd(Sm - Sh - Sl), rnd(g - gh - gl)
                out = out + np.einsum(spec, Sl, gl, optimize=True) + np.einsum(spec, Sh, gl2, optimize=True) + np.einsum(spec, Sl2, gh, optimize=True)
        return out.astype(np.float32)

    def solve(f, mode):
        X, Y, Z = f.shape
        Sx, Sy, Sz = S(X), S(Y), S(Z)
        g = f.astype(np.float32)
        g = mm("ck,abk->abc", Sz, g, mode); g = mm("ai,ijk->ajk", Sx, g, mode); g = mm("bj,ajk->abk", Sy, g, mode)  # noqa: E702
        g = (g / (lam1(X)[:, None, None] + lam1(Y)[None, :, None] + lam1(Z)[None, None, :]).astype(np.float32)).astype(np.float32)
        g = mm("bj,ajk->abk", Sy, g, mode); g = mm("ai,ijk->ajk", Sx, g, mode); g = mm("ck,abk->abc", Sz, g, mode)  # noqa: E702
        return g

    shape = (64, 128, 64)
    f = rhs(shape)
    for mode in ("fp32", "tf32", "tf32x3", "bf16", "bf16x3", "bf16x6"):
        q = solve(f, mode)
        res = np.linalg.norm(f - P.M(q.astype(np.float64))) / np.linalg.norm(f)
        _, it = P.pcg(f, lambda r: solve(r, mode).astype(np.float64), rtol=1e-6)
        print("%-7s one application: residual %.1e | PCG iterations to 1e-6: %d" % (mode, res, it))


def overlap():
    def weights(Y, world, rank, ovl):
        lo = [Y * r // world for r in range(world + 1)]
        j0, j1 = lo[rank], lo[rank + 1]
        je0, je1 = (j0 - ovl if rank > 0 else j0), (j1 + ovl if rank + 1 < world else j1)
        w = np.zeros(Y)
        for j in range(je0, je1):
            v = 1.0
            if rank > 0 and j < j0 + ovl:
                v = min(v, (j - (j0 - ovl) + 0.5) / (2.0 * ovl))
            if rank + 1 < world and j >= j1 - ovl:
                v = min(v, ((j1 + ovl) - j - 0.5) / (2.0 * ovl))
            w[j] = v
        return je0, je1, w

    def B(r, world, ovl):
        z = np.zeros_like(r)
        for rank in range(world):
            je0, je1, w = weights(r.shape[1], world, rank, ovl)
            sw = np.sqrt(w)[None, :, None]
            z[:, je0:je1] += sw[:, je0:je1] * dst_solve((sw * r)[:, je0:je1], np.float64)
        return z

    for shape, world in [((128, 512, 128), 2), ((128, 1024, 128), 4)]:
        f = rhs(shape)
        for ovl in (4, 8, 16, 32):
            _, it = P.pcg(f, lambda r: B(r, world, ovl), rtol=1e-6)
            print(shape, "ranks", world, "overlap", ovl, "PCG iterations", it, flush=True)


if __name__ == "__main__":
    {"accuracy": accuracy, "schwarz": schwarz, "overlap": overlap, "precision": precision}[sys.argv[1] if len(sys.argv) > 1 else "accuracy"]()
