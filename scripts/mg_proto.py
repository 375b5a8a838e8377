#!/usr/bin/env python
"""numpy prototype of the MG-preconditioned CG pressure solve (design exploration for csrc/mg.cuh; CPU only).
M x = 6x - sum(in-grid nbrs) (zero Dirichlet ghosts).  Level operator s_l * M_l, s_l = 2^l.
usage: mg_proto.py X Y Z [variant ...]   variants: const (shipping), tri (trilinear P, R = P^T)"""
import sys
import time

import numpy as np


def M(x):
    p = np.pad(x, 1)
    return 6 * x - (p[2:, 1:-1, 1:-1] + p[:-2, 1:-1, 1:-1] + p[1:-1, 2:, 1:-1] + p[1:-1, :-2, 1:-1] + p[1:-1, 1:-1, 2:] + p[1:-1, 1:-1, :-2])


def restrict_const(r):
    X, Y, Z = r.shape
    p = np.pad(r, ((0, X % 2), (0, Y % 2), (0, Z % 2)))
    return p.reshape((X + 1) // 2, 2, (Y + 1) // 2, 2, (Z + 1) // 2, 2).sum(axis=(1, 3, 5))


def prolong_const(e, shape):
    f = np.repeat(np.repeat(np.repeat(e, 2, 0), 2, 1), 2, 2)
    return f[:shape[0], :shape[1], :shape[2]]


def _p1d(n_f):
    """cell-centred linear interpolation matrix (n_f x n_c), zero ghost beyond both ends."""
    n_c = (n_f + 1) // 2
    P = np.zeros((n_f, n_c))
    for f in range(n_f):
        c = f // 2
        o = c + (1 if f % 2 else -1)
        P[f, c] = 0.75
        if 0 <= o < n_c:
            P[f, o] = 0.25
    return P


def prolong_tri(e, shape):
    Px, Py, Pz = (_p1d(n) for n in shape)
    return np.einsum('ia,jb,kc,abc->ijk', Px, Py, Pz, e, optimize=True)


def restrict_tri(r):
    Px, Py, Pz = (_p1d(n) for n in r.shape)
    return np.einsum('ia,jb,kc,ijk->abc', Px, Py, Pz, r, optimize=True)


class MG:
    def __init__(self, shape, variant, omega=0.8, nu=2, coarse_cells=2, rscale=1.0):
        self.variant, self.omega, self.nu, self.rscale = variant, omega, nu, rscale
        self.shapes = [tuple(shape)]
        while np.prod(self.shapes[-1]) > coarse_cells:
            s = self.shapes[-1]
            self.shapes.append(tuple((n + 1) // 2 for n in s))

    def smooth(self, x, r, s, n):
        c = self.omega / (6 * s)
        for _ in range(n):
            x = c * r if x is None else x + c * (r - s * M(x))
        return x

    def cycle(self, r, l=0):
        s = 2.0 ** l
        if l == len(self.shapes) - 1:
            return self.smooth(None, r, s, 8)
        x = self.smooth(None, r, s, self.nu)
        res = r - s * M(x)
        if self.variant == 'tri':
            rc = restrict_tri(res) * self.rscale
            e = self.cycle(rc, l + 1)
            x = x + prolong_tri(e, r.shape)
        else:
            rc = restrict_const(res)
            e = self.cycle(rc, l + 1)
            x = x + prolong_const(e, r.shape)
        return self.smooth(x, r, s, self.nu)


def pcg(f, B, rtol=1e-6, maxit=200):
    x = np.zeros_like(f)
    r = f.copy()
    z = B(r)
    d = z.copy()
    rz = float(np.vdot(r, z))
    bb = float(np.vdot(f, f))
    for it in range(1, maxit + 1):
        w = M(d)
        a = rz / float(np.vdot(d, w))
        x += a * d
        r -= a * w
        rr = float(np.vdot(r, r))
        if rr <= rtol * rtol * bb:
            return x, it
        z = B(r)
        rz_new = float(np.vdot(r, z))
        d = z + (rz_new / rz) * d
        rz = rz_new
    return x, maxit


if __name__ == '__main__':
    shape = tuple(int(a) for a in sys.argv[1:4])
    variants = sys.argv[4:] or ['const', 'tri']
    rng = np.random.default_rng(0)
    u = rng.standard_normal((3,) + shape)
    f = (np.diff(u[0], axis=0, append=0) + np.diff(u[1], axis=1, append=0) + np.diff(u[2], axis=2, append=0))
    for v in variants:
        name, *kv = v.split(',')
        kw = {k: float(val) if '.' in val else int(val) for k, val in (a.split('=') for a in kv)}
        mg = MG(shape, name, **kw)
        t = time.time()
        x, it = pcg(f, mg.cycle)
        print(f"{v:40s} iters {it:3d}  true rel res {np.linalg.norm(f - M(x)) / np.linalg.norm(f):.2e}  ({time.time() - t:.1f} s)")
