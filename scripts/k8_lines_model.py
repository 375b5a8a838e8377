"""L1 wavefront model of the RK2 particle kernel (K8), the reasoning behind DESIGN.md section 4.

A global gather whose 32 lanes touch n distinct 128-byte lines costs about n L1 wavefronts (B300_MICROARCH.md); K8 issues
48 such gathers per particle-warp (2 samples x 3 components x 8 corners).  This script draws the bench's particle cloud
(uniform, `ppc` particles per cell), orders it the way the library bins it, and counts the distinct lines of every gather
of every warp — for whole-cell bins, for the quarter-column bins the library uses (particle_kernels.cuh: bin_key), and for
the planned shared-memory tile (coalesced tile loads + conflict-free shared loads).

    python scripts/k8_lines_model.py [X Y Z] [ppc] [shift]   # default 32 32 128, 2.4 particles per cell, midpoint shift 0.02 cells

Output of the defaults (measured on B200 with ncu: 3.0 lines per gather for whole-cell bins, 2.6 for quarter-column bins):
    whole-cell bins                3.6 lines per gather
    quarter-column bins (library)  2.8
    z-blocked quarter bins         2.65   (tried here only: 19 particles per (quarter, 32-cell z block) cannot fill a warp)
    tile of 2 columns              66 wavefronts per particle-warp against 135
"""
import sys

import numpy as np


def corner_offsets(px, py, pz, dims):
    """element offsets (internal layout: y slowest, z fastest) of the 8 corners of the three staggered samples"""
    X, Y, Z = dims
    out = []
    for sx, sy, sz in ((0.0, 0.5, 0.5), (0.5, 0.0, 0.5), (0.5, 0.5, 0.0)):     # u, v, w (src/systemHelpers.cpp:110-115)
        i = np.floor(px - sx).astype(np.int64)
        j = np.floor(py - sy).astype(np.int64)
        k = np.floor(pz - sz).astype(np.int64)
        for dj in (0, 1):
            for di in (0, 1):
                for dk in (0, 1):
                    ii, jj, kk = np.clip(i + di, 0, X - 1), np.clip(j + dj, 0, Y - 1), np.clip(k + dk, 0, Z - 1)
                    out.append((jj * X + ii) * Z + kk)
    return out                                                                   # 24 arrays of n offsets


def lines_per_gather(offsets, order):
    """mean number of distinct 128-byte lines (32 floats) per warp-wide gather"""
    tot, cnt = 0, 0
    n = len(order) // 32 * 32
    for off in offsets:
        lines = (off[order][:n] // 32).reshape(-1, 32)
        lines = np.sort(lines, axis=1)
        tot += (1 + (np.diff(lines, axis=1) != 0).sum(axis=1)).sum()
        cnt += lines.shape[0]
    return tot / cnt


def main():
    a = [float(x) for x in sys.argv[1:]]
    dims = tuple(int(x) for x in a[:3]) if len(a) >= 3 else (32, 32, 128)
    ppc = a[3] if len(a) >= 4 else 2.4
    shift = a[4] if len(a) >= 5 else 0.02
    X, Y, Z = dims
    rng = np.random.default_rng(0)
    n = int(ppc * X * Y * Z)
    p = rng.uniform(0.5, np.array(dims) - 0.5, size=(n, 3))
    # RK2 midpoint: |0.5 dt rho u| <= 0.5 * 0.2 * 0.95 = 0.095 cells under the CFL rule; typical values are a fraction of that
    mid = p + rng.uniform(-shift, shift, size=(n, 3))
    i, j, k = (np.floor(p[:, a]).astype(np.int64) for a in range(3))
    hx, hy = (p[:, 0] - i >= 0.5).astype(np.int64), (p[:, 1] - j >= 0.5).astype(np.int64)
    orders = {
        "unsorted": np.arange(n),
        "whole-cell bins": np.argsort((j * X + i) * Z + k, kind="stable"),
        "quarter-column bins (library)": np.argsort(((2 * j + hy) * (2 * X) + (2 * i + hx)) * Z + k, kind="stable"),
        "z-blocked (32) quarter bins": np.argsort((((j * X + i) * (Z // 32) + k // 32) * 4 + (2 * hy + hx)) * 32 + k % 32, kind="stable"),
    }
    offs = corner_offsets(p[:, 0], p[:, 1], p[:, 2], dims) + corner_offsets(mid[:, 0], mid[:, 1], mid[:, 2], dims)
    print(f"grid {dims}, {ppc} particles per cell, {n} particles, 48 gathers per particle-warp")
    for name, order in orders.items():
        l = lines_per_gather(offs, order)
        print(f"  {name:32s} {l:5.2f} lines per gather -> {48 * l:6.1f} L1 wavefronts per particle-warp")
    # shared-memory tile, block = cell columns (i .. i+cols-1, j): rows of Z floats both samples can touch (midpoint within 0.095)
    for cols in (1, 2, 4):
        rows = (cols + 3) * 3 + (cols + 2) * 4 + (cols + 2) * 3                  # u: i-1..i+cols+1 x j-1..j+1; v: i-1..i+cols x j-1..j+2; w: .. x j-1..j+1
        warps = cols * ppc * Z / 32
        ld = rows * Z / 32 / warps                                               # coalesced 128-byte loads, and as many shared stores
        print(f"  tile of {cols} column(s): {rows} rows, {rows * Z * 4 / 1024:5.1f} KB; per particle-warp {ld:4.1f} LDG + {ld:4.1f} STS + 48 LDS"
              f" = {2 * ld + 48:5.1f} wavefronts")


if __name__ == "__main__":
    main()
