"""Seeded synthetic inputs shared by the parity tests (no dependency on /root/reference at run time)."""
import numpy as np

C1 = (8, 25, 8)      # BASELINE config 1 grid
C2 = (20, 45, 20)    # BASELINE config 2 grid
ODD = (7, 9, 7)      # ragged: no dimension divisible by 4 (scalar kernel paths); X == Z as the reference requires (Q1)
RECT = (6, 9, 5)     # X != Z: the reference's grid2mat is not a bijection there (Q1) -> oracle-only cases
THIN = (1, 3, 2)     # degenerate: single-cell extent in x (oracle-only)


def rand_field(dims, seed, scale=0.3, dtype=np.float32):
    """Smooth-ish random staggered velocity (sum of a few sines + noise), |u| ~ scale."""
    rng = np.random.default_rng(seed)
    X, Y, Z = dims
    i, j, k = np.meshgrid(np.arange(X), np.arange(Y), np.arange(Z), indexing="ij")
    out = []
    for c in range(3):
        a = rng.normal(size=3)
        ph = rng.uniform(0, 6.28, size=3)
        f = np.sin(0.7 * i * a[0] + ph[0]) + np.sin(0.5 * j * a[1] + ph[1]) + np.sin(0.9 * k * a[2] + ph[2])
        f = f / 3.0 + 0.25 * rng.normal(size=dims)
        out.append((scale * f).astype(dtype))
    return out


def box_particles(dims, n, seed, dtype=np.float32):
    """Uniform particles in the initParticles box of the reference (src/system.cpp:102-104)."""
    rng = np.random.default_rng(seed)
    X, Y, Z = dims
    lo = np.array([X / 2.0 - X / 5.0, Y - 1.5, Z / 2.0 - Z / 3.0])
    hi = np.array([X / 2.0 + X / 5.0, Y - 1.0, Z / 2.0 + Z / 3.0])
    return (lo + rng.random((n, 3)) * (hi - lo)).astype(dtype)


def uniform_particles(dims, n, seed, dtype=np.float32, margin=0.0):
    rng = np.random.default_rng(seed)
    d = np.array(dims, dtype=np.float64)
    return (margin + rng.random((n, 3)) * (d - 2 * margin)).astype(dtype)


def wild_particles(dims, n, seed, dtype=np.float32):
    """Particles including some outside the grid and some exactly on walls/cell faces (edge cases Q5, Q8, Q13)."""
    rng = np.random.default_rng(seed)
    d = np.array(dims, dtype=np.float64)
    p = -1.5 + rng.random((n, 3)) * (d + 3.0)
    p[: n // 8] = np.floor(p[: n // 8])          # exactly on cell faces
    p[n // 8: n // 6, 0] = d[0]                  # exactly on the high x wall
    p[n // 6: n // 5, 1] = 0.0                   # exactly on the low y wall
    return p.astype(dtype)


def rel_inf(a, b):
    """max |a-b| / max(|b|, tiny): field-scale relative error."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
