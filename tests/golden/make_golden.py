"""Generates tests/golden/ref_<X>x<Y>x<Z>.npz from the COMPILED REFERENCE (oracle/_ref/libinkref.so, built by
oracle/build_ref.sh from /root/reference/src).  Run here (where /root/reference exists):

    python tests/golden/make_golden.py

Each file holds seeded inputs and the reference's outputs phase by phase, each phase starting from the SAME
injected state (never a long free run: the scheme is chaotic).  These are the pins of the oracle:
the reference itself ships no tests or golden vectors.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle  # noqa: E402
import util  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(dims, seed, n_part):
    oracle.build()
    R = oracle.RefSystem(*dims)
    u, v, w = util.rand_field(dims, seed)
    out = {"dims": np.array(dims), "u": u, "v": v, "w": w}

    def fresh(scale=1.0, particles=None):
        R.set_velocity(u * np.float32(scale), v * np.float32(scale), w * np.float32(scale))
        R.set_particles(particles if particles is not None else part)

    part = util.box_particles(dims, n_part, seed + 1)
    # force footprint test: a few particles scattered inside the grid
    part_sparse = util.uniform_particles(dims, 12, seed + 2)
    wild = util.wild_particles(dims, 400, seed + 3)
    out.update(part=part, part_sparse=part_sparse, wild=wild)

    fresh()
    out["sample_wild"] = R.sample_velocity(wild)
    out["dt_035"] = np.float64(R.calc_dt(0.35))
    out["dt_1e9"] = np.float64(R.calc_dt(1e9))
    out["div"] = R.divergence()

    fresh(); R.advect(0.2); out["advect_small"] = np.stack(R.get_velocity())
    fresh(scale=8.0); R.advect(0.9); out["advect_big"] = np.stack(R.get_velocity())   # displacements of several cells
    fresh(particles=part_sparse); R.forces(0.25); out["forces_sparse"] = np.stack(R.get_velocity())
    fresh(); R.forces(0.25); out["forces_box"] = np.stack(R.get_velocity())
    fresh(); R.viscosity(0.17); out["viscosity"] = np.stack(R.get_velocity())
    fresh(); out["pressure_p"] = R.calc_pressure(0.17)
    fresh(); R.pressure(0.17); out["pressure_vel"] = np.stack(R.get_velocity()); out["pressure_div"] = R.divergence()
    fresh(); R.vorticity(0.17); out["vorticity"] = np.stack(R.get_velocity()); out["curl"] = np.stack(R.get_field("curl"))
    fresh(); R.particles(0.2)
    pos, vel = R.get_particles(); out["particles_pos"] = pos; out["particles_vel"] = vel
    fresh(particles=wild); R.particles(0.35)
    pos, vel = R.get_particles(); out["wild_pos"] = pos; out["wild_vel"] = vel
    # three full System::solve steps from the injected state (dt, field and particles after each)
    fresh()
    for s in range(3):
        out[f"solve{s}_dt"] = np.float64(R.solve(0.35))
        out[f"solve{s}_vel"] = np.stack(R.get_velocity())
        out[f"solve{s}_pos"] = R.get_particles()[0]
    R.close()
    path = os.path.join(HERE, "ref_%dx%dx%d.npz" % dims)
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    if not os.path.isdir("/root/reference/src"):
        sys.exit("needs /root/reference (the golden files are committed; regenerate only in the build container)")
    make(util.C1, 11, 2000)
    make(util.ODD, 23, 500)
    make((14, 30, 14), 37, 1500)
    make(util.C2, 53, 100000)   # BASELINE config 2 (20x45x20, 100 k particles): the largest scene the reference itself runs (17 s LU set-up)
