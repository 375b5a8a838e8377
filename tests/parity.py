"""Implementation-agnostic parity checks against the golden vectors of the COMPILED REFERENCE (tests/golden/).

`make(dims)` must return an object with the OracleSystem / ink_simulation_b200.System phase API.  Used twice:
with the CPU oracle (-m "not gpu": pins the oracle) and with the CUDA path through the C ABI (-m gpu).

Tolerances (BASELINE.json north_star / SURVEY §8c):
  * every phase except the pressure solve: BIT-EXACT in fp32 (same operation order, no FMA contraction);
  * pressure: the reference's own fp32 SparseLU is ~4e-6 (relative, in p) away from the exact solution, so
    p is compared at 2e-5 and the projected velocity at 1e-5 of the field scale (north_star: 1e-5 relative);
  * particles after a full solve: 1e-6 relative per step (north_star).
"""
import glob
import os

import numpy as np

from util import rel_inf

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_*.npz")))
P_TOL = 2e-5
VEL_TOL = 1e-5
POS_TOL = 1e-6
# END-OF-STEP velocity (after vorticity confinement).  The north_star's 1e-5 is on the PROJECTED field (check_pressure) and holds
# everywhere.  At the end of a step it holds too, with ONE exception: step 0 of the 20x45x20 fixture.  The confinement force
# N x omega with N = grad|omega| / |omega| is not normalised and amplifies pressure round-off where |omega| is small;
# the reference's own fp32 SparseLU is 6.0e-6 (p) / 3.4e-6 (projected velocity) away from the exact solution there, which the rough
# random injected field turns into 2.09e-5 at the end of step 0 (3.0e-6 / 2.0e-6 on the developed fields of steps 1-2).  That
# number is the REFERENCE's noise: it is what the oracle with the exact fp64 DST solve measures against the fixture
# (MEASURED_STEP_VEL, asserted to within 10 % by the oracle test so that drift shows), and check_solve's `exact` argument holds the
# implementation under test to 1e-5 against that exact-solve oracle on the very same step.
STEP_VEL_TOL = {((20, 45, 20), 0): 3e-5}
MEASURED_STEP_VEL = {((20, 45, 20), 0): 2.09e-5}     # oracle (exact solve) vs compiled reference, tests/test_oracle_golden.py


def load(path):
    return dict(np.load(path))


def _eq(name, got, want):
    got = np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, name
    if not np.array_equal(got.astype(np.float32), want):
        bad = np.argwhere(got.astype(np.float32) != want)
        raise AssertionError(f"{name}: {len(bad)} of {want.size} values differ; first at {bad[0]}: "
                             f"{got[tuple(bad[0])]!r} vs {want[tuple(bad[0])]!r}; max abs diff "
                             f"{np.max(np.abs(got.astype(np.float64) - want))}")


def check_phases_exact(make, G):
    """fp32, strict arithmetic: everything but the pressure solve must equal the reference bit for bit."""
    dims = tuple(int(x) for x in G["dims"])
    u, v, w = G["u"], G["v"], G["w"]

    def fresh(scale=1.0, particles=None):
        S = make(dims)
        S.set_velocity(u * np.float32(scale), v * np.float32(scale), w * np.float32(scale))
        S.set_particles(G["part"] if particles is None else particles)
        return S

    S = fresh()
    _eq("sample_wild", S.sample_velocity(G["wild"]), G["sample_wild"])
    assert S.calc_dt(0.35) == float(G["dt_035"])
    assert S.calc_dt(1e9) == float(G["dt_1e9"])
    _eq("div", S.divergence(), G["div"])
    S = fresh(); S.advect(0.2); _eq("advect_small", np.stack(S.get_velocity()), G["advect_small"])
    S = fresh(8.0); S.advect(0.9); _eq("advect_big", np.stack(S.get_velocity()), G["advect_big"])
    S = fresh(particles=G["part_sparse"]); S.forces(0.25); _eq("forces_sparse", np.stack(S.get_velocity()), G["forces_sparse"])
    S = fresh(); S.forces(0.25); _eq("forces_box", np.stack(S.get_velocity()), G["forces_box"])
    S = fresh(); S.viscosity(0.17); _eq("viscosity", np.stack(S.get_velocity()), G["viscosity"])
    S = fresh(); S.vorticity(0.17)
    _eq("curl", np.stack(S.get_field("curl")), G["curl"])
    _eq("vorticity", np.stack(S.get_velocity()), G["vorticity"])
    S = fresh(); S.particles(0.2)
    pos, vel = S.get_particles()
    _eq("particles_pos", pos, G["particles_pos"]); _eq("particles_vel", vel, G["particles_vel"])
    S = fresh(particles=G["wild"]); S.particles(0.35)
    pos, vel = S.get_particles()
    _eq("wild_pos", pos, G["wild_pos"]); _eq("wild_vel", vel, G["wild_vel"])
    # gradient subtract with the REFERENCE's pressure: isolates getGradient/applyPressure arithmetic from the solver
    S = fresh(); S.pressure_given_p(0.17, G["pressure_p"]); _eq("pressure_given_p", np.stack(S.get_velocity()), G["pressure_vel"])


def check_pressure(make, G, p_tol=P_TOL, vel_tol=VEL_TOL):
    dims = tuple(int(x) for x in G["dims"])
    S = make(dims)
    S.set_velocity(G["u"], G["v"], G["w"])
    S.set_particles(G["part"])
    S.pressure(0.17)
    ep = rel_inf(S.get_pressure(), G["pressure_p"])
    ev = rel_inf(np.stack(S.get_velocity()), G["pressure_vel"])
    assert ep <= p_tol, f"pressure rel err {ep}"
    assert ev <= vel_tol, f"projected velocity rel err {ev}"
    # post-projection divergence (reference's own getDivergence) no worse than the reference's (north_star)
    d = np.abs(S.divergence()).max()
    dref = np.abs(G["pressure_div"]).max()
    assert d <= dref * (1 + 1e-3) + 1e-6, (d, dref)
    return ep, ev


def check_solve(make, G, vel_tol=None, pos_tol=POS_TOL, exact=None):
    """Three System::solve steps.  Step 0 starts from identical state; later steps are re-synchronised to the
    reference's state first (parity is judged per step from identical input).
    exact: optional factory of the oracle with the exact pressure solve; the same step must then ALSO agree with it at VEL_TOL /
    pos_tol, so that a loosened STEP_VEL_TOL entry only ever covers the reference's own solver noise."""
    dims = tuple(int(x) for x in G["dims"])
    errs = []
    for s in range(3):
        tol = vel_tol if vel_tol is not None else STEP_VEL_TOL.get((dims, s), VEL_TOL)
        todo = [make(dims)] + ([exact(dims)] if exact is not None else [])
        for S in todo:
            if s == 0:
                S.set_velocity(G["u"], G["v"], G["w"]); S.set_particles(G["part"])
            else:
                S.set_velocity(*G[f"solve{s-1}_vel"]); S.set_particles(G[f"solve{s-1}_pos"])
        S = todo[0]
        dt = S.solve(0.35)
        assert dt == float(G[f"solve{s}_dt"]), (s, dt, float(G[f"solve{s}_dt"]))
        vel = np.stack(S.get_velocity())
        ev = rel_inf(vel, G[f"solve{s}_vel"])
        pos = S.get_particles()[0]
        ref = G[f"solve{s}_pos"]
        ep = float(np.max(np.abs(pos.astype(np.float64) - ref) / np.maximum(np.abs(ref), 1.0)))
        assert ev <= tol, f"step {s}: velocity rel err {ev}"
        assert ep <= pos_tol, f"step {s}: particle rel err {ep}"
        if exact is not None:
            E = todo[1]
            assert E.solve(0.35) == dt
            ex = rel_inf(vel, np.stack(E.get_velocity()))
            epos = E.get_particles()[0]
            exp = float(np.max(np.abs(pos.astype(np.float64) - epos) / np.maximum(np.abs(epos), 1.0)))
            assert ex <= VEL_TOL, f"step {s}: velocity rel err {ex} against the exact-solve oracle"
            assert exp <= pos_tol, f"step {s}: particle rel err {exp} against the exact-solve oracle"
        errs.append((ev, ep))
    return errs
