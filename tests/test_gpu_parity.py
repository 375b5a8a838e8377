"""-m gpu: the CUDA path, called through the C ABI (libinkb200.so via ctypes), against
  (1) the golden vectors of the compiled reference (tests/golden/), and
  (2) the CPU oracle on fresh seeded inputs, including shapes the reference cannot take (X != Z, Q1).
fp32 + strict arithmetic is BIT-EXACT for every phase but the Poisson solve; tolerances for the rest are the
north_star's (1e-5 field-relative on the projected velocity, 1e-6 relative on particle positions)."""
import os

import numpy as np
import pytest

import parity
import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def gpu(dims, dtype=np.float32, **kw):
    from ink_simulation_b200 import System

    kw.setdefault("cg_rtol", 1e-7 if dtype == np.float32 else 1e-12)
    return System(*dims, dtype=dtype, **kw)


@pytest.mark.parametrize("path", parity.GOLDEN, ids=lambda p: p.split("ref_")[-1][:-4])
def test_gpu_phases_bit_exact_vs_reference_golden(path):
    parity.check_phases_exact(lambda d: gpu(d), parity.load(path))


@pytest.mark.parametrize("path", parity.GOLDEN, ids=lambda p: p.split("ref_")[-1][:-4])
def test_gpu_pressure_and_solve_vs_reference_golden(oracle_mod, path):
    G = parity.load(path)
    parity.check_pressure(lambda d: gpu(d), G)
    # end of step: against the reference's fixture AND, at 1e-5 with no per-fixture allowance, against the exact-solve oracle
    parity.check_solve(lambda d: gpu(d), G, exact=lambda d: oracle_mod.OracleSystem(*d))


@pytest.mark.parametrize("dims", [util.C1, util.RECT, util.THIN, util.C2], ids=str)
@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_phases_bit_exact_vs_oracle(oracle_mod, dims, dtype):
    u, v, w = util.rand_field(dims, 7, scale=0.6, dtype=dtype)
    part = util.wild_particles(dims, 3000, 8, dtype=dtype)
    inside = util.uniform_particles(dims, 50, 9, dtype=dtype)
    for phase, dt, pp in (("advect", 0.15, part), ("advect", 1.7, part), ("forces", 0.3, inside), ("forces", 0.3, part),
                          ("viscosity", 0.2, part), ("vorticity", 0.2, part), ("particles", 0.35, part), ("particles", 0.05, inside)):
        O = oracle_mod.OracleSystem(*dims, dtype=dtype)
        S = gpu(dims, dtype)
        for T in (O, S):
            T.set_velocity(u, v, w)
            T.set_particles(pp)
            getattr(T, phase)(dt)
        assert np.array_equal(np.stack(O.get_velocity()), np.stack(S.get_velocity())), (phase, dt)
        po, vo = O.get_particles()
        ps, vs = S.get_particles()
        assert np.array_equal(po, ps) and np.array_equal(vo, vs), (phase, dt)
        if phase == "forces":
            assert np.array_equal(O.forces(dt, return_mark=True), S.forces(dt, return_mark=True))
        if phase == "vorticity":
            assert np.array_equal(np.stack(O.get_field("curl")), np.stack(S.get_field("curl")))
        # the fused max-speed / next dt
        assert O.calc_dt(0.35) == S.calc_dt(0.35), (phase, dt)
        S.close()
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    S = gpu(dims, dtype)
    for T in (O, S):
        T.set_velocity(u, v, w)
    assert np.array_equal(O.sample_velocity(part), S.sample_velocity(part))
    assert np.array_equal(O.divergence(), S.divergence())
    assert np.array_equal(O.pressure_rhs(0.17), S.pressure_rhs(0.17))


@pytest.mark.parametrize("dims", [util.C1, util.RECT, util.C2, (32, 64, 32)], ids=str)
@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_poisson_cg_vs_exact_dst(oracle_mod, dims, dtype):
    """Matrix-free CG (replacing SparseLU) against the oracle's exact DST-I solve of the same system."""
    b = np.random.default_rng(3).normal(size=dims).astype(dtype)
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    p_ref, _ = O.poisson_solve(b)
    S = gpu(dims, dtype)
    p, res = S.poisson_solve(b)
    tol = 2e-5 if dtype == np.float32 else 1e-9
    assert util.rel_inf(p, p_ref) < tol, (util.rel_inf(p, p_ref), res, S.stats().cg_iters)


@pytest.mark.parametrize("dims", [util.C1, util.RECT, util.C2], ids=str)
@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_projection_vs_oracle(oracle_mod, dims, dtype):
    u, v, w = util.rand_field(dims, 17, scale=0.4, dtype=dtype)
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    S = gpu(dims, dtype)
    for T in (O, S):
        T.set_velocity(u, v, w)
        T.pressure(0.17)
    tol = 1e-5 if dtype == np.float32 else 1e-9
    assert util.rel_inf(np.stack(S.get_velocity()), np.stack(O.get_velocity())) < tol
    assert util.rel_inf(S.get_pressure(), O.get_pressure()) < 2 * tol
    # post-projection divergence no worse than the exact-solve oracle's (up to solver tolerance)
    assert np.abs(S.divergence()).max() <= np.abs(O.divergence()).max() * (1 + 1e-3) + 10 * tol


@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_full_steps_track_oracle_per_step(oracle_mod, dtype):
    """20 System::solve steps of BASELINE config 1 (8x25x8, 10k particles in the initParticles box); after every step the
    GPU state is compared with the oracle stepped from the SAME input state (then the oracle state is adopted).

    End-of-step tolerance in fp32: the vorticity-confinement phase divides grad|curl| by |curl_cell| (Q11); where the
    field is round-off noise this normalisation amplifies 1e-7 differences to ~1e-4 of the field scale — the same
    spread separates the reference arithmetic in fp32 from itself in fp64 (see test_gpu_step_phase_by_phase, which
    shows every phase but the solve is bit-exact and the PROJECTED field is within 1e-5).  fp64 mode is tight."""
    dims = util.C1
    part = util.box_particles(dims, 10000, 1, dtype=dtype)
    S = gpu(dims, dtype)
    S.set_particles(part)
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    O.set_particles(part)
    ptol, vtol = (1e-6, 5e-4) if dtype == np.float32 else (1e-9, 1e-8)
    for step in range(20):
        dt_o = O.solve(0.35)
        dt_s = S.solve(0.35)
        assert dt_s == dt_o, (step, dt_s, dt_o)
        vo, vs = np.stack(O.get_velocity()), np.stack(S.get_velocity())
        assert util.rel_inf(vs, vo) < vtol, (step, util.rel_inf(vs, vo))
        po, ps = O.get_particles()[0], S.get_particles()[0]
        assert np.max(np.abs(ps - po) / np.maximum(np.abs(po), 1.0)) < ptol, step
        # re-synchronise: identical input state for the next step
        S.set_velocity(*O.get_velocity())
        S.set_particles(po, O.get_particles()[1])


@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_step_phase_by_phase(oracle_mod, dtype):
    """One developing scene (config 1), 12 steps, checked PHASE BY PHASE from identical input: convection, forces,
    viscosity bit-exact; projected velocity within 1e-5 of the field scale with divergence no worse than the exact
    solve's (north_star); then — from the oracle's projected field — vorticity and particles bit-exact."""
    dims = util.C1
    part = util.box_particles(dims, 10000, 1, dtype=dtype)
    S = gpu(dims, dtype)
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    for T in (S, O):
        T.set_particles(part)
    for step in range(12):
        dt = O.calc_dt(0.35)
        assert S.calc_dt(0.35) == dt
        for ph in ("advect", "forces", "viscosity"):
            getattr(O, ph)(dt); getattr(S, ph)(dt)
            assert np.array_equal(np.stack(O.get_velocity()), np.stack(S.get_velocity())), (step, ph)
        O.pressure(dt); S.pressure(dt)
        vo, vs = np.stack(O.get_velocity()), np.stack(S.get_velocity())
        assert util.rel_inf(vs, vo) < (1e-5 if dtype == np.float32 else 1e-9), (step, util.rel_inf(vs, vo))
        assert np.abs(S.divergence()).max() <= np.abs(O.divergence()).max() * 1.001 + 1e-5 * np.abs(vo).max()
        S.set_velocity(*O.get_velocity())          # identical input for the phases after the solve
        O.vorticity(dt); S.vorticity(dt)
        assert np.array_equal(np.stack(O.get_velocity()), np.stack(S.get_velocity())), (step, "vorticity")
        O.particles(dt); S.particles(dt)
        assert np.array_equal(O.get_particles()[0], S.get_particles()[0]), (step, "particles")
        assert np.array_equal(O.get_particles()[1], S.get_particles()[1]), (step, "particle velocities")


@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
def test_gpu_fast_arith_particles_within_tolerance(oracle_mod, dtype):
    """INK_ARITH_FAST (FMA lerps, no weight-sum division) vs the strict oracle: <= 1e-6 relative per step (north_star),
    apart from particles whose wall-clamp branch flipped (counted, must be a vanishing fraction)."""
    dims = util.C2
    u, v, w = util.rand_field(dims, 5, scale=0.7, dtype=dtype)
    part = util.uniform_particles(dims, 100000, 6, dtype=dtype)
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    S = gpu(dims, dtype, arith="fast")
    for T in (O, S):
        T.set_velocity(u, v, w)
        T.set_particles(part)
        T.particles(0.3)
    po, ps = O.get_particles()[0], S.get_particles()[0]
    rel = np.max(np.abs(ps.astype(np.float64) - po) / np.maximum(np.abs(po), 1.0), axis=1)
    flipped = int((rel > 1e-6).sum())
    assert flipped <= 2, (flipped, rel.max())
    assert np.median(rel) < 1e-7


def test_gpu_particle_order_is_stable_across_resorts():
    """Binning permutes storage, never the caller-visible order (render.py keys particles by index)."""
    dims = util.C2
    S = gpu(dims, resort_interval=2)
    a = util.uniform_particles(dims, 5001, 1)
    b = util.box_particles(dims, 777, 2)
    S.set_particles(a, drop=0)
    S.set_particles(b, vel=b * 0.5, drop=1)
    assert S.num_drops() == 2 and S.num_particles(0) == 5001 and S.num_particles(1) == 777
    pa, va = S.get_particles(0)
    pb, vb = S.get_particles(1)
    assert np.array_equal(pa, a) and np.array_equal(pb, b) and np.array_equal(vb, b * 0.5) and not va.any()
    # zero velocity field: particles must not move through steps and re-binnings
    for _ in range(5):
        S.particles(0.1)
    assert S.stats().resorts >= 3
    assert np.array_equal(S.get_particles(0)[0], a) and np.array_equal(S.get_particles(1)[0], b)
    # replace drop 0 with a different count; drop 1 must survive untouched
    a2 = util.uniform_particles(dims, 1234, 3)
    S.set_particles(a2, drop=0)
    assert np.array_equal(S.get_particles(0)[0], a2) and np.array_equal(S.get_particles(1)[0], b)


def test_gpu_empty_and_single_particle():
    S = gpu(util.C1)
    assert S.solve(0.35) == pytest.approx(0.005)      # zero field: MIN_TIMESTEP (Q12)
    S.set_particles(np.zeros((0, 3), np.float32))
    S.solve(0.35)
    S.set_particles(np.array([[4.0, 20.0, 4.0]], np.float32))
    for _ in range(3):
        S.solve(0.35)
    S.checkNanAndInf()
    assert np.abs(np.stack(S.get_velocity())).max() > 0   # gravity acted around the particle


def test_gpu_advect_skip_is_exact_identity(oracle_mod):
    """The CFL-bounded convection phase is skipped as a proven identity; check against the oracle doing the work."""
    dims = util.C2
    u, v, w = util.rand_field(dims, 3, scale=0.5)
    O = oracle_mod.OracleSystem(*dims)
    S = gpu(dims)
    for T in (O, S):
        T.set_velocity(u, v, w)
    dt = O.calc_dt(0.35)
    assert S.calc_dt(0.35) == dt
    O.advect(dt); S.advect(dt)
    assert S.stats().advect_skipped == 1
    assert np.array_equal(np.stack(O.get_velocity()), np.stack(S.get_velocity()))


def test_gpu_streaming_io_equals_blocking_io():
    """ink_prefetch_positions / ink_commit_positions / ink_step / ink_fetch_positions_async (copies on their own streams,
    double-buffered) must give exactly what ink_set_particles / ink_step / ink_get_particles give: same positions bit for
    bit (binning differs, per-particle arithmetic does not), same field."""
    import torch

    dims = util.C2
    u, v, w = util.rand_field(dims, 21, scale=0.4)
    A, B = gpu(dims, resort_interval=3), gpu(dims, resort_interval=3)
    n = 20000
    inputs = [util.uniform_particles(dims, n, 30 + s, margin=0.3) for s in range(5)]
    pins = [torch.from_numpy(p).pin_memory() for p in inputs]
    outs = [torch.empty((n, 3), dtype=torch.float32).pin_memory() for _ in range(5)]
    for T in (A, B):
        T.set_velocity(u, v, w)
        T.set_particles(inputs[0])
    B.prefetch_positions(pins[0])
    for s in range(5):
        A.set_particles(inputs[s])
        dta = A.solve(0.35)
        pa = A.get_particles()[0]
        B.commit_positions()
        if s + 1 < 5:
            B.prefetch_positions(pins[s + 1])      # overlaps the step below
        dtb = B.solve(0.35)
        B.fetch_positions_async(outs[s])            # overlaps the next step
        assert dta == dtb
        last = pa
    B.sync_io()
    np.testing.assert_array_equal(outs[4].numpy(), last)
    for fa, fb in zip(A.get_velocity(), B.get_velocity()):
        np.testing.assert_array_equal(fa, fb)
    # every intermediate result arrived too
    C_ = gpu(dims, resort_interval=3)
    C_.set_velocity(u, v, w)
    for s in range(5):
        C_.set_particles(inputs[s])
        C_.solve(0.35)
        np.testing.assert_array_equal(outs[s].numpy(), C_.get_particles()[0])


def test_gpu_two_systems_of_different_size_in_one_process():
    """A System whose whole grid is the multigrid tail (23 KB of shared memory) must keep working after a second System with
    a much smaller tail is created: the shared-memory limit is a property of the kernel, not of the System (regression:
    'invalid argument' at 4 ranks in tests/mgpu_check.py, where a distributed and a single-GPU System share a process)."""
    dims_a, dims_b = (16, 10, 12), (16, 20, 12)
    A = gpu(dims_a)
    ref = gpu(dims_a)
    u, v, w = util.rand_field(dims_a, 21, scale=0.5)
    for T in (A, ref):
        T.set_velocity(u, v, w)
    ref.pressure(0.17)
    B = gpu(dims_b)          # created AFTER A, smaller tail
    B.set_velocity(*util.rand_field(dims_b, 22, scale=0.5))
    B.pressure(0.17)
    A.pressure(0.17)
    assert np.array_equal(np.stack(A.get_velocity()), np.stack(ref.get_velocity()))
    for T in (A, B, ref):
        T.close()


def test_gpu_zero_rhs_after_a_developed_solve():
    """Steady-state solves do not wait for the initial residual on the host (one sync less per step); the f == 0 special
    case (solution exactly 0, whatever the warm start holds) is then recognised at the first poll.  A System with solver
    history must give the same answer as a fresh one."""
    dims = util.C2
    A, B = gpu(dims), gpu(dims)
    u, v, w = util.rand_field(dims, 31, scale=0.5)
    A.set_velocity(u, v, w)
    for _ in range(3):
        A.pressure(0.17)                      # builds history: iteration count > 1, non-zero warm start
    assert A.stats().cg_iters > 1
    zero = [np.zeros(dims, np.float32)] * 3
    for T in (A, B):
        T.set_velocity(*zero)
        T.pressure(0.17)
    assert A.stats().cg_iters == 0 and B.stats().cg_iters == 0
    assert np.array_equal(A.get_pressure(), B.get_pressure()) and np.array_equal(A.get_warm_start(), np.zeros(dims, np.float32))
    assert np.array_equal(np.stack(A.get_velocity()), np.stack(B.get_velocity()))
    # and the next non-trivial solve is unaffected by what the blind iterations left behind
    for T in (A, B):
        T.set_velocity(u, v, w)
        T.pressure(0.17)
    assert util.rel_inf(np.stack(A.get_velocity()), np.stack(B.get_velocity())) < 1e-5
    A.close(); B.close()


def test_gpu_check_finite_is_the_reference_checkNanAndInf(oracle_mod):
    """System::checkNanAndInf (src/systemHelpers.cpp:212-241): no NaN / Inf component, norm() < 10000 — as one device
    reduction returning two counters.  Clean state passes; every kind of violation is counted where the reference asserts."""
    from ink_simulation_b200 import InkError

    dims = util.C1
    S = gpu(dims)
    u, v, w = util.rand_field(dims, 41, scale=0.5)
    S.set_velocity(u, v, w)
    S.set_particles(util.box_particles(dims, 5000, 42))
    S.solve(0.35)
    assert S.check_finite() == (0, 0)
    S.checkNanAndInf()
    bad_u = u.copy(); bad_u[1, 2, 3] = np.nan; bad_u[2, 2, 2] = np.inf
    big_v = v.copy(); big_v[3, 3, 3] = 2e4                      # finite, but norm() >= 10000
    S.set_velocity(bad_u, big_v, w)
    assert S.check_finite() == (2, 3)                           # NaN and Inf norms also fail `< 10000`, as in the reference
    with pytest.raises(InkError, match="checkNanAndInf"):
        S.checkNanAndInf()
    S.set_velocity(u, v, w)
    pos = util.box_particles(dims, 100, 43)
    vel = np.zeros_like(pos)
    pos[7, 1] = np.nan                                          # position: NaN / Inf only
    vel[9, 0] = 10000.0                                         # exactly 10000 is not < 10000
    vel[11, 2] = -np.inf
    S.set_particles(pos, vel)
    assert S.check_finite() == (2, 2)
    S.close()


def test_gpu_solver_stops_on_non_finite_residual():
    """A NaN in the field makes rr NaN, which never satisfies rr <= tol^2 bb: ink_step must fail at the first poll instead of
    running cg_max_iters garbage iterations and returning 0 (the reference aborts in checkNanAndInf)."""
    import time

    from ink_simulation_b200 import InkError

    dims = util.C2
    S = gpu(dims)
    u, v, w = util.rand_field(dims, 44, scale=0.5)
    u[3, 4, 5] = np.nan
    S.set_velocity(u, v, w)
    t0 = time.perf_counter()
    with pytest.raises(InkError, match="non-finite"):
        S.pressure(0.17)
    assert time.perf_counter() - t0 < 2.0
    assert S.stats().cg_converged == 0 and S.stats().cg_iters < 100
    S.close()


def test_gpu_library_runs_on_its_own_non_blocking_stream():
    """the ctx owns a cudaStreamNonBlocking stream (never the NULL stream); device-pointer inputs produced on torch's
    stream are picked up correctly through the Python host, device outputs are complete on return."""
    import torch

    dims = util.C1
    S = gpu(dims)
    assert S.cuda_stream() != 0
    ext = torch.cuda.ExternalStream(S.cuda_stream())
    u, v, w = util.rand_field(dims, 45, scale=0.5)
    host = gpu(dims)
    host.set_velocity(u, v, w)
    # produce the inputs on torch's stream with a long-running chain in front of them
    big = torch.randn(4096, 4096, device="cuda")
    for _ in range(10):
        big = big @ big * 1e-3
    tu, tv, tw = (torch.from_numpy(a).cuda() + 0 * big[0, 0] for a in (u, v, w))
    S.set_velocity(tu.contiguous(), tv.contiguous(), tw.contiguous())
    for a, b in zip(S.get_velocity(), host.get_velocity()):
        np.testing.assert_array_equal(a, b)
    # events recorded on the ctx stream see the library's work
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    S.set_particles(util.box_particles(dims, 10000, 46))
    e0.record(ext)
    for _ in range(5):
        S.solve(0.35)
    e1.record(ext)
    e1.synchronize()
    assert e0.elapsed_time(e1) > 0.05
    S.close(); host.close()


def test_gpu_cuda_array_interface_inputs_are_ordered_before_the_ctx_stream():
    """inputs handed over as __cuda_array_interface__ objects (cupy, numba; played here by a wrapper around a torch
    tensor) are produced on a stream the library's private stream knows nothing about.  The Python host synchronises the stream
    the interface names (or the device when it names none) before it passes the pointer on, so the library never reads them
    half-written."""
    import torch

    class Foreign:                      # exposes nothing but the interface: no data_ptr, no torch attributes
        def __init__(self, t, stream):
            self._keep = t
            self.__cuda_array_interface__ = dict(t.__cuda_array_interface__, stream=stream)

    dims = util.C1
    u, v, w = util.rand_field(dims, 47, scale=0.5)
    pos = util.box_particles(dims, 5000, 48)
    host = gpu(dims)
    host.set_velocity(u, v, w)
    host.set_particles(pos)
    side = torch.cuda.Stream()
    for named in (side.cuda_stream, None):
        S = gpu(dims)
        with torch.cuda.stream(side):
            big = torch.randn(4096, 4096, device="cuda")
            for _ in range(10):
                big = big @ big * 1e-3       # a long chain in front of the inputs on the side stream
            tu, tv, tw = ((torch.from_numpy(a).cuda() + 0 * big[0, 0]).contiguous() for a in (u, v, w))
            tp = (torch.from_numpy(pos).cuda() + 0 * big[0, 0]).contiguous()
        S.set_velocity(Foreign(tu, named), Foreign(tv, named), Foreign(tw, named))
        S.set_particles(Foreign(tp, named))
        for a, b in zip(S.get_velocity(), host.get_velocity()):
            np.testing.assert_array_equal(a, b)
        np.testing.assert_array_equal(S.get_particles()[0], host.get_particles()[0])
        S.close()
    host.close()


@pytest.fixture
def k8_tile_env(monkeypatch):
    """INK_K8_TILE is read when a System is created: 2 = the TMA-staged shared-memory tile kernel wherever it is legal
    (also for clouds the heuristic would leave to the global-gather kernel), 0 = never."""
    def make(mode, dims, dtype=np.float32, **kw):
        monkeypatch.setenv("INK_K8_TILE", str(mode))
        S = gpu(dims, dtype, **kw)
        monkeypatch.delenv("INK_K8_TILE")
        return S
    return make


@pytest.mark.parametrize("path", parity.GOLDEN, ids=lambda p: p.split("ref_")[-1][:-4])
def test_gpu_tile_particle_kernel_vs_reference_golden(k8_tile_env, path):
    """The tile kernel (k_particles_rk2_tile: rows staged by cp.async.bulk, gathers from shared memory) against the COMPILED
    REFERENCE's particle update, bit for bit: the dense box cloud and the wild cloud (out-of-grid, on-wall, on-face particles:
    those leave the tile and take the global sampler)."""
    G = parity.load(path)
    dims = tuple(int(x) for x in G["dims"])
    if (dims[2] * 4) % 16 != 0:
        pytest.skip("rows are not 16-byte multiples: the tile path is not legal here (global-gather kernel runs)")
    for cloud, dt, kp, kv in (("part", 0.2, "particles_pos", "particles_vel"), ("wild", 0.35, "wild_pos", "wild_vel")):
        S = k8_tile_env(2, dims)
        S.set_velocity(G["u"], G["v"], G["w"])
        S.set_particles(G[cloud])
        S.particles(dt)
        pos, vel = S.get_particles()
        np.testing.assert_array_equal(pos, G[kp]); np.testing.assert_array_equal(vel, G[kv])
        S.close()


@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
@pytest.mark.parametrize("arith", ["strict", "fast"])
def test_gpu_tile_particle_kernel_equals_global_gather_kernel(k8_tile_env, dtype, arith):
    """Same inputs through both K8 kernels: positions, velocities and the force mask seeded for the next step are identical,
    bit for bit — uniform and wild clouds, CFL-sized and far too large dt (samples that leave the tile), and many steps
    without re-binning (particles drift out of the cells they were binned in; the tile follows their bounding box)."""
    dims = (32, 24, 64)
    u, v, w = util.rand_field(dims, 51, scale=0.6, dtype=dtype)
    clouds = {"uniform": util.uniform_particles(dims, 60000, 52, dtype=dtype), "wild": util.wild_particles(dims, 5000, 53, dtype=dtype),
              "box": util.box_particles(dims, 30000, 54, dtype=dtype)}
    for name, part in clouds.items():
        for dt in (0.15, 2.5):
            A, B = k8_tile_env(2, dims, dtype, arith=arith, resort_interval=1000), k8_tile_env(0, dims, dtype, arith=arith, resort_interval=1000)
            for T in (A, B):
                T.set_velocity(u, v, w); T.set_particles(part)
            for step in range(6):
                A.particles(dt); B.particles(dt)
                pa, va = A.get_particles(); pb, vb = B.get_particles()
                assert np.array_equal(pa, pb) and np.array_equal(va, vb), (name, dt, step)
            A.forces(0.1); B.forces(0.1)          # consumes the mask K8 seeded from the new positions
            assert np.array_equal(A.get_mark(), B.get_mark()), (name, dt)
            A.close(); B.close()


def test_gpu_tile_particle_kernel_whole_steps(k8_tile_env, oracle_mod):
    """Config-2 scene stepped with the tile kernel forced: the step-by-step oracle comparison of
    test_gpu_full_steps_track_oracle_per_step holds unchanged (re-binning every 4 steps rebuilds the work list)."""
    dims = util.C2
    part = util.box_particles(dims, 100000, 1)
    S = k8_tile_env(2, dims, resort_interval=4)
    O = oracle_mod.OracleSystem(*dims, dtype=np.float32)
    for T in (S, O):
        T.set_particles(part)
    for step in range(10):
        assert S.solve(0.35) == O.solve(0.35)
        po, ps = O.get_particles()[0], S.get_particles()[0]
        assert np.max(np.abs(ps - po) / np.maximum(np.abs(po), 1.0)) < 1e-6, step
        S.set_velocity(*O.get_velocity())
        S.set_particles(po, O.get_particles()[1])
    S.close()


@pytest.mark.parametrize("dims", [(16, 24, 64), (20, 45, 20), (32, 64, 32), (10, 9, 132), (8, 25, 8)], ids=str)
@pytest.mark.parametrize("dtype", [np.float32, np.float64], ids=["fp32", "fp64"])
@pytest.mark.parametrize("switch", ["INK_MG3_UP_TILE", "INK_MG3_DOWN_TILE", "INK_CG_MATVEC_TILE"])
def test_gpu_up_tile_kernel_is_the_same_preconditioner(oracle_mod, monkeypatch, dims, dtype, switch):
    """k_mg3_up_tile (x' staged through a shared-memory tile) against k_mg3_up (one coarse cell per thread) as the up pass of the
    V-cycle: same operator, different rounding order => the PCG takes the same number of iterations (+-1) and both meet the exact
    DST solution; ragged X / Y, Z below and above the tile length, the default-size and the tiny grids."""
    O = oracle_mod.OracleSystem(*dims, dtype=dtype)
    b = np.random.default_rng(61).normal(size=dims).astype(dtype)
    p_exact, _ = O.poisson_solve(b)
    out = {}
    for mode in (2, 0):   # the same holds for the tiled down pass (residual + restriction in one launch) and the tiled CG matvec
        monkeypatch.setenv(switch, str(mode))
        S = gpu(dims, dtype, cg_rtol=1e-7 if dtype == np.float32 else 1e-12)
        monkeypatch.delenv(switch)
        p, res = S.poisson_solve(b)
        out[mode] = (p, S.stats().cg_iters)
        assert util.rel_inf(p, p_exact) < (2e-5 if dtype == np.float32 else 1e-9), (mode, util.rel_inf(p, p_exact))
        S.close()
    assert abs(out[2][1] - out[0][1]) <= 1, (out[2][1], out[0][1])
    assert util.rel_inf(out[2][0], out[0][0]) < (2e-5 if dtype == np.float32 else 1e-9)


@pytest.mark.parametrize("dims,npart", [(util.C1, 10000), ((32, 64, 32), 157000)], ids=["C1-box", "C3-density"])
def test_gpu_fast_arith_tracks_strict_over_several_steps(dims, npart):
    """The bench headline runs INK_ARITH_FAST (FMA lerps, weight-sum division dropped); the bit-exact path is INK_ARITH_STRICT.
    Over 8 whole steps, each from the strict run's state: same dt, particle positions within 1e-6 relative per step apart from a
    vanishing number of wall-clamp branch flips, velocity field within 1e-5 — at config 1 (dense box cloud) and at the bench's
    particle density (2.4 per cell, uniform)."""
    part = util.box_particles(dims, npart, 71) if dims == util.C1 else util.uniform_particles(dims, npart, 71, margin=0.5)
    A, B = gpu(dims, arith="strict", cg_rtol=1e-7), gpu(dims, arith="fast", cg_rtol=1e-7)
    for T in (A, B):
        T.set_particles(part)
    for step in range(8):
        assert A.solve(0.35) == B.solve(0.35), step
        pa, va = A.get_particles(); pb = B.get_particles()[0]
        rel = np.max(np.abs(pb.astype(np.float64) - pa) / np.maximum(np.abs(pa), 1.0), axis=1)
        assert int((rel > 1e-6).sum()) <= max(2, npart // 50000), (step, int((rel > 1e-6).sum()), rel.max())
        assert util.rel_inf(np.stack(B.get_velocity()), np.stack(A.get_velocity())) < 1e-5, step
        B.set_velocity(*A.get_velocity()); B.set_particles(pa, va)
    A.close(); B.close()


@pytest.mark.parametrize("dims", [util.C2, (32, 64, 32)], ids=str)
def test_gpu_bench_solver_tolerance_meets_the_north_star(oracle_mod, dims):
    """bench.py runs the ABI default cg_rtol = 1e-6 (the parity suite mostly 1e-7): the projected velocity must still be within 1e-5
    of the field scale of the exact (DST) projection, step after step on a developing scene, with warm starts as in a real run."""
    part = util.uniform_particles(dims, int(2.4 * np.prod(dims)), 81, margin=0.5)
    S = gpu(dims, cg_rtol=1e-6)
    O = oracle_mod.OracleSystem(*dims, dtype=np.float32)
    u, v, w = util.rand_field(dims, 82, scale=0.4)
    for T in (S, O):
        T.set_velocity(u, v, w); T.set_particles(part)
    worst = 0.0
    for step in range(6):
        dt = O.calc_dt(0.35)
        assert S.calc_dt(0.35) == dt
        for ph in ("advect", "forces", "viscosity"):
            getattr(O, ph)(dt); getattr(S, ph)(dt)
        O.pressure(dt); S.pressure(dt)
        err = util.rel_inf(np.stack(S.get_velocity()), np.stack(O.get_velocity()))
        worst = max(worst, err)
        assert err < 1e-5, (step, err, S.stats().cg_iters)
        S.set_velocity(*O.get_velocity())
        O.vorticity(dt); S.vorticity(dt)
        O.particles(dt); S.particles(dt)
    S.close()


def test_gpu_every_kernel_variant_runs_on_ragged_grids():
    """scripts/sanitize_r2.py: whole steps with wild particles on small ragged grids under every opt-in kernel switch of round 2
    (tiled up / down / matvec, K8 tile, one-pass down, cluster coarse kernel) and the device seeders; every run must stay finite
    (ink_check_finite).  Own process: the switches are environment variables read when a context is created."""
    import subprocess
    import sys

    out = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_r2.py")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "sanitize_r2: done" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_gpu_refused_graph_capture_falls_back_to_plain_launches():
    """a runtime that refuses the capture of the PCG iteration (simulated by INK_TEST_FAIL_CAPTURE=1) must not fail
    ink_step: the ctx drops to plain launches (graph_launches stays 0, one INK_VERBOSE line) and, the kernels and their order
    being the same, produces the same bits as the graph path.  Own processes: the switch is read when the library loads."""
    import subprocess
    import sys

    code = (
        "import sys, hashlib, numpy as np\n"
        f"sys.path.insert(0, {ROOT!r}); sys.path.insert(0, {os.path.join(ROOT, 'tests')!r})\n"
        "import util\n"
        "from ink_simulation_b200 import System\n"
        "dims = util.C2\n"
        "S = System(*dims, cg_rtol=1e-7)\n"
        "S.set_velocity(*util.rand_field(dims, 71, scale=0.5)); S.set_particles(util.box_particles(dims, 20000, 72))\n"
        "for _ in range(3): S.solve(0.35)\n"
        "st = S.stats()\n"
        "h = hashlib.sha1(np.stack(S.get_velocity()).tobytes() + S.get_particles()[0].tobytes()).hexdigest()\n"
        "print('RESULT', h, st.cg_iters, st.graph_launches)\n"
    )
    res = {}
    for fail in ("0", "1"):
        env = dict(os.environ, INK_TEST_FAIL_CAPTURE=fail, INK_VERBOSE="1")
        out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env)
        assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
        line = [l for l in out.stdout.splitlines() if l.startswith("RESULT")][0].split()
        res[fail] = (line[1], int(line[2]), int(line[3]), out.stderr)
    assert res["0"][2] > 0 and res["1"][2] == 0
    assert "plain launches" in res["1"][3] and "plain launches" not in res["0"][3]
    assert res["0"][:2] == res["1"][:2]


@pytest.mark.parametrize("dims,npart,steps", [((128, 256, 128), 10_000_000, 30), ((256, 256, 256), 20_000_000, 8)],
                         ids=["c3_128x256x128_10M", "c4grid_256x256x256_20M"])
def test_gpu_parity_at_bench_size(oracle_mod, dims, npart, steps):
    """VERDICT r2, missing 1: every throughput number is quoted at sizes no test had compared.  The bench's own workload (uniform
    cloud, fast arithmetic, rtol 1e-6) is developed for a few steps on the GPU, then ONE step is taken phase by phase by the
    strict CUDA path and by the CPU oracle from that identical state (bench.full_size_parity, the block BENCH prints as
    `parity`): dt equal, stencil phases and particles bit-exact, projected velocity <= 1e-5, divergence no worse than the exact
    solve's, fast-arithmetic positions <= 1e-6.  256^3 is where Z > 128 (the long-row residual kernel) and the 6-level
    hierarchy first run; 10 M / 20 M particles exercise the binning keys and 32-bit offsets of the particle kernels."""
    import bench

    S = gpu(dims, arith="fast", cg_rtol=1e-6)
    S.set_particles(bench.make_particles(dims, npart, "uniform"))
    for _ in range(steps):
        S.solve(0.35)
    u, v, w = S.get_velocity()
    pos, vel = S.get_particles()
    S.close()
    r = bench.full_size_parity(dims, u, v, w, pos, vel)
    print(r)
    assert r["ok"], r
