"""Multi-GPU parity script (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_check.py

Every rank builds (a) the y-slab / particle-shard DISTRIBUTED System over NCCL and (b) a private single-GPU System of the
whole problem on its own device, steps both from the same seeded state and compares:
  * phases without a solve (forces, viscosity, vorticity, particles): bit-exact (same kernels, redundant-row scheme),
  * pressure / full steps: within the north_star tolerances (the slab-local preconditioner changes the CG iterates,
    not the solution): projected velocity 1e-5 field-relative PER PHASE inside whole steps, particle positions 1e-6
    relative per step.
Exit code 0 = all ranks agree.  Used by tests/test_multi_gpu.py (-m gpu, skipped with < 2 GPUs)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist

    import util
    from ink_simulation_b200 import System, create_distributed, shard_range

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")   # only carries the 128-byte NCCL id; the data path is the library's own NCCL comm
    rank, world = dist.get_rank(), dist.get_world_size()
    failures = []

    def check(name, ok, info=""):
        if not ok:
            failures.append(f"rank {rank}: {name} {info}")

    # the last case has >= 32 rows per rank (>= 2 * ovl = 16) and Z % 4 == 0: the vectorised band kernels, the fused peer-store
    # pushes and interior ranks with TWO neighbours are what runs there
    cases = (((16, 5 * world, 12), 5000), ((16, 40, 12), 20000), ((32, 64, 32), 200000), ((64, 32 * world, 64), 2000000))
    if os.environ.get("MGPU_CASES"):   # e.g. MGPU_CASES=0,2: run a subset (debugging at large world sizes)
        cases = tuple(cases[int(i)] for i in os.environ["MGPU_CASES"].split(","))
    for dims, npart in cases:
        u, v, w = util.rand_field(dims, 11, scale=0.5)
        part = util.uniform_particles(dims, npart, 12)
        lo, hi = shard_range(npart, rank, world)

        def pair(**kw):
            D = create_distributed(*dims, device=local, cg_rtol=1e-7, **kw)
            S = System(*dims, device=local, cg_rtol=1e-7, **kw)
            for T in (D, S):
                T.set_velocity(u, v, w)
                T.set_particles(part)
            return D, S

        # --- phases without a solve: bit-exact
        for phase, dt in (("forces", 0.3), ("viscosity", 0.2), ("vorticity", 0.2), ("advect", 1.7), ("particles", 0.3)):
            D, S = pair()
            getattr(D, phase)(dt); getattr(S, phase)(dt)
            check(f"{dims} {phase} velocity", np.array_equal(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
            pd, vd = D.get_particles(); ps, vs = S.get_particles()
            check(f"{dims} {phase} particles", np.array_equal(pd, ps[lo:hi]) and np.array_equal(vd, vs[lo:hi]))
            check(f"{dims} {phase} next dt", D.calc_dt(0.35) == S.calc_dt(0.35))
            D.close(); S.close()
        # --- projection
        for precond in ("mg", "none"):
            D, S = pair(cg_precond=precond)
            D.pressure(0.17); S.pressure(0.17)
            ev = util.rel_inf(np.stack(D.get_velocity()), np.stack(S.get_velocity()))
            ep = util.rel_inf(D.get_pressure(), S.get_pressure())
            check(f"{dims} pressure[{precond}]", ev < 1e-5 and ep < 2e-5, f"vel {ev:.2e} p {ep:.2e} iters {D.stats().cg_iters}/{S.stats().cg_iters}")
            check(f"{dims} divergence[{precond}]", np.abs(D.divergence()).max() <= np.abs(S.divergence()).max() * 1.01 + 1e-5)
            D.close(); S.close()
        # --- whole steps, PHASE BY PHASE from identical input (as tests/test_gpu_parity.py::test_gpu_step_phase_by_phase does
        # against the oracle): stencil phases bit-exact, projected velocity 1e-5 of the field scale, then — from the single-GPU
        # projected field — vorticity and particles bit-exact
        D, S = pair()
        for step in range(4):
            dt = S.calc_dt(0.35)
            check(f"{dims} step {step} dt", D.calc_dt(0.35) == dt)
            for ph in ("advect", "forces", "viscosity"):
                getattr(D, ph)(dt); getattr(S, ph)(dt)
                check(f"{dims} step {step} {ph}", np.array_equal(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
            D.pressure(dt); S.pressure(dt)
            vd, vs = np.stack(D.get_velocity()), np.stack(S.get_velocity())
            check(f"{dims} step {step} projected velocity", util.rel_inf(vd, vs) < 1e-5, f"{util.rel_inf(vd, vs):.2e} iters {D.stats().cg_iters}/{S.stats().cg_iters}")
            check(f"{dims} step {step} divergence", np.abs(D.divergence()).max() <= np.abs(S.divergence()).max() * 1.01 + 1e-5 * np.abs(vs).max())
            D.set_velocity(*S.get_velocity())
            D.vorticity(dt); S.vorticity(dt)
            check(f"{dims} step {step} vorticity", np.array_equal(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
            D.particles(dt); S.particles(dt)
            pd, vd = D.get_particles(); ps, vs = S.get_particles()
            check(f"{dims} step {step} particles", np.array_equal(pd, ps[lo:hi]) and np.array_equal(vd, vs[lo:hi]))
        # --- and undivided System::solve calls, re-synchronised after each: dt equal, particle positions 1e-6 relative per step
        for step in range(3):
            dtd, dts = D.solve(0.35), S.solve(0.35)
            check(f"{dims} solve {step} dt", dtd == dts, f"{dtd} {dts}")
            pd = D.get_particles()[0]; ps = S.get_particles()[0]
            rel = np.max(np.abs(pd - ps[lo:hi]) / np.maximum(np.abs(ps[lo:hi]), 1.0)) if hi > lo else 0.0
            check(f"{dims} solve {step} particles", rel < 1e-6, f"{rel:.2e}")
            bad, big = D.check_finite()
            check(f"{dims} solve {step} check_finite", bad == 0 and big == 0, f"{bad} {big}")
            D.set_velocity(*S.get_velocity())
            D.set_particles(ps, S.get_particles()[1])
        if rank == 0:
            print(f"mgpu_check {dims}: world={world} cg_iters dist/single = {D.stats().cg_iters}/{S.stats().cg_iters}", flush=True)
        D.close(); S.close()

    allf = [None] * world
    dist.all_gather_object(allf, failures)
    flat = [f for fs in allf for f in fs]
    if rank == 0:
        print("mgpu_check:", "OK" if not flat else "FAILED\n  " + "\n  ".join(flat), flush=True)
    dist.destroy_process_group()
    sys.exit(1 if flat else 0)


if __name__ == "__main__":
    main()
