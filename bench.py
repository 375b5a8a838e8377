#!/usr/bin/env python
"""bench.py — throughput of the Ink-Simulation hot path (System::solve) on B200.

One "step" = one System::solve (MAC water-grid update + RK2 particle update, reference src/inkSolver.cpp:5-9) on
synthetic input.  Default workload at N=1 = BASELINE.json configs[2]: 10M particles (uniform, UNSORTED input order)
on a 128x256x128 MAC grid, fp32; at N>1 every GPU keeps one such y-slab and particle shard (weak scaling).
Prints ONE JSON line (see the contract in the task statement):
  value        : particle RK2 updates/s over whole steps (state resident in HBM)
  e2e          : same, through the host-facing call: pinned-host positions in (H2D) -> solve -> positions out (D2H) every step
  roofline     : the pressure phase (PCG kernels, the largest share of a step): algorithmic bytes / CUDA-event time vs
                 MEASURED_PEAKS.json; `roofline.per_kernel` lists K8 (particles) and the stencil phases the same way
  cpu_baseline : the CPU oracle port (OpenMP, all host cores) on a bounded sample of the same workload
  c4           : BASELINE.json configs[3] (256^3, 100 M particles) on the same N GPUs, STRONG-scaled (y-slabs of the fixed grid)
  parity       : N > 1 only — the distributed path against a single-GPU System on a small problem, phase by phase
  strict       : N = 1 only — the same workload with INK_ARITH_STRICT (the bit-exact arithmetic the parity claims cover)
All device timings are CUDA events recorded on the LIBRARY's own stream (ink_get_stream), max over ranks.
`--impl reference` times the CPU implementation (compiled reference where it can run the config, else the oracle port).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    "c1": dict(dims=(8, 25, 8), particles=10_000, cloud="box"),
    "c2": dict(dims=(20, 45, 20), particles=100_000, cloud="box"),
    "c3": dict(dims=(128, 256, 128), particles=10_000_000, cloud="uniform"),
    "c4": dict(dims=(256, 256, 256), particles=100_000_000, cloud="uniform"),
    # BASELINE config 5: RK2 advection sweep on a FROZEN 512^3 analytic field, 125 M particles per GPU (1 B at 8 GPUs); every
    # rank holds the whole field (broadcast once, no per-step traffic), particles shard by index
    "c5": dict(dims=(512, 512, 512), particles=125_000_000, cloud="uniform", sweep=True),
}
RENDER_DT = 0.35  # src/main.cpp:25
MIN_TIMED_MS = float(os.environ.get("INK_BENCH_MIN_MS", "600"))   # the device-timed loop covers at least this long (clock sampling needs it),
                                                                     # whatever --steps says; profiling passes under ncu set it to 0


def make_particles(dims, n, cloud, seed=0x1234, first=0, count=None):
    """Particles [first, first+count) of the n-particle cloud (a rank only generates its own shard; the cloud is i.i.d.
    uniform, any seeded sample of it is the same workload)."""
    rng = np.random.default_rng([seed, first])
    X, Y, Z = dims
    n = n if count is None else count
    if cloud == "box":  # initParticles box, src/system.cpp:102-104
        lo = np.array([X / 2 - X / 5, Y - 1.5, Z / 2 - Z / 3])
        hi = np.array([X / 2 + X / 5, Y - 1.0, Z / 2 + Z / 3])
    else:
        lo = np.array([0.5, 0.5, 0.5])
        hi = np.array([X - 0.5, Y - 0.5, Z - 0.5])
    out = np.empty((n, 3), np.float32)
    chunk = 4_000_000
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        out[s:e] = (lo + rng.random((e - s, 3)) * (hi - lo)).astype(np.float32)
    return out


def make_particles_device(dims, count, seed, device):
    """uniform cloud generated on the device (torch is plumbing): used for the large nested runs, where 1.2 GB of host RNG
    output per rank would dominate the bench's wall time"""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    hi = torch.tensor([dims[0] - 1.0, dims[1] - 1.0, dims[2] - 1.0], device=device)
    return (torch.rand((count, 3), generator=g, device=device, dtype=torch.float32) * hi + 0.5).contiguous()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback 6650 (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(dev), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_port_sample(dims, pos, steps, warm_state=None):
    """The CPU oracle (OpenMP restatement of the reference, oracle/ink_oracle.cpp) on the same workload: one untimed warm-up
    step, then `steps` timed ones.  Returns (seconds per step, OpenMP threads actually used)."""
    import oracle

    oracle.build()
    threads = oracle.set_threads()
    O = oracle.OracleSystem(*dims, dtype=np.float32)
    O.set_particles(pos)
    if warm_state is not None:
        O.set_velocity(*warm_state)
    O.solve(RENDER_DT)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.solve(RENDER_DT)
    return (time.perf_counter() - t0) / steps, threads


def l2_note(rbytes, dims, n, world):
    """How the timing rule 'inputs larger than L2 or flush between iterations' is met, from the actual sizes."""
    part_mb = 9 * rbytes * (n / world) / 1e6                  # per pass: read pos, write pos, write vel (3 components each)
    cells = dims[0] * dims[1] * dims[2] / world
    vec_mb = rbytes * cells / 1e6
    total = part_mb + 15 * vec_mb
    verdict = "inputs larger than L2" if total > 126 else "inputs FIT in L2: a parity-size case, not a bench line"
    return f"{verdict} (per GPU: particle arrays {part_mb:.0f} MB per pass + 15 grid vectors of {vec_mb:.1f} MB; L2 = 126 MB)"


def workload_config(args, name, cfg, dims, n, world, strong=False):
    """The `config` object both arms print (the reference arm runs the same workload and says which sample of it)."""
    rbytes = 8 if args.dtype == "f64" else 4
    if cfg.get("sweep"):
        X, Y, Z = cfg["dims"]
        per = args.particles or cfg["particles"]
        return {"workload": f"{name}: RK2 advection sweep (System::updateParticles), {per * world} particles ({per} per GPU, uniform, unsorted "
                            f"input), frozen analytic {X}x{Y}x{Z} field replicated on every GPU, dt = 0.2", "arith": args.arith,
                "resort_interval": args.resort, "sharding": "particles by index, no data-path collective",
                "l2": "field 1.6 GB + particle arrays 4.5 GB per pass >> L2 (126 MB)"}
    return {"workload": f"{name}: {n} particles ({cfg['cloud']}, unsorted input), {dims[0]}x{dims[1]}x{dims[2]} MAC grid, "
                        f"full System::solve per step", "arith": args.arith, "cg_rtol": args.cg_rtol,
            "resort_interval": args.resort, "precond": args.precond,
            "sharding": "single GPU" if world == 1 else f"{world} y-slabs of {dims[1] // world} rows ({'strong' if strong else 'weak'} scaling), "
                                                        f"particles by index; NCCL + peer-memory exchanges",
            "l2": l2_note(rbytes, dims, n, world)}


def scaled_workload(cfg, world, strong):
    dims, n = cfg["dims"], cfg["particles"]
    if world > 1 and not strong:   # weak scaling: every GPU keeps one config-sized y-slab and particle shard
        dims, n = (dims[0], dims[1] * world, dims[2]), n * world
    return dims, n


def run_reference_arm(args, cfg, rank, world):
    """--impl reference: the CPU implementation on the host cores, on the SAME workload as the GPU arm at this N (the whole
    N-GPU grid and particle count, not one GPU's share).  The compiled reference (hash-map grid + SparseLU) cannot run grids
    beyond ~32x64x32, so for c3/c4/c5 the OpenMP oracle port stands in (kind: "port") with every CPU the process
    may use — the thread count is set explicitly (torchrun exports OMP_NUM_THREADS=1) and the count printed is the one
    omp_get_max_threads() reports."""
    if rank != 0:
        return
    import oracle

    oracle.build()
    sweep = bool(cfg.get("sweep"))
    dims, n = (cfg["dims"], cfg["particles"] * world) if sweep else scaled_workload(cfg, world, args.strong)
    use_ref = oracle.have_ref() and dims[0] * dims[1] * dims[2] <= 20000 and dims[0] == dims[2] and not sweep
    cap = args.cpu_particles * (1 if args.strong else world)
    sample_n = n if use_ref else min(n, cap)
    pos = make_particles(dims, sample_n, cfg["cloud"])
    if use_ref:
        T = oracle.RefSystem(*dims)
        kind, cores = "reference", 1
    else:
        cores = oracle.set_threads()
        T = oracle.OracleSystem(*dims, dtype=np.float32)
        kind = "port"
    T.set_particles(pos)
    if sweep:   # config 5: particles-only sweep on the frozen analytic field (same field as run_sweep)
        ax = [np.arange(m, dtype=np.float32) * np.float32(2 * np.pi / m) for m in dims]
        gx, gy, gz = ax[0][:, None, None], ax[1][None, :, None], ax[2][None, None, :]
        u = np.broadcast_to(np.sin(gz) + np.cos(gy), dims).astype(np.float32)
        v = np.broadcast_to(np.sin(gx) + np.cos(gz), dims).astype(np.float32)
        w = np.broadcast_to(np.sin(gy) + np.cos(gx), dims).astype(np.float32)
        mx = np.float32(np.sqrt(u * u + v * v + w * w).max())
        T.set_velocity(u / mx, v / mx, w / mx)
        step = lambda: T.particles(0.2)
    else:
        step = lambda: T.solve(RENDER_DT)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    sec = (time.perf_counter() - t0) / args.steps
    scale = n / sample_n    # particle-bound sample of the same grid (only when --cpu-particles caps it); stated below
    val = sample_n / sec
    sample = (f"{args.steps} {'updateParticles sweeps' if sweep else 'System::solve steps'} of the WHOLE {world}-GPU workload: grid "
              f"{dims[0]}x{dims[1]}x{dims[2]}, {sample_n} of {n} particles ({cfg['cloud']})"
              f"{'' if scale == 1 else ' — value counts the sampled particles only'}; "
              f"{'compiled reference, 1 thread (it has no threading)' if use_ref else f'OpenMP oracle port on {cores} threads (omp_get_max_threads), exact DST Poisson solve'}"
              f"; flow steps timed: {args.warmup + 1}..{args.warmup + args.steps} from rest (the GPU arm times a later, developed stretch of the same flow; "
              f"the CPU step cost does not depend on it: direct solve, no iterations)")
    line = {
        "impl": "reference", "metric": "particle_rk2_updates_per_s", "value": val, "unit": "particle-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "strong" if (world > 1 and args.strong) else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.config, cfg, dims, n, world, args.strong),
        "mac_cell_steps_per_s": dims[0] * dims[1] * dims[2] / sec,
        "cpu_baseline": {"value": val, "unit": "particle-updates/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "particle-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if use_ref and hasattr(T, "solve_timed"):   # the reference as shipped, per phase
        ph = []
        for _ in range(max(5, min(args.steps, 20))):
            ph.append(T.solve_timed(RENDER_DT)[1])
        med = np.median(np.array(ph), axis=0)
        line["reference_phase_ms"] = dict(zip(("dt", "advect", "forces", "viscosity", "pressure", "vorticity", "particles"), (float(x) * 1e3 for x in med)))
    print(json.dumps(line), flush=True)


class Harness:
    """rank / world plumbing shared by the bench legs"""

    def __init__(self, rank, world, local):
        import torch

        self.torch, self.rank, self.world, self.local = torch, rank, world, local
        self.dist = None
        if world > 1:
            import torch.distributed as dist

            self.dist = dist

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return float(t.item())

    def events(self, S):
        """(stream, e0, e1): timing events live on the LIBRARY's stream (torch.cuda.Event sees only the stream it is recorded on)"""
        t = self.torch
        ext = t.cuda.ExternalStream(S.cuda_stream(), device=t.device("cuda", self.local))
        return ext, t.cuda.Event(enable_timing=True), t.cuda.Event(enable_timing=True)

    def timed_solves(self, S, n_solves):
        """n_solves whole steps bracketed by barrier + synchronize; device time (ms) as max over ranks, CG iterations per step"""
        ext, e0, e1 = self.events(S)
        iters = []
        self.barrier()
        e0.record(ext)
        for _ in range(n_solves):
            S.solve(RENDER_DT)
            iters.append(S.stats().cg_iters)
        e1.record(ext)
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1)), iters


def run_sweep(args, cfg, H):
    """Config 5: particles-only RK2 sweep (ink_particles = System::updateParticles) on a frozen analytic field.  Every rank is a
    single-GPU System holding the WHOLE field and its index shard of the particles: no data-path collective (weak scaling)."""
    torch = H.torch
    from ink_simulation_b200 import System

    rank, world, local = H.rank, H.world, H.local
    dims, n_local = cfg["dims"], args.particles or cfg["particles"]
    X, Y, Z = dims
    ncell = X * Y * Z
    S = System(*dims, dtype=np.float32, arith=args.arith, device=local, resort_interval=args.resort)
    # ABC-type flow, max |u| = 1  =>  dt = K_CFL * h / maxV = 0.2; built on the device, (X, Y, Z) C-order
    dev = torch.device("cuda", local)
    ax = [torch.arange(m, device=dev, dtype=torch.float32) * (2 * np.pi / m) for m in dims]
    gx, gy, gz = ax[0][:, None, None], ax[1][None, :, None], ax[2][None, None, :]
    u = (torch.sin(gz) + torch.cos(gy)).expand(X, Y, Z).contiguous()
    v = (torch.sin(gx) + torch.cos(gz)).expand(X, Y, Z).contiguous()
    w = (torch.sin(gy) + torch.cos(gx)).expand(X, Y, Z).contiguous()
    mx = float(torch.sqrt(u * u + v * v + w * w).max())
    for f in (u, v, w):
        f.div_(mx)
    S.set_velocity(u, v, w)
    del u, v, w, gx, gy, gz
    torch.cuda.empty_cache()
    pos = make_particles_device(dims, n_local, 0x1234 + rank, dev)
    S.set_particles_shard(pos)
    del pos
    torch.cuda.empty_cache()
    dt = 0.2
    for _ in range(max(3, args.warmup)):
        S.particles(dt)
    S.sync()
    st0 = S.stats()
    sampler = ClockSampler(local)
    ext, e0, e1 = H.events(S)
    steps = max(args.steps, int(math.ceil(MIN_TIMED_MS / 4.5)))
    H.barrier()
    e0.record(ext)
    for _ in range(steps):
        S.particles(dt)
    e1.record(ext)
    H.barrier()
    ms = H.max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop()
    st1 = S.stats()
    ms_step = ms / steps
    # K8 alone (no re-binning in the timed region): dt = 0 keeps the particles binned
    S.resort_particles()
    S.sync()
    e0.record(ext)
    for _ in range(10):
        S.particles(0.0)
    e1.record(ext)
    H.barrier()
    ms_k8 = H.max_over_ranks(e0.elapsed_time(e1)) / 10
    peak, peak_src = peaks()
    alg = 36 * n_local + 12 * ncell
    gbs = alg / ms_step / 1e6
    line = {
        "metric": "particle_rk2_updates_per_s", "value": n_local * world / (ms_step * 1e-3), "unit": "particle-updates/s", "n_gpus": world,
        "steps": args.steps, "timed_solves": steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, args.config, cfg, dims, n_local * world, world),
        "roofline": {"kernel": f"k_particles_rk2<f32,{args.arith}> (+ periodic re-binning inside the timed region)", "bound": "hbm",
                     "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None, "alg_bytes_per_launch": alg,
                     "peak_source": peak_src,
                     "k8_only": {"ms": ms_k8, "gbs": alg / ms_k8 / 1e6, "frac": alg / ms_k8 / 1e6 / peak, "note": "10 launches with dt = 0 right after a re-binning"}},
        "cpu_baseline": None, "e2e": None,
        "gpu_launches": int(st1.kernel_launches - st0.kernel_launches) * world, "resorts_in_timed_region": int(st1.resorts - st0.resorts),
        "clocks": clocks,
    }
    if rank == 0:
        print(json.dumps(line), flush=True)


def rand_field(dims, seed, scale=0.5):
    """smooth random staggered velocity for the parity block (same recipe as tests/util.rand_field)"""
    rng = np.random.default_rng(seed)
    X, Y, Z = dims
    i, j, k = np.meshgrid(np.arange(X), np.arange(Y), np.arange(Z), indexing="ij")
    out = []
    for _ in range(3):
        a = rng.normal(size=3)
        ph = rng.uniform(0, 6.28, size=3)
        f = np.sin(0.7 * i * a[0] + ph[0]) + np.sin(0.5 * j * a[1] + ph[1]) + np.sin(0.9 * k * a[2] + ph[2])
        out.append((scale * (f / 3.0 + 0.25 * rng.normal(size=dims))).astype(np.float32))
    return out


def rel_inf(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def full_size_parity(dims, u, v, w, pos, vel, device=0, label=""):
    """GPU path against the CPU oracle at FULL workload size, from one identical state (u, v, w, pos, vel), ONE step taken phase
    by phase as tests/test_gpu_parity.py::test_gpu_step_phase_by_phase does at config 1: dt equal; convection, forces, viscosity
    bit-exact; projected velocity <= 1e-5 of the field scale with a divergence no worse than the exact solve's; then, from the
    oracle's projected field, vorticity and particles (positions AND velocities) bit-exact in strict arithmetic, and the
    fast-arithmetic particle kernel <= 1e-6 relative apart from wall-clamp flips (counted).  The oracle is the checker here
    (cpu_baseline leg / tests only); nothing timed runs through it."""
    import oracle
    from ink_simulation_b200 import System

    oracle.build()
    oracle.set_threads()
    S = System(*dims, device=device, arith="strict", cg_rtol=1e-7)
    O = oracle.OracleSystem(*dims, dtype=np.float32)
    for T in (S, O):
        T.set_velocity(u, v, w)
        T.set_particles(pos, vel)
    vstack = lambda T: np.stack(T.get_velocity())   # noqa: E731
    dt = O.calc_dt(RENDER_DT)
    out = {"case": f"{label}{dims[0]}x{dims[1]}x{dims[2]}, {len(pos)} particles, strict arithmetic, cg_rtol 1e-7, one step phase by phase "
                   "from the developed state", "dt": dt, "dt_equal": bool(S.calc_dt(RENDER_DT) == dt)}
    exact = {}
    for ph in ("advect", "forces", "viscosity"):
        getattr(O, ph)(dt); getattr(S, ph)(dt)
        exact[ph] = bool(np.array_equal(vstack(O), vstack(S)))
    O.pressure(dt); S.pressure(dt)
    vo = vstack(O)
    out["projected_vel_rel"] = rel_inf(vstack(S), vo)
    out["p_rel"] = rel_inf(S.get_pressure(), O.get_pressure())
    ds, do = float(np.abs(S.divergence()).max()), float(np.abs(O.divergence()).max())
    out["div_max_gpu"], out["div_max_oracle"] = ds, do
    out["div_ok"] = bool(ds <= do * 1.001 + 1e-5 * float(np.abs(vo).max()))
    out["cg_iters"] = int(S.stats().cg_iters)
    S.set_velocity(*O.get_velocity())             # identical input for the phases after the solve
    O.vorticity(dt); S.vorticity(dt)
    exact["vorticity"] = bool(np.array_equal(vstack(O), vstack(S)))
    field = O.get_velocity()
    pos0, vel0 = O.get_particles()
    O.particles(dt); S.particles(dt)
    po, vo_p = O.get_particles()
    ps, vs_p = S.get_particles()
    exact["particles"] = bool(np.array_equal(po, ps) and np.array_equal(vo_p, vs_p))
    out["pos_rel"] = float(np.max(np.abs(ps.astype(np.float64) - po) / np.maximum(np.abs(po), 1.0)))
    S.close()
    del ps, vs_p, vo_p
    F = System(*dims, device=device, arith="fast")
    F.set_velocity(*field); F.set_particles(pos0, vel0)
    F.particles(dt)
    pf = F.get_particles()[0]
    F.close()
    rel = np.max(np.abs(pf.astype(np.float64) - po) / np.maximum(np.abs(po), 1.0), axis=1)
    flips = int((rel > 1e-6).sum())
    out["fast_arith"] = {"pos_rel_median": float(np.median(rel)), "pos_rel_max_without_flips": float(rel[rel <= 1e-6].max()) if flips < len(rel) else None,
                         "clamp_flips": flips, "flip_budget": max(2, len(rel) // 50000)}
    out["bit_exact_phases"] = exact
    out["tolerances"] = {"projected_vel_rel": 1e-5, "pos_rel": 1e-6}
    out["ok"] = bool(out["dt_equal"] and all(exact.values()) and out["projected_vel_rel"] < 1e-5 and out["div_ok"] and out["pos_rel"] <= 1e-6
                     and flips <= out["fast_arith"]["flip_budget"])
    O.close()
    return out


def parity_block(args, H, dims=None, npart=None, fp64_too=True, develop=0):
    """N > 1: the distributed System (y-slabs + particle shards, peer-memory exchanges) against a single-GPU System of the same
    problem, from identical seeded input, phase by phase: stencil phases and particles bit-exact, projected velocity
    <= 1e-5 of the field scale, particle positions <= 1e-6 relative after a whole step.  >= 32 rows per rank and Z % 4 == 0, so
    the vectorised band kernels, both-sided interior ranks and the fused peer pushes are what runs.  Default: a small problem
    (64 x 32N x 64); the nested c4 block calls it with the FULL 256^3 grid (every rank also holds the single-GPU System).
    develop > 0: instead of the rough random field, the state a single-GPU System reaches after `develop` steps from rest with
    these particles (the bench's own kind of flow; on a 256^3 grid the rough random field is so ill-conditioned that two solves
    which both meet rtol 1e-7 differ by 1.7e-5 in the projected velocity, measured at 2 GPUs).
    fp64_too: the whole step once more in fp64 on both, to show that the fp32 end-of-step figure is solver round-off amplified
    by the un-normalised confinement force and not a difference between the two paths."""
    from ink_simulation_b200 import System, create_distributed, shard_range

    rank, world, local = H.rank, H.world, H.local
    if dims is None:
        dims, npart = (64, 32 * world, 64), 100_000 * world
    u, v, w = rand_field(dims, 11) if develop <= 0 else (None, None, None)
    rng = np.random.default_rng(12)
    part = np.empty((npart, 3), np.float32)
    for s0 in range(0, npart, 4_000_000):
        e0 = min(npart, s0 + 4_000_000)
        part[s0:e0] = (rng.random((e0 - s0, 3)) * np.array(dims)).astype(np.float32)
    lo, hi = shard_range(npart, rank, world)
    kw = dict(device=local, cg_rtol=1e-7, arith="strict")
    pvel = None
    if develop > 0:
        T = System(*dims, **kw)
        T.set_particles(part)
        for _ in range(develop):
            T.solve(RENDER_DT)
        u, v, w = T.get_velocity()
        part, pvel = T.get_particles()
        T.close()
    D, S = create_distributed(*dims, **kw), System(*dims, **kw)
    for T in (D, S):
        T.set_velocity(u, v, w)
        T.set_particles(part, pvel)
    exact = {}
    dt = 0.17
    for phase in ("forces", "viscosity"):
        getattr(D, phase)(dt); getattr(S, phase)(dt)
        exact[phase] = bool(np.array_equal(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
    D.pressure(dt); S.pressure(dt)
    vel_rel = rel_inf(np.stack(D.get_velocity()), np.stack(S.get_velocity()))
    p_rel = rel_inf(D.get_pressure(), S.get_pressure())
    it = (int(D.stats().cg_iters), int(S.stats().cg_iters))
    D.set_velocity(*S.get_velocity())
    D.vorticity(dt); S.vorticity(dt)
    exact["vorticity"] = bool(np.array_equal(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
    D.particles(dt); S.particles(dt)
    pd, vd = D.get_particles(); ps, vs = S.get_particles()
    exact["particles"] = bool(np.array_equal(pd, ps[lo:hi]) and np.array_equal(vd, vs[lo:hi]))
    # a whole step from identical state
    D.set_velocity(*S.get_velocity()); D.set_particles(ps, vs)
    dtd, dts = D.solve(RENDER_DT), S.solve(RENDER_DT)
    pd = D.get_particles()[0]; ps = S.get_particles()[0]
    pos_rel = float(np.max(np.abs(pd - ps[lo:hi]) / np.maximum(np.abs(ps[lo:hi]), 1.0))) if hi > lo else 0.0
    step_vel_rel = rel_inf(np.stack(D.get_velocity()), np.stack(S.get_velocity()))
    ok = all(exact.values()) and vel_rel < 1e-5 and p_rel < 2e-5 and pos_rel < 1e-6 and dtd == dts
    D.close(); S.close()
    step_vel_rel_fp64 = None
    if fp64_too:
        kw64 = dict(kw, dtype=np.float64, cg_rtol=1e-12)
        D, S = create_distributed(*dims, **kw64), System(*dims, **kw64)
        for T in (D, S):
            T.set_velocity(u, v, w)
            T.set_particles(part)
        ok = ok and D.solve(RENDER_DT) == S.solve(RENDER_DT)
        step_vel_rel_fp64 = H.max_over_ranks(rel_inf(np.stack(D.get_velocity()), np.stack(S.get_velocity())))
        ok = ok and step_vel_rel_fp64 < 1e-9
        D.close(); S.close()
    ok_all = H.min_over_ranks(1.0 if ok else 0.0) > 0.5
    return {"world": world, "case": f"{dims[0]}x{dims[1]}x{dims[2]}, {npart} particles, strict arithmetic, cg_rtol 1e-7, "
                                    + (f"state developed over {develop} steps from rest" if develop > 0 else "random rough field"),
            "bit_exact_phases": exact, "vel_rel": H.max_over_ranks(vel_rel), "p_rel": H.max_over_ranks(p_rel),
            "pos_rel": H.max_over_ranks(pos_rel), "step_vel_rel": H.max_over_ranks(step_vel_rel), "step_vel_rel_fp64": step_vel_rel_fp64,
            "cg_iters_dist_single": it,
            "tolerances": {"vel_rel": 1e-5, "p_rel": 2e-5, "pos_rel": 1e-6, "step_vel_rel_fp64": 1e-9}, "ok": bool(ok_all)}


def nested_solve_block(args, H, name, strong, steps, arith=None, particles_from=None):
    """A second workload on the same N GPUs, device-resident only (no e2e / CPU legs): used for BASELINE configs[3] (c4,
    strong-scaled) and for the strict-arithmetic line.  Returns a dict, identical on every rank."""
    torch = H.torch
    from ink_simulation_b200 import System, create_distributed, shard_range

    cfg = CONFIGS[name]
    dims, n = scaled_workload(cfg, H.world, strong)
    lo, hi = shard_range(n, H.rank, H.world)
    dtype = np.float32 if args.dtype == "f32" else np.float64
    kw = dict(dtype=dtype, arith=arith or args.arith, device=H.local, cg_rtol=args.cg_rtol, resort_interval=args.resort, cg_precond=args.precond)
    S = create_distributed(*dims, **kw) if H.world > 1 else System(*dims, **kw)
    pos = particles_from if particles_from is not None else make_particles_device(dims, hi - lo, 0x1234 + 977 * H.rank, torch.device("cuda", H.local))
    S.set_particles_shard(pos)
    del pos
    torch.cuda.empty_cache()
    for _ in range(max(3, args.warmup)):
        S.solve(RENDER_DT)
    S.sync()
    st0 = S.stats()
    ms_total, iters = H.timed_solves(S, steps)
    st1 = S.stats()
    phase = S.phase_ms()
    ms_step = ms_total / steps
    ncell = dims[0] * dims[1] * dims[2]
    out = {"config": workload_config(args, name, cfg, dims, n, H.world, strong), "scaling": "strong" if (strong and H.world > 1) else "weak",
           "arith": arith or args.arith, "steps": steps, "ms_per_step": ms_step, "particle_updates_per_s": n / (ms_step * 1e-3),
           "cell_steps_per_s": ncell / (ms_step * 1e-3), "cg_iters": float(np.mean(iters)), "phase_ms_last_step": phase,
           "launches_per_step": (int(st1.kernel_launches - st0.kernel_launches)) / steps,
           "graph_launches_per_step": (int(st1.graph_launches - st0.graph_launches)) / steps}
    S.close()
    del S
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--particles", type=int, default=None)
    ap.add_argument("--arith", default="fast", choices=["strict", "fast"],
                    help="fast: FMA lerps in the particle sampler (<= 1e-6 relative vs the reference, tested); strict: bit-exact "
                         "(reported beside it as the `strict` block)")
    ap.add_argument("--precond", default="mg", choices=["mg", "mg_const", "none"])
    ap.add_argument("--strong", action="store_true", help="N > 1: keep the total problem fixed instead of the per-GPU problem")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cg-rtol", type=float, default=1e-6)
    ap.add_argument("--resort", type=int, default=16)
    ap.add_argument("--cpu-steps", type=int, default=5, help="oracle steps for cpu_baseline after one warm-up step (0 = skip)")
    ap.add_argument("--cpu-particles", type=int, default=10_000_000, help="particle cap PER GPU of the CPU legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-blocking", action="store_true", help="e2e through the blocking set/solve/get calls instead of the streaming ones")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel roofline micro-timings")
    ap.add_argument("--no-c4", action="store_true", help="skip the nested BASELINE configs[3] (256^3 / 100 M, strong-scaled) block")
    ap.add_argument("--c4-steps", type=int, default=10)
    ap.add_argument("--no-strict", action="store_true", help="skip the nested strict-arithmetic block (N = 1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the N > 1 parity block")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cfg = dict(CONFIGS[args.config])
    if args.particles:
        cfg["particles"] = args.particles
    if args.impl == "reference":
        run_reference_arm(args, cfg, rank, world)
        return

    import torch

    from ink_simulation_b200 import System

    torch.cuda.set_device(local)
    # One process per GPU: keep this rank's pinned staging buffers and copy-issuing thread on the GPU's own NUMA node (every
    # step moves 120 MB each way per GPU).  N = 1 is left alone: the cpu_baseline leg uses all host cores.
    host_affinity = "unchanged (single process)"
    if world > 1 and not os.environ.get("INK_NO_AFFINITY"):
        from ink_simulation_b200 import bind_host_to_gpu

        host_affinity = bind_host_to_gpu(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    H = Harness(rank, world, local)
    if cfg.get("sweep"):
        run_sweep(args, cfg, H)
        if world > 1:
            dist.destroy_process_group()
        return
    dims, n = scaled_workload(cfg, world, args.strong)
    ncell = dims[0] * dims[1] * dims[2]
    ncl = ncell // world
    dtype = np.float32 if args.dtype == "f32" else np.float64
    rbytes = 4 if args.dtype == "f32" else 8
    from ink_simulation_b200 import create_distributed, shard_range

    lo, hi = shard_range(n, rank, world)
    n_local = hi - lo
    pos_host = torch.from_numpy(make_particles(dims, n, cfg["cloud"], first=lo, count=n_local)).pin_memory()
    out_host = torch.empty_like(pos_host).pin_memory()

    kw = dict(dtype=dtype, arith=args.arith, device=local, cg_rtol=args.cg_rtol, resort_interval=args.resort, cg_precond=args.precond)
    S = create_distributed(*dims, **kw) if world > 1 else System(*dims, **kw)
    S.set_particles_shard(pos_host)

    for _ in range(max(3, args.warmup)):
        S.solve(RENDER_DT)
    S.sync()

    # ---- timed region: whole steps, state resident in HBM; at least --steps of them and at least MIN_TIMED_MS of device time --
    est_ms, _ = H.timed_solves(S, 3)
    n_solves = max(args.steps, int(math.ceil(MIN_TIMED_MS / max(est_ms / 3, 1e-3))))
    n_solves = int(H.max_over_ranks(float(n_solves)))
    st0 = S.stats()
    sampler = ClockSampler(local)
    ms_total, iters = H.timed_solves(S, n_solves)
    clocks = sampler.stop()
    st1 = S.stats()
    ms_step = ms_total / n_solves
    value = n / (ms_step * 1e-3)
    # phase split (CUDA events recorded inside ink_step on the library's stream): median over a few extra steps
    ph_runs, ph_iters = [], []
    for _ in range(9):
        S.solve(RENDER_DT)
        ph_runs.append(S.phase_ms())
        ph_iters.append(int(S.stats().cg_iters))
    phase_acc = {k: float(np.median([p[k] for p in ph_runs])) for k in ph_runs[0]}
    # the pressure roofline pairs iterations and time of the SAME nine steps (means of both), not the last step's count with a median
    last_iters = float(np.mean(ph_iters))
    cg_ms_mean = float(np.mean([p["pressure"] for p in ph_runs])) if "pressure" in ph_runs[0] else 0.0

    # ---- e2e: host positions in -> solve -> host positions out, every step ------------------------------
    e2e = None
    if not args.no_e2e:
        # Every step: positions come from pinned host memory (H2D, 12 B/particle), System::solve runs, the new positions go
        # back to pinned host memory (D2H, 12 B/particle).  Streaming calls of the C ABI (ink_prefetch_positions /
        # ink_commit_positions / ink_step / ink_fetch_positions_async): the copies run on their own streams, so the upload of
        # step i+1's input and the download of step i's result overlap the compute in between (inputs are independent host
        # buffers, as with a producer feeding frames; results land one step later).  All copies of the k2 steps start and
        # finish inside the timed region (sync_io before the closing event).  --e2e-blocking: the serial set/solve/get calls.
        S.get_particles(0, out_pos=out_host)
        in_bufs = [out_host.clone().pin_memory(), out_host.clone().pin_memory()]   # two producer buffers holding the current cloud
        out_bufs = [out_host, torch.empty_like(out_host).pin_memory()]
        k2 = max(args.steps, 50)
        ext, e0, e1 = H.events(S)
        if args.e2e_blocking:
            H.barrier()
            e0.record(ext)
            for i in range(k2):
                S.set_particles_shard(in_bufs[i & 1])
                S.solve(RENDER_DT)
                S.get_particles(0, out_pos=out_bufs[i & 1])
            e1.record(ext)
        else:
            for i in range(2):   # warm the staging buffers / streams (untimed)
                S.prefetch_positions(in_bufs[i & 1]); S.commit_positions(); S.solve(RENDER_DT); S.fetch_positions_async(out_bufs[i & 1])
            S.sync_io()
            H.barrier()
            e0.record(ext)
            S.prefetch_positions(in_bufs[0])
            for i in range(k2):
                S.commit_positions()
                if i + 1 < k2:
                    S.prefetch_positions(in_bufs[(i + 1) & 1])
                S.solve(RENDER_DT)
                S.fetch_positions_async(out_bufs[i & 1])
            S.sync_io()          # every copy of the k2 steps has landed ...
            e1.record(ext)       # ... before the closing event is recorded
        H.barrier()
        ms_e2e = H.max_over_ranks(e0.elapsed_time(e1)) / k2
        e2e = {"value": n / (ms_e2e * 1e-3), "unit": "particle-updates/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": n_local * 12, "d2h_bytes_per_step": n_local * 12,
               "mode": "blocking calls" if args.e2e_blocking else "streaming calls: H2D(i+1) and D2H(i-1) overlap step i", "steps": k2}
        if True:
            # copy-only loop (no ink_step): the PCIe / host-memory floor of this e2e pattern with N ranks copying at once
            S.sync_io()
            H.barrier()
            e0.record(ext)
            S.prefetch_positions(in_bufs[0])
            for i in range(k2):
                S.commit_positions()
                if i + 1 < k2:
                    S.prefetch_positions(in_bufs[(i + 1) & 1])
                S.fetch_positions_async(out_bufs[i & 1])
            S.sync_io()
            e1.record(ext)
            H.barrier()
            ms_copy = H.max_over_ranks(e0.elapsed_time(e1)) / k2
            e2e["copy_only_ms_per_step"] = ms_copy
            e2e["copy_only_aggregate_gbs"] = 24 * n / ms_copy / 1e6
            e2e["note"] = ("copy_only = the same H2D + D2H traffic with no compute in between: the host-memory / PCIe floor of the box with "
                           f"{world} rank(s) copying at once; e2e cannot be faster than that")

    # ---- per-kernel micro-timings (CUDA events on the library stream, same process, after the timed region) -------------
    peak, peak_src = peaks()
    kernels = {}
    ext, e0, e1 = H.events(S)
    if not args.no_kernels and world == 1:
        # K8 particles_rk2 on the current (developed) field, re-binned first; 36 B/particle (fp32) + field once
        S.resort_particles()
        S.sync()
        reps = 20
        e0.record(ext)
        for _ in range(reps):
            S._ck(S.L.ink_particles(S.h, 0.0))  # dt = 0: same memory traffic and arithmetic, particles stay binned
        e1.record(ext)
        e1.synchronize()
        ms = e0.elapsed_time(e1) / reps
        alg = (9 * rbytes) * n + 3 * rbytes * ncell
        kernels["k_particles_rk2"] = {"ms": ms, "alg_bytes": alg, "gbs": alg / ms / 1e6, "frac": alg / ms / 1e6 / peak,
                                      "note": "20 launches with dt = 0 right after a re-binning (cache-warm field)"}
        # CG iteration pair ALONE (k_cg_dir_matvec + k_cg_update), no preconditioner: 10*R B/cell/iteration; exactly 100
        # iterations (cg_max_iters = 100, tolerance 1e-30 on a random right-hand side is never met)
        from ink_simulation_b200.system import _ptr

        S2 = System(*dims, dtype=dtype, device=local, cg_rtol=1e-30 if rbytes == 8 else 0.0, cg_max_iters=100, cg_check_every=100, cg_precond="none")
        b = torch.randn(dims, dtype=torch.float32 if rbytes == 4 else torch.float64, device="cuda")
        pout = torch.empty_like(b)
        torch.cuda.synchronize()
        hd = 32 if rbytes == 4 else 64
        S2._ck(S2.L.ink_poisson_solve(S2.h, _ptr(b), _ptr(pout), hd))   # _ptr orders the device inputs before the ctx stream
        ext2, f0, f1 = H.events(S2)
        f0.record(ext2)
        S2._ck(S2.L.ink_poisson_solve(S2.h, _ptr(b), _ptr(pout), hd))   # _ptr orders the device inputs before the ctx stream
        f1.record(ext2)
        f1.synchronize()
        ms = f0.elapsed_time(f1)
        it = S2.stats().cg_iters
        alg = 10 * rbytes * ncell * it
        kernels["cg_pair_unpreconditioned"] = {"ms_per_iter": ms / max(it, 1), "iters": it, "alg_bytes_per_iter": 10 * rbytes * ncell,
                                               "gbs": alg / ms / 1e6, "frac": alg / ms / 1e6 / peak,
                                               "note": f"{it} iterations of the CG pair without preconditioner; includes the conversion / transposes of b and p (2 passes) around them"}
        S2.close()
        del b, pout

    # ---- roofline ------------------------------------------------------------------------------------------------------
    # Lead entry: the pressure phase (PCG: k_cg_* + k_mg3_*), the largest share of a step.  Algorithmic bytes = what the
    # ALGORITHM needs on the fine level per iteration (CG pair 10 vector passes + V-cycle 5.25 [trilinear V(1,1)] or 11.5
    # [constant V(2,2)]) x iterations + 16 passes around the solve (RHS, warm start, gradient subtract);
    # the launch-latency-bound coarse levels count against it.  K8 and the stencil phases follow in per_kernel.
    cg_ms = cg_ms_mean
    part_ms = phase_acc.get("particles", 0.0)
    alg_k8 = (9 * rbytes) * n_local + 3 * rbytes * ncell
    traffic_note = None
    traffic = None
    pcg_iter_traffic = None
    if world == 1 and args.config == "c3" and args.dtype == "f32":
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            traffic = tj.get(f"k_particles_rk2:{args.config}:{args.arith}:{args.dtype}")
            traffic_note = tj.get("_capture", "ncu --set full capture (profiles/), not this run")
            pcg_iter_traffic = tj.get("pcg_iteration:c3:f32")
        except Exception:
            pass
    per_kernel = [{"kernel": f"k_particles_rk2<{args.dtype},{args.arith}>", "bound": "hbm", "achieved": alg_k8 / max(part_ms, 1e-9) / 1e6,
                   "frac": alg_k8 / max(part_ms, 1e-9) / 1e6 / peak, "alg_bytes_per_launch": alg_k8, "launch_ms": part_ms, "traffic": traffic,
                   "traffic_source": traffic_note,
                   "note": "36 B per particle-update + the velocity field once; 48 trilinear corner gathers per particle"}]
    for name, bpc in (("forces", 25), ("viscosity", 24), ("vorticity", 60)):
        ms = phase_acc.get(name, 0.0)
        if ms > 0 and world == 1:   # at N > 1 these phases also hold exchanges: only meaningful at N = 1
            b = bpc * (rbytes / 4) * ncl
            per_kernel.append({"kernel": {"forces": "k_dilate_add_gravity_v4", "viscosity": "k_viscosity_v4", "vorticity": "k_curl_v4 + k_vorticity_v4"}[name],
                               "bound": "hbm", "achieved": b / ms / 1e6, "frac": b / ms / 1e6 / peak, "alg_bytes_per_launch": b, "launch_ms": ms})
    vec_passes = (10 + {"mg": 5.25, "mg_const": 11.5, "none": 0}[args.precond]) * max(last_iters, 0) + 16
    alg_p = vec_passes * rbytes * ncl
    roof = {"kernel": "pressure phase: PCG (k_cg_dir_matvec, k_cg_update, k_mg3_* V-cycle) + RHS + gradient subtract", "bound": "hbm",
            "achieved": alg_p / max(cg_ms, 1e-9) / 1e6, "peak": peak, "unit": "GB/s", "frac": alg_p / max(cg_ms, 1e-9) / 1e6 / peak,
            "traffic": (pcg_iter_traffic * max(last_iters, 0)) if pcg_iter_traffic else None,
            "traffic_source": "DRAM bytes of one PCG iteration (ncu --set full, profiles/r2_h_top_kernels.txt, C3 on one GPU; not this run) x iterations: "
                              "below the algorithmic bytes because most of the working set stays in L2" if pcg_iter_traffic else None,
            "peak_source": peak_src, "alg_bytes": alg_p, "ms": cg_ms, "cg_iters": last_iters,
            "pairing": "mean iterations and mean pressure-phase time of the same 9 steps after the timed loop",
            "share_of_step": cg_ms / max(sum(phase_acc.values()), 1e-9),
            "survey_8d": {"bytes_per_cell_per_iter": 44, "frac": (44 * last_iters + 44) * (rbytes / 4) * ncl / max(cg_ms, 1e-9) / 1e6 / peak},
            "per_kernel": per_kernel}

    cpu = None
    parity = None
    if args.cpu_steps > 0 and world == 1:
        sample_n = min(n, args.cpu_particles)
        u, v, w = S.get_velocity()
        sec, threads = cpu_port_sample(dims, pos_host.numpy()[:sample_n], args.cpu_steps, warm_state=(u, v, w))
        cpu = {"value": sample_n / sec, "unit": "particle-updates/s", "cores": threads, "kind": "port",
               "sample": f"{args.cpu_steps} System::solve steps after 1 warm-up step of the same workload ({sample_n} particles, grid {dims}) "
                         f"from the GPU's developed velocity field; OpenMP oracle port on {threads} threads, exact DST Poisson solve",
               "ms_per_step": sec * 1e3, "mac_cell_steps_per_s": ncell / sec}
        if not args.no_parity:
            # parity AT THE SIZE THAT IS TIMED, from the very state the timed loop left behind
            pp, pv = S.get_particles()
            parity = full_size_parity(dims, u, v, w, pp, pv, device=local, label=f"{args.config}: ")
            del pp, pv
        del u, v, w

    launches_per_step = (int(st1.kernel_launches - st0.kernel_launches)) / n_solves
    graphs_per_step = (int(st1.graph_launches - st0.graph_launches)) / n_solves
    water_ms = sum(phase_acc.get(k, 0.0) for k in ("dt", "advect", "forces", "viscosity", "pressure", "vorticity"))
    line = {
        "metric": "particle_rk2_updates_per_s", "value": value, "unit": "particle-updates/s", "n_gpus": world, "steps": args.steps,
        "timed_solves": n_solves, "timed_region_ms": ms_total, "timed_region_s": ms_total * 1e-3,
        "steps_note": "`steps` echoes --steps; the device-timed loop ran `timed_solves` whole steps (>= --steps and >= 0.6 s, for clock sampling) "
                      "and ms_per_step = timed_region_ms / timed_solves",
        "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if (world > 1 and args.strong) else "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": workload_config(args, args.config, cfg, dims, n, world, args.strong), "host_affinity": host_affinity,
        "mac_cell_steps_per_s": ncell / (ms_step * 1e-3),
        "mac_only_cell_steps_per_s": ncell / (water_ms * 1e-3) if water_ms > 0 else None,
        "particles_only_updates_per_s": n / (part_ms * 1e-3) if part_ms > 0 else None,
        "cg_iters_per_step": {"mean": float(np.mean(iters)), "min": int(min(iters)), "max": int(max(iters))},
        "phase_ms_median": phase_acc,
        "roofline": roof, "kernels": kernels, "cpu_baseline": cpu, "e2e": e2e,
        "gpu_launches": int(st1.kernel_launches - st0.kernel_launches) * world, "launches_per_step": launches_per_step,
        "graph_launches_per_step": graphs_per_step, "resorts_in_timed_region": int(st1.resorts - st0.resorts),
        "clocks": clocks,
    }
    # ---- nested blocks: free the main System first ------------------------------------------------------------------------
    strict_pos = None
    if world == 1 and not args.no_strict and args.arith != "strict":
        strict_pos = pos_host.to(torch.device("cuda", local))
    S.close()
    del S, pos_host, out_host
    if e2e is not None:
        del in_bufs, out_bufs
    torch.cuda.empty_cache()
    if strict_pos is not None:
        # the bit-exact arithmetic (the ABI default, what the parity suite pins) on the same workload, beside the headline
        blk = nested_solve_block(args, H, args.config, args.strong, n_solves, arith="strict", particles_from=strict_pos)   # same protocol: same number of timed solves
        line["strict"] = {k: blk[k] for k in ("arith", "steps", "ms_per_step", "particle_updates_per_s", "cell_steps_per_s", "cg_iters", "phase_ms_last_step")}
        del strict_pos
        torch.cuda.empty_cache()
    if world > 1 and not args.no_parity:
        line["parity"] = parity_block(args, H)
    elif parity is not None:
        line["parity"] = parity
    if args.config == "c3" and not args.no_c4 and not args.particles:
        # BASELINE configs[3]: 256^3 grid, 100 M particles, y-slab sharded over the same N GPUs, total problem FIXED
        line["c4"] = nested_solve_block(args, H, "c4", True, max(args.c4_steps, 3))
        if world > 1 and not args.no_parity:
            # the distributed path against the single-GPU one on the FULL 256^3 grid (16 M particles: every rank generates and
            # holds the whole cloud for the single-GPU side), one step phase by phase
            line["c4"]["parity"] = parity_block(args, H, dims=CONFIGS["c4"]["dims"], npart=16_000_000, fp64_too=False, develop=5)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
