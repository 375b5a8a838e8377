#!/usr/bin/env python
"""bench.py — throughput of the Ink-Simulation hot path (System::solve) on B200.

One "step" = one System::solve (MAC water-grid update + RK2 particle update, reference src/inkSolver.cpp:5-9) on
synthetic input.  Default workload at N=1 = BASELINE.json configs[2]: 10M particles (uniform, UNSORTED input order)
on a 128x256x128 MAC grid, fp32.  Prints ONE JSON line (see the contract in the task statement):
  value      : particle RK2 updates/s over whole steps (state resident in HBM)
  e2e        : same, through the host-facing call: pinned-host positions in (H2D) -> solve -> positions out (D2H) every step
  roofline   : dominant kernel's algorithmic bytes / CUDA-event time vs MEASURED_PEAKS.json
  cpu_baseline: the CPU oracle port (OpenMP, all host cores) on a bounded sample of the same workload
`--impl reference` times the CPU implementation (compiled reference where it can run the config, else the oracle port).
"""
import argparse
import json
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


def make_particles(dims, n, cloud, seed=0x1234, first=0, count=None):
    """Particles [first, first+count) of the n-particle cloud (a rank only generates its own shard; chunk-seeded so the
    cloud does not depend on how it is sharded as long as shards start on chunk boundaries — they need not: the cloud
    is i.i.d. uniform, any seeded sample of it is the same workload)."""
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


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, dev):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(dev), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
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
    """The CPU oracle (OpenMP restatement of the reference, oracle/ink_oracle.cpp) on the same workload; returns
    (seconds per step, threads)."""
    import oracle

    oracle.build()
    O = oracle.OracleSystem(*dims, dtype=np.float32)
    O.set_particles(pos)
    if warm_state is not None:
        O.set_velocity(*warm_state)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.solve(RENDER_DT)
    dt = (time.perf_counter() - t0) / steps
    return dt, os.cpu_count()


def l2_note(args, dims, n, world):
    """How the timing rule 'inputs larger than L2 or flush between iterations' is met, from the actual sizes."""
    rb = 8 if args.dtype == "f64" else 4
    part_mb = 9 * rb * (n / world) / 1e6                      # per pass: read pos, write pos, write vel (3 components each)
    cells = dims[0] * dims[1] * dims[2] / world
    vec_mb = rb * cells / 1e6
    total = part_mb + 15 * vec_mb
    verdict = "inputs larger than L2" if total > 126 else "inputs FIT in L2: a parity-size case, not a bench line"
    return f"{verdict} (per GPU: particle arrays {part_mb:.0f} MB per pass + 15 grid vectors of {vec_mb:.1f} MB; L2 = 126 MB)"


def workload_config(args, cfg, dims, n, world):
    """The `config` object both arms print (the reference arm runs a bounded sample of it and says which)."""
    if cfg.get("sweep"):
        X, Y, Z = cfg["dims"]
        per = args.particles or cfg["particles"]
        return {"workload": f"{args.config}: RK2 advection sweep (System::updateParticles), {per * world} particles ({per} per GPU, uniform, unsorted "
                            f"input), frozen analytic {X}x{Y}x{Z} field replicated on every GPU, dt = 0.2", "arith": args.arith,
                "resort_interval": args.resort, "sharding": "particles by index, no data-path collective",
                "l2": "field 1.6 GB + particle arrays 4.5 GB per pass >> L2 (126 MB)"}
    return {"workload": f"{args.config}: {n} particles ({cfg['cloud']}, unsorted input), {dims[0]}x{dims[1]}x{dims[2]} MAC grid, "
                        f"full System::solve per step", "arith": args.arith, "cg_rtol": args.cg_rtol,
            "resort_interval": args.resort, "precond": args.precond,
            "sharding": "single GPU" if world == 1 else f"{world} y-slabs of {dims[1] // world} rows, particles by index; NCCL + peer-memory exchanges",
            "l2": l2_note(args, dims, n, world)}


def run_reference_arm(args, cfg, rank, world):
    """--impl reference: the CPU implementation on the host cores.  The compiled reference (hash-map grid + SparseLU)
    cannot run grids beyond ~32x64x32, so for c3/c4 the OpenMP oracle port stands in (kind: "port")."""
    if rank != 0:
        return
    import oracle

    oracle.build()
    dims, n = cfg["dims"], cfg["particles"]
    use_ref = oracle.have_ref() and dims[0] * dims[1] * dims[2] <= 20000 and dims[0] == dims[2]
    sample_n = n if use_ref else min(n, args.cpu_particles)
    pos = make_particles(dims, sample_n, cfg["cloud"])
    if use_ref:
        T = oracle.RefSystem(*dims)
        kind, cores = "reference", 1
    else:
        T = oracle.OracleSystem(*dims, dtype=np.float32)
        kind, cores = "port", os.cpu_count()
    T.set_particles(pos)
    if cfg.get("sweep"):   # config 5: particles-only sweep on the frozen analytic field (same field as run_sweep)
        X, Y, Z = dims
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
    val = sample_n / sec
    sample = (f"{args.steps} {'updateParticles sweeps' if cfg.get('sweep') else 'System::solve steps'} of one GPU's share of the workload: grid {dims[0]}x{dims[1]}x{dims[2]}, {sample_n} particles "
              f"({cfg['cloud']}); {'compiled reference, 1 thread (it has no threading)' if use_ref else 'OpenMP oracle port, exact DST Poisson solve'}")
    wdims, wn = dims, n
    if world > 1 and not args.strong:
        wdims, wn = (dims[0], dims[1] * world, dims[2]), n * world
    line = {
        "impl": "reference", "metric": "particle_rk2_updates_per_s", "value": val, "unit": "particle-updates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, cfg, wdims, wn, world),
        "mac_cell_steps_per_s": dims[0] * dims[1] * dims[2] / sec,
        "cpu_baseline": {"value": val, "unit": "particle-updates/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "particle-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_sweep(args, cfg, rank, world, local):
    """Config 5: particles-only RK2 sweep (ink_particles = System::updateParticles) on a frozen analytic field.  Every rank is a
    single-GPU System holding the WHOLE field and its index shard of the particles: no data-path collective (weak scaling)."""
    import torch

    from ink_simulation_b200 import System

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
    pos = torch.from_numpy(make_particles(dims, n_local * world, "uniform", first=rank * n_local, count=n_local))
    S.set_particles_shard(pos)
    del pos
    dt = 0.2
    for _ in range(max(3, args.warmup)):
        S.particles(dt)
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.barrier(); torch.cuda.synchronize()
    st0 = S.stats()
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        S.particles(dt)
    S.sync()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop()
    st1 = S.stats()
    ms_step = ms / args.steps
    peak, peak_src = peaks()
    alg = 36 * n_local + 12 * ncell
    gbs = alg / ms_step / 1e6
    line = {
        "metric": "particle_rk2_updates_per_s", "value": n_local * world / (ms_step * 1e-3), "unit": "particle-updates/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, cfg, dims, n_local * world, world),
        "roofline": {"kernel": f"k_particles_rk2<f32,{args.arith}> (+ periodic re-binning inside the timed region)", "bound": "hbm",
                     "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None, "alg_bytes_per_launch": alg,
                     "peak_source": peak_src},
        "cpu_baseline": None, "e2e": None,
        "gpu_launches": int(st1.kernel_launches - st0.kernel_launches) * world, "resorts_in_timed_region": int(st1.resorts - st0.resorts),
        "clocks": clocks,
    }
    if rank == 0:
        print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS))
    ap.add_argument("--particles", type=int, default=None)
    ap.add_argument("--arith", default="fast", choices=["strict", "fast"],
                    help="fast: FMA lerps in the particle sampler (<= 1e-6 relative vs the reference, tested); strict: bit-exact")
    ap.add_argument("--precond", default="mg", choices=["mg", "mg_const", "none"])
    ap.add_argument("--strong", action="store_true", help="N > 1: keep the total problem fixed instead of the per-GPU problem")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cg-rtol", type=float, default=1e-6)
    ap.add_argument("--resort", type=int, default=16)
    ap.add_argument("--cpu-steps", type=int, default=1, help="oracle steps for cpu_baseline (0 = skip)")
    ap.add_argument("--cpu-particles", type=int, default=10_000_000)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-blocking", action="store_true", help="e2e through the blocking set/solve/get calls instead of the streaming ones")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel roofline micro-timings")
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
    # step moves 120 MB each way per GPU; round 1 measured the e2e of 8 unbound ranks at ~100 GB/s for the whole box).
    # N = 1 is left alone: the cpu_baseline leg uses all host cores.  INK_NO_AFFINITY=1 switches it off for A/B runs.
    host_affinity = "unchanged (single process)"
    if world > 1 and not os.environ.get("INK_NO_AFFINITY"):
        from ink_simulation_b200 import bind_host_to_gpu

        host_affinity = bind_host_to_gpu(local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dims, n = cfg["dims"], cfg["particles"]
    if world > 1 and not args.strong:   # weak scaling: every GPU keeps one config-sized y-slab and particle shard
        dims = (dims[0], dims[1] * world, dims[2])
        n = n * world
    if cfg.get("sweep"):
        run_sweep(args, cfg, rank, world, local)
        if world > 1:
            dist.destroy_process_group()
        return
    ncell = dims[0] * dims[1] * dims[2]
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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(3, args.warmup)):
        S.solve(RENDER_DT)
    barrier()

    # ---- timed region: K whole steps, state resident in HBM -------------------------------------------
    st0 = S.stats()
    sampler = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phase_acc = {}
    iters = []
    barrier()
    e0.record()
    for _ in range(args.steps):
        S.solve(RENDER_DT)
        iters.append(S.stats().cg_iters)
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop()
    st1 = S.stats()
    ms_step = ms_total / args.steps
    value = n / (ms_step * 1e-3)
    # phase split of the LAST timed step (CUDA events recorded inside ink_step on the launching stream)
    phase_acc = S.phase_ms()

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
        k2 = max(3, args.steps)
        if args.e2e_blocking:
            barrier()
            e0.record()
            for i in range(k2):
                S.set_particles_shard(in_bufs[i & 1])
                S.solve(RENDER_DT)
                S.get_particles(0, out_pos=out_bufs[i & 1])
            e1.record()
        else:
            for i in range(2):   # warm the staging buffers / streams (untimed)
                S.prefetch_positions(in_bufs[i & 1]); S.commit_positions(); S.solve(RENDER_DT); S.fetch_positions_async(out_bufs[i & 1])
            S.sync_io()
            barrier()
            e0.record()
            S.prefetch_positions(in_bufs[0])
            for i in range(k2):
                S.commit_positions()
                if i + 1 < k2:
                    S.prefetch_positions(in_bufs[(i + 1) & 1])
                S.solve(RENDER_DT)
                S.fetch_positions_async(out_bufs[i & 1])
            S.sync_io()
            e1.record()
        barrier()
        ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / k2
        e2e = {"value": n / (ms_e2e * 1e-3), "unit": "particle-updates/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": n_local * 12, "d2h_bytes_per_step": n_local * 12,
               "mode": "blocking calls" if args.e2e_blocking else "streaming calls: H2D(i+1) and D2H(i-1) overlap step i", "steps": k2}

    # ---- per-kernel roofline micro-timings (CUDA events, same process, after the timed region) -------------
    peak, peak_src = peaks()
    kernels = {}
    if not args.no_kernels and world == 1:
        # K8 particles_rk2 on the current (developed) field, re-binned first; 36 B/particle (fp32) + field once
        S.resort_particles()
        dt = S.stats().dt
        torch.cuda.synchronize()
        reps = 10
        e0.record()
        for _ in range(reps):
            S._ck(S.L.ink_particles(S.h, 0.0))  # dt = 0: same memory traffic and arithmetic, particles stay binned
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        alg = (9 * rbytes) * n + 3 * rbytes * ncell
        kernels["k_particles_rk2"] = {"ms": ms, "alg_bytes": alg, "gbs": alg / ms / 1e6, "frac": alg / ms / 1e6 / peak}
        # CG iteration pair (k_cg_dir_matvec + k_cg_update): 10*R B/cell/iteration; fixed 100 iterations, tolerance off
        S2 = System(*dims, dtype=dtype, device=local, cg_rtol=0.0, cg_max_iters=100, cg_check_every=100)
        b = torch.randn(dims, dtype=torch.float32 if rbytes == 4 else torch.float64, device="cuda")
        pout = torch.empty_like(b)
        import ctypes as C

        S2._ck(S2.L.ink_poisson_solve(S2.h, C.c_void_p(b.data_ptr()), C.c_void_p(pout.data_ptr()), 32 if rbytes == 4 else 64))
        torch.cuda.synchronize()
        e0.record()
        S2._ck(S2.L.ink_poisson_solve(S2.h, C.c_void_p(b.data_ptr()), C.c_void_p(pout.data_ptr()), 32 if rbytes == 4 else 64))
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        it = S2.stats().cg_iters
        alg = 10 * rbytes * ncell * it
        kernels["cg_iteration"] = {"ms_per_iter": ms / max(it, 1), "iters": it, "alg_bytes_per_iter": 10 * rbytes * ncell,
                                   "gbs": alg / ms / 1e6, "frac": alg / ms / 1e6 / peak,
                                   "note": "includes conversion/transposes of b and p (2 passes) around 100 iterations"}
        S2.close()

    # ---- roofline of the dominant kernel -------------------------------------------------------------------------------
    # k_particles_rk2 is the longest single launch of a step and the kernel BASELINE.json's headline metric counts.
    # achieved = algorithmic bytes per launch (36 B per particle-update + the velocity field read once) /
    # average launch duration from the CUDA events ink_step records around the particle phase on ITS stream (timed region).
    cg_ms = phase_acc.get("pressure", 0.0)
    part_ms = phase_acc.get("particles", 0.0)
    last_iters = iters[-1] if iters else 0
    alg_k8 = (9 * rbytes) * n_local + 3 * rbytes * ncell
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"k_particles_rk2:{args.config}:{args.arith}:{args.dtype}")
    except Exception:
        pass
    roof = {"kernel": f"k_particles_rk2<{args.dtype},{args.arith}>", "bound": "hbm", "achieved": alg_k8 / max(part_ms, 1e-9) / 1e6,
            "peak": peak, "unit": "GB/s", "frac": alg_k8 / max(part_ms, 1e-9) / 1e6 / peak, "traffic": traffic,
            "peak_source": peak_src + " (MEASURED_PEAKS.json hbm_gbs)" if peak_src == "measured" else "fallback 6650",
            "alg_bytes_per_launch": alg_k8, "launch_ms": part_ms,
            "note": "issue/L1-gather bound, not HBM bound: 48 trilinear corner gathers per particle (profiles/)"}
    # the pressure phase as a whole (CG pair + V-cycle kernels): bytes the ALGORITHM needs per iteration on the fine level
    # (CG 10 + V-cycle 5.25 [trilinear V(1,1)] or 11.5 [constant V(2,2)] vector passes) vs time; launch-latency-bound coarse levels count against it
    if last_iters > 0 and cg_ms > 0:
        ncl = ncell // world
        vec_passes = (10 + {"mg": 5.25, "mg_const": 11.5, "none": 0}[args.precond]) * last_iters + 16
        roof_p = {"what": "pressure phase (PCG: k_cg_* + k_mg_*)", "cg_iters": last_iters, "ms": cg_ms,
                  "alg_bytes": vec_passes * rbytes * ncl, "achieved": vec_passes * rbytes * ncl / cg_ms / 1e6,
                  "frac": vec_passes * rbytes * ncl / cg_ms / 1e6 / peak}
    else:
        roof_p = None

    # MAC stencil phases of the last timed step (CUDA events inside ink_step): algorithmic bytes per cell vs time.
    # forces = K2b dilate + gravity (25 B/cell), viscosity = K3 (24), vorticity = K7a curl + K7b confinement (24 + 36; at N > 1 the
    # phase also holds the velocity all-gather, so the fraction is only meaningful at N = 1)
    mac_phases = {}
    for name, bpc in (("forces", 25), ("viscosity", 24), ("vorticity", 60)):
        ms = phase_acc.get(name, 0.0)
        if ms > 0:
            b = bpc * (rbytes / 4) * (ncell // world)
            mac_phases[name] = {"ms": ms, "alg_bytes": b, "gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak}

    cpu = None
    if args.cpu_steps > 0 and world == 1:
        sample_n = min(n, args.cpu_particles)
        u, v, w = S.get_velocity()
        sec, cores = cpu_port_sample(dims, pos_host.numpy()[:sample_n], args.cpu_steps, warm_state=(u, v, w))
        cpu = {"value": sample_n / sec, "unit": "particle-updates/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_steps} System::solve step(s) of the same workload ({sample_n} particles, grid {dims}) "
                         f"from the GPU's developed velocity field; OpenMP oracle port, exact DST Poisson solve",
               "ms_per_step": sec * 1e3, "mac_cell_steps_per_s": ncell / sec}

    water_ms = sum(phase_acc.get(k, 0.0) for k in ("dt", "advect", "forces", "viscosity", "pressure", "vorticity"))
    line = {
        "metric": "particle_rk2_updates_per_s", "value": value, "unit": "particle-updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if (world > 1 and args.strong) else "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": dict(workload_config(args, cfg, dims, n, world), host_affinity=host_affinity),
        "mac_cell_steps_per_s": ncell / (ms_step * 1e-3),
        "mac_only_cell_steps_per_s": ncell / (water_ms * 1e-3) if water_ms > 0 else None,
        "particles_only_updates_per_s": n / (part_ms * 1e-3) if part_ms > 0 else None,
        "cg_iters_per_step": {"mean": float(np.mean(iters)), "min": int(min(iters)), "max": int(max(iters))},
        "phase_ms_last_step": phase_acc,
        "roofline": roof, "pressure_phase": roof_p, "mac_phases": mac_phases, "kernels": kernels, "cpu_baseline": cpu, "e2e": e2e,
        "gpu_launches": int(st1.kernel_launches - st0.kernel_launches) * world, "resorts_in_timed_region": int(st1.resorts - st0.resorts),
        "clocks": clocks,
    }
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
