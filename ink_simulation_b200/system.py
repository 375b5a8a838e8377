"""Host-side mirror of the reference's `System` (src/system.h:70-138) over the C ABI.

`System` keeps the reference's method names (`init`, `solve`, `getInkParticles`, `getWaterGrid`, `checkNanAndInf`)
and adds the per-phase hooks the parity tests need (same names as oracle.OracleSystem / RefSystem so a test reads
the same against all three).  All compute happens in libinkb200.so on the GPU; this file only marshals numpy arrays.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import _capi


@dataclass
class Params:
    """The global constants of reference src/system.h:13-38 (defaults = committed values)."""

    cell_dim: float = 1.0
    density: float = 0.95
    k_vort: float = 1.0
    viscosity: float = 1.0016
    p_atm: float = 1.0
    gravity: tuple = (0.0, -0.1, 0.0)
    k_cfl: float = 0.2
    dt_min: float = 0.005
    dt_max: float = 1.0
    buffer_size: int = 3


class InkError(RuntimeError):
    pass


def _ptr(a):
    """Pointer of `a` for the C ABI, with the ordering the ABI asks for: the ctx runs on its own non-blocking stream, so a DEVICE
    input must be complete before the call (include/ink_capi.h).  numpy arrays and host tensors need nothing.  A torch CUDA tensor:
    torch's current stream of its device is synchronised.  Any other object with __cuda_array_interface__ (cupy, numba, ...): the
    stream the interface names is synchronised (v3 `stream`: 1 = legacy default, 2 = per-thread default, else a cudaStream_t), the
    whole device when it names none.  A bare integer address carries no such information: the caller orders it (synchronise, or
    make the ctx stream from cuda_stream() wait on an event)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):  # torch tensor (host pinned or device) — plumbing only
        if getattr(a, "is_cuda", False):
            import torch

            torch.cuda.current_stream(a.device).synchronize()
        return C.c_void_p(a.data_ptr())
    cai = getattr(a, "__cuda_array_interface__", None)
    if cai is not None:
        import torch

        st = cai.get("stream")
        if st is None:
            torch.cuda.synchronize()
        elif st == 1:
            torch.cuda.default_stream().synchronize()     # torch's default stream is the legacy default stream
        elif st == 2:
            torch.cuda.synchronize()                      # per-thread default stream: no handle to wait on, take the device
        else:
            torch.cuda.ExternalStream(int(st)).synchronize()
        return C.c_void_p(int(cai["data"][0]))
    return C.c_void_p(int(a))


def _dev_meta(a):
    """(shape, itemsize, contiguous) of a torch tensor or of any __cuda_array_interface__ object"""
    if hasattr(a, "data_ptr"):
        return tuple(a.shape), a.element_size(), a.is_contiguous()
    cai = a.__cuda_array_interface__
    return tuple(cai["shape"]), np.dtype(cai["typestr"]).itemsize, cai.get("strides") is None


def _is_dev_like(a):
    return hasattr(a, "data_ptr") or hasattr(a, "__cuda_array_interface__")


def nccl_unique_id() -> bytes:
    """ncclGetUniqueId on the calling rank (rank 0); broadcast the 128 bytes to the other ranks yourself."""
    buf = (C.c_uint8 * 128)()
    L = _capi.load()
    if L.ink_nccl_unique_id(buf) != 0:
        raise InkError(L.ink_last_error(None).decode())
    return bytes(buf)


def shard_range(n, rank, world_size):
    """Index range [first, last) of an n-particle drop that `rank` keeps (same rule as ink_set_particles)."""
    return n * rank // world_size, n * (rank + 1) // world_size


def slab_range(Y, rank, world_size):
    """Grid rows [j0, j1) that `rank` owns (same rule as ink_create)."""
    return Y * rank // world_size, Y * (rank + 1) // world_size


def bind_host_to_gpu(device=0):
    """Restrict the calling process to the CPUs NVML reports as local to CUDA device `device`, so that pinned staging buffers
    allocated afterwards (first touch) and the copy-issuing thread sit on that GPU's NUMA node.  With one process per GPU and
    120 MB of positions each way per step and per GPU (bench.py e2e, the .ply frame path), host buffers on the wrong socket
    send every byte across the inter-socket link: 8 ranks then share ~100 GB/s instead of having a PCIe link each.
    Purely a placement hint: any failure (no NVML, container cpuset that excludes those CPUs, single NUMA node) leaves the
    process as it was.  Returns a one-line description of what was done."""
    try:
        import pynvml

        pynvml.nvmlInit()
        handle = None
        try:
            import torch

            uuid = str(torch.cuda.get_device_properties(device).uuid)
            handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        except Exception:
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v for v in vis.split(",") if v.strip()]
            idx = int(ids[device]) if ids and all(v.strip().isdigit() for v in ids) and device < len(ids) else device
            handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (ncpu + 63) // 64)
        local = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        target = local & allowed
        if not target:
            return "unchanged (the GPU's local CPUs are outside this process's cpuset)"
        if target == allowed:
            return f"unchanged (all {len(allowed)} allowed CPUs are local to the GPU)"
        if len(target) < 4:   # NCCL's proxy and torch's helper threads inherit the mask: never squeeze a rank onto a couple of cores
            return f"unchanged (only {len(target)} of the allowed CPUs are local to the GPU)"
        os.sched_setaffinity(0, target)
        return f"bound to the {len(target)} CPUs local to GPU {device} (of {len(allowed)} allowed)"
    except Exception as e:  # noqa: BLE001 - a hint must never take the run down
        return f"unchanged ({type(e).__name__})"


def create_distributed(X, Y, Z, **kw):
    """One `System` per torch.distributed rank (launched by torchrun, backend nccl or gloo for the id exchange):
    rank 0 creates the NCCL id, everyone receives it, every rank opens its own GPU (LOCAL_RANK)."""
    import os

    import torch.distributed as dist

    rank, world = dist.get_rank(), dist.get_world_size()
    box = [nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    kw.setdefault("device", int(os.environ.get("LOCAL_RANK", rank)))
    return System(X, Y, Z, rank=rank, world_size=world, nccl_id=box[0], **kw)


class System:
    """Drop-in for the reference `System`: one MAC water grid + ink particle drops, stepped on one B200."""

    def __init__(self, X=20, Y=45, Z=20, dtype=np.float32, params: Params | None = None, arith="strict", device=0,
                 cg_rtol=None, cg_max_iters=None, cg_check_every=None, cg_warm_start=None, cg_precond=None,
                 resort_interval=None, rank=0, world_size=1, nccl_id=None):
        self.L = _capi.load()
        self.dims = (int(X), int(Y), int(Z))
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
            raise ValueError("dtype must be float32 or float64")
        self.hd = 32 if self.dtype == np.float32 else 64
        self.params = params or Params()
        p = _capi.default_params()
        q = self.params
        p.X, p.Y, p.Z = self.dims
        p.cell_dim, p.density, p.k_vort, p.viscosity, p.p_atm = q.cell_dim, q.density, q.k_vort, q.viscosity, q.p_atm
        p.gravity[0], p.gravity[1], p.gravity[2] = q.gravity
        p.k_cfl, p.dt_min, p.dt_max, p.buffer_size = q.k_cfl, q.dt_min, q.dt_max, q.buffer_size
        p.precision = self.hd
        p.arith = {"strict": _capi.ARITH_STRICT, "fast": _capi.ARITH_FAST}[arith]
        p.device = device
        if cg_rtol is not None:
            p.cg_rtol = cg_rtol
        if cg_max_iters is not None:
            p.cg_max_iters = cg_max_iters
        if cg_check_every is not None:
            p.cg_check_every = cg_check_every
        if cg_warm_start is not None:
            p.cg_warm_start = int(cg_warm_start)
        if cg_precond is not None:
            p.cg_precond = {"none": _capi.PCG_NONE, "mg_const": _capi.PCG_MG, "mg": _capi.PCG_MG_TRI}[cg_precond]
        if resort_interval is not None:
            p.resort_interval = resort_interval
        # multi-GPU: one process per GPU; the grid is split in y-slabs, particles by index (include/ink_capi.h)
        self.rank, self.world_size = int(rank), int(world_size)
        p.rank, p.world_size = self.rank, self.world_size
        if self.world_size > 1:
            if nccl_id is None or len(nccl_id) != 128:
                raise ValueError("world_size > 1 needs the 128-byte nccl_id from nccl_unique_id() of rank 0")
            for i, b in enumerate(bytes(nccl_id)):
                p.nccl_id[i] = b
        self._p = p
        h = C.c_void_p()
        rc = self.L.ink_create(C.byref(p), C.byref(h))
        if rc != 0:
            raise InkError(f"ink_create failed ({rc}): {self.L.ink_last_error(None).decode()}")
        self.h = h

    # ---- lifetime
    def close(self):
        if getattr(self, "h", None):
            self.L.ink_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise InkError(self.L.ink_last_error(self.h).decode())

    def _a(self, x):
        return np.ascontiguousarray(x, dtype=self.dtype)

    # ---- reference API (src/system.h:75-81)
    def init(self):
        """System::init (src/system.cpp:10-26): the grid is created zeroed by the constructor; particle seeding is the
        caller's job here (set_particles) — the emitter/box seeders are host-side helpers in ink_simulation_b200.seeding."""
        return self

    def solve(self, timeToNextRender):
        dt = C.c_double()
        self._ck(self.L.ink_step(self.h, float(timeToNextRender), C.byref(dt)))
        return dt.value

    def getInkParticles(self):
        """list over drops of (positions (N,3), velocities (N,3)) in the caller's particle order."""
        return [self.get_particles(d) for d in range(self.num_drops())]

    def getWaterGrid(self):
        """currVelocity of every cell as three dense (X,Y,Z) arrays."""
        return self.get_field("curr")

    def check_finite(self):
        """(non-finite count, norm >= 10000 count) over grid velocities and particles: the tests of System::checkNanAndInf
        (src/systemHelpers.cpp:212-241) as one device reduction (ink_check_finite)."""
        out = (C.c_int64 * 2)()
        self._ck(self.L.ink_check_finite(self.h, out))
        return int(out[0]), int(out[1])

    def checkNanAndInf(self):
        bad, big = self.check_finite()
        if bad or big:
            raise InkError(f"checkNanAndInf: {bad} grid cells / particles hold a NaN or Inf, {big} velocities with norm() >= 10000")

    def get_phantom(self):
        have, v = C.c_int32(), (C.c_double * 3)()
        self._ck(self.L.ink_get_phantom(self.h, C.byref(have), v))
        return bool(have.value), tuple(v)

    def set_phantom(self, have, v):
        self._ck(self.L.ink_set_phantom(self.h, int(bool(have)), (C.c_double * 3)(*[float(x) for x in v])))

    def cuda_stream(self):
        """the ctx's own non-blocking cudaStream_t as an int (ink_get_stream); wrap it with torch.cuda.ExternalStream to record events"""
        s = C.c_void_p()
        self._ck(self.L.ink_get_stream(self.h, C.byref(s)))
        return s.value or 0

    # ---- state
    def set_particles(self, pos, vel=None, drop=0):
        if _is_dev_like(pos):  # torch tensor or __cuda_array_interface__ object: float32/float64, contiguous (N,3)
            shape, isz, contig = _dev_meta(pos)
            assert len(shape) == 2 and shape[1] == 3 and contig
            self._ck(self.L.ink_set_particles(self.h, drop, _ptr(pos), _ptr(vel), shape[0], 32 if isz == 4 else 64))
            return
        pos = self._a(pos).reshape(-1, 3)
        v = None if vel is None else self._a(vel).reshape(-1, 3)
        self._ck(self.L.ink_set_particles(self.h, drop, _ptr(pos), _ptr(v), len(pos), self.hd))

    def set_particles_shard(self, pos, vel=None, drop=0):
        """Multi-GPU: hand over only this rank's particles (same as set_particles on one GPU)."""
        if hasattr(pos, "data_ptr"):
            hd = 32 if pos.element_size() == 4 else 64
            self._ck(self.L.ink_set_particles_shard(self.h, drop, _ptr(pos), _ptr(vel), pos.shape[0], hd))
            return
        pos = self._a(pos).reshape(-1, 3)
        v = None if vel is None else self._a(vel).reshape(-1, 3)
        self._ck(self.L.ink_set_particles_shard(self.h, drop, _ptr(pos), _ptr(v), len(pos), self.hd))

    def num_drops(self):
        return int(self.L.ink_num_drops(self.h))

    def num_particles(self, drop=0):
        return int(self.L.ink_num_particles(self.h, drop))

    def get_particles(self, drop=0, out_pos=None, out_vel=None):
        if out_pos is not None or out_vel is not None:  # caller-provided buffers (numpy or torch, host or device)
            t = out_pos if out_pos is not None else out_vel
            hd = (32 if t.element_size() == 4 else 64) if hasattr(t, "data_ptr") else (32 if t.dtype == np.float32 else 64)
            self._ck(self.L.ink_get_particles(self.h, drop, _ptr(out_pos), _ptr(out_vel), hd))
            return out_pos, out_vel
        n = self.num_particles(drop)
        pos = np.empty((n, 3), self.dtype)
        vel = np.empty((n, 3), self.dtype)
        self._ck(self.L.ink_get_particles(self.h, drop, _ptr(pos), _ptr(vel), self.hd))
        return pos, vel

    def set_velocity(self, u, v, w):
        if _is_dev_like(u):   # torch tensors (host or DEVICE) or __cuda_array_interface__ objects, contiguous, (X, Y, Z): pointers go straight to the C ABI
            metas = [_dev_meta(a) for a in (u, v, w)]
            assert all(m[0] == self.dims and m[2] for m in metas) and len({m[1] for m in metas}) == 1
            self._ck(self.L.ink_set_velocity(self.h, _ptr(u), _ptr(v), _ptr(w), 32 if metas[0][1] == 4 else 64))
            return
        u, v, w = self._a(u), self._a(v), self._a(w)
        assert u.shape == self.dims and v.shape == self.dims and w.shape == self.dims
        self._ck(self.L.ink_set_velocity(self.h, _ptr(u), _ptr(v), _ptr(w), self.hd))

    def get_field(self, which="curr"):
        idx = {"old": 0, "curr": 1, "curl": 2}[which]
        out = [np.empty(self.dims, self.dtype) for _ in range(3)]
        self._ck(self.L.ink_get_field(self.h, idx, *[_ptr(a) for a in out], self.hd))
        return out

    def get_velocity(self):
        return self.get_field("curr")

    def get_pressure(self):
        p = np.empty(self.dims, self.dtype)
        self._ck(self.L.ink_get_pressure(self.h, _ptr(p), self.hd))
        return p

    def set_pressure(self, p):
        p = self._a(p)
        assert p.shape == self.dims
        self._ck(self.L.ink_set_pressure(self.h, _ptr(p), self.hd))

    def get_warm_start(self):
        """q = p - p_atm exactly as the Poisson solver holds it (ink_get_warm_start; checkpoints)."""
        q = np.empty(self.dims, self.dtype)
        self._ck(self.L.ink_get_warm_start(self.h, _ptr(q), self.hd))
        return q

    def set_warm_start(self, q):
        q = self._a(q)
        assert q.shape == self.dims
        self._ck(self.L.ink_set_warm_start(self.h, _ptr(q), self.hd))

    def get_mark(self):
        m = np.empty(self.dims, np.uint8)
        self._ck(self.L.ink_get_mark(self.h, _ptr(m)))
        return m

    # ---- phases (each followed by the reference's old = curr commit)
    def calc_dt(self, ttnr):
        dt = C.c_double()
        self._ck(self.L.ink_calc_dt(self.h, float(ttnr), C.byref(dt)))
        return dt.value

    def advect(self, dt):
        self._ck(self.L.ink_advect(self.h, float(dt)))

    def forces(self, dt, return_mark=False):
        self._ck(self.L.ink_forces(self.h, float(dt)))
        return self.get_mark() if return_mark else None

    def viscosity(self, dt):
        self._ck(self.L.ink_viscosity(self.h, float(dt)))

    def pressure_rhs(self, dt):
        b = np.empty(self.dims, self.dtype)
        self._ck(self.L.ink_pressure_rhs(self.h, float(dt), _ptr(b), self.hd))
        return b

    def pressure(self, dt):
        self._ck(self.L.ink_pressure(self.h, float(dt)))

    def pressure_given_p(self, dt, p):
        self.set_pressure(p)
        self._ck(self.L.ink_pressure_given_p(self.h, float(dt)))

    def poisson_solve(self, b):
        b = self._a(b)
        p = np.empty(self.dims, self.dtype)
        self._ck(self.L.ink_poisson_solve(self.h, _ptr(b), _ptr(p), self.hd))
        return p, self.stats().cg_rel_residual

    def vorticity(self, dt):
        self._ck(self.L.ink_vorticity(self.h, float(dt)))

    def divergence(self):
        d = np.empty(self.dims, self.dtype)
        self._ck(self.L.ink_get_divergence(self.h, _ptr(d), self.hd))
        return d

    def sample_velocity(self, pos):
        pos = self._a(pos).reshape(-1, 3)
        out = np.empty_like(pos)
        self._ck(self.L.ink_sample_velocity(self.h, _ptr(pos), len(pos), _ptr(out), self.hd))
        return out

    def velocity_field_update(self, dt):
        self._ck(self.L.ink_velocity_field_update(self.h, float(dt)))

    def particles(self, dt):
        self._ck(self.L.ink_particles(self.h, float(dt)))

    # ---- streaming particle I/O (include/ink_capi.h): copies overlap the step in between ------------------------
    @staticmethod
    def _hd_of(t):
        return (32 if t.element_size() == 4 else 64) if hasattr(t, "data_ptr") else (32 if t.dtype == np.float32 else 64)

    def prefetch_positions(self, pos, drop=0):
        """start the H2D copy of the next positions of `drop` (pinned torch tensor or numpy array, (N,3)); returns at once"""
        self._ck(self.L.ink_prefetch_positions(self.h, drop, _ptr(pos), pos.shape[0], self._hd_of(pos)))

    def commit_positions(self, drop=0):
        self._ck(self.L.ink_commit_positions(self.h, drop))

    def fetch_positions_async(self, out_pos, drop=0):
        """export + D2H of the positions of `drop` into out_pos on the output copy stream; complete after sync_io()"""
        self._ck(self.L.ink_fetch_positions_async(self.h, drop, _ptr(out_pos), self._hd_of(out_pos)))

    def sync_io(self):
        self._ck(self.L.ink_sync_io(self.h))

    def resort_particles(self):
        self._ck(self.L.ink_resort_particles(self.h))

    def sync(self):
        self._ck(self.L.ink_sync(self.h))

    def stats(self) -> _capi.InkStats:
        s = _capi.InkStats()
        self._ck(self.L.ink_get_stats(self.h, C.byref(s)))
        return s

    def phase_ms(self):
        out = (C.c_float * 8)()
        self._ck(self.L.ink_get_phase_ms(self.h, out))
        return dict(zip(("dt", "advect", "forces", "viscosity", "pressure", "vorticity", "particles", "resort"), list(out)))
