"""TEST INFRASTRUCTURE — ctypes front-ends for the CPU oracle and the compiled reference.

* ``OracleSystem``  — oracle/ink_oracle.cpp (flat-array restatement, float or double).
* ``RefSystem``     — oracle/_ref/libinkref.so, the reference's own sources compiled by
  oracle/build_ref.sh (only available where that library was built / travelled to).

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs
may import this package.  The product (``ink_simulation_b200``) never does.

Array convention everywhere: velocity components are dense ``(X, Y, Z)`` C-order arrays (z fastest);
particles are ``(N, 3)`` xyz arrays.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "libinkoracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libinkref.so")


@dataclass
class RefParams:
    """The global constants of the reference's src/system.h:13-38 (defaults = committed values)."""

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


def build(force: bool = False) -> None:
    """Compile the oracle (and the reference when /root/reference exists).  Building the checker is not using it."""
    src = os.path.join(_HERE, "ink_oracle.cpp")
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libinkoracle.so"], stdout=subprocess.DEVNULL)
    subprocess.check_call([os.path.join(_HERE, "build_ref.sh")], stdout=subprocess.DEVNULL)


def set_threads(n=None) -> int:
    """OpenMP threads of the oracle port: n, or every CPU this process may run on (sched_getaffinity) — NOT what
    OMP_NUM_THREADS says (torchrun exports 1).  Returns the team size omp_get_max_threads() then reports."""
    L = _load_oracle()
    L.orc_set_threads(C.c_int(int(n) if n else len(os.sched_getaffinity(0))))
    return int(L.orc_max_threads())


def have_ref() -> bool:
    return os.path.exists(REF_SO)


_orc = None
_ref = None


def _load_oracle():
    global _orc
    if _orc is None:
        if not os.path.exists(ORACLE_SO):
            build()
        L = C.CDLL(ORACLE_SO)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int] * 4 + [C.c_float] * 11 + [C.c_int]
        for name in ("orc_calc_dt", "orc_solve", "orc_poisson_solve", "orc_solve_timed"):
            getattr(L, name).restype = C.c_double
        L.orc_num_particles.restype = C.c_long
        _orc = L
    return _orc


def _load_ref():
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(f"{REF_SO} not built (run oracle/build_ref.sh where /root/reference exists)")
        L = C.CDLL(REF_SO)
        L.ref_create.restype = C.c_void_p
        L.ref_create.argtypes = [C.c_int] * 3
        L.ref_num_particles.restype = C.c_long
        L.ref_grid_entries.restype = C.c_long
        L.ref_max_speed_all_entries.restype = C.c_float
        for name in ("ref_calc_dt", "ref_solve", "ref_solve_timed"):
            getattr(L, name).restype = C.c_double
        _ref = L
    return _ref


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleSystem:
    """CPU oracle; ``dtype`` np.float32 (bit-compatible with the reference) or np.float64."""

    def __init__(self, X, Y, Z, dtype=np.float32, params: RefParams | None = None):
        self.L = _load_oracle()
        self.dims = (X, Y, Z)
        self.dtype = np.dtype(dtype)
        self.params = params or RefParams()
        q = self.params
        self.h = C.c_void_p(
            self.L.orc_create(
                32 if self.dtype == np.float32 else 64, X, Y, Z, q.cell_dim, q.density, q.k_vort, q.viscosity, q.p_atm,
                q.gravity[0], q.gravity[1], q.gravity[2], q.k_cfl, q.dt_min, q.dt_max, q.buffer_size,
            )
        )

    def close(self):
        if self.h:
            self.L.orc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _a(self, x):
        return np.ascontiguousarray(x, dtype=self.dtype)

    # ---- state
    def set_particles(self, pos, vel=None, drop=0):
        pos = self._a(pos).reshape(-1, 3)
        v = None if vel is None else self._a(vel).reshape(-1, 3)
        self.L.orc_set_particles(self.h, C.c_int(drop), _p(pos), None if v is None else _p(v), C.c_long(len(pos)))

    def num_particles(self, drop=0):
        return int(self.L.orc_num_particles(self.h, C.c_int(drop)))

    def get_particles(self, drop=0):
        n = self.num_particles(drop)
        pos = np.empty((n, 3), self.dtype)
        vel = np.empty((n, 3), self.dtype)
        self.L.orc_get_particles(self.h, C.c_int(drop), _p(pos), _p(vel))
        return pos, vel

    def set_velocity(self, u, v, w):
        u, v, w = self._a(u), self._a(v), self._a(w)
        assert u.shape == self.dims
        self.L.orc_set_velocity(self.h, _p(u), _p(v), _p(w))

    def get_field(self, which="curr"):
        idx = {"old": 0, "curr": 1, "curl": 2}[which]
        out = [np.empty(self.dims, self.dtype) for _ in range(3)]
        self.L.orc_get_field(self.h, C.c_int(idx), *[_p(a) for a in out])
        return out

    def get_velocity(self):
        return self.get_field("curr")

    def get_pressure(self):
        p = np.empty(self.dims, self.dtype)
        self.L.orc_get_pressure(self.h, _p(p))
        return p

    # ---- phases (each followed by the reference's old=curr sweep)
    def calc_dt(self, ttnr):
        return float(self.L.orc_calc_dt(self.h, C.c_double(ttnr)))

    def advect(self, dt):
        self.L.orc_advect(self.h, C.c_double(dt))

    def forces(self, dt, return_mark=False):
        mark = np.zeros(self.dims, np.uint8) if return_mark else None
        self.L.orc_forces(self.h, C.c_double(dt), None if mark is None else _p(mark))
        return mark

    def viscosity(self, dt):
        self.L.orc_viscosity(self.h, C.c_double(dt))

    def pressure_rhs(self, dt):
        b = np.empty(self.dims, self.dtype)
        self.L.orc_pressure_rhs(self.h, C.c_double(dt), _p(b))
        return b

    def pressure(self, dt):
        self.L.orc_pressure(self.h, C.c_double(dt))

    def pressure_given_p(self, dt, p):
        p = self._a(p)
        self.L.orc_pressure_given_p(self.h, C.c_double(dt), _p(p))

    def poisson_solve(self, b):
        b = self._a(b)
        p = np.empty(self.dims, self.dtype)
        res = float(self.L.orc_poisson_solve(self.h, _p(b), _p(p)))
        return p, res

    def vorticity(self, dt):
        self.L.orc_vorticity(self.h, C.c_double(dt))

    def divergence(self):
        d = np.empty(self.dims, self.dtype)
        self.L.orc_divergence(self.h, _p(d))
        return d

    def sample_velocity(self, pos):
        pos = self._a(pos).reshape(-1, 3)
        out = np.empty_like(pos)
        self.L.orc_sample_velocity(self.h, _p(pos), C.c_long(len(pos)), _p(out))
        return out

    def velocity_field_update(self, dt):
        self.L.orc_velocity_field_update(self.h, C.c_double(dt))

    def particles(self, dt):
        self.L.orc_particles(self.h, C.c_double(dt))

    def solve(self, ttnr):
        return float(self.L.orc_solve(self.h, C.c_double(ttnr)))

    def solve_timed(self, ttnr):
        t = (C.c_double * 7)()
        dt = float(self.L.orc_solve_timed(self.h, C.c_double(ttnr), t))
        return dt, list(t)

    def num_phantom(self):
        return int(self.L.orc_num_phantom(self.h))


class RefSystem:
    """The compiled reference (fp32, constants of src/system.h as committed; grid size run-time)."""

    dtype = np.dtype(np.float32)

    def __init__(self, X, Y, Z):
        self.L = _load_ref()
        self.dims = (X, Y, Z)
        self.h = C.c_void_p(self.L.ref_create(X, Y, Z))

    def close(self):
        if self.h:
            self.L.ref_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _a(x):
        return np.ascontiguousarray(x, dtype=np.float32)

    def set_particles(self, pos, vel=None, drop=0):
        pos = self._a(pos).reshape(-1, 3)
        v = None if vel is None else self._a(vel).reshape(-1, 3)
        self.L.ref_set_particles(self.h, C.c_int(drop), _p(pos), None if v is None else _p(v), C.c_long(len(pos)))

    def num_drops(self):
        return int(self.L.ref_num_drops(self.h))

    def num_particles(self, drop=0):
        return int(self.L.ref_num_particles(self.h, C.c_int(drop)))

    def get_particles(self, drop=0):
        n = self.num_particles(drop)
        pos = np.empty((n, 3), np.float32)
        vel = np.empty((n, 3), np.float32)
        self.L.ref_get_particles(self.h, C.c_int(drop), _p(pos), _p(vel))
        return pos, vel

    def init_particles_box(self, seed=1):
        self.L.ref_init_particles_box(self.h, C.c_uint(seed))

    def init_emitter(self):
        self.L.ref_init_emitter(self.h)

    def set_velocity(self, u, v, w):
        u, v, w = self._a(u), self._a(v), self._a(w)
        assert u.shape == self.dims
        self.L.ref_set_velocity(self.h, _p(u), _p(v), _p(w))

    def get_field(self, which="curr"):
        idx = {"old": 0, "curr": 1, "curl": 2}[which]
        out = [np.empty(self.dims, np.float32) for _ in range(3)]
        self.L.ref_get_field(self.h, C.c_int(idx), *[_p(a) for a in out])
        return out

    def get_velocity(self):
        return self.get_field("curr")

    def grid_entries(self):
        return int(self.L.ref_grid_entries(self.h))

    def max_speed_all_entries(self):
        return float(self.L.ref_max_speed_all_entries(self.h))

    def calc_dt(self, ttnr):
        return float(self.L.ref_calc_dt(self.h, C.c_double(ttnr)))

    def advect(self, dt):
        self.L.ref_advect(self.h, C.c_float(dt))

    def forces(self, dt):
        self.L.ref_forces(self.h, C.c_float(dt))

    def viscosity(self, dt):
        self.L.ref_viscosity(self.h, C.c_float(dt))

    def calc_pressure(self, dt):
        p = np.empty(self.dims, np.float32)
        self.L.ref_calc_pressure(self.h, C.c_float(dt), _p(p))
        return p

    def pressure(self, dt):
        self.L.ref_pressure(self.h, C.c_float(dt))

    def vorticity(self, dt):
        self.L.ref_vorticity(self.h, C.c_float(dt))

    def divergence(self):
        d = np.empty(self.dims, np.float32)
        self.L.ref_divergence(self.h, _p(d))
        return d

    def sample_velocity(self, pos):
        pos = self._a(pos).reshape(-1, 3)
        out = np.empty_like(pos)
        self.L.ref_sample_velocity(self.h, _p(pos), C.c_long(len(pos)), _p(out))
        return out

    def velocity_field_update(self, dt):
        self.L.ref_velocity_field_update(self.h, C.c_float(dt))

    def particles(self, dt):
        self.L.ref_particles(self.h, C.c_float(dt))

    def solve(self, ttnr):
        return float(self.L.ref_solve(self.h, C.c_double(ttnr)))

    def solve_timed(self, ttnr):
        t = (C.c_double * 7)()
        dt = float(self.L.ref_solve_timed(self.h, C.c_double(ttnr), t))
        return dt, list(t)


def write_ply_with_reference(directory: str, frame: int, drops) -> None:
    """InkSim::writeToFile (src/inkSim.cpp:43-96) run by the compiled reference; ``drops`` = list of (N,3) arrays.
    Note: grid dims must have been set (any RefSystem created) and should be small (the InkSim ctor runs init())."""
    L = _load_ref()
    arrs = [np.ascontiguousarray(d, np.float32).reshape(-1, 3) for d in drops]
    for i in range(len(arrs)):
        os.makedirs(os.path.join(directory, f"drop{i}"), exist_ok=True)
    ptrs = (C.POINTER(C.c_float) * len(arrs))(*[a.ctypes.data_as(C.POINTER(C.c_float)) for a in arrs])
    ns = (C.c_long * len(arrs))(*[len(a) for a in arrs])
    L.ref_write_ply(directory.encode(), C.c_int(frame), C.c_int(len(arrs)), ptrs, ns)
