"""ctypes binding of include/ink_capi.h (libinkb200.so).  No torch types cross this boundary; there is no
CPU fallback — if the CUDA library is missing or no B200 is present, creation fails loudly."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libinkb200.so")

ARITH_STRICT, ARITH_FAST = 0, 1
FIELD_OLD, FIELD_CURR, FIELD_CURL = 0, 1, 2
PCG_NONE, PCG_MG, PCG_MG_TRI = 0, 1, 2
NOALLOC_MAX_N = 1 << 28   # INK_NOALLOC_MAX_N


class InkParams(C.Structure):
    """struct ink_params (include/ink_capi.h) — run-time copy of the globals in reference src/system.h:13-38."""

    _fields_ = [
        ("X", C.c_int32), ("Y", C.c_int32), ("Z", C.c_int32),
        ("cell_dim", C.c_float), ("density", C.c_float), ("k_vort", C.c_float), ("viscosity", C.c_float),
        ("p_atm", C.c_float), ("gravity", C.c_float * 3), ("k_cfl", C.c_float), ("dt_min", C.c_float),
        ("dt_max", C.c_float), ("buffer_size", C.c_int32), ("precision", C.c_int32), ("arith", C.c_int32),
        ("device", C.c_int32), ("cg_rtol", C.c_double), ("cg_max_iters", C.c_int32), ("cg_check_every", C.c_int32),
        ("cg_warm_start", C.c_int32), ("cg_precond", C.c_int32), ("resort_interval", C.c_int32),
        ("rank", C.c_int32), ("world_size", C.c_int32), ("nccl_id", C.c_uint8 * 128),
    ]


class InkStats(C.Structure):
    _fields_ = [
        ("dt", C.c_double), ("cg_iters", C.c_int32), ("cg_rel_residual", C.c_double), ("advect_skipped", C.c_int32),
        ("steps", C.c_int64), ("resorts", C.c_int64), ("kernel_launches", C.c_int64),
        ("cg_converged", C.c_int32), ("graph_launches", C.c_int32),
    ]


# every symbol include/ink_capi.h declares: name -> (restype, argtypes)
_VP, _I, _D, _L = C.c_void_p, C.c_int, C.c_double, C.c_int64
SYMBOLS = {
    "ink_default_params": (None, [C.POINTER(InkParams)]),
    "ink_nccl_unique_id": (_I, [C.POINTER(C.c_uint8)]),
    "ink_create": (_I, [C.POINTER(InkParams), C.POINTER(_VP)]),
    "ink_destroy": (None, [_VP]),
    "ink_last_error": (C.c_char_p, [_VP]),
    "ink_sync": (_I, [_VP]),
    "ink_set_particles": (_I, [_VP, _I, _VP, _VP, _L, _I]),
    "ink_set_particles_shard": (_I, [_VP, _I, _VP, _VP, _L, _I]),
    "ink_seed_emitter": (_I, [_VP, _I, _VP, _I, _VP, _VP, _I, _VP, _VP, _I]),
    "ink_seed_box": (_I, [_VP, _I, _VP, _L]),
    "ink_num_drops": (_I, [_VP]),
    "ink_num_particles": (_L, [_VP, _I]),
    "ink_get_particles": (_I, [_VP, _I, _VP, _VP, _I]),
    "ink_set_velocity": (_I, [_VP, _VP, _VP, _VP, _I]),
    "ink_get_field": (_I, [_VP, _I, _VP, _VP, _VP, _I]),
    "ink_get_pressure": (_I, [_VP, _VP, _I]),
    "ink_set_pressure": (_I, [_VP, _VP, _I]),
    "ink_get_warm_start": (_I, [_VP, _VP, _I]),
    "ink_set_warm_start": (_I, [_VP, _VP, _I]),
    "ink_get_mark": (_I, [_VP, _VP]),
    "ink_get_divergence": (_I, [_VP, _VP, _I]),
    "ink_sample_velocity": (_I, [_VP, _VP, _L, _VP, _I]),
    "ink_step": (_I, [_VP, _D, C.POINTER(_D)]),
    "ink_calc_dt": (_I, [_VP, _D, C.POINTER(_D)]),
    "ink_advect": (_I, [_VP, _D]),
    "ink_forces": (_I, [_VP, _D]),
    "ink_viscosity": (_I, [_VP, _D]),
    "ink_pressure_rhs": (_I, [_VP, _D, _VP, _I]),
    "ink_pressure": (_I, [_VP, _D]),
    "ink_pressure_given_p": (_I, [_VP, _D]),
    "ink_poisson_solve": (_I, [_VP, _VP, _VP, _I]),
    "ink_vorticity": (_I, [_VP, _D]),
    "ink_velocity_field_update": (_I, [_VP, _D]),
    "ink_particles": (_I, [_VP, _D]),
    "ink_resort_particles": (_I, [_VP]),
    "ink_prefetch_positions": (_I, [_VP, _I, _VP, _L, _I]),
    "ink_commit_positions": (_I, [_VP, _I]),
    "ink_fetch_positions_async": (_I, [_VP, _I, _VP, _I]),
    "ink_sync_io": (_I, [_VP]),
    "ink_check_finite": (_I, [_VP, C.POINTER(C.c_int64)]),
    "ink_get_phantom": (_I, [_VP, C.POINTER(C.c_int32), C.POINTER(C.c_double)]),
    "ink_set_phantom": (_I, [_VP, C.c_int32, C.POINTER(C.c_double)]),
    "ink_get_stream": (_I, [_VP, C.POINTER(_VP)]),
    "ink_get_stats": (_I, [_VP, C.POINTER(InkStats)]),
    "ink_get_phase_ms": (_I, [_VP, C.POINTER(C.c_float)]),
    "ink_check_finite_noalloc": (_I, [_VP, _VP, _VP, _L, _L, _I, _I, _VP, _VP]),
}

_lib = None


def load():
    """dlopen libinkb200.so and type every entry point.  Raises if the extension was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C ink_simulation_b200/csrc). There is no CPU fallback."
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def default_params() -> InkParams:
    p = InkParams()
    load().ink_default_params(C.byref(p))
    return p
