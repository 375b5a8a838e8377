"""seeding.py — the reference's particle seeders (System::init, src/system.cpp:10-26) for the Python host.

The arithmetic lives once, in C++ (host/seeding.cpp, checked against the compiled reference: tests/test_host_cpp.py), and is
reached here through libinkhost.so's C hooks; nothing in this file computes positions itself.
    emitter_scene(dims)      the committed scene: 120 hemispheres x 73 x 37 float-loop iterations = 324 120 particles
                             (emitParticleHemisphere, src/inkSolver.cpp:11-42)                       -> (pos, vel)
    box_scene(dims, n)       initParticles (src/system.cpp:93-115) with glibc's unseeded rand()     -> pos
    from_file(path)          initFromFile (src/system.cpp:117-148): "x y z vx vy vz" per line       -> (pos, vel)
"""
import ctypes as C
import os

import numpy as np

_HOST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host", "libinkhost.so")
_lib = None


def host_lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_HOST):
            raise OSError(f"{_HOST} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`")
        L = C.CDLL(_HOST)
        for name in ("inkhost_emitter_scene", "inkhost_box_scene", "inkhost_from_file", "inkhost_read_ply"):
            getattr(L, name).restype = C.c_long
        _lib = L
    return _lib


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def emitter_scene(dims):
    cap = 400000
    pos, vel = np.empty((cap, 3), np.float32), np.empty((cap, 3), np.float32)
    n = host_lib().inkhost_emitter_scene(C.c_int(dims[0]), C.c_int(dims[1]), C.c_int(dims[2]), _vp(pos), _vp(vel), C.c_long(cap))
    return pos[:n].copy(), vel[:n].copy()


def box_scene(dims, n):
    pos = np.empty((n, 3), np.float32)
    m = host_lib().inkhost_box_scene(C.c_int(dims[0]), C.c_int(dims[1]), C.c_int(dims[2]), C.c_int(n), _vp(pos), C.c_long(n))
    assert m == n
    return pos


def from_file(path):
    L = host_lib()
    n = L.inkhost_from_file(str(path).encode(), None, None, C.c_long(0))
    if n < 0:
        raise OSError(f"PART_FILE cannot be read: {path}")
    pos, vel = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32)
    L.inkhost_from_file(str(path).encode(), _vp(pos), _vp(vel), C.c_long(n))
    return pos, vel


# ---- device seeding (ink_seed_emitter / ink_seed_box): the host supplies only what is sequential or libm-dependent ----------
def emitter_tables():
    """(radius, sin_i, cos_i, sin_j, cos_j): the float loop counters of the emitter scene and glibc's sinf / cosf of them"""
    cap = 256
    arrs = [np.empty(cap, np.float32) for _ in range(5)]
    nr, ni, nj = C.c_int(), C.c_int(), C.c_int()
    host_lib().inkhost_emitter_tables(_vp(arrs[0]), C.byref(nr), _vp(arrs[1]), _vp(arrs[2]), C.byref(ni), _vp(arrs[3]), _vp(arrs[4]), C.byref(nj),
                                      C.c_int(cap))
    assert max(nr.value, ni.value, nj.value) <= cap
    return (arrs[0][:nr.value].copy(), arrs[1][:ni.value].copy(), arrs[2][:ni.value].copy(), arrs[3][:nj.value].copy(), arrs[4][:nj.value].copy())


def rand_stream(n):
    """the first n draws of glibc's unseeded rand() (what initParticles consumes: x, y, z per particle)"""
    out = np.empty(n, np.int32)
    host_lib().inkhost_rand_stream(_vp(out), C.c_long(n))
    return out


def seed_emitter_on_device(system, drop=0):
    """System::init's emitter scene produced by a kernel (ink_seed_emitter); returns the particle count"""
    r, si, ci, sj, cj = emitter_tables()
    system._ck(system.L.ink_seed_emitter(system.h, drop, _vp(r), len(r), _vp(si), _vp(ci), len(si), _vp(sj), _vp(cj), len(sj)))
    return len(r) * len(si) * len(sj)


def seed_box_on_device(system, n, drop=0, stream=None):
    """initParticles produced by a kernel (ink_seed_box) from the rand() stream"""
    rnd = rand_stream(3 * n) if stream is None else np.ascontiguousarray(stream, np.int32)
    system._ck(system.L.ink_seed_box(system.h, drop, _vp(rnd), n))
    return n
