"""The C++ host layer (ink_simulation_b200/host): drop-in `System` / `InkSim` over the C ABI.

-m "not gpu": particle seeding and the .ply frame writer, checked against fixtures generated from the COMPILED REFERENCE
              (tests/golden/make_host_golden.py) and, where oracle/_ref exists, against the reference run live.
-m gpu      : the Qt-free `main` (inksim_b200_c1: 8x25x8, 10000 particles of the initParticles box) run end to end;
              its frames must equal, byte for byte, what the Python host produces through the same C ABI from the same
              glibc rand() seeding, and its headers must be the reference's."""
import ctypes as C
import hashlib
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "ink_simulation_b200", "host")
GOLD = os.path.join(ROOT, "tests", "golden")


def host_lib():
    path = os.path.join(HOST, "libinkhost.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "ink_simulation_b200", "csrc")])
        subprocess.check_call(["make", "-C", HOST])
    L = C.CDLL(path)
    L.inkhost_emitter_scene.restype = C.c_long
    return L


def emitter(L, dims):
    cap = 400000
    pos = np.empty((cap, 3), np.float32)
    vel = np.empty((cap, 3), np.float32)
    n = L.inkhost_emitter_scene(C.c_int(dims[0]), C.c_int(dims[1]), C.c_int(dims[2]), pos.ctypes.data_as(C.c_void_p),
                                vel.ctypes.data_as(C.c_void_p), C.c_long(cap))
    return pos[:n].copy(), vel[:n].copy()


def test_emitter_scene_matches_reference_golden():
    G = np.load(os.path.join(GOLD, "host_emitter.npz"))
    pos, vel = emitter(host_lib(), tuple(int(x) for x in G["dims"]))
    assert len(pos) == int(G["n"]) == 324120          # 120 radii x 73 x 37 float-loop iterations
    assert np.array_equal(pos[::997], G["pos_sample"]) and np.array_equal(vel[::997], G["vel_sample"])
    assert hashlib.sha256(pos.tobytes()).hexdigest() == str(G["pos_sha"])
    assert hashlib.sha256(vel.tobytes()).hexdigest() == str(G["vel_sha"])


LIVE = """
import sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
import oracle
dims = (6, 11, 6)   # X == Z: the reference pressure index is only a bijection then (SURVEY Q1)
R = oracle.RefSystem(*dims); R.init_emitter(); rp, rv = R.get_particles()
import test_host_cpp as t
pos, vel = t.emitter(t.host_lib(), dims)
assert pos.tobytes() == rp.tobytes() and vel.tobytes() == rv.tobytes()
print("LIVE-OK")
"""


@pytest.mark.skipif(not __import__("oracle").have_ref(), reason="compiled reference not built here")
def test_emitter_scene_matches_live_reference():
    """Own process: the reference library and libinkhost.so both define `System`/`InkSim` (that is the point of a drop-in),
    so the reference must be loaded first and never share a process with other tests' handles."""
    out = subprocess.run([os.sys.executable, "-c", LIVE % (ROOT, os.path.join(ROOT, "tests"))], capture_output=True, text=True)
    assert "LIVE-OK" in out.stdout, out.stdout + out.stderr


def test_frame_writer_bytes_equal_reference(tmp_path):
    drop = np.load(os.path.join(GOLD, "host_ply_drop.npy"))
    want = open(os.path.join(GOLD, "host_ply.bin"), "rb").read()
    out = str(tmp_path / "f.ply")
    host_lib().inkhost_write_particles_ply(out.encode(), drop.ctypes.data_as(C.c_void_p), C.c_long(len(drop)))
    assert open(out, "rb").read() == want
    assert want.startswith(b"ply\nformat binary_small_endian 1.0\nelement vertex 257\nproperty float x\n")


def read_ply(path, props):
    data = open(path, "rb").read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    header = data[:end].decode()
    n = int([ln for ln in header.splitlines() if ln.startswith("element vertex")][0].split()[-1])
    return header, np.frombuffer(data[end:], np.float32).reshape(n, props)


def glibc_box_particles(dims, n):
    """initParticles (reference src/system.cpp:62-115) with glibc's unseeded rand() (== srand(1))."""
    libc = C.CDLL("libc.so.6")
    libc.srand(1)
    X, Y, Z = dims
    f = np.float32
    lo = [float(f(X) / f(2) - f(X) / f(5)), Y - 1.5, float(f(Z) / f(2) - f(Z) / f(3))]
    hi = [float(f(X) / f(2) + f(X) / f(5)), Y - 1.0, float(f(Z) / f(2) + f(Z) / f(3))]
    out = np.empty((n, 3), np.float32)
    RAND_MAX = 2147483647
    for p in range(n):
        for a in range(3):
            out[p, a] = lo[a] + (libc.rand() / RAND_MAX) * (hi[a] - lo[a])
    return out


def test_rand_stream_is_the_unseeded_rand_sequence_and_leaves_the_global_generator_alone():
    """inkhost_rand_stream used to call srand(1) on the process-global generator.  It now draws from a private
    random_r state: same numbers as a fresh process's rand() (what the reference's initParticles consumes), and a
    caller's own rand() sequence is not disturbed by it."""
    import ctypes

    from ink_simulation_b200 import seeding

    libc = ctypes.CDLL(None)
    libc.srand(1)
    want = np.array([libc.rand() for _ in range(1000)], dtype=np.int32)
    assert np.array_equal(seeding.rand_stream(1000), want)
    libc.srand(7)
    a = [libc.rand() for _ in range(5)]
    libc.srand(7)
    first = libc.rand()
    seeding.rand_stream(100)
    assert [first] + [libc.rand() for _ in range(4)] == a


@pytest.mark.gpu
def test_inksim_binary_end_to_end(tmp_path):
    from ink_simulation_b200 import System

    exe = os.path.join(HOST, "inksim_b200_c1")
    assert os.path.exists(exe), "build with __graft_entry__.build()"
    steps, render = 60, 0.05
    out = str(tmp_path / "output")
    env = dict(os.environ, INK_ARITH="strict")
    subprocess.check_call([exe, out, str(render), str(steps)], env=env, stdout=subprocess.DEVNULL)
    dims = (8, 25, 8)
    S = System(*dims, arith="strict")
    S.set_particles(glibc_box_particles(dims, 10000))
    frame, since = 1, np.float32(0)
    for _ in range(steps):
        dt = np.float32(S.solve(float(np.float32(render) - since)))   # float timeStep / timeSinceLastRender, src/inkSim.cpp:21-24
        since = np.float32(since + dt)
        if since >= np.float32(render):
            header, pos = read_ply(os.path.join(out, "drop0", f"{frame}.ply"), 3)
            assert header == ("ply\nformat binary_small_endian 1.0\nelement vertex 10000\nproperty float x\nproperty float y\n"
                              "property float z\nend_header\n")
            assert np.array_equal(pos, S.get_particles()[0]), frame
            frame += 1
            since = np.float32(0)
    assert frame > 3
    assert not os.path.exists(os.path.join(out, "drop0", f"{frame}.ply"))
    header, rec = read_ply(os.path.join(out, "velocityField.ply"), 6)
    assert "element vertex 1600\n" in header and header.count("property float") == 6
    u, v, w = S.get_velocity()
    assert np.array_equal(rec[:, 3:], np.stack([u.ravel(), v.ravel(), w.ravel()], axis=1))
    i, j, k = np.meshgrid(np.arange(8), np.arange(25), np.arange(8), indexing="ij")
    assert np.array_equal(rec[:, :3], np.stack([i.ravel(), j.ravel(), k.ravel()], axis=1).astype(np.float32) + 0.5)


@pytest.mark.gpu
def test_gpu_device_seeding_equals_the_reference_seeders():
    """the emitter scene and the initParticles box produced by KERNELS (ink_seed_emitter / ink_seed_box; the host only
    supplies the float loop counters with their sinf / cosf, or the rand() stream) equal the compiled reference's arrays bit for
    bit — the emitter against the committed golden hashes, the box against the host seeder that the CPU tests pin."""
    from ink_simulation_b200 import System, seeding

    G = np.load(os.path.join(GOLD, "host_emitter.npz"))
    dims = tuple(int(x) for x in G["dims"])
    S = System(*dims)
    assert seeding.seed_emitter_on_device(S) == int(G["n"]) == 324120
    pos, vel = S.get_particles()
    assert np.array_equal(pos[::997], G["pos_sample"]) and np.array_equal(vel[::997], G["vel_sample"])
    assert hashlib.sha256(pos.tobytes()).hexdigest() == str(G["pos_sha"])
    assert hashlib.sha256(vel.tobytes()).hexdigest() == str(G["vel_sha"])
    S.close()
    for dims, n in (((8, 25, 8), 10000), ((20, 45, 20), 100000), ((7, 9, 5), 333)):
        S = System(*dims)
        seeding.seed_box_on_device(S, n)
        pos, vel = S.get_particles()
        assert np.array_equal(pos, seeding.box_scene(dims, n)) and not vel.any()
        # a second drop continues the same rand() stream, as initParticles does for NUM_PARTICLES = {n, m}
        stream = seeding.rand_stream(3 * (n + 50))
        seeding.seed_box_on_device(S, 50, drop=1, stream=stream[3 * n:])
        assert np.array_equal(S.get_particles(1)[0], seeding.box_scene(dims, n + 50)[n:])
        S.close()
