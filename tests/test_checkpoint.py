"""Checkpoint / resume and the on-disk formats either side of the hot path (the reference can dump
velocityField.ply but never read anything back; "matching load" is ours).

-m "not gpu": the C++ writer/reader (libinkhost.so hooks) and the Python mirror (ink_simulation_b200/inksim.py) produce and
              accept the same bytes; malformed files are refused by both.
-m gpu      : a run stopped at a checkpoint and resumed — in a fresh System (Python) or a fresh process (C++ binary), by
              either host from either host's files — continues BIT FOR BIT like the run that never stopped."""
import ctypes as C
import filecmp
import os
import shutil
import subprocess

import numpy as np
import pytest

from ink_simulation_b200 import inksim, seeding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "ink_simulation_b200", "host")
GOLD = os.path.join(ROOT, "tests", "golden")


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


def host_dims():
    d = (C.c_int * 3)()
    seeding.host_lib().inkhost_grid_dims(d)
    return tuple(d)


def cpp_read(path):
    L = seeding.host_lib()
    props = C.c_int(0)
    n = L.inkhost_read_ply(path.encode(), None, C.c_long(0), C.byref(props))
    if n < 0:
        return None
    out = np.empty((n, props.value), np.float32)
    assert L.inkhost_read_ply(path.encode(), _vp(out), C.c_long(out.size), C.byref(props)) == n
    return out


def test_grid_dump_python_and_cpp_write_the_same_bytes(tmp_path):
    dims = host_dims()
    assert dims == (20, 45, 20)                      # libinkhost.so is built for the reference's committed scene
    rng = np.random.default_rng(5)
    u, v, w = (rng.standard_normal(dims).astype(np.float32) for _ in range(3))
    a, b = str(tmp_path / "cpp.ply"), str(tmp_path / "py.ply")
    seeding.host_lib().inkhost_write_grid_ply(a.encode(), C.c_int(3), _vp(u), _vp(v), _vp(w))
    inksim.write_velocity_ply(b, dims, u, v, w)
    assert open(a, "rb").read() == open(b, "rb").read()
    head = open(a, "rb").read(200)
    assert head.startswith(b"ply\nformat binary_small_endian 1.0\nelement vertex 18000\nproperty float x\nproperty float y\nproperty float z\n"
                           b"property float u\nproperty float v\nproperty float w\nend_header\n")   # src/inkSim.cpp:109-118
    # both readers give back the fields and the cell centres of src/inkSim.cpp:137-141
    for got in (cpp_read(b), inksim.read_ply(a)[1]):
        assert np.array_equal(got[:, 3], u.ravel()) and np.array_equal(got[:, 4], v.ravel()) and np.array_equal(got[:, 5], w.ravel())
        assert np.array_equal(got[0, :3], [0.5, 0.5, 0.5]) and np.array_equal(got[-1, :3], [19.5, 44.5, 19.5])
        assert np.array_equal(got[1, :3], [0.5, 0.5, 1.5])                      # k fastest
    ru, rv, rw = inksim.read_velocity_ply(a, dims)
    assert np.array_equal(ru, u) and np.array_equal(rv, v) and np.array_equal(rw, w)
    # the one-field variant used for the Poisson warm start
    seeding.host_lib().inkhost_write_grid_ply(a.encode(), C.c_int(1), _vp(u), None, None)
    inksim.write_ply(b, inksim.grid_records(dims, (u,)), ("x", "y", "z", "q"))
    assert open(a, "rb").read() == open(b, "rb").read()


def test_readers_accept_the_reference_frame_and_refuse_damaged_files(tmp_path):
    want = np.load(os.path.join(GOLD, "host_ply_drop.npy"))
    gold = os.path.join(GOLD, "host_ply.bin")          # a frame written by the COMPILED REFERENCE (make_host_golden.py)
    props, rec = inksim.read_ply(gold)
    assert props == 3 and np.array_equal(rec, want) and np.array_equal(cpp_read(gold), want)
    data = open(gold, "rb").read()
    for name, blob in (("short", data[:-4]), ("long", data + b"\0\0\0\0"), ("nohdr", data.replace(b"end_header\n", b"end_headr\n", 1)),
                       ("noprops", data.replace(b"property float", b"property doubl"))):
        bad = str(tmp_path / f"{name}.ply")
        open(bad, "wb").write(blob)
        assert cpp_read(bad) is None, name
        with pytest.raises(ValueError):
            inksim.read_ply(bad)
    assert cpp_read(str(tmp_path / "missing.ply")) is None


def test_empty_drop_round_trips(tmp_path):
    p = str(tmp_path / "e.ply")
    inksim.write_ply(p, np.empty((0, 6), np.float32), ("x", "y", "z", "vx", "vy", "vz"))
    props, rec = inksim.read_ply(p)
    assert props == 6 and rec.shape == (0, 6) and cpp_read(p).shape == (0, 6)


# ---------------------------------------------------------------------------------------------------------------------------
C1 = (8, 25, 8)


def c1_system(**kw):
    from ink_simulation_b200 import System

    S = System(*C1, arith="strict", **kw)
    S.set_particles(seeding.box_scene(C1, 10000))
    return S


def state(S):
    pos, vel = S.get_particles()
    return [pos, vel, *S.get_velocity(), S.get_warm_start(), S.get_pressure()]


@pytest.mark.gpu
def test_gpu_python_checkpoint_resume_is_bit_exact(tmp_path):
    render, n1, n2 = 0.05, 23, 17
    straight = inksim.InkSim(tmp_path / "a", c1_system())
    straight.simulate(render, n1)
    straight.writeCheckpoint()
    shutil.copytree(tmp_path / "a", tmp_path / "b")
    # the uninterrupted run goes on in the same System, device state untouched; only the frame pacing is told to continue
    # (a plain second simulate() restarts at frame 1, like the reference)
    straight._resumed = True
    straight.simulate(render, n2)
    again = inksim.InkSim(tmp_path / "b", c1_system())          # fresh System, fresh bins, no solver history
    assert again.resume()
    again.simulate(render, n2)
    for x, y in zip(state(straight.ink_system), state(again.ink_system)):
        assert np.array_equal(x, y)
    cmp = filecmp.dircmp(tmp_path / "a", tmp_path / "b")
    frames = sorted(os.listdir(tmp_path / "a" / "drop0"), key=lambda s: int(s.split(".")[0]))
    assert len(frames) > 4 and frames == sorted(os.listdir(tmp_path / "b" / "drop0"), key=lambda s: int(s.split(".")[0]))
    assert filecmp.cmpfiles(tmp_path / "a" / "drop0", tmp_path / "b" / "drop0", frames, shallow=False)[0] == frames
    assert "velocityField.ply" in cmp.common_files and filecmp.cmp(tmp_path / "a" / "velocityField.ply", tmp_path / "b" / "velocityField.ply", shallow=False)
    assert not inksim.InkSim(tmp_path / "none", c1_system()).resume()


@pytest.mark.gpu
def test_gpu_checkpoint_carries_the_phantom_cell_state(tmp_path):
    """out-of-grid particles create phantom cells whose accumulated dt*g enters calcTimeStep's maximum (Q8).  With
    every particle above the grid at step 0 the grid itself is still at rest after that step, so the phantom term IS the
    maximum: a resume that dropped it would take dt_min instead.  The checkpoint carries it (meta.txt, double bits):
    resumed run == uninterrupted run, bit for bit, dt included."""
    from ink_simulation_b200 import System

    def make():
        S = System(*C1, arith="strict")
        pos = seeding.box_scene(C1, 2000)
        pos[:, 1] = C1[1] + 0.75                   # above the grid: phantom cells at step 0, clamped inside by the particle update
        S.set_particles(pos)
        return S

    render, n1, n2 = 0.05, 1, 12
    straight = inksim.InkSim(tmp_path / "a", make())
    straight.simulate(render, n1)
    have, pv = straight.ink_system.get_phantom()
    assert have and pv[1] < 0                      # the phantom cells received dt*g at step 0
    straight.writeCheckpoint()
    assert "phantom 1 " in open(tmp_path / "a" / "checkpoint" / "meta.txt").read()
    shutil.copytree(tmp_path / "a", tmp_path / "b")
    again = inksim.InkSim(tmp_path / "b", make())
    assert again.resume()
    assert again.ink_system.get_phantom() == (have, pv)
    want = straight.ink_system.calc_dt(0.045)
    assert again.ink_system.calc_dt(0.045) == want
    again.ink_system.set_phantom(False, (0, 0, 0))  # the term matters here: without it the resumed run would take another step
    assert again.ink_system.calc_dt(0.045) != want
    again.ink_system.set_phantom(have, pv)
    straight._resumed = True
    straight.simulate(render, n2)
    again.simulate(render, n2)
    assert straight.ink_system.stats().dt == again.ink_system.stats().dt
    for x, y in zip(state(straight.ink_system), state(again.ink_system)):
        assert np.array_equal(x, y)


@pytest.mark.gpu
def test_gpu_cpp_binary_checkpoint_resume_and_cross_host(tmp_path):
    exe = os.path.join(HOST, "inksim_b200_c1")
    assert os.path.exists(exe), "build with __graft_entry__.build()"
    env = dict(os.environ, INK_ARITH="strict")
    render, n1, n2 = "0.05", 23, 17
    a, b, c = str(tmp_path / "a"), str(tmp_path / "b"), str(tmp_path / "c")
    run = lambda *args: subprocess.check_call([exe, *args], env=env, stdout=subprocess.DEVNULL)
    run(a, render, str(n1 + n2))                               # never stopped
    run(b, render, str(n1), "--checkpoint")                    # stopped ...
    shutil.copytree(b, c)
    run(b, render, str(n2), "--resume")                        # ... and continued by a fresh process
    frames = sorted(os.listdir(os.path.join(a, "drop0")), key=lambda s: int(s.split(".")[0]))
    assert len(frames) > 4
    assert sorted(os.listdir(os.path.join(b, "drop0")), key=lambda s: int(s.split(".")[0])) == frames
    assert filecmp.cmpfiles(os.path.join(a, "drop0"), os.path.join(b, "drop0"), frames, shallow=False)[0] == frames
    assert filecmp.cmp(os.path.join(a, "velocityField.ply"), os.path.join(b, "velocityField.ply"), shallow=False)
    # cross-host: the Python host continues the C++ host's checkpoint and lands on the same bytes
    from ink_simulation_b200 import System

    P = inksim.InkSim(c, System(*C1, arith="strict"))
    assert P.resume()
    P.simulate(float(np.float32(0.05)), n2)
    assert filecmp.cmpfiles(os.path.join(a, "drop0"), os.path.join(c, "drop0"), frames, shallow=False)[0] == frames
    assert filecmp.cmp(os.path.join(a, "velocityField.ply"), os.path.join(c, "velocityField.ply"), shallow=False)


# ---------------------------------------------------------------------------------------------------------------------------
class FakeSystem:
    """Stands in for System on a CPU-only machine: deterministic state that depends on every solve() so far, and the
    accessors InkSim uses.  Lets the run loop, frame pacing and checkpoint bookkeeping of inksim.py be tested without a GPU."""

    class _P:
        cell_dim = 1.0

    def __init__(self, dims=(3, 4, 2), n=5):
        self.dims, self.params = dims, self._P()
        self.pos = np.arange(3 * n, dtype=np.float32).reshape(n, 3)
        self.vel = np.zeros((n, 3), np.float32)
        self.u = [np.full(dims, a, np.float32) for a in (1, 2, 3)]
        self.q = np.zeros(dims, np.float32)
        self.asked = []

    def solve(self, ttnr):
        self.asked.append(ttnr)
        dt = min(ttnr, 0.02 + 0.013 * (int(self.pos[0, 0] * 1000) % 3))   # a function of the STATE only (as in the real solver):
        # sometimes the CFL step, sometimes the time to the next frame
        self.pos = (self.pos + np.float32(dt)).astype(np.float32)
        self.vel = (self.vel * np.float32(0.5) + np.float32(dt)).astype(np.float32)
        self.u = [(a * np.float32(1.01) + np.float32(dt)).astype(np.float32) for a in self.u]
        self.q = (self.q + self.u[0] * np.float32(1e-3)).astype(np.float32)
        return dt

    def num_drops(self): return 1
    def get_particles(self, d=0): return self.pos.copy(), self.vel.copy()
    def get_velocity(self): return [a.copy() for a in self.u]
    def get_warm_start(self): return self.q.copy()
    def get_phantom(self): return getattr(self, "phantom", (False, (0.0, 0.0, 0.0)))
    def set_phantom(self, have, v): self.phantom = (bool(have), tuple(float(x) for x in v))
    def set_velocity(self, u, v, w): self.u = [np.array(a, np.float32) for a in (u, v, w)]
    def set_warm_start(self, q): self.q = np.array(q, np.float32)
    def set_particles(self, pos, vel=None, drop=0): self.pos, self.vel = np.array(pos, np.float32), np.array(vel, np.float32)


def tree(root):
    out = {}
    for d, _, files in os.walk(root):
        for f in files:
            if "checkpoint" not in d:
                out[os.path.relpath(os.path.join(d, f), root)] = open(os.path.join(d, f), "rb").read()
    return out


def test_run_loop_pacing_and_checkpoint_bookkeeping_without_a_gpu(tmp_path):
    render, n1, n2 = 0.05, 11, 9
    a = inksim.InkSim(tmp_path / "a", FakeSystem())
    a.simulate(render, n1 + n2)
    # pacing of src/inkSim.cpp:17-31 in float arithmetic: solve() is asked for the time left to the next frame, a frame is
    # written when the accumulated time reaches renderTimestep, numbering starts at 1
    asked = a.ink_system.asked
    assert asked[0] == float(np.float32(render)) and all(0 < t <= asked[0] for t in asked)
    frames = sorted(int(f.split(".")[0]) for f in os.listdir(tmp_path / "a" / "drop0"))
    assert frames == list(range(1, len(frames) + 1)) and 3 < len(frames) < n1 + n2
    b = inksim.InkSim(tmp_path / "b", FakeSystem())
    b.simulate(render, n1)
    b.writeCheckpoint()
    meta = open(tmp_path / "b" / "checkpoint" / "meta.txt").read().split()
    assert meta[:4] == ["grid", "3", "4", "2"] and meta[4:6] == ["drops", "1"] and meta[6] == "next_frame" and meta[8] == "time_since_last_render_bits"
    c = inksim.InkSim(tmp_path / "b", FakeSystem())          # fresh object, fresh "device" state
    assert c.resume()
    assert np.array_equal(c.ink_system.pos, b.ink_system.pos) and np.array_equal(c.ink_system.q, b.ink_system.q)
    c.simulate(render, n2)
    assert tree(tmp_path / "a") == tree(tmp_path / "b")      # every frame and velocityField.ply, byte for byte
    assert c.ink_system.asked == asked[n1:]                  # the resumed run asked solve() for exactly the same times
    # a plain second simulate() starts again at frame 1, like the reference; a checkpoint of another grid is refused
    c.simulate(render, 3)
    assert c.ink_system.asked[n2] == float(np.float32(render))
    with pytest.raises(ValueError):
        inksim.InkSim(tmp_path / "b", FakeSystem(dims=(3, 4, 3))).resume()
