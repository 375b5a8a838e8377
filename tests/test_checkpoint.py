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
