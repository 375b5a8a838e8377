"""inksim.py — Python mirror of the reference's run loop `InkSim` (src/inkSim.{h,cpp}) over `System`, plus the on-disk
formats either side of the hot path.  The C++ host (host/inkSim.cpp) is the drop-in; this mirror drives
the SAME C ABI from Python and writes/reads the SAME bytes, so either host can continue the other's run.

Files (all "ply / format binary_small_endian 1.0 / element vertex N / property float ... / end_header" + raw host-endian fp32,
the reference's non-standard format token included — src/inkSim.cpp:62-68, :109-118; render.py:290-294 only looks for
"element vertex" and "end_header"):
    <dir>/drop<i>/<frame>.ply          x y z                 one frame of one drop, particle index order (src/inkSim.cpp:45-96)
    <dir>/velocityField.ply            x y z u v w           cell centre + currVelocity, i-j-k order (exactly X*Y*Z records)
    <dir>/checkpoint/velocityField.ply x y z u v w           same format, state between two steps
    <dir>/checkpoint/warmstart.ply     x y z q               start vector of the next pressure solve (ink_get_warm_start)
    <dir>/checkpoint/drop<i>.ply       x y z vx vy vz        particles of drop i
    <dir>/checkpoint/meta.txt          grid, drops, next frame number, time since the last frame (float bits), phantom-cell velocity (Q8, double bits)
"""
import os
import struct

import numpy as np

from .system import System


def ply_header(n, props):
    return ("ply\nformat binary_small_endian 1.0\nelement vertex %d\n" % n + "".join("property float %s\n" % p for p in props)
            + "end_header\n").encode()


def write_ply(path, records, props):
    rec = np.ascontiguousarray(records, np.float32).reshape(-1, len(props))
    with open(path, "wb") as f:
        f.write(ply_header(len(rec), props))
        f.write(rec.tobytes())


def read_ply(path):
    """-> (property count, (n, props) float32 array).  Same acceptance rule as host/inkSim.cpp: read_ply."""
    data = open(path, "rb").read()
    end = data.find(b"end_header\n")
    if end < 0:
        raise ValueError(f"not a ply dump of this program: {path}")
    end += len(b"end_header\n")
    lines = data[:end].decode().splitlines()
    nv = [ln for ln in lines if ln.startswith("element vertex ")]
    props = sum(ln.startswith("property float ") for ln in lines)
    if not nv or props == 0:
        raise ValueError(f"not a ply dump of this program: {path}")
    n = int(nv[0].split()[-1])
    if len(data) - end != 4 * n * props:
        raise ValueError(f"payload size does not match the header: {path}")
    return props, np.frombuffer(data, np.float32, offset=end).reshape(n, props).copy()


def grid_records(dims, fields, h=1.0):
    """(cx, cy, cz, f0, ...) per cell, i-j-k order, centre = index + h/2 (src/inkSim.cpp:137-141)."""
    X, Y, Z = dims
    i, j, k = np.meshgrid(np.arange(X), np.arange(Y), np.arange(Z), indexing="ij")
    off = np.float32(h) / np.float32(2)
    cols = [(a.ravel().astype(np.float32) + off) for a in (i, j, k)] + [np.asarray(f, np.float32).ravel() for f in fields]
    return np.stack(cols, axis=1)


def write_velocity_ply(path, dims, u, v, w, h=1.0):
    write_ply(path, grid_records(dims, (u, v, w), h), ("x", "y", "z", "u", "v", "w"))


def read_velocity_ply(path, dims):
    props, rec = read_ply(path)
    if props != 6 or len(rec) != int(np.prod(dims)):
        raise ValueError(f"{path}: not a velocity dump of a {dims} grid")
    return tuple(rec[:, 3 + a].reshape(dims).copy() for a in range(3))


class InkSim:
    """InkSim(writeDirectory) / simulate(renderTimestep, totalTimesteps) as in src/inkSim.h:11,19.  `system` is a ready
    System (grid size, seeding are the caller's: the reference fixes them at compile time in src/system.h:15-19)."""

    def __init__(self, writeDirectory, system: System):
        self.ink_system = system
        self.writeDirectory = str(writeDirectory)
        self._next_frame, self._since, self._resumed = 1, np.float32(0), False

    # src/inkSim.cpp:13-43; timeStep and timeSinceLastRender are floats there
    def simulate(self, renderTimestep, totalTimesteps):
        render = np.float32(renderTimestep)
        frame = self._next_frame if self._resumed else 1
        since = self._since if self._resumed else np.float32(0)
        self._resumed = False
        for _ in range(int(totalTimesteps)):
            dt = np.float32(self.ink_system.solve(float(render - since)))
            since = np.float32(since + dt)
            if since >= render:
                self.writeToFile(frame)
                frame += 1
                since = np.float32(0)
        self.writeWaterGridVelocities()
        self._next_frame, self._since = frame, since

    def writeToFile(self, frameNum):
        S = self.ink_system
        for d in range(S.num_drops()):
            path = os.path.join(self.writeDirectory, f"drop{d}")
            os.makedirs(path, exist_ok=True)
            write_ply(os.path.join(path, f"{frameNum}.ply"), S.get_particles(d)[0], ("x", "y", "z"))

    def writeWaterGridVelocities(self):
        S = self.ink_system
        os.makedirs(self.writeDirectory, exist_ok=True)
        write_velocity_ply(os.path.join(self.writeDirectory, "velocityField.ply"), S.dims, *S.get_velocity(), h=S.params.cell_dim)

    # ---- checkpoint / resume (host/inkSim.cpp: writeCheckpoint / resume; same files) ------------------------------------
    def writeCheckpoint(self):
        S = self.ink_system
        ck = os.path.join(self.writeDirectory, "checkpoint")
        os.makedirs(ck, exist_ok=True)
        h = S.params.cell_dim
        write_velocity_ply(os.path.join(ck, "velocityField.ply"), S.dims, *S.get_velocity(), h=h)
        write_ply(os.path.join(ck, "warmstart.ply"), grid_records(S.dims, (S.get_warm_start(),), h), ("x", "y", "z", "q"))
        for d in range(S.num_drops()):
            pos, vel = S.get_particles(d)
            write_ply(os.path.join(ck, f"drop{d}.ply"), np.concatenate([pos, vel], axis=1), ("x", "y", "z", "vx", "vy", "vz"))
        bits = struct.unpack("<I", struct.pack("<f", float(self._since)))[0]
        with open(os.path.join(ck, "meta.txt"), "w") as f:
            f.write("grid %d %d %d\ndrops %d\nnext_frame %d\ntime_since_last_render_bits 0x%08x\n"
                    % (*S.dims, S.num_drops(), self._next_frame, bits))
            have, pv = S.get_phantom()   # phantom cells of out-of-grid particles (Q8) enter calcTimeStep's maximum
            f.write("phantom %d %s\n" % (int(have), " ".join("0x%016x" % struct.unpack("<Q", struct.pack("<d", x))[0] for x in pv)))

    def resume(self):
        S = self.ink_system
        ck = os.path.join(self.writeDirectory, "checkpoint")
        meta = os.path.join(ck, "meta.txt")
        if not os.path.exists(meta):
            return False
        kv = {ln.split()[0]: ln.split()[1:] for ln in open(meta).read().splitlines() if ln.strip()}
        if tuple(int(x) for x in kv["grid"]) != tuple(S.dims):
            raise ValueError("checkpoint is for another grid size")
        drops = int(kv["drops"][0])
        u, v, w = read_velocity_ply(os.path.join(ck, "velocityField.ply"), S.dims)
        props, q = read_ply(os.path.join(ck, "warmstart.ply"))
        if props != 4 or len(q) != int(np.prod(S.dims)):
            raise ValueError("checkpoint warmstart.ply: wrong shape")
        parts = []
        for d in range(drops):   # read everything before touching the state
            props, rec = read_ply(os.path.join(ck, f"drop{d}.ply"))
            if props != 6:
                raise ValueError("checkpoint drop file: wrong shape")
            parts.append(rec)
        S.set_velocity(u, v, w)
        S.set_warm_start(q[:, 3].reshape(S.dims))
        for d, rec in enumerate(parts):
            S.set_particles(rec[:, :3], rec[:, 3:], drop=d)
        if "phantom" in kv:   # absent in checkpoints written before the phantom state was saved
            S.set_phantom(int(kv["phantom"][0]), [struct.unpack("<d", struct.pack("<Q", int(x, 16)))[0] for x in kv["phantom"][1:4]])
        self._next_frame = int(kv["next_frame"][0])
        self._since = np.float32(struct.unpack("<f", struct.pack("<I", int(kv["time_since_last_render_bits"][0], 16)))[0])
        self._resumed = True
        return True
