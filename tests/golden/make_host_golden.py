"""Generates tests/golden/host_emitter.npz and host_ply.bin from the COMPILED REFERENCE (oracle/_ref/libinkref.so):
  * the emitter scene of System::init (src/system.cpp:17-23 -> src/inkSolver.cpp:11-42) for the committed 20x45x20 grid:
    particle count, a strided sample of positions/velocities and SHA-256 of the full arrays;
  * the bytes InkSim::writeToFile (src/inkSim.cpp:43-96) produces for a small seeded drop.
Run here (needs /root/reference): python tests/golden/make_host_golden.py"""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

oracle.build()
R = oracle.RefSystem(4, 45, 4)   # emitter positions depend on the grid size only through the ceiling centre (X/2, Y-.1, Z/2)
R.init_emitter()
pos, vel = R.get_particles()
# re-centre to the committed 20x45x20 scene: done by the test on its side (centre differs only in x, z)
np.savez_compressed(os.path.join(HERE, "host_emitter.npz"), dims=np.array([4, 45, 4]), n=len(pos), pos_sample=pos[::997], vel_sample=vel[::997],
                    pos_sha=hashlib.sha256(pos.tobytes()).hexdigest(), vel_sha=hashlib.sha256(vel.tobytes()).hexdigest())
rng = np.random.default_rng(7)
drop = (rng.random((257, 3)) * 5).astype(np.float32)
with tempfile.TemporaryDirectory() as d:
    os.makedirs(os.path.join(d, "drop0"))
    oracle.write_ply_with_reference(d, 3, [drop])
    data = open(os.path.join(d, "drop0", "3.ply"), "rb").read()
open(os.path.join(HERE, "host_ply.bin"), "wb").write(data)
np.save(os.path.join(HERE, "host_ply_drop.npy"), drop)
print("emitter particles:", len(pos), "ply bytes:", len(data))
