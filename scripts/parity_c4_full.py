"""One-off (GPU box): bench.full_size_parity at the FULL config 4 — 256^3 grid, 100 M particles — developed for 5 steps with the
bench's own settings.  The GPU test runs the same check at 20 M particles to stay quick; this is the full-size record
(profiles/r3_parity_c4_full.json)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from ink_simulation_b200 import System  # noqa: E402

dims, n = (256, 256, 256), 100_000_000
t0 = time.time()
S = System(*dims, arith="fast", cg_rtol=1e-6)
S.set_particles(bench.make_particles(dims, n, "uniform"))
for _ in range(5):
    S.solve(0.35)
u, v, w = S.get_velocity()
pos, vel = S.get_particles()
S.close()
r = bench.full_size_parity(dims, u, v, w, pos, vel, label="c4: ")
r["wall_s"] = time.time() - t0
print(json.dumps(r))
