"""Small driver for compute-sanitizer (memcheck / racecheck): every new kernel of round 2 on small ragged grids.
usage (GPU box): compute-sanitizer --tool memcheck python scripts/sanitize_r2.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from ink_simulation_b200 import System, seeding  # noqa: E402

SWITCHES = [{}, {"INK_MG3_UP_TILE": "2"}, {"INK_MG3_DOWN_TILE": "2"}, {"INK_CG_MATVEC_TILE": "1"}, {"INK_K8_TILE": "2"}, {"INK_MG3_FUSED_DOWN": "1"},
            {"INK_MG_COARSE": "2"}]
for env in SWITCHES:
    for k, v in env.items():
        os.environ[k] = v
    for dims in ((16, 24, 64), (20, 45, 20), (10, 9, 132), (40, 24, 260)):
        S = System(*dims, cg_rtol=1e-6)
        u, v, w = util.rand_field(dims, 3, scale=0.4)
        S.set_velocity(u, v, w)
        S.set_particles(util.wild_particles(dims, 4000, 4))
        for _ in range(3):
            S.solve(0.35)
        assert S.check_finite() == (0, 0)
        S.close()
    for k in env:
        del os.environ[k]
    print("ok", env, flush=True)
S = System(20, 45, 20)
seeding.seed_emitter_on_device(S)
seeding.seed_box_on_device(S, 5000, drop=1)
S.solve(0.35)
S.close()
print("sanitize_r2: done")
