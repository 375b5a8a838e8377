"""ink_simulation_b200 — B200-native (sm_100a) implementation of Ink-Simulation's per-timestep hot path
(System::solve: MAC water-grid update + RK2 ink-particle update) behind the reference's System / InkSim interface.

Layout:  csrc/  CUDA kernels + the C ABI (include/ink_capi.h -> libinkb200.so)
         host/  C++ drop-in `System` / `InkSim` over the C ABI (the reference is C++)
         system.py / inksim.py / seeding.py  Python mirror of the same interface (System, InkSim + .ply formats and
                checkpoints, the reference's seeders) for tests and bench.py
"""
from ._capi import LIB_PATH, load  # noqa: F401
from .system import InkError, Params, System, bind_host_to_gpu, create_distributed, nccl_unique_id, shard_range, slab_range  # noqa: F401
from .inksim import InkSim  # noqa: F401,E402
