"""Multi-GPU coverage.

-m gpu (needs >= 2 B200s on the box, skipped otherwise): tests/mgpu_check.py under torchrun — the NCCL y-slab / particle
shard path against a single-GPU run of the same problem (bit-exact phases, solver within tolerance).
-m "not gpu": the host-side sharding rules on 2 gloo ranks (slab rows, particle index shards, the NCCL-id broadcast
plumbing of create_distributed up to — not including — the GPU call)."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_multi_gpu_parity_under_torchrun():
    """every GPU of the box (up to 8): ranks with two neighbours only exist from 3 ranks on"""
    import torch

    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "mgpu_check.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=1500)
    assert out.returncode == 0 and "mgpu_check: OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


GLOO_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import numpy as np
import torch.distributed as dist
from ink_simulation_b200 import shard_range, slab_range
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# the id exchange create_distributed() performs (rank 0's 128 bytes reach everyone)
box = [bytes(range(128)) if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
assert box[0] == bytes(range(128))
# shards tile the index range / the grid rows exactly, in rank order, for awkward sizes too
for n in (0, 1, 7, 10_000_001):
    parts = [None] * world
    dist.all_gather_object(parts, shard_range(n, rank, world))
    assert parts[0][0] == 0 and parts[-1][1] == n and all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
for Y in (8, 45, 256):
    rows = [None] * world
    dist.all_gather_object(rows, slab_range(Y, rank, world))
    assert rows[0][0] == 0 and rows[-1][1] == Y and all(rows[i][1] == rows[i + 1][0] for i in range(world - 1))
# a rank-sharded gather reproduces the caller-order array (what get_particles shards concatenate to)
full = np.arange(30, dtype=np.float32).reshape(10, 3)
lo, hi = shard_range(10, rank, world)
got = [None] * world
dist.all_gather_object(got, full[lo:hi])
assert np.array_equal(np.concatenate(got), full)
dist.destroy_process_group()
print("GLOO-OK", rank)
"""


def test_sharding_rules_on_two_gloo_ranks(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(GLOO_WORKER % ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29534", str(script)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and out.stdout.count("GLOO-OK") == 2, out.stdout[-2000:] + out.stderr[-2000:]


def test_bind_host_to_gpu_is_a_safe_hint(monkeypatch):
    """bench.py calls bind_host_to_gpu(LOCAL_RANK) under torchrun so that each rank's pinned buffers sit on its GPU's NUMA
    node.  It must never raise, must only ever SHRINK the affinity to CPUs the process already has, and must do nothing when
    NVML is missing or its answer is useless."""
    import types

    from ink_simulation_b200 import bind_host_to_gpu

    before = os.sched_getaffinity(0)
    try:
        monkeypatch.setitem(sys.modules, "pynvml", None)                 # import fails
        assert bind_host_to_gpu(0).startswith("unchanged") and os.sched_getaffinity(0) == before
        fake = types.ModuleType("pynvml")
        fake.nvmlInit = lambda: None
        fake.nvmlDeviceGetHandleByUUID = lambda u: (_ for _ in ()).throw(RuntimeError("no such uuid"))
        fake.nvmlDeviceGetHandleByIndex = lambda i: ("handle", i)
        monkeypatch.setitem(sys.modules, "pynvml", fake)
        mask_all = sum(1 << c for c in before)
        fake.nvmlDeviceGetCpuAffinity = lambda h, n: [(mask_all >> (64 * w)) & (2**64 - 1) for w in range(n)]
        assert bind_host_to_gpu(0).startswith("unchanged") and os.sched_getaffinity(0) == before      # everything is local already
        fake.nvmlDeviceGetCpuAffinity = lambda h, n: [0] * n
        assert bind_host_to_gpu(0).startswith("unchanged") and os.sched_getaffinity(0) == before      # empty intersection
        one = 1 << min(before)
        fake.nvmlDeviceGetCpuAffinity = lambda h, n: [(one >> (64 * w)) & (2**64 - 1) for w in range(n)]
        assert bind_host_to_gpu(0).startswith("unchanged") and os.sched_getaffinity(0) == before      # would squeeze the rank onto 1 core
        if len(before) >= 8:
            keep = sorted(before)[: len(before) // 2]
            sub = sum(1 << c for c in keep) | (1 << (max(before) + 7))                                 # plus a CPU we do not own
            fake.nvmlDeviceGetCpuAffinity = lambda h, n: [(sub >> (64 * w)) & (2**64 - 1) for w in range(n)]
            assert bind_host_to_gpu(0).startswith("bound to the") and os.sched_getaffinity(0) == set(keep)
    finally:
        os.sched_setaffinity(0, before)
