"""Which part of the streaming e2e loop fails to overlap?  (GPU diagnostic, not a bench.)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ink_simulation_b200 import System
import bench

dims = (128, 256, 128); n = 10_000_000
pos = torch.from_numpy(bench.make_particles(dims, n, "uniform", first=0, count=n)).pin_memory()
S = System(*dims, arith="fast", cg_rtol=1e-6, resort_interval=16)
S.set_particles_shard(pos)
for _ in range(10): S.solve(0.35)
out = [torch.empty_like(pos).pin_memory() for _ in range(2)]
S.get_particles(0, out_pos=out[0])
inb = [out[0].clone().pin_memory(), out[0].clone().pin_memory()]

def run(name, body, k=16):
    S.sync_io(); torch.cuda.synchronize(); t0 = time.time()
    body(k)
    S.sync_io(); torch.cuda.synchronize()
    print(f"{name:34s} {(time.time() - t0) / k * 1e3:7.3f} ms/step", flush=True)

def solve_only(k):
    for i in range(k): S.solve(0.35)
def in_only(k):
    S.prefetch_positions(inb[0])
    for i in range(k):
        S.commit_positions()
        if i + 1 < k: S.prefetch_positions(inb[(i + 1) & 1])
        S.solve(0.35)
def out_only(k):
    for i in range(k):
        S.solve(0.35); S.fetch_positions_async(out[i & 1])
def both(k):
    S.prefetch_positions(inb[0])
    for i in range(k):
        S.commit_positions()
        if i + 1 < k: S.prefetch_positions(inb[(i + 1) & 1])
        S.solve(0.35); S.fetch_positions_async(out[i & 1])
def commit_no_copy(k):   # prefetch from a DEVICE buffer: D2D copy on the copy stream (no PCIe)
    d = inb[0].cuda()
    S.prefetch_positions(d)
    for i in range(k):
        S.commit_positions()
        if i + 1 < k: S.prefetch_positions(d)
        S.solve(0.35)
for rep in range(2):
    run("solve only", solve_only); run("prefetch+commit+solve", in_only); run("solve+fetch_async", out_only)
    run("both", both); run("commit from device buffer + solve", commit_no_copy)
