# round-1 (e) GPU pass: parity tests, smoke, full bench line (+c4), reference arm, ncu launch list, ncu --set full of K8
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err || tail -5 gpurun_out/bench_c3.err
python bench.py --config c4 --steps 10 --warmup 5 --cpu-steps 0 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err || tail -5 gpurun_out/bench_c4.err
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r1e.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_particles_rk2 -s 2 -c 1 -o gpurun_out/prof_k8_e $B > gpurun_out/ncu_k8.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_mg3_up|k_cg_dir|k_cg_upd|k_vorticity|k_viscosity" -s 40 -c 12 -o gpurun_out/prof_top_e $B > gpurun_out/ncu_top.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err || tail -3 gpurun_out/bench_ref.err
INK_TRACE_ITER=5 python bench.py --steps 3 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels 2>&1 >/dev/null | grep trace | tail -1 > gpurun_out/trace_e.txt
python - <<'PY'
import torch, time
n = 120_000_000
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(f, reps=5):
    torch.cuda.synchronize(); t0 = time.time()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.time() - t0) / reps * 1e3
def h2d():
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
def both(): h2d(); d2h()
for f in (h2d, d2h, both): f()
print("PCIe 120 MB: h2d %.2f ms, d2h %.2f ms, both concurrently %.2f ms" % (t(h2d), t(d2h), t(both)))
PY
ls -la gpurun_out | tail -5
