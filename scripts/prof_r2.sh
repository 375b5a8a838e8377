# round-2 opening GPU pass (one B200, ~4 GPU-minutes): everything round 1 could no longer afford to re-capture.
#   gpurun --timeout 900 -- 'bash scripts/prof_r2.sh'
# parity tests, smoke, headline bench line, ncu launch list of the same command (incl. the new k_dilate_add_gravity_v4),
# ncu --set full of K8 and of the dilation kernel, iteration trace.  ncu only after the plain command has exited 0.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/r2_a_bench_c3.json 2> gpurun_out/r2_a_bench_c3.err || tail -5 gpurun_out/r2_a_bench_c3.err
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2_a_launches.csv $B > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_summary.py gpurun_out/r2_a_launches.csv > gpurun_out/r2_a_launches_summary.txt 2>&1; head -30 gpurun_out/r2_a_launches_summary.txt
ncu --set full --clock-control none --import-source on -k regex:"k_particles_rk2|k_dilate_add_gravity" -s 4 -c 2 -o gpurun_out/r2_a_k8_dilate $B > gpurun_out/ncu_k8.log 2>&1
INK_TRACE_ITER=5 python bench.py --steps 3 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels 2>&1 >/dev/null | grep trace | tail -1 > gpurun_out/r2_a_trace.txt; cat gpurun_out/r2_a_trace.txt
ls -la gpurun_out | tail -8
