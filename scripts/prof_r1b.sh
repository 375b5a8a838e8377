# round-1 (b) GPU pass: parity tests, full bench line, reference arm, ncu launch list, ncu --set full of K8 / MG
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err || tail -5 gpurun_out/bench_c3.err
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err || tail -5 gpurun_out/bench_ref.err
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r1b.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_particles_rk2 -s 2 -c 1 -o gpurun_out/prof_k8 $B > gpurun_out/ncu_k8.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_mg_ -s 200 -c 24 -o gpurun_out/prof_mg $B > gpurun_out/ncu_mg.log 2>&1
ls -la gpurun_out
