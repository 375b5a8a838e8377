set -x
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 4200 --csv --log-file gpurun_out/launches_r1.csv $B > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_particles_rk2 -s 2 -c 2 -o gpurun_out/prof_k8 $B > gpurun_out/ncu_k8.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_cg_ -s 300 -c 4 -o gpurun_out/prof_cg $B > gpurun_out/ncu_cg.log 2>&1
ls -la gpurun_out
