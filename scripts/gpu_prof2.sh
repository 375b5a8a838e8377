set -x
B="python bench.py --steps 2 --warmup 3 --cpu-steps 0 --no-kernels"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"import|export|mark|bin_|scan" -c 80 --csv --log-file gpurun_out/launches_io.csv $B > gpurun_out/ncu_io.log 2>&1
python scripts/launch_summary.py gpurun_out/launches_io.csv
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-kernels --no-e2e"
ncu --set full --clock-control none --import-source on -k regex:k_mg3_up -s 12 -c 4 -o gpurun_out/prof_mg3_up $B > gpurun_out/ncu_up.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_mg3_res" -s 16 -c 4 -o gpurun_out/prof_mg3_down $B > gpurun_out/ncu_down.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_cg_dir|k_cg_upd" -s 8 -c 2 -o gpurun_out/prof_cg $B > gpurun_out/ncu_cg.log 2>&1
ls -la gpurun_out/*.ncu-rep
