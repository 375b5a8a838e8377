# round-2 final single-GPU evidence pass, part 2 (the first pass produced > 64 MiB of .ncu-rep and nothing came back but stdout):
# bench lines of both arms, then ONE --set full capture pass summarised to text ON THE BOX (the report itself is deleted).
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench_c3.json 2> gpurun_out/r2h_bench_c3.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c3.json 2> gpurun_out/r2h_bench_ref_c3.err; echo "ref c3 rc=$?"
timeout 300 python bench.py --impl reference --config c1 --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c1.json 2>/dev/null; echo "ref c1 rc=$?"
timeout 600 python bench.py --impl reference --config c2 --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c2.json 2>/dev/null; echo "ref c2 rc=$?"
for c in c1 c2; do timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-c4 --cpu-steps 0 > gpurun_out/r2h_bench_$c.json 2>/dev/null; echo "gpu $c rc=$?"; done
B="python bench.py --steps 20 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels --no-c4 --no-strict"
export INK_NO_GRAPH=1 INK_BENCH_MIN_MS=0
timeout 900 ncu --set full --clock-control none -k regex:"k_cg_dir_matvec|k_cg_update|k_mg3_up_tile|k_mg3_residual_rz_v4|k_mg3_restrict_xy_v4|k_particles_rk2|k_viscosity_v4|k_vorticity_v4|k_curl_v4|k_dilate_add_gravity_v4" -s 600 -c 34 -o gpurun_out/r2h_top -f $B > gpurun_out/r2h_ncu_top.log 2>&1; echo "ncu top rc=$?"
python scripts/ncu_summary.py gpurun_out/r2h_top.ncu-rep > gpurun_out/r2h_top_kernels.txt 2>&1; rm -f gpurun_out/r2h_top.ncu-rep
grep -E "^==|duration|dram read|dram write|dram % of peak|issue slots" gpurun_out/r2h_top_kernels.txt | head -80
du -sh gpurun_out
