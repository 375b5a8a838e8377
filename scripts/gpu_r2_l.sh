# C4 (256^3) on one GPU: ncu --set full of the level-0 PCG kernels, summarised on the box
mkdir -p gpurun_out
B="python bench.py --config c4 --steps 3 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels --no-c4 --no-strict"
export INK_NO_GRAPH=1 INK_BENCH_MIN_MS=0
timeout 900 ncu --set full --clock-control none -k regex:"k_cg_dir_matvec|k_cg_update|k_mg3_up_tile|k_mg3_residual_rz_v4|k_mg3_restrict_xy_v4" -s 300 -c 17 -o gpurun_out/r2l_c4 -f $B > gpurun_out/r2l_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/ncu_summary.py gpurun_out/r2l_c4.ncu-rep > gpurun_out/r2l_c4_kernels.txt 2>&1
ncu -i gpurun_out/r2l_c4.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
ki=h.index('Kernel Name'); di=h.index('gpu__time_duration.sum')
want=[c for c in h if c.startswith('smsp__average_warps_issue_stalled') and c.endswith('per_issue_active.ratio')]
for r in rows[2:]:
    if float(r[di].replace(',',''))<20000: continue
    print(r[ki][:40], r[di], ' '.join('%s=%s'%(c.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio',''), r[h.index(c)]) for c in want if float(r[h.index(c)] or 0)>0.5))
" > gpurun_out/r2l_c4_stalls.txt 2>&1
rm -f gpurun_out/r2l_c4.ncu-rep
grep -E "^==|duration|dram read|dram write|dram % of peak|issue slots|L1/TEX|L2 hit|occupancy|warp instr" gpurun_out/r2l_c4_kernels.txt | awk '/^==/{k=$0} /duration/{d=$2} /warp instr/{if (d>20) print k, d}' | sort | uniq | head
cat gpurun_out/r2l_c4_stalls.txt | head -20
