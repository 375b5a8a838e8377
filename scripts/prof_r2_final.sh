# round-2 final single-GPU evidence pass:   gpurun --timeout 1500 -- 'bash scripts/prof_r2_final.sh'
# GPU test suite, smoke, the headline bench line (both arms, plus the reference AS SHIPPED at C1 / C2), then — only after the plain
# command has exited 0 — the ncu launch list of the same step and one --set full capture of each top kernel, and an iteration trace.
# The ncu passes run with INK_NO_GRAPH=1 (direct launches: same kernels, same order) so the launch list is one row per kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2h_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/r2h_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench_c3.json 2> gpurun_out/r2h_bench_c3.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c3.json 2> gpurun_out/r2h_bench_ref_c3.err; echo "ref c3 rc=$?"
timeout 300 python bench.py --impl reference --config c1 --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c1.json 2>/dev/null; echo "ref c1 rc=$?"
timeout 600 python bench.py --impl reference --config c2 --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref_c2.json 2>/dev/null; echo "ref c2 rc=$?"
for c in c1 c2; do timeout 300 python bench.py --config $c --steps 20 --warmup 5 --no-c4 --cpu-steps 0 > gpurun_out/r2h_bench_$c.json 2>/dev/null; echo "gpu $c rc=$?"; done
B="python bench.py --steps 20 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels --no-c4 --no-strict"
export INK_NO_GRAPH=1 INK_BENCH_MIN_MS=0
timeout 300 $B > gpurun_out/r2h_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 3000 --csv --log-file gpurun_out/r2h_launches.csv $B > gpurun_out/r2h_ncu_launch.log 2>&1; echo "ncu list rc=$?"
python scripts/launch_summary.py gpurun_out/r2h_launches.csv > gpurun_out/r2h_launches_summary.txt 2>&1; head -24 gpurun_out/r2h_launches_summary.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_cg_dir_matvec|k_cg_update|k_mg3_up_tile|k_mg3_residual_rz_v4|k_mg3_restrict_xy_v4|k_particles_rk2|k_viscosity_v4|k_vorticity_v4|k_dilate_add_gravity_v4" -s 600 -c 90 -o gpurun_out/r2h_top -f $B > gpurun_out/r2h_ncu_top.log 2>&1; echo "ncu top rc=$?"
unset INK_NO_GRAPH INK_BENCH_MIN_MS
INK_TRACE_ITER=5 timeout 200 $B 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2h_trace_c3.txt
python - <<PY
import json
d = json.loads(open("gpurun_out/r2h_bench_c3.json").read().strip().splitlines()[-1])
print("C3 N=1: %.3f ms/step %.3g upd/s e2e %.3f ms; pressure frac %.3f share %.2f; K8 frac %.3f; cpu %.1f ms (%d thr); c4 %.3f ms; strict %.3f ms" % (
    d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["roofline"]["share_of_step"], d["roofline"]["per_kernel"][0]["frac"],
    d["cpu_baseline"]["ms_per_step"], d["cpu_baseline"]["cores"], d["c4"]["ms_per_step"], d["strict"]["ms_per_step"]))
for f in ("ref_c3", "ref_c1", "ref_c2", "c1", "c2"):
    try:
        r = json.loads(open(f"gpurun_out/r2h_bench_{f}.json").read().strip().splitlines()[-1])
        print(f, "%.3f ms/step" % r["ms_per_step"], r.get("cpu_baseline") and r["cpu_baseline"].get("kind"), r.get("reference_phase_ms"))
    except Exception as e:
        print(f, "unreadable", e)
PY
