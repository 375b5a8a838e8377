# final state of the tree: GPU suite, smoke, headline bench line, ncu launch list of the same command (CSV kept, gzip)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2n_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 | tee gpurun_out/r2n_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2n_bench_c3.json 2> gpurun_out/r2n_bench_c3.err; echo "bench rc=$?"
B="python bench.py --steps 20 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels --no-c4 --no-strict"
export INK_NO_GRAPH=1 INK_BENCH_MIN_MS=0
timeout 300 $B > gpurun_out/r2n_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1500 -c 3000 --csv --log-file gpurun_out/r2n_launches.csv $B > gpurun_out/r2n_ncu_launch.log 2>&1; echo "ncu list rc=$?"
python scripts/launch_summary.py gpurun_out/r2n_launches.csv > gpurun_out/r2n_launches_summary.txt 2>&1; head -12 gpurun_out/r2n_launches_summary.txt
gzip -f gpurun_out/r2n_launches.csv
python - <<PY
import json
d = json.loads(open("gpurun_out/r2n_bench_c3.json").read().strip().splitlines()[-1])
print("C3 N=1: %.3f ms/step %.3g upd/s e2e %.3f ms (copy-only %.3f); pressure %.3f ms frac %.3f share %.2f; K8 frac %.3f; cpu %.1f ms (%d thr); c4 %.3f ms; strict %.3f ms" % (
    d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["e2e"]["copy_only_ms_per_step"], d["roofline"]["ms"], d["roofline"]["frac"], d["roofline"]["share_of_step"], d["roofline"]["per_kernel"][0]["frac"],
    d["cpu_baseline"]["ms_per_step"], d["cpu_baseline"]["cores"], d["c4"]["ms_per_step"], d["strict"]["ms_per_step"]))
PY
du -sh gpurun_out
