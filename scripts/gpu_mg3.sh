# trilinear V(1,1) preconditioner: parity tests, bench old vs new, launch list
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
for p in mg mg_const; do
  python bench.py --steps 10 --warmup 5 --cpu-steps 0 --no-e2e --precond $p > gpurun_out/b_$p.json 2> gpurun_out/b_$p.err || tail -5 gpurun_out/b_$p.err
done
python - <<'PY'
import json
for f in ("b_mg", "b_mg_const"):
    try:
        d = json.load(open("gpurun_out/" + f + ".json"))
        print(f, "ms/step", round(d["ms_per_step"], 3), "iters", d["cg_iters_per_step"])
        print("  kernels", {k: {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a != "note"} for k, v in d["kernels"].items()})
        print("  phases", {k: round(v, 3) for k, v in d["phase_ms_last_step"].items()})
    except Exception as e:
        print(f, "failed", e)
PY
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mg3.csv $B > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_summary.py gpurun_out/launches_mg3.csv | head -30
