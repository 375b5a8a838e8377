# trilinear V(1,1) preconditioner: parity tests, bench variants (env knobs), launch list
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
run() {  # name, env...
  name=$1; shift
  env "$@" python bench.py --steps 10 --warmup 5 --cpu-steps 0 --no-e2e --no-kernels > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err || tail -5 gpurun_out/b_$name.err
  python - $name <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open("gpurun_out/b_" + f + ".json"))
    print(f, "ms/step", round(d["ms_per_step"], 3), "iters", d["cg_iters_per_step"]["mean"], "pressure", round(d["phase_ms_last_step"]["pressure"], 3))
except Exception as e:
    print(f, "failed", e)
PY
}
run default X=1
run unfused INK_MG3_UNFUSED=1
run tail16 INK_MG_TAIL_KB=16
run tail48 INK_MG_TAIL_KB=48
B="python bench.py --steps 1 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mg3.csv $B > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_summary.py gpurun_out/launches_mg3.csv | head -16
INK_MG_TAIL_KB=16 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mg3_t16.csv $B > gpurun_out/ncu_launch.log 2>&1
python scripts/launch_summary.py gpurun_out/launches_mg3_t16.csv | head -16
