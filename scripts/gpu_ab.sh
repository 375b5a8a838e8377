# A/B of env knobs on the C3 bench (no CPU baseline); usage: gpu_ab.sh "name ENV=.." ...
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
run() {
  name=$1; shift
  env "$@" python bench.py --steps 20 --warmup 5 --cpu-steps 0 $BENCH_EXTRA > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err || tail -5 gpurun_out/b_$name.err
  python - $name <<'PY'
import json, sys
f = sys.argv[1]
try:
    d = json.load(open("gpurun_out/b_" + f + ".json"))
    e = d.get("e2e") or {}
    print(f, "ms/step", round(d["ms_per_step"], 3), "iters", d["cg_iters_per_step"]["mean"], "e2e ms", round(e.get("ms_per_step", 0), 3),
          "phases", {k: round(v, 3) for k, v in d["phase_ms_last_step"].items()})
    if d.get("kernels"): print("   kernels", {k: round(v["ms"] if "ms" in v else v["ms_per_iter"], 4) for k, v in d["kernels"].items()})
except Exception as ex:
    print(f, "failed", ex)
PY
}
for spec in "$@"; do run $spec; done
