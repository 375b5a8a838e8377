# quick GPU check: parity tests + short strict/fast bench (no CPU baseline, no e2e)
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for a in strict fast; do
  python bench.py --steps 5 --warmup 5 --cpu-steps 0 --no-e2e --arith $a > gpurun_out/b_$a.json 2> gpurun_out/b_$a.err || tail -5 gpurun_out/b_$a.err
done
python - <<'PY'
import json
for f in ("b_strict", "b_fast"):
    try:
        d = json.load(open("gpurun_out/" + f + ".json"))
        print(f, "ms/step", round(d["ms_per_step"], 3), "iters", d["cg_iters_per_step"]["mean"])
        print("  kernels", {k: {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items() if a != "note"} for k, v in d["kernels"].items()})
        print("  phases", {k: round(v, 3) for k, v in d["phase_ms_last_step"].items()})
    except Exception as e:
        print(f, "failed", e)
PY
