# tiled down kernel A/B (one GPU)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "same_preconditioner" > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j_pytest.log
INK_MG3_DOWN_TILE=2 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "poisson or projection or golden or full_steps or zero_rhs" > gpurun_out/r2j_pytest2.log 2>&1; echo "pytest(down tile everywhere) rc=$?"; tail -3 gpurun_out/r2j_pytest2.log
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --cpu-steps 0"
for mode in "default" "INK_MG3_DOWN_TILE=1" "INK_MG3_DOWN_TILE=2" "INK_MG3_DOWN_TILE=1 INK_MG3_UP_TILE=2"; do
  tag=$(echo "$mode" | tr ' =' '__')
  env $( [ "$mode" = default ] || echo $mode ) timeout 300 python bench.py $F 2>gpurun_out/r2j_$tag.err | tail -1 > gpurun_out/r2j_$tag.json
  python - "$mode" gpurun_out/r2j_$tag.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    c4 = d.get("c4") or {}
    print("%-40s C3 %.3f ms/step pressure %.3f CG %.1f launches %.0f | C4 %.3f ms pressure %.2f CG %.1f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], d["launches_per_step"], c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("pressure", 0), c4.get("cg_iters", 0)))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(sys.argv[2].replace(".json", ".err")).read()[-1500:])
PY
done
G="--steps 3 --warmup 5 --no-e2e --no-kernels --no-c4 --no-parity --no-strict --cpu-steps 0"
INK_MG3_DOWN_TILE=2 INK_TRACE_ITER=5 timeout 200 python bench.py $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2j_trace_c3.txt
INK_MG3_DOWN_TILE=2 INK_TRACE_ITER=5 timeout 200 python bench.py --config c4 $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2j_trace_c4.txt
