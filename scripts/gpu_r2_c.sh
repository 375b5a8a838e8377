# graphs + cluster coarse kernel: parity subset, then A/B bench lines on one GPU
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_checkpoint.py -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2c_pytest.log
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --cpu-steps 0"
for mode in "default" "INK_NO_GRAPH=1" "INK_MG_COARSE=0" "INK_MG_COARSE=0 INK_NO_GRAPH=1" "INK_MG_CLUSTER_CTAS=8" "INK_MG_COARSE_CELLS=600000"; do
  tag=$(echo "$mode" | tr ' =' '__')
  env $( [ "$mode" = default ] || echo $mode ) timeout 300 python bench.py $F 2>gpurun_out/r2c_$tag.err | tail -1 > gpurun_out/r2c_$tag.json
  python - "$mode" gpurun_out/r2c_$tag.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    c4 = d.get("c4") or {}
    print("%-34s C3 %.3f ms/step pressure %.3f CG %.1f launches %.0f graphs %.0f | C4 %.3f ms pressure %.2f CG %.1f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], d["launches_per_step"], d["graph_launches_per_step"], c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("pressure", 0), c4.get("cg_iters", 0)))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(sys.argv[2].replace(".json", ".err")).read()[-1500:])
PY
done
