# INK_MG_NODONE A/B (one GPU)
mkdir -p gpurun_out
INK_MG_NODONE=1 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_checkpoint.py -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; echo "pytest(nodone) rc=$?"; tail -3 gpurun_out/r2k_pytest.log
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --cpu-steps 0"
for mode in "default" "INK_MG_NODONE=1" "default" "INK_MG_NODONE=1"; do
  env $( [ "$mode" = default ] || echo $mode ) timeout 300 python bench.py $F 2>gpurun_out/r2k.err | tail -1 > gpurun_out/r2k.json
  python - "$mode" gpurun_out/r2k.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    c4 = d.get("c4") or {}
    print("%-20s C3 %.3f ms/step pressure %.3f CG %.1f | C4 %.3f ms pressure %.2f CG %.1f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("pressure", 0), c4.get("cg_iters", 0)))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("gpurun_out/r2k.err").read()[-1500:])
PY
done
