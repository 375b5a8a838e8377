# quick check + bench of the current library (one GPU)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "poisson or projection or golden or full_steps or zero_rhs or same_preconditioner or north_star" > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2m_pytest.log
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --cpu-steps 0"
for rep in 1 2; do
  timeout 300 python bench.py $F 2>gpurun_out/r2m.err | tail -1 > gpurun_out/r2m.json
  python - gpurun_out/r2m.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    c4 = d.get("c4") or {}
    print("C3 %.3f ms/step pressure %.3f CG %.1f | C4 %.3f ms pressure %.2f CG %.1f" % (d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("pressure", 0), c4.get("cg_iters", 0)))
except Exception as e:
    print("FAILED", e); print(open("gpurun_out/r2m.err").read()[-1500:])
PY
done
G="--steps 3 --warmup 5 --no-e2e --no-kernels --no-c4 --no-parity --no-strict --cpu-steps 0"
INK_TRACE_ITER=5 timeout 200 python bench.py --config c4 $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2m_trace_c4.txt
