# 2-GPU box: folded all-reduce — mgpu_check (default and with the tile up kernel forced), A/B bench lines, one trace
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $T --master-port 29511 tests/mgpu_check.py > gpurun_out/r2g_mgpu2.log 2>&1; echo "mgpu rc=$?"; grep -E "^mgpu_check|FAILED|rank [0-9]+:|Error" gpurun_out/r2g_mgpu2.log | head -20
INK_MG3_UP_TILE=2 MGPU_CASES=2,3 timeout 400 $T --master-port 29515 tests/mgpu_check.py > gpurun_out/r2g_mgpu2_uptile.log 2>&1; echo "mgpu uptile rc=$?"; grep -E "^mgpu_check|FAILED|rank [0-9]+:|Error" gpurun_out/r2g_mgpu2_uptile.log | head -20
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --cpu-steps 0"
for mode in "default" "INK_NO_FOLD=1"; do
  tag=$(echo "$mode" | tr ' =' '__')
  env $( [ "$mode" = default ] || echo $mode ) timeout 300 $T --master-port 29512 bench.py --gpus 2 $F 2>gpurun_out/r2g_$tag.err | tail -1 > gpurun_out/r2g_$tag.json
  python - "$mode" gpurun_out/r2g_$tag.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    c4 = d.get("c4") or {}
    print("%-16s N=2 C3w %.3f ms/step pressure %.3f CG %.1f launches %.0f parity %s | C4s %.3f ms pressure %.2f CG %.1f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], d["launches_per_step"], d.get("parity", {}).get("ok"), c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("pressure", 0), c4.get("cg_iters", 0)))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(sys.argv[2].replace(".json", ".err")).read()[-1500:])
PY
done
G="--steps 3 --warmup 5 --no-e2e --no-kernels --no-c4 --no-parity --no-strict --cpu-steps 0"
INK_TRACE_ITER=5 timeout 200 $T --master-port 29513 bench.py --gpus 2 $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2g_trace_c3_2gpu.txt
