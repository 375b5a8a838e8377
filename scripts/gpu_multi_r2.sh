# round-2 multi-GPU check (N = 2, 4 or 8; charged N x):   gpurun --gpus N --timeout 600 -- 'bash scripts/gpu_multi_r2.sh N'
# tests/mgpu_check.py (all three cases; it had only been run at 2 ranks after the shared-memory-limit fix), then the weak-
# scaling bench line and one iteration trace.  Everything under `timeout`: a peer-memory spin that never completes must not
# hang the box.
N=${1:-4}
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 240 $T --master-port 29511 tests/mgpu_check.py > gpurun_out/mgpu_check_$N.log 2>&1; grep -E "^mgpu_check|FAILED|rank [0-9]+:|InkError" gpurun_out/mgpu_check_$N.log | head -20
timeout 200 $T --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/r2_bench_${N}gpu.err | tail -1 > gpurun_out/r2_bench_${N}gpu.json
INK_TRACE_ITER=5 timeout 120 $T --master-port 29513 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e 2>&1 >/dev/null | grep trace | tail -1 > gpurun_out/r2_trace_${N}gpu.txt
python - <<PY
import json
d = json.loads(open("gpurun_out/r2_bench_${N}gpu.json").read())
print("N=%d: %.3f ms/step, %.3g particle-updates/s, CG iters %s" % (d["n_gpus"], d["ms_per_step"], d["value"], d["cg_iters_per_step"]["mean"]))
print({k: round(v, 3) for k, v in d["phase_ms_last_step"].items()})
PY
cat gpurun_out/r2_trace_${N}gpu.txt
