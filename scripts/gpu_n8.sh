N=${1:-8}
INK_VERBOSE=1 timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/b${N}.json 2> gpurun_out/b${N}.err || tail -8 gpurun_out/b${N}.err
grep inkb200 gpurun_out/b${N}.err | head -2
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/b${N}.json").read().strip().splitlines()[-1])
    print("b${N} value %.3e ms/step %.3f iters %s e2e %s" % (d["value"], d["ms_per_step"], d["cg_iters_per_step"]["mean"], d["e2e"] and d["e2e"]["ms_per_step"]))
    print("  phases", {k: round(v,3) for k,v in d["phase_ms_last_step"].items()})
except Exception as e: print("failed", e)
PY
