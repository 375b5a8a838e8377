N=${1:-2}
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --config c4 --strong --steps 10 --warmup 5 --no-e2e > gpurun_out/c4s_$N.json 2> gpurun_out/c4s_$N.err || tail -8 gpurun_out/c4s_$N.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/c4s_$N.json").read().strip().splitlines()[-1])
    print("c4 strong N=$N value %.3e ms/step %.3f iters %s scaling %s" % (d["value"], d["ms_per_step"], d["cg_iters_per_step"]["mean"], d["scaling"]))
    print("  phases", {k: round(v,3) for k,v in d["phase_ms_last_step"].items()})
except Exception as e: print("failed", e)
PY
