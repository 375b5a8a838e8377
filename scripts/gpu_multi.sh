# usage: gpu_multi.sh N  — multi-GPU parity script, then bench at 1 and N GPUs
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_check.py 2>&1 | grep -E "mgpu_check|Error|error|rank" | tail -20
python bench.py --steps 10 --warmup 5 --cpu-steps 0 > gpurun_out/b1.json 2> gpurun_out/b1.err || tail -3 gpurun_out/b1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/b$N.json 2> gpurun_out/b$N.err || tail -5 gpurun_out/b$N.err
python - <<PY
import json
for f in ("b1","b$N"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, "value %.3e ms/step %.3f iters %s e2e %.3e" % (d["value"], d["ms_per_step"], d["cg_iters_per_step"], d["e2e"]["value"]))
        print("  phases", {k: round(v,3) for k,v in d["phase_ms_last_step"].items()})
    except Exception as e: print(f, "failed", e)
PY
