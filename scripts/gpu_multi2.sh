# usage: gpu_multi2.sh N — bench at N GPUs with peer-memory exchanges on and off (INK_NO_P2P=1)
N=${1:-2}
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export INK_NO_P2P=1; else unset INK_NO_P2P; fi
  INK_VERBOSE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 5 --no-e2e > gpurun_out/b${N}_$mode.json 2> gpurun_out/b${N}_$mode.err || tail -5 gpurun_out/b${N}_$mode.err
  grep inkb200 gpurun_out/b${N}_$mode.err | head -2
done
python - <<PY
import json
for f in ("b${N}_p2p","b${N}_nccl"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1])
        print(f, "value %.3e ms/step %.3f iters %s" % (d["value"], d["ms_per_step"], d["cg_iters_per_step"]["mean"]))
        print("  phases", {k: round(v,3) for k,v in d["phase_ms_last_step"].items()})
    except Exception as e: print(f, "failed", e)
PY
