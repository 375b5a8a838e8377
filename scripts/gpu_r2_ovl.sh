# Schwarz overlap A/B on an N-GPU box: mgpu_check with the wider overlap, then C3-weak and C4-strong lines for overlap 8 / 16
N=${1:-4}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
INK_SCHWARZ_OVL=16 MGPU_CASES=2,3 timeout 400 $T --master-port 29511 tests/mgpu_check.py > gpurun_out/r2p_mgpu_ovl16.log 2>&1; echo "mgpu(ovl 16) rc=$?"; grep -E "^mgpu_check|FAILED|rank [0-9]+:" gpurun_out/r2p_mgpu_ovl16.log | head
F="--steps 20 --warmup 5 --no-e2e --no-kernels --no-strict --no-parity --cpu-steps 0"
for o in 8 16; do
  INK_SCHWARZ_OVL=$o timeout 400 $T --master-port 29512 bench.py --gpus $N $F 2>gpurun_out/r2p_$o.err | tail -1 > gpurun_out/r2p_$o.json
  python - $o <<PY
import json, sys
try:
    d = json.loads(open(f"gpurun_out/r2p_{sys.argv[1]}.json").read()); c4 = d["c4"]
    print("ovl %s: C3 weak %.3f ms/step pressure %.3f CG %.1f | C4 strong %.3f ms pressure %.2f CG %.1f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["pressure"], d["cg_iters_per_step"]["mean"], c4["ms_per_step"], c4["phase_ms_last_step"]["pressure"], c4["cg_iters"]))
except Exception as e:
    print("FAILED", e); print(open(f"gpurun_out/r2p_{sys.argv[1]}.err").read()[-1500:])
PY
done
