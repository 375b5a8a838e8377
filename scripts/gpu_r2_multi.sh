# multi-GPU evidence pass on an N-GPU box (N = 2, 4, 8; charged N x):   gpurun --gpus N --timeout 1500 -- 'bash scripts/gpu_r2_multi.sh N [c5]'
# tests/mgpu_check.py at N ranks (all cases, then the vector case with the tile up kernel forced), the weak-scaling bench line with its
# nested parity / C4-strong blocks, one iteration trace; with "c5" also BASELINE configs[4] (125 M particles per GPU on a frozen 512^3 field).
# Everything under `timeout`: a peer-memory spin that never completes must not hang the box.
N=${1:-4}
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 800 $T --master-port 29511 tests/mgpu_check.py > gpurun_out/r2_mgpu_$N.log 2>&1; echo "mgpu rc=$?"; grep -E "^mgpu_check|FAILED|rank [0-9]+:|Error" gpurun_out/r2_mgpu_$N.log | head -20
INK_MG3_UP_TILE=2 MGPU_CASES=3 timeout 300 $T --master-port 29515 tests/mgpu_check.py > gpurun_out/r2_mgpu_${N}_uptile.log 2>&1; echo "mgpu uptile rc=$?"; grep -E "^mgpu_check|FAILED|rank [0-9]+:|Error" gpurun_out/r2_mgpu_${N}_uptile.log | head -8
timeout 500 $T --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 2>gpurun_out/r2_bench_${N}gpu.err | tail -1 > gpurun_out/r2_bench_${N}gpu.json; echo "bench rc=$?"
python - $N <<PY
import json, sys
N = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/r2_bench_{N}gpu.json").read())
    c4 = d.get("c4") or {}
    print("N=%s C3 weak %.3f ms/step %.3g upd/s CG %.1f e2e %.3f ms (copy-only %.3f ms, %.0f GB/s) parity %s" % (N, d["ms_per_step"], d["value"], d["cg_iters_per_step"]["mean"], d["e2e"]["ms_per_step"], d["e2e"].get("copy_only_ms_per_step", 0), d["e2e"].get("copy_only_aggregate_gbs", 0), d.get("parity", {}).get("ok")))
    print("  phases", {k: round(v, 3) for k, v in d["phase_ms_median"].items()})
    print("  C4 strong %.3f ms/step CG %.1f" % (c4.get("ms_per_step", 0), c4.get("cg_iters", 0)), {k: round(v, 2) for k, v in c4.get("phase_ms_last_step", {}).items()})
    print("  parity", d.get("parity"))
except Exception as e:
    print("bench line unreadable:", e); print(open(f"gpurun_out/r2_bench_{N}gpu.err").read()[-2000:])
PY
G="--steps 3 --warmup 5 --no-e2e --no-kernels --no-c4 --no-parity --no-strict --cpu-steps 0"
INK_TRACE_ITER=5 timeout 200 $T --master-port 29513 bench.py --gpus $N $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2_trace_c3_${N}gpu.txt
INK_TRACE_ITER=5 timeout 300 $T --master-port 29514 bench.py --gpus $N --config c4 --strong $G 2>&1 >/dev/null | grep trace | tail -1 | tee gpurun_out/r2_trace_c4_${N}gpu.txt
if [ "$2" = "c5" ]; then
  timeout 500 $T --master-port 29516 bench.py --gpus $N --config c5 --steps 20 --warmup 5 2>gpurun_out/r2_bench_c5_${N}gpu.err | tail -1 > gpurun_out/r2_bench_c5_${N}gpu.json; echo "c5 rc=$?"
  python -c "
import json; d=json.loads(open('gpurun_out/r2_bench_c5_${N}gpu.json').read()); print('C5 N=%d: %.3f ms/sweep %.3g upd/s frac %.3f (K8 only %.3f ms frac %.3f)' % (d['n_gpus'], d['ms_per_step'], d['value'], d['roofline']['frac'], d['roofline']['k8_only']['ms'], d['roofline']['k8_only']['frac']))" || tail -5 gpurun_out/r2_bench_c5_${N}gpu.err
fi
nvidia-smi topo -m > gpurun_out/r2_topo_${N}gpu.txt 2>&1
