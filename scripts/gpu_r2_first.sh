# round-2 opening pass on a 2-GPU box:  gpurun --gpus 2 --timeout 1500 -- 'bash scripts/gpu_r2_first.sh'
# GPU test suite (includes tests/mgpu_check.py at 2 ranks), smoke, the 1-GPU bench line and the 2-GPU one.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2a_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py --steps 20 --warmup 5 2>gpurun_out/r2a_bench_c3.err | tail -1 > gpurun_out/r2a_bench_c3.json; echo "bench1 rc=$?"; tail -3 gpurun_out/r2a_bench_c3.err
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $T --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 2>gpurun_out/r2a_bench_2gpu.err | tail -1 > gpurun_out/r2a_bench_2gpu.json; echo "bench2 rc=$?"; tail -3 gpurun_out/r2a_bench_2gpu.err
python - <<PY
import json
for f in ("r2a_bench_c3", "r2a_bench_2gpu"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read())
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f, "N=%d: %.3f ms/step, %.3g upd/s, CG %.1f, launches/step %.0f" % (d["n_gpus"], d["ms_per_step"], d["value"], d["cg_iters_per_step"]["mean"], d["launches_per_step"]))
    print("  phases", {k: round(v, 3) for k, v in d["phase_ms_median"].items()})
    print("  e2e", d["e2e"] and round(d["e2e"]["ms_per_step"], 3), "roofline", round(d["roofline"]["frac"], 3), [(k["kernel"][:20], round(k["frac"], 3)) for k in d["roofline"]["per_kernel"]])
    print("  c4", d.get("c4") and (round(d["c4"]["ms_per_step"], 3), d["c4"]["cg_iters"], {k: round(v, 2) for k, v in d["c4"]["phase_ms_last_step"].items()}))
    print("  strict", d.get("strict") and round(d["strict"]["ms_per_step"], 3), "parity", d.get("parity"))
    print("  cpu", d.get("cpu_baseline") and (round(d["cpu_baseline"]["ms_per_step"], 1), d["cpu_baseline"]["cores"]))
PY
