python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -8 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 5 --cpu-steps 0 --no-kernels > gpurun_out/b_stream.json 2> gpurun_out/b_stream.err || tail -5 gpurun_out/b_stream.err
python bench.py --steps 10 --warmup 5 --cpu-steps 0 --no-kernels --e2e-blocking > gpurun_out/b_block.json 2> gpurun_out/b_block.err || tail -5 gpurun_out/b_block.err
python - <<'PY'
import json
for f in ("b_stream", "b_block"):
    try:
        d = json.load(open("gpurun_out/" + f + ".json"))
        print(f, "ms/step", round(d["ms_per_step"], 3), "e2e", d["e2e"])
    except Exception as e:
        print(f, "failed", e)
PY
