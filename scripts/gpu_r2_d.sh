# K8 tile kernel: parity tests, A/B bench lines, ncu --set full of the tile kernel and of the cluster coarse kernel (one GPU)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tile or golden or phases_bit_exact" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2d_pytest.log
F="--steps 20 --warmup 5 --no-e2e --no-strict --cpu-steps 0"
for mode in "INK_K8_TILE=0" "INK_K8_TILE=1"; do
  tag=$(echo "$mode" | tr ' =' '__')
  env $mode timeout 300 python bench.py $F 2>gpurun_out/r2d_$tag.err | tail -1 > gpurun_out/r2d_$tag.json
  python - "$mode" gpurun_out/r2d_$tag.json <<PY
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read())
    c4 = d.get("c4") or {}
    k8 = d["kernels"].get("k_particles_rk2", {})
    print("%-16s C3 %.3f ms/step particles %.3f (frac %.3f; alone %.3f ms frac %.3f) | C4 %.3f ms particles %.3f" % (sys.argv[1], d["ms_per_step"], d["phase_ms_median"]["particles"], d["roofline"]["per_kernel"][0]["frac"], k8.get("ms", 0), k8.get("frac", 0), c4.get("ms_per_step", 0), c4.get("phase_ms_last_step", {}).get("particles", 0)))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(sys.argv[2].replace(".json", ".err")).read()[-1500:])
PY
done
G="--steps 3 --warmup 3 --no-e2e --no-strict --no-kernels --no-c4 --cpu-steps 0"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_particles_rk2_tile -s 20 -c 1 -o gpurun_out/r2d_k8_tile -f python bench.py $G > gpurun_out/r2d_ncu_k8.log 2>&1; echo "ncu k8 rc=$?"
INK_MG_COARSE=2 timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mg3_coarse -s 40 -c 1 -o gpurun_out/r2d_mg3_cluster -f python bench.py $G > gpurun_out/r2d_ncu_cluster.log 2>&1; echo "ncu cluster rc=$?"
ls -la gpurun_out/*.ncu-rep
