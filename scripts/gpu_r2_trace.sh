# per-iteration CUDA-event traces (INK_TRACE_ITER) of the PCG on a 2-GPU box: C4 on one GPU, C4 strong-scaled over 2, C3 weak over 2
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
F="--steps 3 --warmup 5 --no-e2e --no-kernels --no-c4 --no-parity --no-strict --cpu-steps 0"
INK_TRACE_ITER=5 timeout 200 python bench.py --config c4 $F 2>&1 >/dev/null | grep trace | tail -2 > gpurun_out/r2b_trace_c4_1gpu.txt
INK_TRACE_ITER=5 timeout 200 $T --master-port 29513 bench.py --gpus 2 --config c4 --strong $F 2>&1 >/dev/null | grep trace | tail -2 > gpurun_out/r2b_trace_c4_2gpu.txt
INK_TRACE_ITER=5 timeout 200 $T --master-port 29514 bench.py --gpus 2 $F 2>&1 >/dev/null | grep trace | tail -2 > gpurun_out/r2b_trace_c3_2gpu.txt
INK_TRACE_ITER=5 timeout 200 python bench.py $F 2>&1 >/dev/null | grep trace | tail -2 > gpurun_out/r2b_trace_c3_1gpu.txt
INK_HOSTTIME=1 timeout 200 python bench.py $F 2>&1 >/dev/null | grep "inkb200 host" | tail -4 > gpurun_out/r2b_hosttime_c3_1gpu.txt
tail -n 3 gpurun_out/r2b_*.txt
