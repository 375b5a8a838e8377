N=${1:-2}
export INK_TRACE_ITER=5 INK_VERBOSE=1
python bench.py --steps 3 --warmup 3 --cpu-steps 0 --no-e2e --no-kernels 2>&1 >/dev/null | grep trace | tail -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e 2>&1 >/dev/null | grep -E "trace|inkb200\]" | tail -3
INK_NO_P2P=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-e2e 2>&1 >/dev/null | grep -E "trace|inkb200\]" | tail -3
