#!/bin/bash
# Scaling pass (gpurun --gpus N): bench at N with the C++ step (default) and with the Python-orchestrated transports.
set -u
N=${1:-8}
TAG=${2:-r2f}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err
echo "bench N=$N rc=$?"; tail -3 $OUT/bench_${TAG}_n$N.err
GSX_DP_IMPL=python timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${TAG}_n${N}_pythondp.json 2> $OUT/bench_${TAG}_n${N}_pythondp.err
echo "bench python dp N=$N rc=$?"
