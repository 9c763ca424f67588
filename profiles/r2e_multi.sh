#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): C++ smoke at world 2, torchrun worker (gsx_dp_* + the Python transports), bench at N.
set -u
N=${1:-2}
TAG=${2:-r2e}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L > $OUT/gpus_$TAG.txt 2>&1
timeout 300 python -m pytest tests/test_dp_native.py -m gpu -q -rs -s > $OUT/pytest_dp_native_$TAG.log 2>&1
echo "dp native tests rc=$?"; grep -E "passed|failed|error|SKIP" $OUT/pytest_dp_native_$TAG.log | tail -4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tests/dp_worker.py > $OUT/dp_worker_${TAG}_n$N.log 2>&1
echo "dp_worker rc=$?"; grep -E "DP_OK|Error|error" $OUT/dp_worker_${TAG}_n$N.log | tail -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${TAG}_n$N.json 2> $OUT/bench_${TAG}_n$N.err
echo "bench N=$N rc=$?"; tail -c 600 $OUT/bench_${TAG}_n$N.json; tail -3 $OUT/bench_${TAG}_n$N.err
GSX_DP_IMPL=python timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 10 --warmup 3 > $OUT/bench_${TAG}_n${N}_pythondp.json 2> $OUT/bench_${TAG}_n${N}_pythondp.err
echo "bench python dp N=$N rc=$?"; tail -c 300 $OUT/bench_${TAG}_n${N}_pythondp.json
