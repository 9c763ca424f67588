#!/bin/bash
# Round-2 GPU pass h: the no-allocation entry point (gsx_fused_loss_static) -- its test alone in a fresh process, then the
# whole GPU suite with it collected last, then smoke and the bench line on the same build.
set -u
TAG=${1:-r2h}
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rs > $OUT/pytest_no_alloc_alone_$TAG.log 2>&1
echo "no-alloc test alone rc=$?"; tail -3 $OUT/pytest_no_alloc_alone_$TAG.log
timeout 1500 python -m pytest tests -x -m gpu -q -rs > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
timeout 700 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 300 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
