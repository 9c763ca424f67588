#!/bin/bash
# Round-1 closing GPU pass (run under gpurun): new-row tests, bench line, full GPU suite, ncu evidence.
# usage: bash profiles/final_r1.sh <tag>
set -u
TAG=${1:-r1v14}
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python -m pytest tests -m gpu -x -q -k "adam or optimizer or fused_loss or gtests" > $OUT/pytest_adam_$TAG.log 2>&1
echo "adam tests rc=$?"; tail -4 $OUT/pytest_adam_$TAG.log
timeout 400 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 600 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
timeout 600 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; tail -4 $OUT/pytest_gpu_$TAG.log
timeout 600 bash profiles/prof.sh $TAG
echo "prof rc=$?"
timeout 120 python profiles/adam_bw.py $TAG > $OUT/adam_bw_$TAG.json 2> $OUT/adam_bw_$TAG.err; cat $OUT/adam_bw_$TAG.json
