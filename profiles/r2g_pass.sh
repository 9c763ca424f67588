#!/bin/bash
# Round-2 GPU pass g: full suite on HEAD (two-phase backward default, fused tile ranges, graph replay), bench line, A/B of the
# fused ranges, launch list.
set -u
TAG=${1:-r2g}
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -q -rs -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
timeout 700 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 200 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
GSX_RANGES=separate timeout 300 python bench.py --steps 10 --no-cpu-baseline > $OUT/bench_${TAG}_n1_separate_ranges.json 2> $OUT/bench_${TAG}_sr.err; echo "bench separate ranges rc=$?"
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_$TAG.csv \
    python profiles/prof_driver.py 2 2 > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
