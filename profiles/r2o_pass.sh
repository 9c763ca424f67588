#!/bin/bash
# Round-2 GPU pass o: final build of the round (occupancy caps re-swept) (tile records, staged emission with fused tile-sort histograms, 16-bit keys, onesweep ranking through shared memory)
# -- full suite, bench line, A/B without the records, launch list.
set -u
TAG=${1:-r2o}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -x -m gpu -q -rs > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
timeout 700 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; python - <<PY
import json
d=json.load(open("$OUT/bench_${TAG}_n1.json"))
print(d["value"], d["ms_per_view_per_gpu"], d["e2e"]["value"], {k:v["ms_per_view"] for k,v in d["stages"].items()}, d["parity_c4"]["ok"])
PY
tail -3 $OUT/bench_$TAG.err
GSX_TILE_LISTS=0 timeout 300 python bench.py --steps 10 --no-cpu-baseline > $OUT/bench_${TAG}_n1_no_tile_records.json 2> $OUT/bench_${TAG}_nl.err; echo "bench no-records rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_${TAG}_n1_no_tile_records.json"))
print(d["value"], d["ms_per_view_per_gpu"], {k:v["ms_per_view"] for k,v in d["stages"].items()})
PY
timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_$TAG.csv \
    python profiles/prof_driver.py 2 2 > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
