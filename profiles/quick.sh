#!/bin/bash
# quick GPU check: list tests + short bench (stage times) [+ extra env as arguments]
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sorted_list or tile_binning or view_forward_chain or depth_ties or batch_path or empty_and_error or capacity_overflow or contributing" 2>&1 | tail -3
timeout 300 python bench.py --steps 10 --no-cpu-baseline > $OUT/bench_quick.json 2> $OUT/bench_quick.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("$OUT/bench_quick.json"))
print(d["value"], d["ms_per_view_per_gpu"], d["e2e"]["value"], {k:v["ms_per_view"] for k,v in d["stages"].items()})
PY
