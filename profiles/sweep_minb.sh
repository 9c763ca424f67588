#!/bin/bash
# generic occupancy-cap sweep: bash profiles/sweep_minb.sh <file.cu> "<-D flags>" ...  (rebuilds that object on the box, benches each)
set -u
F=$1; shift
cd 3dgs_b200/csrc
for v in "$@"; do
  rm -f build/${F%.cu}.o; make -s EXTRA="$v" >/dev/null 2>&1
  cd ../..
  python bench.py --steps 10 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), {k:v['ms_per_view'] for k,v in d['stages'].items()})"
  cd 3dgs_b200/csrc
done
rm -f build/${F%.cu}.o; make -s >/dev/null 2>&1
