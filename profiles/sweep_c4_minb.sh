#!/bin/bash
# occupancy cap of the two-phase compositing backward (GSX_C4_BWD_MINB) and of the forward (GSX_C3_FWD_MINB): rebuild on the box, bench each
set -u
cd 3dgs_b200/csrc
for v in "$@"; do
  rm -f build/composite3.o; make -s EXTRA="$v" >/dev/null 2>&1
  cuobjdump -res-usage build/composite3.o 2>/dev/null | grep -A1 "k_composite_bwd4\|k_composite_fwd3ILi128" | grep -o "REG:[0-9]* STACK:[0-9]* SHARED:[0-9]*" | tr '\n' ' '
  cd ../..
  python bench.py --steps 10 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', round(d['value'],1), {k:v['ms_per_view'] for k,v in d['stages'].items() if k.startswith('composite')})"
  cd 3dgs_b200/csrc
done
rm -f build/composite3.o; make -s >/dev/null 2>&1
