#!/bin/bash
# A/B of build variants on ONE box (run under gpurun): profiles/ab.sh <variant.so> ... ; "base" = the in-tree libgsx.so.
# Variants are built here beforehand:  make -C 3dgs_b200/csrc BUILD=build_x OUT=../variants/x.so EXTRA=-DGSX_...=...
cp 3dgs_b200/libgsx.so /tmp/libgsx_base.so
for rep in 1 2; do
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so; else cp 3dgs_b200/variants/$v.so 3dgs_b200/libgsx.so; fi
  timeout 300 python bench.py --no-cpu-baseline > gpurun_out/ab.log 2> gpurun_out/ab.err || { echo "$v failed"; tail -3 gpurun_out/ab.err; continue; }
  python -c "
import json
d=json.loads(open('gpurun_out/ab.log').read().strip().splitlines()[-1])
print('$v', round(d['ms_per_view_per_gpu'],4), {k:v['ms_per_view'] for k,v in d['stages'].items()})
"
done; done
cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so
