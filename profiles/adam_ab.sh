#!/bin/bash
# Same-box A/B of the fused Adam kernel's build variants (run under gpurun): profiles/adam_ab.sh base b1m4 b2m4 ...
# Variants: make -C 3dgs_b200/csrc BUILD=build_x OUT=../variants/adam_x.so EXTRA="-DGSX_ADAM_BATCH=.. -DGSX_ADAM_MINB=.."
cp 3dgs_b200/libgsx.so /tmp/libgsx_base.so
mkdir -p gpurun_out
: > gpurun_out/adam_ab.jsonl
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so; else cp 3dgs_b200/variants/adam_$v.so 3dgs_b200/libgsx.so; fi
  timeout 120 python -m pytest tests -m gpu -x -q -k "adam or optimizer" > gpurun_out/adam_ab_test_$v.log 2>&1; echo "$v tests rc=$? $(tail -1 gpurun_out/adam_ab_test_$v.log)"
  timeout 120 python profiles/adam_bw.py $v >> gpurun_out/adam_ab.jsonl 2> gpurun_out/adam_ab_$v.err || { echo "$v bw failed"; tail -3 gpurun_out/adam_ab_$v.err; }
  tail -1 gpurun_out/adam_ab.jsonl
done
cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so
