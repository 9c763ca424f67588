#!/bin/bash
# Same-box A/B of the loss kernels' build variants (run under gpurun): profiles/loss_ab.sh base loss_reg
cp 3dgs_b200/libgsx.so /tmp/libgsx_base.so
mkdir -p gpurun_out
: > gpurun_out/loss_ab.jsonl
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so; else cp 3dgs_b200/variants/$v.so 3dgs_b200/libgsx.so; fi
  timeout 100 python -m pytest tests -m gpu -x -q -k "fused_loss or gtests" > gpurun_out/loss_ab_test_$v.log 2>&1; echo "$v tests rc=$? $(tail -1 gpurun_out/loss_ab_test_$v.log)"
  timeout 60 python profiles/loss_bw.py $v >> gpurun_out/loss_ab.jsonl 2> gpurun_out/loss_ab_$v.err || { echo "$v bw failed"; tail -3 gpurun_out/loss_ab_$v.err; }
  tail -1 gpurun_out/loss_ab.jsonl
done
cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so
