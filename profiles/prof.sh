#!/bin/bash
# ncu evidence pass (run under gpurun): launch list of one bench step + full captures of the hot kernels.
# usage: bash profiles/prof.sh <tag>
set -u
TAG=${1:-rX}
OUT=gpurun_out
python profiles/prof_driver.py 2 2 > $OUT/prof_plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/prof_plain_$TAG.log; exit 1; }
# second batch only: first batch launches = 2 (batch kernels) + 2 views * per-view launches; skip generously by kernel-id filter
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_$TAG.csv \
    python profiles/prof_driver.py 2 2 > $OUT/ncu_launch_$TAG.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:'k_batch_bwd|k_batch_fwd|k_composite_bwd2|k_composite_fwd2|k_emit|k_onesweep|k_scan_apply|k_histogram|k_depth_keys|k_tile_ranges|k_adam_fused|k_loss_fwd|k_loss_bwd' \
    -s 28 -c 32 -o $OUT/prof_$TAG -f python profiles/prof_driver.py 2 2 > $OUT/ncu_full_$TAG.log 2>&1
tail -3 $OUT/ncu_full_$TAG.log
