#!/bin/bash
# Round-2 third GPU pass: suite on the current build, single-warp-CTA compositing (GSX_COMPOSITE_CTA=32) parity + A/B,
# occupancy variants (3dgs_b200/variants/*.so), fused tile-sort histogram A/B, sort vs CUB, ncu of both CTA shapes.
set -u
TAG=${1:-r2c}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -rs -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
GSX_COMPOSITE_CTA=32 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "view_ or batch_path or config_full or contributing or config1 or host_buffer" > $OUT/pytest_cta32_$TAG.log 2>&1
echo "cta32 tests rc=$?"; grep -E "passed|failed|error" $OUT/pytest_cta32_$TAG.log | tail -3
timeout 600 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 200 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
short() { timeout 300 python bench.py --steps 8 --no-cpu-baseline > $OUT/bench_${TAG}_n1_$1.json 2> $OUT/bench_${TAG}_$1.err; echo "bench $1 rc=$?"; }
GSX_COMPOSITE_CTA=32 short cta32
GSX_NO_FUSED_HIST=1 short nofusedhist
cp 3dgs_b200/libgsx.so /tmp/libgsx_base.so
for v in m9 m10; do cp 3dgs_b200/variants/$v.so 3dgs_b200/libgsx.so; short $v; GSX_COMPOSITE_CTA=32 short ${v}_cta32; done
cp /tmp/libgsx_base.so 3dgs_b200/libgsx.so
timeout 300 python profiles/sort_vs_cub.py C4 10 > $OUT/sort_vs_cub_$TAG.json 2> $OUT/sort_vs_cub_$TAG.err; echo "sort_vs_cub rc=$?"
timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
    --metrics smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_fmaheavy.sum,smsp__inst_executed_pipe_fmalite.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,smsp__inst_executed_pipe_uniform.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed_op_shared_atom.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active \
    -k regex:'k_composite_bwd3|k_composite_fwd3|k_emit|k_onesweep|k_batch_fwd|k_batch_bwd' \
    -s 20 -c 20 -o $OUT/prof_$TAG -f python profiles/prof_driver.py 2 2 > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
GSX_COMPOSITE_CTA=32 timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_cta32_$TAG.log 2>&1 && \
GSX_COMPOSITE_CTA=32 timeout 900 ncu --set full --clock-control none --import-source on \
    -k regex:'k_composite_bwd3|k_composite_fwd3' -s 4 -c 4 -o $OUT/prof_cta32_$TAG -f python profiles/prof_driver.py 2 2 > $OUT/ncu_full_cta32_$TAG.log 2>&1
echo "ncu cta32 rc=$?"
