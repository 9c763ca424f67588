#!/bin/bash
# Round-2 fourth GPU pass: two-phase compositing backward (GSX_COMPOSITE_BWD=two_phase): parity subset, bench A/B, ncu.
set -u
TAG=${1:-r2d}
OUT=gpurun_out
mkdir -p $OUT
GSX_COMPOSITE_BWD=two_phase timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_dp_native.py -m gpu -q -s -x \
  -k "view_ or batch_path or config_full or contributing or config1 or host_buffer or capacity or pipeline or dp_native" > $OUT/pytest_twophase_$TAG.log 2>&1
echo "two-phase tests rc=$?"; grep -E "passed|failed|error" $OUT/pytest_twophase_$TAG.log | tail -3
short() { timeout 300 python bench.py --steps 10 --no-cpu-baseline > $OUT/bench_${TAG}_n1_$1.json 2> $OUT/bench_${TAG}_$1.err; echo "bench $1 rc=$?"; }
short base
GSX_COMPOSITE_BWD=two_phase short twophase
short base2
GSX_COMPOSITE_BWD=two_phase short twophase2
GSX_COMPOSITE_BWD=two_phase timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_twophase_$TAG.log 2>&1 && \
GSX_COMPOSITE_BWD=two_phase timeout 900 ncu --set full --clock-control none --import-source on \
    --metrics smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,lts__t_sectors_op_red.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active \
    -k regex:'k_composite_bwd4' -s 2 -c 2 -o $OUT/prof_twophase_$TAG -f python profiles/prof_driver.py 2 2 > $OUT/ncu_full_twophase_$TAG.log 2>&1
echo "ncu rc=$?"
