#!/bin/bash
# Round-2 GPU pass (run under gpurun): GPU tests, bench line (v3 and v2 compositing), sort vs CUB, micro-benchmarks, ncu.
# usage: bash profiles/r2_pass.sh <tag> [skip_tests]
set -u
TAG=${1:-r2a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi_$TAG.csv 2>&1
if [ "${2:-}" != "skip_tests" ]; then
  timeout 900 python -m pytest tests -m gpu -q -rs -s > $OUT/pytest_gpu_$TAG.log 2>&1
  echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
fi
timeout 600 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 400 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
GSX_COMPOSITE=v2 timeout 300 python bench.py --steps 5 --no-cpu-baseline > $OUT/bench_${TAG}_n1_compositev2.json 2> $OUT/bench_${TAG}_v2.err
echo "bench v2 rc=$?"
timeout 120 profiles/bin/ubench > $OUT/ubench_$TAG.json 2>&1; echo "ubench rc=$?"
timeout 300 python profiles/sort_vs_cub.py C4 10 > $OUT/sort_vs_cub_$TAG.json 2> $OUT/sort_vs_cub_$TAG.err; echo "sort_vs_cub rc=$?"; tail -2 $OUT/sort_vs_cub_$TAG.err
timeout 200 python profiles/hits.py > $OUT/hits_$TAG.json 2> $OUT/hits_$TAG.err; echo "hits rc=$?"
# ncu: launch list, then full captures (+ the counters north_star names) of the hot kernels of the second batch
timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $OUT/launches_$TAG.csv \
    python profiles/prof_driver.py 2 2 > $OUT/ncu_launch_$TAG.log 2>&1
echo "launch list rc=$?"
timeout 300 python profiles/prof_driver.py 2 2 > $OUT/prof_plain2_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on \
    --metrics smsp__inst_executed_pipe_fma.sum,smsp__inst_executed_pipe_fmaheavy.sum,smsp__inst_executed_pipe_fmalite.sum,smsp__inst_executed_pipe_alu.sum,smsp__inst_executed_pipe_xu.sum,smsp__inst_executed_pipe_lsu.sum,smsp__inst_executed_pipe_uniform.sum,smsp__inst_executed_pipe_cbu.sum,smsp__inst_executed_pipe_adu.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed_op_shared_atom.sum,lts__t_sectors_op_red.sum,lts__t_sectors_op_atom.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active \
    -k regex:'k_composite_bwd3|k_composite_fwd3|k_emit|k_onesweep|k_batch_fwd|k_batch_bwd' \
    -s 20 -c 20 -o $OUT/prof_$TAG -f python profiles/prof_driver.py 2 2 > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full_$TAG.log
