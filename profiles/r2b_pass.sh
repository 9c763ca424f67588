#!/bin/bash
# Round-2 second GPU pass: full GPU suite on the new binning / multi-GPU C ABI / densification, bench line, binning v1 vs v2,
# onesweep tile shapes vs CUB, compute-sanitizer memcheck of the small parity tests (one sanitizer tool per call).
set -u
TAG=${1:-r2b}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -rs -s > $OUT/pytest_gpu_$TAG.log 2>&1
echo "gpu suite rc=$?"; grep -E "passed|failed|error" $OUT/pytest_gpu_$TAG.log | tail -3
timeout 600 python bench.py > $OUT/bench_${TAG}_n1.json 2> $OUT/bench_$TAG.err
echo "bench rc=$?"; tail -c 300 $OUT/bench_${TAG}_n1.json; tail -3 $OUT/bench_$TAG.err
GSX_BIN=v1 timeout 300 python bench.py --steps 5 --no-cpu-baseline > $OUT/bench_${TAG}_n1_binv1.json 2> $OUT/bench_${TAG}_binv1.err
echo "bench bin v1 rc=$?"
for c in 0 1 2 3; do
  GSX_RS_CFG=$c timeout 300 python profiles/sort_vs_cub.py C4 10 > $OUT/sort_vs_cub_${TAG}_cfg$c.json 2> $OUT/sort_vs_cub_${TAG}_cfg$c.err; echo "sort_vs_cub cfg $c rc=$?"
done
for c in 1 2 3; do
  GSX_RS_CFG=$c timeout 300 python bench.py --steps 5 --no-cpu-baseline > $OUT/bench_${TAG}_n1_rscfg$c.json 2> $OUT/bench_${TAG}_rscfg$c.err; echo "bench rs cfg $c rc=$?"
done
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py tests/test_dp_native.py -q -x \
  -k "kats or pointwise or sorted_list_bit_exact or render_forward_parity or render_backward_parity or view_backward_parity or batch_path or capacity_overflow or densify_step or clone_and_split or adam_step_parity or fused_loss_parity or dp_native_world1 or depth_ties" \
  > $OUT/sanitizer_memcheck_$TAG.log 2>&1
echo "memcheck rc=$?"; tail -5 $OUT/sanitizer_memcheck_$TAG.log
