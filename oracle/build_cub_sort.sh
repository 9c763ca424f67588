#!/bin/bash
# Builds oracle/_ref/libcubsort.so: cub::DeviceRadixSort::SortPairs from the reference's vendored CCCL (headers used where
# they lie under /root/reference/submodules/cccl; build container only).  Baseline of profiles/sort_vs_cub.py.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
CCCL=/root/reference/submodules/cccl
[ -d "$CCCL/cub" ] || { echo "no reference CCCL tree: skipping"; exit 0; }
mkdir -p "$HERE/_ref"
OUT="$HERE/_ref/libcubsort.so"
if [ -f "$OUT" ] && [ "$OUT" -nt "$HERE/cub_sort_shim.cu" ] && [ "$OUT" -nt "$HERE/build_cub_sort.sh" ]; then echo "[build_cub_sort] up to date"; exit 0; fi
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
# sm_100 (not 100a): the arch the reference's CMake "native" setting produces on a B200, so CUB dispatches Policy1000
$NVCC -O3 -std=c++17 -gencode arch=compute_100,code=sm_100 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda \
  -ccbin /usr/bin/g++ -I"$CCCL/cub" -I"$CCCL/thrust" -I"$CCCL/libcudacxx/include" -shared --cudart shared \
  -Xlinker -rpath,/usr/local/cuda/lib64 -o "$OUT" "$HERE/cub_sort_shim.cu"
echo "[build_cub_sort] built $OUT"
