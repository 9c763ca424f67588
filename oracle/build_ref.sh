#!/bin/bash
# Builds the REFERENCE's own CUDA kernels for sm_100 from the sources where they lie under /root/reference (build
# container only; nothing is copied into the repo) -> oracle/_ref/libref3dgs.so.  Second oracle + same-box baseline.
# The two SH files are left out: they need sphericart's CMake-generated CUDA source + NVRTC.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference
[ -d "$REF/cuda" ] || { echo "no reference tree: skipping"; exit 0; }
mkdir -p "$HERE/_ref"
OUT="$HERE/_ref/libref3dgs.so"
SRCS="projection culling gaussian render projection_backward gaussian_backward render_backward"
NEWEST=$(ls -t $(for s in $SRCS; do echo $REF/cuda/$s.cu; done) "$HERE/ref_shim.cu" | head -1)
if [ -f "$OUT" ] && [ "$OUT" -nt "$NEWEST" ]; then echo "[build_ref] up to date"; exit 0; fi
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -O3 -std=c++17 -gencode arch=compute_100,code=sm_100 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda \
  -ccbin /usr/bin/g++ -I"$REF/include" -I"$REF/cuda" -shared --cudart shared -Xlinker -rpath,/usr/local/cuda/lib64 \
  -o "$OUT" $(for s in $SRCS; do echo $REF/cuda/$s.cu; done) "$HERE/ref_shim.cu"
echo "[build_ref] built $OUT"
