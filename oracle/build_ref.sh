#!/bin/bash
# Builds the REFERENCE's own CUDA kernels for sm_100 from the sources where they lie under /root/reference (build
# container only; nothing is copied into the repo) -> oracle/_ref/libref3dgs.so.  Second oracle + same-box baseline B1.
#
# The two SH files (cuda/spherical_harmonics*.cu) call sphericart's CUDA backend, which JIT-compiles its kernel through
# NVRTC at run time from a string that sphericart's CMake step generates (CMakeLists.txt:12-48 there: the three headers +
# src/sphericart_impl.cu concatenated and wrapped in a raw string literal) plus a generated export header.  Both are
# produced here by the few lines of shell below, into oracle/_ref/sphericart_gen/ (git-ignored build output), so F6/F10
# of the reference are part of the library too.  At run time sphericart dlopen()s libcuda.so / libcudart.so / libnvrtc.so
# by their unversioned names; oracle/ref_cuda.py drops a libcuda.so symlink into oracle/_ref/shlib (on this library's
# RUNPATH) when the driver only ships libcuda.so.1.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference
SP="$REF/submodules/sphericart/sphericart"
[ -d "$REF/cuda" ] || { echo "no reference tree: skipping"; exit 0; }
mkdir -p "$HERE/_ref/shlib"
OUT="$HERE/_ref/libref3dgs.so"
SRCS="projection culling gaussian render projection_backward gaussian_backward render_backward spherical_harmonics spherical_harmonics_backward"
NEWEST=$(ls -t $(for s in $SRCS; do echo $REF/cuda/$s.cu; done) "$HERE/ref_shim.cu" "$HERE/build_ref.sh" | head -1)
if [ -f "$OUT" ] && [ "$OUT" -nt "$NEWEST" ]; then echo "[build_ref] up to date"; exit 0; fi
GEN="$HERE/_ref/sphericart_gen"
mkdir -p "$GEN/generated" "$GEN/include/sphericart"
{ printf 'R"======(\n'
  for h in include/sphericart_impl.cuh include/templates_core.hpp include/macros.hpp; do printf '\n'; cat "$SP/$h"; done
  printf '\n'; cat "$SP/src/sphericart_impl.cu"; printf '\n)======"'; } > "$GEN/generated/wrapped_sphericart_impl.cu"
cat > "$GEN/include/sphericart/exports.h" <<'H'
#ifndef SPHERICART_EXPORT_H
#define SPHERICART_EXPORT_H
#define SPHERICART_EXPORT __attribute__((visibility("default")))
#define SPHERICART_NO_EXPORT __attribute__((visibility("hidden")))
#endif
H
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -O3 -std=c++17 -gencode arch=compute_100,code=sm_100 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda \
  -ccbin /usr/bin/g++ -I"$REF/include" -I"$REF/cuda" -I"$SP/include" -I"$GEN" -I"$GEN/include" -shared --cudart shared \
  -Xlinker -rpath,/usr/local/cuda/lib64 -Xlinker -rpath,'$ORIGIN/shlib' -Xlinker --enable-new-dtags \
  -o "$OUT" $(for s in $SRCS; do echo $REF/cuda/$s.cu; done) "$SP/src/cuda_base.cpp" "$SP/src/sphericart_cuda.cpp" \
  "$HERE/ref_shim.cu" -ldl
echo "[build_ref] built $OUT"
