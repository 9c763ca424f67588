#!/bin/bash
# Drop-in proof: compiles the REFERENCE's own gtest files (tests/cuda_forward_test.cpp, cuda_backward_test.cpp,
# optimizer_test.cpp, adaptive_density_test.cpp, from
# /root/reference, against the reference's own headers) and links them with libgsx.so instead of the reference's
# gsplat_cuda library.  googletest 1.13.0 comes from the reference tree (vendored inside submodules/yaml-cpp/test).
# Outputs: oracle/_ref/ref_gtest_forward, oracle/_ref/ref_gtest_backward, oracle/_ref/ref_gtest_optimizer (git-ignored, travel to the GPU box).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=/root/reference
GT=$REF/submodules/yaml-cpp/test/googletest-1.13.0/googletest
[ -d "$REF/tests" ] && [ -d "$GT" ] || { echo "no reference tree / googletest: skipping"; exit 0; }
LIB="$HERE/../3dgs_b200"
OUT="$HERE/_ref"
mkdir -p "$OUT"
if [ -f "$OUT/ref_gtest_density" ] && [ "$OUT/ref_gtest_density" -nt "$LIB/libgsx.so" ]; then echo "[build_ref_gtests] up to date"; exit 0; fi
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
[ -f "$OUT/gtest-all.o" ] || $NVCC -O2 -std=c++17 -ccbin /usr/bin/g++ -I$GT/include -I$GT -c $GT/src/gtest-all.cc -o "$OUT/gtest-all.o"
[ -f "$OUT/gtest_main.o" ] || $NVCC -O2 -std=c++17 -ccbin /usr/bin/g++ -I$GT/include -I$GT -c $GT/src/gtest_main.cc -o "$OUT/gtest_main.o"
CFLAGS="-O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -I$REF/include -I$GT/include -I$GT"
LFLAGS="-gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ --cudart shared -Xlinker -rpath,\$ORIGIN/../../3dgs_b200 -Xlinker -rpath,/usr/local/cuda/lib64 -lpthread -L$LIB -lgsx"
$NVCC $CFLAGS -x cu -c $REF/tests/cuda_forward_test.cpp -o "$OUT/fwd_test.o"
$NVCC $CFLAGS -x cu -c $REF/tests/cuda_backward_test.cpp -o "$OUT/bwd_test.o"
$NVCC $CFLAGS -x cu -c $REF/tests/optimizer_test.cpp -o "$OUT/opt_test.o"
$NVCC $CFLAGS -x cu -c $REF/tests/adaptive_density_test.cpp -o "$OUT/dens_test.o"
$NVCC "$OUT/fwd_test.o" "$OUT/gtest-all.o" "$OUT/gtest_main.o" $LFLAGS -o "$OUT/ref_gtest_forward"
$NVCC "$OUT/bwd_test.o" "$OUT/gtest-all.o" $LFLAGS -o "$OUT/ref_gtest_backward"
$NVCC "$OUT/opt_test.o" "$OUT/gtest-all.o" $LFLAGS -o "$OUT/ref_gtest_optimizer"
$NVCC "$OUT/dens_test.o" "$OUT/gtest-all.o" "$OUT/gtest_main.o" $LFLAGS -o "$OUT/ref_gtest_density"
echo "[build_ref_gtests] built $OUT/ref_gtest_forward $OUT/ref_gtest_backward $OUT/ref_gtest_optimizer $OUT/ref_gtest_density"
