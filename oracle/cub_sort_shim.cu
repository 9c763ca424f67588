// oracle/cub_sort_shim.cu -- extern "C" doorway to cub::DeviceRadixSort::SortPairs of the reference's vendored CCCL
// (/root/reference/submodules/cccl, 3.2.0) -- the one kernel of the reference with an sm_100 tuning (Policy1000,
// cub/device/dispatch/tuning/tuning_radix_sort.cuh:1054-1060; call site cuda/culling.cu:455-469).
//
// TEST INFRASTRUCTURE ONLY: the head-to-head baseline of libgsx.so's hand-written onesweep (profiles/sort_vs_cub.py),
// compiled by oracle/build_cub_sort.sh against the headers WHERE THEY LIE into oracle/_ref/libcubsort.so.  Never linked
// into libgsx.so.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

#define CS_API extern "C" __attribute__((visibility("default")))

template <typename K, typename V>
static int sort_pairs(const K *kin, K *kout, const V *vin, V *vout, long long n, int begin_bit, int end_bit, void *temp,
                      size_t *temp_bytes, cudaStream_t st) {
  size_t need = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, need, kin, kout, vin, vout, n, begin_bit, end_bit, st);
  if (e != cudaSuccess) return (int)e;
  if (temp == nullptr || *temp_bytes < need) { *temp_bytes = need; return temp == nullptr ? 0 : -1; }
  return (int)cub::DeviceRadixSort::SortPairs(temp, need, kin, kout, vin, vout, n, begin_bit, end_bit, st);
}

// temp == NULL: size query (*temp_bytes receives the requirement)
CS_API int cub_sort_pairs_u32(const uint32_t *kin, uint32_t *kout, const uint32_t *vin, uint32_t *vout, long long n,
                              int begin_bit, int end_bit, void *temp, size_t *temp_bytes, void *stream) {
  return sort_pairs<uint32_t, uint32_t>(kin, kout, vin, vout, n, begin_bit, end_bit, temp, temp_bytes, (cudaStream_t)stream);
}
CS_API int cub_sort_pairs_u64(const uint64_t *kin, uint64_t *kout, const uint32_t *vin, uint32_t *vout, long long n,
                              int begin_bit, int end_bit, void *temp, size_t *temp_bytes, void *stream) {
  return sort_pairs<uint64_t, uint32_t>(kin, kout, vin, vout, n, begin_bit, end_bit, temp, temp_bytes, (cudaStream_t)stream);
}
// exactly what the reference calls: double keys, int values, all 64 bits (cuda/culling.cu:455-469)
CS_API int cub_sort_pairs_f64(const double *kin, double *kout, const int *vin, int *vout, long long n, void *temp,
                              size_t *temp_bytes, void *stream) {
  return sort_pairs<double, int>(kin, kout, vin, vout, n, 0, 64, temp, temp_bytes, (cudaStream_t)stream);
}
CS_API int cub_version() { return CUB_VERSION; }
