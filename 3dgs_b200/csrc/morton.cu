// morton.cu -- SURVEY 8f rank 4 (first half): Morton re-layout of the Gaussian arrays, the on-device order the splatting
// path reads its parameters in.
//
// Reference: compute_morton_codes, cuda/culling.cu:14-68,345-359 (63-bit codes: 21 bits per axis, quantised against a
// caller-provided bounding box, z|y|x interleaved) and TrainerImpl::sort_gaussians, cuda/trainer.cu:871-924
// (thrust::sequence + thrust::sort_by_key, then one gather per parameter / optimiser array).  Here the codes and the
// identity permutation come out of one kernel, the sort is the library's own onesweep radix sort (stable: equal codes
// keep index order, where thrust's order is unspecified), and a row gather serves every array width.
#include <algorithm>

#include "gsx_common.cuh"

namespace gsx {
namespace {

constexpr uint32_t kBitsPerCoord = 21;                      // culling.cu:15-16
constexpr uint32_t kMaxCoord = (1u << kBitsPerCoord) - 1;   // 2^21 - 1

// Bit spreading.  TRUE = false reproduces the reference's spread_bits mask sequence (culling.cu:18-40) bit for bit: its
// first, third and fourth masks are not the ones of a 3-way interleave, so most coordinate bits are dropped (only bits 0, 2, 3,
// 9, 10 and 12 of a coordinate reach the code) -- SURVEY quirk, kept as the default because the reference's own test and trainer see
// exactly these codes.  TRUE = true is the real interleave (bit b -> bit 3b), opt-in through GSX_MORTON_TRUE_INTERLEAVE.
template <bool TRUE>
__device__ __forceinline__ uint64_t spread3(uint64_t n) {
  n &= kMaxCoord;
  n = (n | (n << 32)) & (TRUE ? 0x1F00000000FFFFull : 0x1F000000FFFFull);
  n = (n | (n << 16)) & 0x1F0000FF0000FFull;
  n = (n | (n << 8)) & (TRUE ? 0x100F00F00F00F00Full : 0x100F807C0F807C0Full);
  n = (n | (n << 4)) & (TRUE ? 0x10C30C30C30C30C3ull : 0x1084210842108421ull);
  n = (n | (n << 2)) & 0x1249249249249249ull;
  return n;
}

template <bool TRUE>
__global__ void __launch_bounds__(256) k_morton_codes(int N, const float *__restrict__ xyz, float x_max, float y_max,
                                                      float z_max, float x_min, float y_min, float z_min,
                                                      uint64_t *__restrict__ codes, uint32_t *__restrict__ iota) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= N) return;
  const size_t b = (size_t)3 * i;
  const float x = xyz[b], y = xyz[b + 1], z = xyz[b + 2];
  // quantisation exactly as culling.cu:55-57: (v - min) * (2097151 / (max - min)), truncated; float -> uint64
  // conversion saturates (negative and NaN give 0)
  const uint64_t xq = __float2ull_rz(__fmul_rn(__fsub_rn(x, x_min), __fdiv_rn((float)kMaxCoord, __fsub_rn(x_max, x_min))));
  const uint64_t yq = __float2ull_rz(__fmul_rn(__fsub_rn(y, y_min), __fdiv_rn((float)kMaxCoord, __fsub_rn(y_max, y_min))));
  const uint64_t zq = __float2ull_rz(__fmul_rn(__fsub_rn(z, z_min), __fdiv_rn((float)kMaxCoord, __fsub_rn(z_max, z_min))));
  codes[i] = (spread3<TRUE>(zq) << 2) | (spread3<TRUE>(yq) << 1) | spread3<TRUE>(xq);
  if (iota) iota[i] = (uint32_t)i;
}

// dst[i][:] = src[order[i]][:], one thread per float; rows are `width` floats
__global__ void __launch_bounds__(256) k_gather_rows(const float *__restrict__ src, const int *__restrict__ order,
                                                     long long total, int width, float *__restrict__ dst) {
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const long long i = e / width;
    const int k = (int)(e - i * width);
    dst[e] = src[(long long)order[i] * width + k];
  }
}

// every array of a Gaussian set in ONE launch (blockIdx.y = array): the reference's sort_gaussians issues 17 gathers
// (cuda/trainer.cu:897-923), each re-reading the permutation
struct GatherArgs { const float *src[24]; float *dst[24]; int width[24]; int n; };
__global__ void __launch_bounds__(256) k_gather_rows_multi(const __grid_constant__ GatherArgs A, const int *__restrict__ order,
                                                           long long N) {
  const float *__restrict__ src = A.src[blockIdx.y];
  float *__restrict__ dst = A.dst[blockIdx.y];
  const int width = A.width[blockIdx.y];
  const long long total = N * width;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total; e += (long long)gridDim.x * 256) {
    const long long i = e / width;
    const int k = (int)(e - i * width);
    dst[e] = src[(long long)order[i] * width + k];
  }
}

}  // namespace
}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_compute_morton_codes(int N, const float *xyz, float x_max, float y_max, float z_max, float x_min, float y_min,
                             float z_min, int flags, uint64_t *codes, void *stream) {
  GSX_REQUIRE_DEV(xyz); GSX_REQUIRE_DEV(codes);
  if (N <= 0) return GSX_OK;
  if (flags & GSX_MORTON_TRUE_INTERLEAVE)
    k_morton_codes<true><<<ceil_div(N, 256), 256, 0, (cudaStream_t)stream>>>(N, xyz, x_max, y_max, z_max, x_min, y_min,
                                                                            z_min, codes, nullptr);
  else
    k_morton_codes<false><<<ceil_div(N, 256), 256, 0, (cudaStream_t)stream>>>(N, xyz, x_max, y_max, z_max, x_min, y_min,
                                                                             z_min, codes, nullptr);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

size_t gsx_morton_order_workspace_bytes(int N) {
  const size_t n = (size_t)(N > 0 ? N : 1);
  return align_up(n * 8, 256) * 2 + align_up(n * 4, 256) + align_up(gsx_sort_workspace_bytes((long long)n), 256);
}

int gsx_morton_order(int N, const float *xyz, float x_max, float y_max, float z_max, float x_min, float y_min,
                     float z_min, int flags, int *order, uint64_t *sorted_codes, void *workspace,
                     size_t workspace_bytes, void *stream) {
  GSX_REQUIRE_DEV(xyz); GSX_REQUIRE_DEV(order); GSX_REQUIRE_DEV(workspace);
  if (sorted_codes) GSX_REQUIRE_DEV(sorted_codes);
  if (N <= 0) return GSX_OK;
  if (workspace_bytes < gsx_morton_order_workspace_bytes(N)) {
    set_error("workspace %zu < %zu bytes", workspace_bytes, gsx_morton_order_workspace_bytes(N));
    return GSX_E_WORKSPACE;
  }
  char *w = reinterpret_cast<char *>(workspace);
  uint64_t *keys_a = reinterpret_cast<uint64_t *>(w); w += align_up((size_t)N * 8, 256);
  uint64_t *keys_b = reinterpret_cast<uint64_t *>(w); w += align_up((size_t)N * 8, 256);
  uint32_t *vals_a = reinterpret_cast<uint32_t *>(w); w += align_up((size_t)N * 4, 256);
  const size_t sort_bytes = gsx_sort_workspace_bytes(N);
  cudaStream_t st = (cudaStream_t)stream;
  if (flags & GSX_MORTON_TRUE_INTERLEAVE)
    k_morton_codes<true><<<ceil_div(N, 256), 256, 0, st>>>(N, xyz, x_max, y_max, z_max, x_min, y_min, z_min, keys_a, vals_a);
  else
    k_morton_codes<false><<<ceil_div(N, 256), 256, 0, st>>>(N, xyz, x_max, y_max, z_max, x_min, y_min, z_min, keys_a, vals_a);
  GSX_LAUNCH_CHECK();
  int rc = gsx_sort_pairs_u64(keys_a, keys_b, vals_a, reinterpret_cast<uint32_t *>(order), N, nullptr, 3 * kBitsPerCoord, w,
                              sort_bytes, stream);
  if (rc != GSX_OK) return rc;
  if (sorted_codes) GSX_CUDA(cudaMemcpyAsync(sorted_codes, keys_b, (size_t)N * 8, cudaMemcpyDeviceToDevice, st));
  return GSX_OK;
}

int gsx_gather_rows(const float *src, const int *order, int N, int width, float *dst, void *stream) {
  GSX_REQUIRE_DEV(src); GSX_REQUIRE_DEV(order); GSX_REQUIRE_DEV(dst);
  if (width <= 0) { set_error("width %d", width); return GSX_E_ARG; }
  if (src == dst) { set_error("gather_rows is out of place"); return GSX_E_ARG; }
  if (N <= 0) return GSX_OK;
  const long long total = (long long)N * width;
  long long blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  k_gather_rows<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(src, order, total, width, dst);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_gather_rows_multi(const float *const *src_host, float *const *dst_host, const int *width_host, int n_arrays,
                          const int *order, int N, void *stream) {
  if (!src_host || !dst_host || !width_host) { set_error("null array table"); return GSX_E_NULL; }
  if (n_arrays < 1 || n_arrays > 24) { set_error("n_arrays %d out of range 1..24", n_arrays); return GSX_E_ARG; }
  GSX_REQUIRE_DEV(order);
  if (N <= 0) return GSX_OK;
  GatherArgs A;
  A.n = n_arrays;
  long long widest = 1;
  for (int a = 0; a < n_arrays; ++a) {
    if (!src_host[a] || !dst_host[a] || src_host[a] == dst_host[a] || width_host[a] <= 0) {
      set_error("array %d: null, aliased or width <= 0 (the gather is out of place)", a);
      return GSX_E_ARG;
    }
    A.src[a] = src_host[a]; A.dst[a] = dst_host[a]; A.width[a] = width_host[a];
    widest = std::max<long long>(widest, width_host[a]);
  }
  long long blocks = ((long long)N * widest + 255) / 256;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  k_gather_rows_multi<<<dim3((unsigned)blocks, (unsigned)n_arrays), 256, 0, (cudaStream_t)stream>>>(A, order, (long long)N);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // extern "C"
