// gsx_common.cuh -- host-side helpers of libgsx.so (error codes, pointer checks, launch helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <nvtx3/nvToolsExt.h>   // header-only: ranges cost a no-op call unless a profiler injected the NVTX library

#include "../../include/gsx.h"

namespace gsx {

void set_error(const char *fmt, ...);

// The reference asserts every pointer is device memory (cuda/checks.cuh:17-38) and exits; the C ABI returns a
// status instead and the C++ compat shim turns it back into exit(EXIT_FAILURE).
int check_device_ptr(const void *p, const char *name);

#define GSX_REQUIRE_DEV(p)                                   \
  do {                                                       \
    int _e = ::gsx::check_device_ptr((p), #p);               \
    if (_e != GSX_OK) return _e;                             \
  } while (0)

// vector accesses on caller-owned arrays need the natural alignment of the vector type (a cudaMalloc'd array has it; a
// slice of a flat buffer at an odd offset does not): rejected with GSX_E_ARG instead of a misaligned-address fault
#define GSX_REQUIRE_ALIGNED(p, bytes)                                                   \
  do {                                                                                  \
    if (((uintptr_t)(p)) & ((bytes)-1)) {                                               \
      ::gsx::set_error("pointer %s must be %d-byte aligned", #p, (int)(bytes));         \
      return GSX_E_ARG;                                                                 \
    }                                                                                   \
  } while (0)

#define GSX_CUDA(call)                                                                            \
  do {                                                                                            \
    cudaError_t _e = (call);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      ::gsx::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e));      \
      return GSX_E_CUDA;                                                                          \
    }                                                                                             \
  } while (0)

#define GSX_LAUNCH_CHECK() GSX_CUDA(cudaGetLastError())

static inline int ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// NVTX range over the host-side enqueue of a stage (tracing); shows up in Nsight Systems / ncu --nvtx
struct NvtxRange {
  explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

constexpr int kNumSMs = 148;  // B200

// grow-only cached device workspace, one per (slot, device), mutex-guarded (binning.cu)
int internal_workspace_slot(int which, size_t bytes, char **out);

// workspace carve-up of the binning stage (binning.cu), shared with the context in api.cu
struct BinLayout {
  size_t touched, offsets, block_sums, total, dk_a, dk_b, dv_a, dv_b, keys_a, keys_b, vals_a, vals_b, sort_ws, sort_ws_bytes, bytes;
};

// Buffers and products of a view's pre-binning (depth sort of the Gaussians that touch a tile + scan of their tile
// counts).  It depends only on the per-Gaussian forward, so the context runs it for the later views of a batch on a side
// stream while earlier views are being composited (api.cu); the per-function F7 path carves one out of its workspace.
struct PreBin {
  uint32_t *key_a = nullptr, *val_a = nullptr;   // ping-pong partners of the depth sort (compact mode: a = the slot's lists)
  uint32_t *key_b = nullptr, *val_b = nullptr;
  int *offsets = nullptr, *block_sums = nullptr;
  int *total = nullptr;      // after the scan: total[0] = K, total[1] = live count (K, or 0 when K exceeds the capacity)
  int *over_flag = nullptr;  // optional mapped host int: receives K when the capacity is exceeded
  char *sort_ws = nullptr;
  size_t sort_ws_bytes = 0;
  const uint32_t *perm = nullptr, *dkeys = nullptr;   // results: ids in depth order and their keys
  const int *count = nullptr;                          // device count of sorted entries (compact mode) or null
};

}  // namespace gsx
