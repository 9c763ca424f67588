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
  // record mode (compact mode with the per-Gaussian forward's tile records, binning.cu): the scan gathers each record once
  // and leaves it in depth order; the emission kernel rebuilds the offsets from block_sums itself (no `offsets` array)
  uint2 *srec = nullptr;     // [N] buffer for the records in depth order
  bool rec_mode = false;     // set by the pre-binning when srec / block_sums are what the emission must use
};

// direct tile binning (tilebin.cu): one stable counting sort on the tile id instead of emission + two radix passes
struct TileBinPlan {
  int warps = 0;            // warps per CTA = segments per CTA
  int ctas = 0;             // G
  int bands = 0;            // tile-row bands (grid.y)
  int rows_per_band = 0;    // tile rows per band
  int tb = 0;               // tiles per band (rows_per_band * ntx, rounded up to even)
  size_t smem_count = 0, smem_emit = 0;
  size_t off_W = 0;         // byte offset of the u16 matrix inside the region (the u32 matrix C is at 0)
  size_t off_seg = 0;       // segment starts [rows + 1]
  size_t bytes = 0;
};
bool tilebin_plan(int ntx, int nty, int n_max, long long capacity, size_t region_bytes, TileBinPlan *p);
int tilebin_run(const TileBinPlan &p, const float *uv, int uv_stride, const float4 *radius, const uint32_t *dkeys,
                const uint32_t *perm, const int *offsets, int n_host, const int *n_dev, const int *total, int ntx, int nty,
                int tight, char *region, int *ranges, uint32_t *out_ids, cudaStream_t st, cudaEvent_t *ev);

}  // namespace gsx
