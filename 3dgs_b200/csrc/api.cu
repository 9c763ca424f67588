// api.cu -- host side of libgsx.so: error plumbing, the persistent per-device context that replaces the
// reference's ForwardPassData (include/gsplat_cuda/cuda_data.cuh:70-86, re-allocated every view there), and the
// view-level entry points that replace rasterize_image (cuda/raster.cu:12-136) and backward_pass
// (cuda/trainer.cu:926-1015).  No host<->device round trip happens inside gsx_forward/gsx_backward: the instance
// count K stays on the device and every kernel downstream of the scan reads it from there.
#include <algorithm>
#include <stdarg.h>
#include <stdlib.h>

#include "gsx_common.cuh"
#include "gsx_math.cuh"
#include "gsx_view.cuh"

namespace gsx {

static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_device_ptr(const void *p, const char *name) {
  if (p == nullptr) { set_error("pointer %s is null", name); return GSX_E_NULL; }
  cudaPointerAttributes a;
  cudaError_t e = cudaPointerGetAttributes(&a, p);
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("cudaPointerGetAttributes(%s) -> %s", name, cudaGetErrorString(e));
    return GSX_E_NOTDEV;
  }
  if (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged) {
    set_error("pointer %s is not a device pointer", name);
    return GSX_E_NOTDEV;
  }
  return GSX_OK;
}

// from binning.cu / composite.cu / batch.cu
BinLayout bin_layout(int N, long long capacity);
int bin_prebin_impl(const float *depth, int depth_stride, const int *touched, int N, bool compact, const int *ccount,
                    PreBin *pb, cudaStream_t st, long long capacity, const uint2 *trec);
int bin_emit_sort_impl(const float *uv, int uv_stride, const float4 *radius, const int *touched, int ntx, int nty, int N,
                       long long capacity, char *ws, const BinLayout &L, const PreBin &pb, int *ranges,
                       const uint32_t **sorted_tiles, const uint32_t **sorted_vals, cudaStream_t st, cudaEvent_t *ev,
                       int tight, const uint2 *trec, const uint16_t *tlist);
size_t prebin_sort_ws_bytes(int N);
size_t prebin_block_sums_count(int N);
int batch_fwd_launch(const BatchConst &bc, const gsx_params_t &p, int N, int l_max, cudaStream_t st);
int batch_bwd_launch(const BatchConst &bc, const gsx_params_t &p, int N, int l_max, const gsx_grads_t &g, int flags,
                     cudaStream_t st);
int extract_view_rgb_launch(const float4 *gacc, const int *touched, int N, float *drgb, cudaStream_t st);
int extract_view_rgb_peers_launch(const float4 *gacc, const int *touched, const float *src_dense, int N, float *mc,
                                  float *const *peers, int n_peers, cudaStream_t st);
int allreduce_multimem_launch(float *mc, size_t n_floats, int rank, int world, cudaStream_t st);
int sh_from_view_rgb_launch(const float *xyz, int N, int l_max, const float *campos_host, int n_views, const float *drgb,
                            size_t view_stride, float *sh_grad, int sh_stride, float *band0_grad, int flags,
                            cudaStream_t st);
int composite_fwd_packed(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges,
                         int W, int H, float bg, int *spp, float *wpp, float *image, cudaStream_t st, const int *order);
int tile_order_launch(const int *ranges, int T, int *order, cudaStream_t st);
int composite_bwd_packed(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges,
                         int W, int H, float bg, const int *spp, const float *wpp, const float *grad_image,
                         float *packed_grads, cudaStream_t st, const int *order);
int count_work_launch(const float4 *r0, const float4 *r1, const int *sorted, const int *ranges, int W, int H,
                      const int *spp, unsigned long long *out_dev, cudaStream_t st);
int composite_hits_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, const int *spp, const float *wpp, const float *grad_image, unsigned long long *out_dev,
                          cudaStream_t st);

}  // namespace gsx

using namespace gsx;

struct gsx_slot {          // per-view state that must outlive the view inside a batch
  SlotPtrs p{};
  gsx_view_t view{};
  bool allocated = false;
  // pre-binning (depth sort + scan) of the view: own buffers so that it can run ahead on the side stream
  PreBin pb{};
  cudaEvent_t ev_pre = nullptr;
  bool pre_async = false;     // issued on the side stream by gsx_batch_forward_params; ev_pre marks its end
  cudaEvent_t ev_bin = nullptr;
  bool bin_async = false;     // emission + tile sort + ranges issued ahead on the binning stream; ev_bin marks their end
};

// Instance buffers + ranges of one binned view.  Two sets: view k uses set k & 1, so that view k+1 can be binned on the
// binning stream while view k is composited out of the other set.
struct BinSet {
  char *ws = nullptr;
  int *ranges = nullptr;
  const uint32_t *sorted_tiles = nullptr, *sorted_vals = nullptr;
  int owner = -1;             // slot whose sorted list the set holds (-1: none)
};

struct gsx_ctx {
  int max_n = 0, max_w = 0, max_h = 0;
  long long capacity = 0;
  gsx_slot slot[GSX_MAX_BATCH];   // slot 0 is the single-view path's state
  int n_slots = 0, cur = 0;       // cur: slot of the last forward
  // binning workspace (depends on capacity), shared by the slots: a view's sorted list lives until the next forward
  char *bin_ws = nullptr;
  BinLayout L{};
  int *ranges = nullptr;
  // per-pixel state of the last forward
  float *image = nullptr, *wpp = nullptr, *grad_image = nullptr;
  int *spp = nullptr;
  const uint32_t *sorted_tiles = nullptr;
  uint64_t *key64 = nullptr;        // tile<<32 | depth bits, rebuilt on demand by gsx_ctx_state (inspection only)
  long long key64_cap = 0;
  const uint32_t *sorted_vals = nullptr;
  int last_n = 0, last_w = 0, last_h = 0;
  int batch_n = 0, batch_nv = 0, batch_lmax = 0;
  int *h_stats = nullptr;           // pinned: [0] = K, [1] = P, [4 + k] = K of slot k
  int *h_over = nullptr;            // pinned + mapped: largest K that exceeded the capacity since the last reserve (0: none)
  int *d_P = nullptr;
  unsigned long long *d_work = nullptr;   // gsx_ctx_count_work's two counters
  int *d_ccount = nullptr;          // [GSX_MAX_BATCH] compacted-Gaussian counters of the slots, [GSX_MAX_BATCH] = tile-list cursor
  uint16_t *tlist = nullptr;        // tile ids of every (view, Gaussian, tile) instance of the current batch, as the per-Gaussian
  unsigned tl_cap = 0;              // forward enumerated them (batch.cu); tl_cap entries, re-sized with the instance capacity
  // optional stage timing (gsx_ctx_enable_timing): 0,1 per-Gaussian forward; 2 view start, 3 emit, 4 sort, 5 ranges,
  // 6 compositing forward; 7,8 compositing backward; 9,10 per-Gaussian backward
  bool timing = false;
  cudaEvent_t ev[11] = {};
  // host pipeline of gsx_batch_render_host: the gradient images of a batch are uploaded on their own stream while the
  // per-Gaussian forward runs, rendered images are downloaded on a third stream under the following backward
  cudaStream_t cs_pre = nullptr;     // pre-binning of the later views of a batch, under the earlier views' compositing
  cudaStream_t cs_bin = nullptr;     // emission + tile sort + ranges of view k+1 under the compositing of view k
  cudaEvent_t ev_fwd = nullptr, ev_free = nullptr;
  BinSet bset[2];                    // bin_ws / ranges / sorted_* below point into the current view's set
  int n_px_tiles = 0;
  const int *cur_total = nullptr;    // device K of the last forward (the slot's PreBin::total)
  cudaStream_t cs_in = nullptr, cs_out = nullptr;
  cudaEvent_t ev_in[GSX_MAX_BATCH] = {}, ev_img = nullptr, ev_out = nullptr, ev_go = nullptr;
  float *grad_stage[GSX_MAX_BATCH] = {};
};

namespace gsx {
static int prebin(gsx_ctx *c, gsx_slot &s, int N, cudaStream_t st);
// the 64-bit keys the two-level sort never materialises, for inspection: tile << 32 | order-preserving depth bits
__global__ void k_compose_keys(const uint32_t *__restrict__ tiles, const uint32_t *__restrict__ ids,
                               const float *__restrict__ depth, const int *__restrict__ n_dev, long long cap,
                               uint64_t *__restrict__ out) {
  const long long n = min((long long)*n_dev, cap);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = ((uint64_t)tiles[i] << 32) | float_sort_bits(depth[ids[i]]);
}

// same when the per-instance tile ids do not exist (direct tile binning): one warp per tile walks its range
__global__ void k_compose_keys_ranges(const int *__restrict__ ranges, int n_tiles, const uint32_t *__restrict__ ids,
                                      const float *__restrict__ depth, uint64_t *__restrict__ out) {
  const int lane = threadIdx.x & 31, warps = (gridDim.x * blockDim.x) >> 5;
  for (int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < n_tiles; t += warps)
    for (int i = ranges[t] + lane; i < ranges[t + 1]; i += 32) out[i] = ((uint64_t)t << 32) | float_sort_bits(depth[ids[i]]);
}

__global__ void k_count_mask(const uint8_t *__restrict__ mask, int N, int *__restrict__ out) {
  int c = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < N; i += gridDim.x * blockDim.x) c += mask[i] ? 1 : 0;
  for (int off = 16; off > 0; off >>= 1) c += __shfl_down_sync(0xffffffffu, c, off);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}

// GSX_TILE_ORDER=1: the compositing grids take the tiles longest list first.  Measured on B200 at C4 (r2k): no gain -- forward
// 0.260 -> 0.271 ms (incl. the 0.01 ms ordering kernel), backward 0.486 -> 0.488 ms; in tile order the long lists already sit
// in the middle of the grid and neighbouring tiles share their records in L2.  Off by default.
static bool tile_order_enabled() { static const bool v = [] { const char *e = getenv("GSX_TILE_ORDER"); return e && e[0] == '1'; }(); return v; }
static const int *tile_order_of(const gsx_ctx *c);
static bool lists_enabled() { const char *e = getenv("GSX_TILE_LISTS"); return !(e && e[0] == '0'); }

// The batch's tile-list buffer (tile ids of the rectangles of more than COOP_CELLS cells; each reserves its cell count):
// room for the instance capacity of every reserved slot.  A rectangle that does not fit gets no list and the emission
// falls back to the SAT for it.  GSX_TILE_LISTS=0: the emission kernel enumerates everything itself (round-1 behaviour).
static int ensure_tile_lists(gsx_ctx *c) {
  const bool off = !lists_enabled() || c->n_px_tiles > 65536;   // 16-bit tile ids
  unsigned long long want = off ? 0ull : std::min<unsigned long long>(0xffffff00ull,
      (unsigned long long)c->capacity * (unsigned long long)std::max(1, c->n_slots));
  if (const char *cap = getenv("GSX_TILE_LISTS_CAP")) want = std::min<unsigned long long>(want, strtoull(cap, nullptr, 10));   // tests: force the fallback
  if (want <= c->tl_cap && !(off && c->tlist)) return GSX_OK;
  if (c->tlist) { cudaFree(c->tlist); c->tlist = nullptr; c->tl_cap = 0; }
  if (want == 0) return GSX_OK;
  GSX_CUDA(cudaMalloc(&c->tlist, want * sizeof(uint16_t)));
  c->tl_cap = (unsigned)want;
  return GSX_OK;
}

static int alloc_bin(gsx_ctx *c, long long capacity) {
  for (BinSet &b : c->bset) {
    if (b.ws) { cudaFree(b.ws); b.ws = nullptr; }
    b.sorted_tiles = b.sorted_vals = nullptr; b.owner = -1;
  }
  c->sorted_tiles = nullptr; c->sorted_vals = nullptr;
  if (capacity >= (1LL << 30)) { set_error("instance capacity %lld >= 2^30", capacity); return GSX_E_ARG; }
  c->capacity = std::max<long long>(capacity, 1024);
  c->L = bin_layout(c->max_n, c->capacity);
  const int sets = c->n_slots > 1 ? 2 : 1;     // the second set exists once a batch of views was reserved
  for (int i = 0; i < sets; ++i) {
    GSX_CUDA(cudaMalloc(&c->bset[i].ws, c->L.bytes));
    // tests poison the instance buffers (GSX_DEBUG_POISON=1) to prove that an overflowed view never reads them
    if (getenv("GSX_DEBUG_POISON")) GSX_CUDA(cudaMemset(c->bset[i].ws, 0xFF, c->L.bytes));
    if (!c->bset[i].ranges) GSX_CUDA(cudaMalloc(&c->bset[i].ranges, sizeof(int) * (size_t)(2 * c->n_px_tiles + 1)));   // ranges [T + 1] | tile order [T]
  }
  c->bin_ws = c->bset[0].ws; c->ranges = c->bset[0].ranges;
  return ensure_tile_lists(c);
}

static void release_slot(gsx_slot &s);
static int alloc_slot_buffers(gsx_ctx *c, int k);
static int alloc_slot(gsx_ctx *c, int k) {
  if (c->slot[k].allocated) return GSX_OK;
  const int rc = alloc_slot_buffers(c, k);
  if (rc != GSX_OK) release_slot(c->slot[k]);   // a failed cudaMalloc must not leak the buffers obtained before it
  return rc;
}
static int alloc_slot_buffers(gsx_ctx *c, int k) {
  gsx_slot &s = c->slot[k];
  const size_t n = (size_t)c->max_n;
  GSX_CUDA(cudaMalloc(&s.p.r0, n * sizeof(float4)));
  GSX_CUDA(cudaMalloc(&s.p.r1, n * sizeof(float4)));
  GSX_CUDA(cudaMalloc(&s.p.r2, n * sizeof(float4)));
  GSX_CUDA(cudaMalloc(&s.p.radius, n * sizeof(float4)));
  GSX_CUDA(cudaMalloc(&s.p.depth, n * sizeof(float)));
  GSX_CUDA(cudaMalloc(&s.p.mask, n));
  GSX_CUDA(cudaMalloc(&s.p.touched, n * sizeof(int)));
  GSX_CUDA(cudaMalloc(&s.p.gacc, n * 12 * sizeof(float)));
  GSX_CUDA(cudaMalloc(&s.p.ckey, n * sizeof(uint32_t)));
  GSX_CUDA(cudaMalloc(&s.p.cid, n * sizeof(uint32_t)));
  GSX_CUDA(cudaMalloc(&s.p.trec, n * sizeof(uint2)));
  s.p.ccount = c->d_ccount + k;
  s.pb.key_a = s.p.ckey; s.pb.val_a = s.p.cid;
  s.pb.over_flag = c->h_over;
  GSX_CUDA(cudaMalloc(&s.pb.key_b, n * sizeof(uint32_t)));
  GSX_CUDA(cudaMalloc(&s.pb.val_b, n * sizeof(uint32_t)));
  GSX_CUDA(cudaMalloc(&s.pb.offsets, n * sizeof(int)));
  GSX_CUDA(cudaMalloc(&s.pb.srec, n * sizeof(uint2)));
  GSX_CUDA(cudaMalloc(&s.pb.block_sums, prebin_block_sums_count(c->max_n) * sizeof(int)));
  GSX_CUDA(cudaMalloc(&s.pb.total, 4 * sizeof(int)));
  GSX_CUDA(cudaMemset(s.pb.total, 0, 4 * sizeof(int)));
  s.pb.sort_ws_bytes = prebin_sort_ws_bytes(c->max_n);
  GSX_CUDA(cudaMalloc(&s.pb.sort_ws, s.pb.sort_ws_bytes));
  GSX_CUDA(cudaEventCreateWithFlags(&s.ev_pre, cudaEventDisableTiming));
  GSX_CUDA(cudaEventCreateWithFlags(&s.ev_bin, cudaEventDisableTiming));
  GSX_CUDA(cudaMemset(s.p.gacc, 0, n * 12 * sizeof(float)));
  GSX_CUDA(cudaMemset(s.p.mask, 0, n));
  GSX_CUDA(cudaMemset(s.p.touched, 0, n * sizeof(int)));
  s.allocated = true;
  if (k + 1 > c->n_slots) c->n_slots = k + 1;
  return GSX_OK;
}
static void release_slot(gsx_slot &s) {   // cudaFree(nullptr) is a no-op
  cudaFree(s.p.r0); cudaFree(s.p.r1); cudaFree(s.p.r2); cudaFree(s.p.radius); cudaFree(s.p.depth); cudaFree(s.p.mask);
  cudaFree(s.p.touched); cudaFree(s.p.gacc); cudaFree(s.p.ckey); cudaFree(s.p.cid); cudaFree(s.p.trec);
  cudaFree(s.pb.key_b); cudaFree(s.pb.val_b); cudaFree(s.pb.offsets); cudaFree(s.pb.srec); cudaFree(s.pb.block_sums); cudaFree(s.pb.total);
  cudaFree(s.pb.sort_ws);
  if (s.ev_pre) cudaEventDestroy(s.ev_pre);
  if (s.ev_bin) cudaEventDestroy(s.ev_bin);
  s = gsx_slot();
}
static void free_slot(gsx_slot &s) {
  if (s.allocated) release_slot(s);
}

static const int *tile_order_of(const gsx_ctx *c) { return tile_order_enabled() ? c->ranges + c->n_px_tiles + 1 : nullptr; }

// Work issued ahead on the side streams reads the slots' per-Gaussian outputs and writes the instance-buffer sets: before
// either is overwritten (next per-Gaussian forward, capacity growth) the caller's stream waits for all of it.
static int drain_async(gsx_ctx *c, cudaStream_t st) {
  for (int k = 0; k < c->n_slots; ++k) {
    gsx_slot &s = c->slot[k];
    if (s.pre_async) { GSX_CUDA(cudaStreamWaitEvent(st, s.ev_pre, 0)); s.pre_async = false; }
    if (s.bin_async) { GSX_CUDA(cudaStreamWaitEvent(st, s.ev_bin, 0)); s.bin_async = false; }
  }
  for (BinSet &b : c->bset) b.owner = -1;
  return GSX_OK;
}

// the batch's tile-list buffer (GSX_TILE_LISTS=0: none -- the emission then repeats the SAT, the round-1 behaviour)
static void set_tile_lists(gsx_ctx *c, BatchConst *bc) {
  bc->tlist = c->tlist; bc->tcursor = (unsigned *)(c->d_ccount + GSX_MAX_BATCH); bc->tl_cap = c->tlist ? c->tl_cap : 0;
}

// emission + tile sort + ranges of slot k's view into instance-buffer set k & 1, on stream `st`
static int issue_emit_sort(gsx_ctx *c, int k, int N, cudaStream_t st, cudaEvent_t *ev) {
  gsx_slot &s = c->slot[k];
  BinSet &b = c->bset[c->bset[1].ws ? (k & 1) : 0];
  const gsx_view_t &v = s.view;
  const int ntx = (v.width + 15) / 16, nty = (v.height + 15) / 16;
  b.owner = k;
  int rc = bin_emit_sort_impl((const float *)s.p.r0, 4, s.p.radius, s.p.touched, ntx, nty, N, c->capacity, b.ws, c->L, s.pb, b.ranges,
                              &b.sorted_tiles, &b.sorted_vals, st, ev, (v.flags & GSX_LISTS_CONTRIBUTING) ? 1 : 0,
                              (lists_enabled() && c->n_px_tiles <= 65536) ? s.p.trec : nullptr, c->tlist);
  if (rc != GSX_OK) return rc;
  // optional: the order in which the compositing grids take the tiles, longest lists first (GSX_TILE_ORDER=1)
  if (tile_order_enabled()) rc = tile_order_launch(b.ranges, ntx * nty, b.ranges + c->n_px_tiles + 1, st);
  return rc;
}

// view k+... binned ahead on the binning stream: after its pre-binning and after everything queued so far on `st` (the
// compositing of the view that used the same instance-buffer set)
static int issue_ahead(gsx_ctx *c, int k, int N, cudaStream_t st) {
  gsx_slot &s = c->slot[k];
  if (!s.pre_async || s.bin_async) return GSX_OK;
  GSX_CUDA(cudaEventRecord(c->ev_free, st));
  GSX_CUDA(cudaStreamWaitEvent(c->cs_bin, c->ev_free, 0));
  GSX_CUDA(cudaStreamWaitEvent(c->cs_bin, s.ev_pre, 0));
  int rc = issue_emit_sort(c, k, N, c->cs_bin, nullptr);
  if (rc != GSX_OK) return rc;
  GSX_CUDA(cudaEventRecord(s.ev_bin, c->cs_bin));
  s.bin_async = true;
  return GSX_OK;
}

// bin + sort + ranges + compositing forward of the view whose per-Gaussian records are in slot k
static int forward_view(gsx_ctx *c, int k, int N, cudaStream_t st) {
  NvtxRange nvtx_("gsx:view_forward(bin+sort+composite)");
  gsx_slot &s = c->slot[k];
  const gsx_view_t &v = s.view;
  c->last_n = N; c->last_w = v.width; c->last_h = v.height; c->cur = k;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[2], st));
  BinSet &b = c->bset[c->bset[1].ws ? (k & 1) : 0];
  int rc;
  if (s.bin_async && b.owner == k) {
    GSX_CUDA(cudaStreamWaitEvent(st, s.ev_bin, 0));    // binned ahead (which waited for the pre-binning)
    s.bin_async = false; s.pre_async = false;
  } else {
    for (int j = 0; j < c->n_slots; ++j)               // out-of-order call: nothing may still be writing this set
      if (c->slot[j].bin_async) { GSX_CUDA(cudaStreamWaitEvent(st, c->slot[j].ev_bin, 0)); c->slot[j].bin_async = false; }
    if (s.pre_async) {
      GSX_CUDA(cudaStreamWaitEvent(st, s.ev_pre, 0));  // issued by gsx_batch_forward_params on the side stream
      s.pre_async = false;
    } else {
      rc = prebin(c, s, N, st);
      if (rc != GSX_OK) return rc;
    }
    rc = issue_emit_sort(c, k, N, st, c->timing ? &c->ev[3] : nullptr);
    if (rc != GSX_OK) return rc;
  }
  c->cur_total = s.pb.total;
  c->bin_ws = b.ws; c->ranges = b.ranges; c->sorted_tiles = b.sorted_tiles; c->sorted_vals = b.sorted_vals;
  // the next view's instance buffers are the other set, last read by the compositing backward of view k-1 (already
  // queued on st): its emission and tile sort start now and run under this view's compositing
  if (c->cs_bin && k + 1 < c->batch_nv) {
    rc = issue_ahead(c, k + 1, N, st);
    if (rc != GSX_OK) return rc;
  }
  rc = composite_fwd_packed(s.p.r0, s.p.r1, s.p.r2, (const int *)c->sorted_vals, c->ranges, v.width, v.height, v.background,
                            c->spp, c->wpp, c->image, st, tile_order_of(c));
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[6], st));
  return rc;
}

// depth sort (+ tie fix-up) of the slot's compacted list and the scan of the tile counts in depth order
static int prebin(gsx_ctx *c, gsx_slot &s, int N, cudaStream_t st) {
  return bin_prebin_impl(nullptr, 0, s.p.touched, N, true, s.p.ccount, &s.pb, st, c->capacity,
                         (lists_enabled() && c->n_px_tiles <= 65536) ? s.p.trec : nullptr);
}

static int check_view(gsx_ctx *c, const gsx_view_t *v, const gsx_params_t *p, int N) {
  if (N < 0 || N > c->max_n || v->width <= 0 || v->height <= 0 || v->width > c->max_w || v->height > c->max_h ||
      (size_t)v->width * v->height > (size_t)c->max_w * c->max_h) {
    set_error("view %dx%d / N %d exceeds the context (%d, %dx%d)", v->width, v->height, N, c->max_n, c->max_w, c->max_h);
    return GSX_E_ARG;
  }
  if (v->l_max < 0 || v->l_max > 3) { set_error("l_max %d out of range 0..3", v->l_max); return GSX_E_ARG; }
  const int shf = ((v->l_max + 1) * (v->l_max + 1) - 1) * 3;
  if (v->l_max > 0 && p->sh_stride < shf) { set_error("sh_stride %d < %d", p->sh_stride, shf); return GSX_E_ARG; }
  if (((uintptr_t)p->quaternion) & 15) { set_error("params.quaternion must be 16-byte aligned (float4 loads)"); return GSX_E_ARG; }
  return GSX_OK;
}
static int check_grads(const gsx_grads_t *g) {
  if (((uintptr_t)g->quaternion) & 15) { set_error("grads.quaternion must be 16-byte aligned (float4 stores)"); return GSX_E_ARG; }
  if (((uintptr_t)g->uv) & 7) { set_error("grads.uv must be 8-byte aligned"); return GSX_E_ARG; }
  return GSX_OK;
}
}  // namespace gsx

extern "C" {

int gsx_version(void) { return 1; }
const char *gsx_last_error(void) { return g_err; }

int gsx_ctx_create(gsx_ctx_t **out, int max_gaussians, int max_width, int max_height, long long instance_capacity) {
  if (!out || max_gaussians <= 0 || max_width <= 0 || max_height <= 0) { set_error("bad ctx arguments"); return GSX_E_ARG; }
  gsx_ctx *c = new gsx_ctx();
  c->max_n = max_gaussians; c->max_w = max_width; c->max_h = max_height;
  const size_t px = (size_t)max_width * max_height;
  const int T = ((max_width + 15) / 16) * ((max_height + 15) / 16);
  if (instance_capacity >= (1LL << 30)) { set_error("instance capacity %lld >= 2^30", instance_capacity); delete c; return GSX_E_ARG; }
  GSX_CUDA(cudaMalloc(&c->d_ccount, (GSX_MAX_BATCH + 1) * sizeof(int)));
  GSX_CUDA(cudaMemset(c->d_ccount, 0, (GSX_MAX_BATCH + 1) * sizeof(int)));
  GSX_CUDA(cudaHostAlloc(&c->h_over, 4 * sizeof(int), cudaHostAllocMapped));
  c->h_over[0] = 0;
  GSX_CUDA(cudaMalloc(&c->d_work, 2 * sizeof(unsigned long long)));
  {
    int rc0 = alloc_slot(c, 0);
    if (rc0 != GSX_OK) return rc0;
    c->cur_total = c->slot[0].pb.total;   // zero until the first forward
  }
  c->n_px_tiles = T;
  GSX_CUDA(cudaMalloc(&c->image, px * 3 * sizeof(float)));
  GSX_CUDA(cudaMalloc(&c->wpp, px * sizeof(float)));
  GSX_CUDA(cudaMalloc(&c->spp, px * sizeof(int)));
  GSX_CUDA(cudaMalloc(&c->grad_image, px * 3 * sizeof(float)));
  GSX_CUDA(cudaMalloc(&c->d_P, sizeof(int)));
  GSX_CUDA(cudaMallocHost(&c->h_stats, (4 + GSX_MAX_BATCH) * sizeof(int)));
  if (instance_capacity <= 0) instance_capacity = std::min<long long>(8LL * max_gaussians, (1LL << 30) - 1);
  int rc = alloc_bin(c, instance_capacity);
  if (rc != GSX_OK) return rc;
  *out = c;
  return GSX_OK;
}

int gsx_ctx_destroy(gsx_ctx_t *c) {
  if (!c) return GSX_OK;
  for (int k = 0; k < GSX_MAX_BATCH; ++k) free_slot(c->slot[k]);
  cudaDeviceSynchronize();
  for (BinSet &b : c->bset) { if (b.ws) cudaFree(b.ws); if (b.ranges) cudaFree(b.ranges); }
  if (c->key64) cudaFree(c->key64); cudaFree(c->image); cudaFree(c->wpp); cudaFree(c->spp);
  for (int k = 0; k < GSX_MAX_BATCH; ++k) {
    if (c->grad_stage[k]) cudaFree(c->grad_stage[k]);
    if (c->ev_in[k]) cudaEventDestroy(c->ev_in[k]);
  }
  if (c->ev_img) cudaEventDestroy(c->ev_img);
  if (c->ev_out) cudaEventDestroy(c->ev_out);
  if (c->ev_go) cudaEventDestroy(c->ev_go);
  if (c->cs_pre) { cudaStreamSynchronize(c->cs_pre); cudaStreamDestroy(c->cs_pre); }
  if (c->cs_bin) cudaStreamDestroy(c->cs_bin);
  if (c->ev_fwd) cudaEventDestroy(c->ev_fwd);
  if (c->ev_free) cudaEventDestroy(c->ev_free);
  if (c->cs_in) cudaStreamDestroy(c->cs_in);
  if (c->cs_out) cudaStreamDestroy(c->cs_out);
  cudaFree(c->grad_image); cudaFree(c->d_P); cudaFree(c->d_ccount); cudaFree(c->tlist); cudaFreeHost(c->h_stats);
  cudaFree(c->d_work); cudaFreeHost(c->h_over);
  delete c;
  return GSX_OK;
}

int gsx_ctx_reserve(gsx_ctx_t *c, long long instance_capacity) {
  if (!c) return GSX_E_NULL;
  if (instance_capacity <= c->capacity) return GSX_OK;
  if (instance_capacity >= (1LL << 30)) { set_error("instance capacity %lld >= 2^30", instance_capacity); return GSX_E_ARG; }
  GSX_CUDA(cudaDeviceSynchronize());
  for (int k = 0; k < c->n_slots; ++k) c->slot[k].pre_async = c->slot[k].bin_async = false;
  c->h_over[0] = 0;
  return alloc_bin(c, instance_capacity);
}

// Single view = a batch of one: the same pair-compacted per-Gaussian kernels (batch.cu) run with one slot.
int gsx_forward(gsx_ctx_t *c, const gsx_view_t *v, const gsx_params_t *p, int N, void *stream) {
  NvtxRange nvtx_("gsx_forward");
  if (!c || !v || !p) { set_error("null ctx/view/params"); return GSX_E_NULL; }
  int rc = check_view(c, v, p, N);
  if (rc != GSX_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  rc = drain_async(c, st);
  if (rc != GSX_OK) return rc;
  c->slot[0].view = *v;
  c->batch_n = N; c->batch_nv = 1; c->batch_lmax = v->l_max;
  BatchConst bc;
  bc.nv = 1; bc.v[0] = make_view_const(*v); bc.s[0] = c->slot[0].p;
  set_tile_lists(c, &bc);
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[0], st));
  GSX_CUDA(cudaMemsetAsync(c->d_ccount, 0, sizeof(int), st));
  GSX_CUDA(cudaMemsetAsync(c->d_ccount + GSX_MAX_BATCH, 0, sizeof(int), st));
  rc = batch_fwd_launch(bc, *p, N, v->l_max, st);
  if (rc != GSX_OK) return rc;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[1], st));
  return forward_view(c, 0, N, st);
}

int gsx_backward(gsx_ctx_t *c, const gsx_view_t *v, const gsx_params_t *p, int N, const float *grad_image,
                 const gsx_grads_t *g, void *stream) {
  NvtxRange nvtx_("gsx_backward");
  if (!c || !v || !p || !g || !grad_image) { set_error("null argument"); return GSX_E_NULL; }
  if (c->cur != 0 || c->batch_nv != 1 || N != c->last_n || v->width != c->last_w || v->height != c->last_h) {
    set_error("gsx_backward does not match the last gsx_forward on this context");
    return GSX_E_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const gsx_slot &s = c->slot[0];
  int rc = check_grads(g);
  if (rc != GSX_OK) return rc;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[7], st));
  rc = composite_bwd_packed(s.p.r0, s.p.r1, s.p.r2, (const int *)c->sorted_vals, c->ranges, v->width, v->height,
                                v->background, c->spp, c->wpp, grad_image, (float *)s.p.gacc, st, tile_order_of(c));
  if (rc != GSX_OK) return rc;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[8], st));
  BatchConst bc;
  bc.nv = 1; bc.v[0] = make_view_const(*v); bc.s[0] = s.p;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[9], st));
  rc = batch_bwd_launch(bc, *p, N, v->l_max, *g, v->flags, st);
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[10], st));
  return rc;
}

// ---- batch path: the views of one training batch share one pass over the parameters (batch.cu) ----------------
int gsx_batch_reserve(gsx_ctx_t *c, int n_slots) {
  if (!c) return GSX_E_NULL;
  if (n_slots < 1 || n_slots > GSX_MAX_BATCH) { set_error("n_slots %d out of range 1..%d", n_slots, GSX_MAX_BATCH); return GSX_E_ARG; }
  for (int k = 0; k < n_slots; ++k) {
    int rc = alloc_slot(c, k);
    if (rc != GSX_OK) return rc;
  }
  if (c->n_slots > 1 && !c->bset[1].ws) {        // second instance-buffer set for the binning pipeline
    GSX_CUDA(cudaMalloc(&c->bset[1].ws, c->L.bytes));
    GSX_CUDA(cudaMalloc(&c->bset[1].ranges, sizeof(int) * (size_t)(2 * c->n_px_tiles + 1)));
  }
  return ensure_tile_lists(c);
}

static int fill_batch_const(gsx_ctx *c, int nv, BatchConst *bc) {
  set_tile_lists(c, bc);
  bc->nv = nv;
  for (int k = 0; k < nv; ++k) { bc->v[k] = make_view_const(c->slot[k].view); bc->s[k] = c->slot[k].p; }
  return GSX_OK;
}

int gsx_batch_forward_params(gsx_ctx_t *c, const gsx_view_t *views, int n_views, const gsx_params_t *p, int N,
                             void *stream) {
  NvtxRange nvtx_("gsx_batch_forward_params(per-Gaussian fwd)");
  if (!c || !views || !p) { set_error("null ctx/views/params"); return GSX_E_NULL; }
  if (n_views < 1 || n_views > c->n_slots) { set_error("n_views %d exceeds the reserved slots (%d)", n_views, c->n_slots); return GSX_E_ARG; }
  for (int k = 0; k < n_views; ++k) {
    int rc = check_view(c, &views[k], p, N);
    if (rc != GSX_OK) return rc;
    if (views[k].l_max != views[0].l_max) { set_error("all views of a batch must share l_max"); return GSX_E_ARG; }
    c->slot[k].view = views[k];
  }
  c->batch_n = N; c->batch_nv = n_views; c->batch_lmax = views[0].l_max;
  BatchConst bc;
  fill_batch_const(c, n_views, &bc);
  cudaStream_t st = (cudaStream_t)stream;
  {
    int rcd = drain_async(c, st);          // side-stream work of a previous batch whose views were never consumed
    if (rcd != GSX_OK) return rcd;
  }
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[0], st));
  GSX_CUDA(cudaMemsetAsync(c->d_ccount, 0, (GSX_MAX_BATCH + 1) * sizeof(int), st));
  int rc = batch_fwd_launch(bc, *p, N, views[0].l_max, st);
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[1], st));
  if (rc != GSX_OK) return rc;
  // Pre-binning (depth sort + scan: ten short latency-bound launches per view) depends only on the kernel above: with
  // more than one view it runs on a side stream, so views 1.. are pre-binned under the compositing of the views before
  // them.  With stage timing on everything stays on the caller's stream so that the stages remain attributable.
  if (n_views > 1 && !c->timing) {
    if (!c->cs_pre) {
      // high priority: their short kernels take CTA slots as the (thousands of queued) compositing CTAs retire, instead
      // of waiting behind the whole compositing grid
      int lo = 0, hi = 0;
      GSX_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
      GSX_CUDA(cudaStreamCreateWithPriority(&c->cs_pre, cudaStreamNonBlocking, hi));
      GSX_CUDA(cudaStreamCreateWithPriority(&c->cs_bin, cudaStreamNonBlocking, hi));
      GSX_CUDA(cudaEventCreateWithFlags(&c->ev_fwd, cudaEventDisableTiming));
      GSX_CUDA(cudaEventCreateWithFlags(&c->ev_free, cudaEventDisableTiming));
    }
    GSX_CUDA(cudaEventRecord(c->ev_fwd, st));
    GSX_CUDA(cudaStreamWaitEvent(c->cs_pre, c->ev_fwd, 0));
    for (int k = 0; k < n_views; ++k) {
      gsx_slot &s = c->slot[k];
      rc = prebin(c, s, N, c->cs_pre);
      if (rc != GSX_OK) return rc;
      GSX_CUDA(cudaEventRecord(s.ev_pre, c->cs_pre));
      s.pre_async = true;
    }
    // both instance-buffer sets are free here (everything of the previous batch is ahead of this call on st)
    for (int k = 0; k < 2 && k < n_views && c->bset[1].ws; ++k) {
      rc = issue_ahead(c, k, N, st);
      if (rc != GSX_OK) return rc;
    }
  }
  return GSX_OK;
}

int gsx_batch_forward_view(gsx_ctx_t *c, int slot, void *stream) {
  if (!c) return GSX_E_NULL;
  if (slot < 0 || slot >= c->batch_nv) { set_error("slot %d outside the current batch (%d views)", slot, c->batch_nv); return GSX_E_ARG; }
  return forward_view(c, slot, c->batch_n, (cudaStream_t)stream);
}

int gsx_batch_backward_view(gsx_ctx_t *c, int slot, const float *grad_image, void *stream) {
  NvtxRange nvtx_("gsx_batch_backward_view(composite bwd)");
  if (!c || !grad_image) { set_error("null argument"); return GSX_E_NULL; }
  if (slot != c->cur || slot >= c->batch_nv) { set_error("gsx_batch_backward_view(%d) must follow gsx_batch_forward_view(%d)", slot, slot); return GSX_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  const gsx_slot &s = c->slot[slot];
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[7], st));
  int rc = composite_bwd_packed(s.p.r0, s.p.r1, s.p.r2, (const int *)c->sorted_vals, c->ranges, s.view.width, s.view.height,
                                s.view.background, c->spp, c->wpp, grad_image, (float *)s.p.gacc, st, tile_order_of(c));
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[8], st));
  return rc;
}

int gsx_batch_backward_params(gsx_ctx_t *c, const gsx_params_t *p, int N, const gsx_grads_t *g, int flags, void *stream) {
  NvtxRange nvtx_("gsx_batch_backward_params(per-Gaussian bwd)");
  if (!c || !p || !g) { set_error("null argument"); return GSX_E_NULL; }
  if (N != c->batch_n || c->batch_nv < 1) { set_error("gsx_batch_backward_params does not match the current batch"); return GSX_E_ARG; }
  if (((uintptr_t)p->quaternion) & 15) { set_error("params.quaternion must be 16-byte aligned (float4 loads)"); return GSX_E_ARG; }
  {
    int rcg = check_grads(g);
    if (rcg != GSX_OK) return rcg;
  }
  BatchConst bc;
  fill_batch_const(c, c->batch_nv, &bc);
  cudaStream_t st = (cudaStream_t)stream;
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[9], st));
  int rc = batch_bwd_launch(bc, *p, N, c->batch_lmax, *g, flags, st);
  if (c->timing) GSX_CUDA(cudaEventRecord(c->ev[10], st));
  return rc;
}

int gsx_batch_extract_view_rgb_grad(gsx_ctx_t *c, int slot, float *drgb, void *stream) {
  if (!c || !drgb) { set_error("null argument"); return GSX_E_NULL; }
  if (slot < 0 || slot >= c->batch_nv) { set_error("slot %d outside the current batch (%d views)", slot, c->batch_nv); return GSX_E_ARG; }
  return extract_view_rgb_launch(c->slot[slot].p.gacc, c->slot[slot].p.touched, c->batch_n, drgb, (cudaStream_t)stream);
}

int gsx_batch_extract_view_rgb_grad_peers(gsx_ctx_t *c, int slot, float *multicast_drgb, float *const *peer_drgb_host,
                                          int n_peers, void *stream) {
  if (!c || (!multicast_drgb && !peer_drgb_host)) { set_error("null argument"); return GSX_E_NULL; }
  if (slot < 0 || slot >= c->batch_nv) { set_error("slot %d outside the current batch (%d views)", slot, c->batch_nv); return GSX_E_ARG; }
  return extract_view_rgb_peers_launch(c->slot[slot].p.gacc, c->slot[slot].p.touched, nullptr, c->batch_n, multicast_drgb,
                                       peer_drgb_host, n_peers, (cudaStream_t)stream);
}

int gsx_push_view_rgb_peers(const float *drgb_local, int N, float *multicast_drgb, float *const *peer_drgb_host, int n_peers,
                            void *stream) {
  if (!drgb_local || (!multicast_drgb && !peer_drgb_host)) { set_error("null argument"); return GSX_E_NULL; }
  return extract_view_rgb_peers_launch(nullptr, nullptr, drgb_local, N, multicast_drgb, peer_drgb_host, n_peers,
                                       (cudaStream_t)stream);
}

int gsx_allreduce_multimem_f32(float *multicast_ptr, size_t n_floats, int rank, int world, void *stream) {
  return allreduce_multimem_launch(multicast_ptr, n_floats, rank, world, (cudaStream_t)stream);
}

int gsx_sh_grads_from_view_rgb(const float *xyz_w, int N, int l_max, const float *campos_host, int n_views,
                               const float *drgb, size_t view_stride, float *sh_grad, int sh_stride, float *band0_grad,
                               int flags, void *stream) {
  if (!campos_host) { set_error("null campos"); return GSX_E_NULL; }
  GSX_REQUIRE_DEV(xyz_w); GSX_REQUIRE_DEV(drgb); GSX_REQUIRE_DEV(band0_grad);
  if (l_max > 0) GSX_REQUIRE_DEV(sh_grad);
  return sh_from_view_rgb_launch(xyz_w, N, l_max, campos_host, n_views, drgb, view_stride, sh_grad, sh_stride, band0_grad,
                                 flags, (cudaStream_t)stream);
}

// Host-buffer batch step: per view H2D of the loss-gradient image and D2H of the rendered image (pinned host memory),
// forward + backward of every view, one pass of parameter gradients.  Synchronises.
static int ensure_host_pipeline(gsx_ctx_t *c, int n_views) {
  if (!c->cs_in) {
    GSX_CUDA(cudaStreamCreateWithFlags(&c->cs_in, cudaStreamNonBlocking));
    GSX_CUDA(cudaStreamCreateWithFlags(&c->cs_out, cudaStreamNonBlocking));
    GSX_CUDA(cudaEventCreateWithFlags(&c->ev_img, cudaEventDisableTiming));
    GSX_CUDA(cudaEventCreateWithFlags(&c->ev_out, cudaEventDisableTiming));
    GSX_CUDA(cudaEventCreateWithFlags(&c->ev_go, cudaEventDisableTiming));
  }
  const size_t bytes = (size_t)c->max_w * c->max_h * 3 * sizeof(float);
  for (int k = 0; k < n_views; ++k) {
    if (!c->ev_in[k]) GSX_CUDA(cudaEventCreateWithFlags(&c->ev_in[k], cudaEventDisableTiming));
    if (!c->grad_stage[k]) GSX_CUDA(cudaMalloc(&c->grad_stage[k], bytes));
  }
  return GSX_OK;
}

// Whole batch from host buffers (pinned for the copies to overlap).  Three streams: uploads of the n gradient images
// start at once and hide under the per-Gaussian forward, view k's image goes back to the host while its backward runs;
// the compute stream waits only for the upload it is about to read and for the image buffer it is about to overwrite.
int gsx_batch_render_host(gsx_ctx_t *c, const gsx_view_t *views, int n_views, const gsx_params_t *p, int N,
                          const float *const *grad_images_host, float *const *images_host, const gsx_grads_t *g, int flags,
                          void *stream) {
  NvtxRange nvtx_("gsx_batch_render_host");
  if (!c || !views || !p || !g || !grad_images_host) { set_error("null argument"); return GSX_E_NULL; }
  if (n_views < 1 || n_views > GSX_MAX_BATCH) { set_error("n_views out of range"); return GSX_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  int rc = ensure_host_pipeline(c, n_views);
  if (rc != GSX_OK) return rc;
  GSX_CUDA(cudaEventRecord(c->ev_go, st));            // the staging buffers may still be read by work queued on st
  GSX_CUDA(cudaStreamWaitEvent(c->cs_in, c->ev_go, 0));
  for (int k = 0; k < n_views; ++k) {
    if (views[k].width > c->max_w || views[k].height > c->max_h) { set_error("view larger than the context"); return GSX_E_ARG; }
    if (!grad_images_host[k]) { set_error("null gradient image"); return GSX_E_NULL; }
    const size_t bytes = (size_t)views[k].width * views[k].height * 3 * sizeof(float);
    GSX_CUDA(cudaMemcpyAsync(c->grad_stage[k], grad_images_host[k], bytes, cudaMemcpyHostToDevice, c->cs_in));
    GSX_CUDA(cudaEventRecord(c->ev_in[k], c->cs_in));
  }
  rc = gsx_batch_forward_params(c, views, n_views, p, N, stream);
  bool downloading = false;
  for (int k = 0; k < n_views && rc == GSX_OK; ++k) {
    const size_t bytes = (size_t)views[k].width * views[k].height * 3 * sizeof(float);
    if (downloading) GSX_CUDA(cudaStreamWaitEvent(st, c->ev_out, 0));   // c->image is about to be overwritten
    rc = gsx_batch_forward_view(c, k, stream);
    if (rc != GSX_OK) break;
    if (images_host && images_host[k]) {
      GSX_CUDA(cudaEventRecord(c->ev_img, st));
      GSX_CUDA(cudaStreamWaitEvent(c->cs_out, c->ev_img, 0));
      GSX_CUDA(cudaMemcpyAsync(images_host[k], c->image, bytes, cudaMemcpyDeviceToHost, c->cs_out));
      GSX_CUDA(cudaEventRecord(c->ev_out, c->cs_out));
      downloading = true;
    }
    GSX_CUDA(cudaStreamWaitEvent(st, c->ev_in[k], 0));
    rc = gsx_batch_backward_view(c, k, c->grad_stage[k], stream);
  }
  if (rc == GSX_OK) rc = gsx_batch_backward_params(c, p, N, g, flags, stream);
  // never return with copies in flight, error or not
  cudaStreamSynchronize(c->cs_in);
  cudaStreamSynchronize(c->cs_out);
  if (rc != GSX_OK) { cudaStreamSynchronize(st); return rc; }
  GSX_CUDA(cudaStreamSynchronize(st));
  return GSX_OK;
}

// out_host[2] = {S, S+} of the last forward (work counters for the roofline arithmetic).  Synchronises.
int gsx_ctx_count_work(gsx_ctx_t *c, unsigned long long *out_host, void *stream) {
  if (!c || !out_host) return GSX_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long *d = c->d_work;
  GSX_CUDA(cudaMemsetAsync(d, 0, 2 * sizeof(unsigned long long), st));
  int rc = count_work_launch(c->slot[c->cur].p.r0, c->slot[c->cur].p.r1, (const int *)c->sorted_vals, c->ranges, c->last_w, c->last_h, c->spp, d, st);
  if (rc == GSX_OK) {
    GSX_CUDA(cudaMemcpyAsync(out_host, d, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    GSX_CUDA(cudaStreamSynchronize(st));
  }
  return rc;
}

// out_host[2] = {warp-hits, contributing warp-hits} of the compositing backward over the last forward's lists: the
// number of (8x8 pixel block, entry) pairs that survive the extent-box cull, and of those with a contributing pixel --
// the units the compositing kernels' instruction counts scale with.  Diagnostics; synchronises.
int gsx_ctx_count_hits(gsx_ctx_t *c, const float *grad_image, unsigned long long *out_host, void *stream) {
  if (!c || !out_host || !grad_image) return GSX_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  const gsx_slot &s = c->slot[c->cur];
  GSX_CUDA(cudaMemsetAsync(c->d_work, 0, 2 * sizeof(unsigned long long), st));
  int rc = composite_hits_launch(s.p.r0, s.p.r1, s.p.r2, (const int *)c->sorted_vals, c->ranges, c->last_w, c->last_h, c->spp,
                                 c->wpp, grad_image, c->d_work, st);
  if (rc != GSX_OK) return rc;
  GSX_CUDA(cudaMemcpyAsync(out_host, c->d_work, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  GSX_CUDA(cudaStreamSynchronize(st));
  return GSX_OK;
}

int gsx_ctx_enable_timing(gsx_ctx_t *c, int on) {
  if (!c) return GSX_E_NULL;
  if (on && !c->ev[0])
    for (int i = 0; i < 11; ++i) GSX_CUDA(cudaEventCreate(&c->ev[i]));
  c->timing = on != 0;
  return GSX_OK;
}

// ms_out[7]: preprocess_fwd, scan+emit, sort, ranges, composite_fwd, composite_bwd, preprocess_bwd of the last
// forward + backward issued with timing enabled.  Synchronises the device.
int gsx_ctx_timing(gsx_ctx_t *c, float *ms_out) {
  if (!c || !ms_out) return GSX_E_NULL;
  if (!c->timing) { set_error("timing not enabled"); return GSX_E_ARG; }
  GSX_CUDA(cudaDeviceSynchronize());
  const int a[7] = {0, 2, 3, 4, 5, 7, 9}, b[7] = {1, 3, 4, 5, 6, 8, 10};
  for (int i = 0; i < 7; ++i) {   // a stage whose events were never recorded reports -1
    if (cudaEventElapsedTime(&ms_out[i], c->ev[a[i]], c->ev[b[i]]) != cudaSuccess) { cudaGetLastError(); ms_out[i] = -1.0f; }
  }
  return GSX_OK;
}

namespace gsx {
__global__ void __launch_bounds__(256) k_fma_peak(float *out, int iters) {
  float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f,
        a7 = a0 + 7.f;
  const float m = 0.999f, c = 1e-4f * blockIdx.x;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
      a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
  }
  const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
  if (s == 123.456f) out[0] = s;
}
}  // namespace gsx

// Measured FP32 FMA peak of this device (TFLOP/s, FMA = 2 FLOP): the roofline denominator for the compositing
// kernels (MEASURED_PEAKS.json only carries HBM and bf16 tensor numbers).
int gsx_measure_fp32_tflops(float *tflops_host) {
  if (!tflops_host) return GSX_E_NULL;
  float *d = nullptr;
  GSX_CUDA(cudaMalloc(&d, 4));
  cudaEvent_t e0, e1;
  GSX_CUDA(cudaEventCreate(&e0)); GSX_CUDA(cudaEventCreate(&e1));
  const int blocks = kNumSMs * 8, iters = 4096;
  k_fma_peak<<<blocks, 256>>>(d, 64);
  float best = 0.f;
  for (int rep = 0; rep < 5; ++rep) {
    GSX_CUDA(cudaEventRecord(e0));
    k_fma_peak<<<blocks, 256>>>(d, iters);
    GSX_CUDA(cudaEventRecord(e1));
    GSX_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    GSX_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double flop = 2.0 * 64.0 * iters * 256.0 * blocks;
    best = fmaxf(best, (float)(flop / (ms * 1e-3) / 1e12));
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops_host = best;
  return GSX_OK;
}

int gsx_ctx_stats(gsx_ctx_t *c, gsx_stats_t *out, void *stream) {
  if (!c || !out) return GSX_E_NULL;
  cudaStream_t st = (cudaStream_t)stream;
  GSX_CUDA(cudaMemsetAsync(c->d_P, 0, sizeof(int), st));
  if (c->last_n > 0) {
    k_count_mask<<<kNumSMs * 4, 256, 0, st>>>(c->slot[c->cur].p.mask, c->last_n, c->d_P);
    GSX_LAUNCH_CHECK();
  }
  GSX_CUDA(cudaMemcpyAsync(&c->h_stats[0], (const char *)c->cur_total, sizeof(int), cudaMemcpyDeviceToHost, st));
  GSX_CUDA(cudaMemcpyAsync(&c->h_stats[1], c->d_P, sizeof(int), cudaMemcpyDeviceToHost, st));
  // every view of the current batch is checked, not only the last one forwarded: an overflowed view composites nothing
  // (binning.cu k_scan_block_sums), so its image and gradients are missing until the caller reserves and re-runs
  const int nv = std::max(1, std::min(c->batch_nv, c->n_slots));
  for (int k = 0; k < nv; ++k) {
    c->h_stats[4 + k] = 0;
    if (c->slot[k].allocated && c->slot[k].pb.total)
      GSX_CUDA(cudaMemcpyAsync(&c->h_stats[4 + k], c->slot[k].pb.total, sizeof(int), cudaMemcpyDeviceToHost, st));
  }
  GSX_CUDA(cudaStreamSynchronize(st));
  out->N = c->last_n; out->P = c->h_stats[1]; out->K = c->h_stats[0]; out->capacity = c->capacity;
  long long worst = out->K;
  for (int k = 0; k < nv; ++k) worst = std::max<long long>(worst, c->h_stats[4 + k]);
  if (worst > c->capacity) {
    if (out->K <= c->capacity) out->K = worst;   // report what has to be reserved
    set_error("instance capacity exceeded: K=%lld > %lld (gsx_ctx_reserve and re-run)", worst, c->capacity);
    return GSX_E_CAPACITY;
  }
  return GSX_OK;
}

// No synchronisation: reads the pinned flag the scan kernels raise.  GSX_E_CAPACITY (and the K to reserve for) if any view
// forwarded since the last gsx_ctx_reserve exceeded the instance capacity; such a view rendered background only and
// contributed no gradients.  Definitive once the stream that ran the views is synchronised; a training loop polls it once per step.
int gsx_ctx_overflow(gsx_ctx_t *c, long long *required_capacity_host) {
  if (!c) return GSX_E_NULL;
  const int k = *(volatile int *)c->h_over;
  if (required_capacity_host) *required_capacity_host = k;
  if (k > 0) {
    set_error("instance capacity exceeded: K=%d > %lld (gsx_ctx_reserve and re-run)", k, c->capacity);
    return GSX_E_CAPACITY;
  }
  return GSX_OK;
}

int gsx_ctx_state(gsx_ctx_t *c, gsx_state_t *o) {
  if (!c || !o) return GSX_E_NULL;
  o->image = c->image; o->weight_per_pixel = c->wpp; o->splats_per_pixel = c->spp;
  const SlotPtrs &sp = c->slot[c->cur].p;
  o->rec0 = (const float *)sp.r0; o->rec1 = (const float *)sp.r1; o->rec2 = (const float *)sp.r2;
  o->radius = (const float *)sp.radius; o->depth = sp.depth; o->mask = sp.mask;
  o->tiles_touched = sp.touched;
  o->sorted_gaussians = (const int *)c->sorted_vals;
  o->sorted_keys = nullptr;   /* never materialised by the two-level sort: gsx_ctx_build_sorted_keys */
  o->ranges = c->ranges; o->num_instances = (const int *)((const char *)c->cur_total);
  o->packed_grads = (const float *)sp.gacc;
  return GSX_OK;
}

// The 64-bit tile|depth keys of the last forward, rebuilt for inspection (tests): synchronous, never on the hot path.
int gsx_ctx_build_sorted_keys(gsx_ctx_t *c, const uint64_t **out) {
  if (!c || !out) return GSX_E_NULL;
  *out = nullptr;
  if (!c->sorted_vals) return GSX_OK;
  if (c->key64_cap < c->capacity) {
    if (c->key64) cudaFree(c->key64);
    c->key64 = nullptr; c->key64_cap = 0;
    GSX_CUDA(cudaMalloc(&c->key64, sizeof(uint64_t) * (size_t)c->capacity));
    c->key64_cap = c->capacity;
  }
  GSX_CUDA(cudaDeviceSynchronize());
  if (c->sorted_tiles)
    k_compose_keys<<<kNumSMs * 8, 256>>>(c->sorted_tiles, c->sorted_vals, c->slot[c->cur].p.depth,
                                         c->cur_total + 1, c->capacity, c->key64);
  else
    k_compose_keys_ranges<<<kNumSMs * 8, 256>>>(c->ranges, ((c->last_w + 15) / 16) * ((c->last_h + 15) / 16), c->sorted_vals,
                                                c->slot[c->cur].p.depth, c->key64);
  GSX_LAUNCH_CHECK();
  GSX_CUDA(cudaDeviceSynchronize());
  *out = c->key64;
  return GSX_OK;
}

// ---- CUDA-graph replay of a whole batch step ("persistent workspace, CUDA-graph replay") ---------------
// The ~20 launches per view of the batch path (latency-bound sort / scan kernels, side streams, events) are captured once
// -- forward of every view, the loss gradient of each view (given, or computed in the graph from a ground-truth image by
// gsx_fused_loss), backward, parameter gradients -- and replayed with one cudaGraphLaunch per training step.  Everything the
// step reads is behind fixed device pointers (parameters, ground truth / loss gradients, gradients), so the caller only
// updates their contents between launches.  The view set and N are baked in: capture again when they change (or when
// gsx_ctx_reserve re-allocated the instance buffers).
struct gsx_graph {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  void *loss_ws = nullptr;
  float *loss_dev = nullptr;
};

int gsx_graph_destroy(gsx_graph_t *g) {
  if (!g) return GSX_OK;
  if (g->exec) cudaGraphExecDestroy(g->exec);
  if (g->graph) cudaGraphDestroy(g->graph);
  cudaFree(g->loss_ws); cudaFree(g->loss_dev);
  delete g;
  return GSX_OK;
}

int gsx_graph_launch(gsx_graph_t *g, void *stream) {
  if (!g || !g->exec) { set_error("null graph"); return GSX_E_NULL; }
  GSX_CUDA(cudaGraphLaunch(g->exec, (cudaStream_t)stream));
  return GSX_OK;
}

int gsx_batch_graph_capture(gsx_ctx_t *c, const gsx_view_t *views, int n_views, const gsx_params_t *p, int N,
                            const float *const *grad_images_dev, const float *const *gt_images_dev, float ssim_weight,
                            float *const *images_out_dev, float *losses_dev, const gsx_grads_t *g, int flags, void *stream,
                            gsx_graph_t **out) {
  if (!c || !views || !p || !g || !out) { set_error("null argument"); return GSX_E_NULL; }
  if (!grad_images_dev && !gt_images_dev) { set_error("either loss-gradient images or ground-truth images are needed"); return GSX_E_NULL; }
  if (n_views < 1 || n_views > c->n_slots) { set_error("n_views %d exceeds the reserved slots (%d)", n_views, c->n_slots); return GSX_E_ARG; }
  if (c->timing) { set_error("stage timing must be off while capturing"); return GSX_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  if (st == nullptr) { set_error("capture needs a non-default stream"); return GSX_E_ARG; }
  gsx_graph *gr = new gsx_graph();
  size_t lws = 0;
  if (gt_images_dev) {
    for (int k = 0; k < n_views; ++k) lws = std::max(lws, gsx_fused_loss_workspace_bytes(views[k].height, views[k].width));
    if (cudaMalloc(&gr->loss_ws, lws) != cudaSuccess || cudaMalloc(&gr->loss_dev, sizeof(float)) != cudaSuccess) {
      gsx_graph_destroy(gr); set_error("cannot allocate the loss workspace"); return GSX_E_CUDA;
    }
  }
  // one eager step first: creates the side streams / events lazily created by the batch path, sizes nothing new inside
  int rc = gsx_batch_forward_params(c, views, n_views, p, N, stream);
  for (int k = 0; k < n_views && rc == GSX_OK; ++k) rc = gsx_batch_forward_view(c, k, stream);
  if (rc == GSX_OK) rc = drain_async(c, st);
  if (rc != GSX_OK) { gsx_graph_destroy(gr); return rc; }
  cudaError_t e = cudaStreamSynchronize(st);
  if (e == cudaSuccess) e = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed);
  if (e != cudaSuccess) { gsx_graph_destroy(gr); set_error("cudaStreamBeginCapture -> %s", cudaGetErrorString(e)); return GSX_E_CUDA; }
  rc = gsx_batch_forward_params(c, views, n_views, p, N, stream);
  for (int k = 0; k < n_views && rc == GSX_OK; ++k) {
    rc = gsx_batch_forward_view(c, k, stream);
    if (rc != GSX_OK) break;
    const size_t bytes = (size_t)views[k].width * views[k].height * 3 * sizeof(float);
    if (images_out_dev && images_out_dev[k] &&
        cudaMemcpyAsync(images_out_dev[k], c->image, bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess) { rc = GSX_E_CUDA; break; }
    const float *gi = grad_images_dev ? grad_images_dev[k] : c->grad_image;
    if (gt_images_dev) {
      rc = gsx_fused_loss(c->image, gt_images_dev[k], views[k].height, views[k].width, ssim_weight, c->grad_image, gr->loss_ws, lws,
                          losses_dev ? losses_dev + k : gr->loss_dev, nullptr, stream);
      if (rc != GSX_OK) break;
    }
    rc = gsx_batch_backward_view(c, k, gi, stream);
  }
  if (rc == GSX_OK) rc = gsx_batch_backward_params(c, p, N, g, flags, stream);
  e = cudaStreamEndCapture(st, &gr->graph);   // always end the capture, error or not
  if (rc == GSX_OK && e != cudaSuccess) { set_error("cudaStreamEndCapture -> %s", cudaGetErrorString(e)); rc = GSX_E_CUDA; }
  if (rc == GSX_OK) {
    e = cudaGraphInstantiate(&gr->exec, gr->graph, 0);
    if (e != cudaSuccess) { set_error("cudaGraphInstantiate -> %s", cudaGetErrorString(e)); rc = GSX_E_CUDA; }
  }
  // the capture enqueued nothing: the bookkeeping of work "in flight" on the side streams must not outlive it
  for (int k = 0; k < c->n_slots; ++k) c->slot[k].pre_async = c->slot[k].bin_async = false;
  for (BinSet &b : c->bset) b.owner = -1;
  if (rc != GSX_OK) { cudaGetLastError(); gsx_graph_destroy(gr); return rc; }
  *out = gr;
  return GSX_OK;
}

int gsx_render_view_host(gsx_ctx_t *c, const gsx_view_t *v, const gsx_params_t *p, int N,
                         const float *grad_image_host, float *image_host, const gsx_grads_t *g, void *stream) {
  if (!c || !v || !p) { set_error("null argument"); return GSX_E_NULL; }
  cudaStream_t st = (cudaStream_t)stream;
  int rc = check_view(c, v, p, N);   // before any copy: an oversize view must not overrun the staging buffer
  if (rc != GSX_OK) return rc;
  const size_t bytes = (size_t)v->width * v->height * 3 * sizeof(float);
  if (grad_image_host) GSX_CUDA(cudaMemcpyAsync(c->grad_image, grad_image_host, bytes, cudaMemcpyHostToDevice, st));
  rc = gsx_forward(c, v, p, N, stream);
  if (rc != GSX_OK) return rc;
  if (image_host) GSX_CUDA(cudaMemcpyAsync(image_host, c->image, bytes, cudaMemcpyDeviceToHost, st));
  if (grad_image_host && g) {
    rc = gsx_backward(c, v, p, N, c->grad_image, g, stream);
    if (rc != GSX_OK) return rc;
  }
  GSX_CUDA(cudaStreamSynchronize(st));
  return GSX_OK;
}

}  // extern "C"
