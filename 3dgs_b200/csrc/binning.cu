// binning.cu -- tile binning: exact per-Gaussian instance counts, depth sort of the Gaussians, exclusive scan,
// deterministic instance emission in depth order, stable tile sort, tile ranges.
//
// Two-level sort (B200 design; the reference sorts one 64-bit double key per instance with 8 CUB passes,
// cuda/culling.cu:455-469): the order (tile, depth bits, Gaussian id) is produced by
//   1. a stable 4-pass radix sort of the N Gaussians on their 32 order-preserving depth bits (ties keep ascending id),
//   2. emitting each Gaussian's instances in that order (exclusive scan of the counts through the permutation),
//   3. a stable radix sort of the K instances on the tile id alone: ceil(log2 T) bits = 2 passes up to 65536 tiles.
// Round 2: steps 2 and 3 of the view-level path run in "record mode" -- the per-Gaussian forward's exact tile count leaves
// an 8-byte tile record per Gaussian (hit mask of a small rectangle / where a large one's tile ids are, batch.cu), the scan
// gathers the records once into depth order, the emission expands them without any SAT and writes 16-bit tile ids, and the
// tile sort runs on 16-bit keys (k_scan_reduce_rec, k_emit_lists below; radix_sort.cu).
// Per instance this moves 2 x 16 B instead of 6 x 24 B; the 64-bit tile|depth keys are never materialised (they can
// be rebuilt for inspection, api.cu).  Result is bit-identical to sorting the 64-bit keys.  Replaces the reference's coarse_binning_kernel (run twice) + generate_splats_kernel
// + thrust temporaries + 4 blocking D2H reads (cuda/culling.cu:197-300, :386-475): the candidate window and the
// 4-axis SAT are evaluated in registers per Gaussian, so no coarse pair list (>= 25 per Gaussian) is
// ever materialised, and emission order is Gaussian-major so equal (tile, depth) keys come out in ascending
// Gaussian id after the stable sort.
#include <algorithm>
#include <climits>
#include <mutex>

#include "gsx_common.cuh"
#include "gsx_math.cuh"

namespace gsx {

int sort_pairs_impl(uint64_t *keys_a, uint64_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                    const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st,
                    int *result_in_b);
int sort_pairs_u32_impl(uint32_t *keys_a, uint32_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                        const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st,
                        int *result_in_b, int flags);
int sort_prepare_impl(long long n, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st);
int sort_tiles_with_ranges_impl(uint32_t *keys_a, uint32_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                                const int *n_dev, int key_bits, int n_tiles, int *ranges, void *workspace,
                                size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags);
int sort_tiles_with_ranges_u16_impl(uint16_t *keys_a, uint16_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                                    const int *n_dev, int key_bits, int n_tiles, int *ranges, void *workspace,
                                    size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags);
unsigned *sort_hist_impl(void *workspace);
int tile_ranges_impl(const uint64_t *sorted_keys, long long n, const int *n_dev, int n_tiles, int *ranges,
                     cudaStream_t st);
int tile_ranges_u32_impl(const uint32_t *sorted_tiles, long long n, const int *n_dev, int n_tiles, int *ranges,
                         cudaStream_t st);
size_t sort_workspace_bytes_impl(long long n);

namespace {

constexpr int kBlock = 256;
constexpr int SCAN_IPT = 8;
constexpr int SCAN_TILE = kBlock * SCAN_IPT;
// record mode (k_scan_reduce_rec / k_emit_lists): entries per thread and per block.  Measured on B200 at C4 (r2k): 4 instead
// of 8 (2600 emission blocks instead of 1300: a fuller last wave) shortens the stage timed alone, 0.364 -> 0.305 ms, but the
// pipelined step -- where the emission of view k + 1 runs under the compositing of view k and idle issue slots are filled
// anyway -- gets slower, 575 -> 567 views/s: what counts there is the instruction total, and twice the blocks repeat the
// per-block prologue.  Stays 8.
#ifndef GSX_REC_IPT
#define GSX_REC_IPT 8
#endif
constexpr int REC_IPT = GSX_REC_IPT;
constexpr int REC_TILE = kBlock * REC_IPT;

// exact instance count of one Gaussian (+ its coarse candidate count, what the reference's size query returns)
__device__ __forceinline__ int count_tiles(float u, float v, float4 r, int ntx, int nty, int *coarse) {
  const TileWindow w = candidate_window(u, v, r.x, ntx, nty);
  const int nx = max(0, w.x1 - w.x0), ny = max(0, w.y1 - w.y0);
  *coarse = nx * ny;
  if (nx * ny == 0) return 0;
  const ObbSat s = obb_sat_setup(compute_obb(u, v, r.x, r.y, r.z, r.w));
  const TileRange tr = sat_aabb_range(s, w);
  int c = 0;
  for (int tx = tr.x0; tx <= tr.x1; ++tx) {
    const SatCol col = sat_col(s, tx);
    for (int ty = tr.y0; ty <= tr.y1; ++ty) c += sat_cell(s, col, ty) ? 1 : 0;
  }
  return c;
}

__global__ void __launch_bounds__(kBlock) k_count(const float *__restrict__ uv, int uv_stride,
                                                  const float4 *__restrict__ radius,
                                                  const uint8_t *__restrict__ mask, int ntx, int nty, int N,
                                                  int *__restrict__ touched, unsigned long long *__restrict__ coarse_total) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  int coarse = 0, c = 0;
  if (i < N && (!mask || mask[i])) c = count_tiles(uv[(size_t)i * uv_stride], uv[(size_t)i * uv_stride + 1], radius[i], ntx, nty, &coarse);
  if (i < N) touched[i] = c;
  if (coarse_total) {
    unsigned long long s = (unsigned long long)coarse;
    for (int off = 16; off > 0; off >>= 1) s += __shfl_down_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(coarse_total, s);
  }
}

// depth keys of the Gaussians that touch a tile (the others sort to the end), ids as values
__global__ void __launch_bounds__(kBlock) k_depth_keys(const float *__restrict__ depth, int depth_stride,
                                                       const int *__restrict__ touched, int N,
                                                       uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  keys[i] = touched[i] > 0 ? min(float_sort_bits(depth[(size_t)i * depth_stride]), 0xfffffffeu) : 0xffffffffu;   // ~0 = "no instances"
  vals[i] = (uint32_t)i;
}

// instance emission in depth order: thread j handles Gaussian perm[j]; its instances go to [offsets[j], +touched).
// Gaussians without instances were sorted to the end (key 0xffffffff): they exit on the coalesced key read without
// touching the permutation.  Rectangles of more than COOP_CELLS tiles are expanded by the whole warp.
__global__ void __launch_bounds__(kBlock) k_emit(const float *__restrict__ uv, int uv_stride,
                                                 const float4 *__restrict__ radius, const uint32_t *__restrict__ dkeys,
                                                 const uint32_t *__restrict__ perm,
                                                 const int *__restrict__ touched, const int *__restrict__ offsets,
                                                 int ntx, int nty, int N_host, const int *__restrict__ n_dev,
                                                 const int *__restrict__ total, long long capacity,
                                                 uint32_t *__restrict__ tiles, uint32_t *__restrict__ vals, int tight,
                                                 unsigned *__restrict__ hist, int bits0, int bits1) {
  // hist (optional): the per-pass digit histograms of the tile sort that follows ([pass][256], pass p = tile-id bits
  // [8p, 8p + 8)), accumulated here so that the sort does not re-read the 4 K bytes of keys.  The low digit of the tile ids
  // of a warp's instances is spread over the bins (shared-memory atomics); the HIGH digit is the same for almost every
  // instance of a Gaussian (it changes every 256 / n_tiles_x tile rows), so it is run-length counted in registers.
  __shared__ unsigned s_hist[2 * 256];
  if (hist) { s_hist[threadIdx.x] = 0; s_hist[threadIdx.x + 256] = 0; __syncthreads(); }
  const unsigned m0 = (1u << bits0) - 1u;
  unsigned run_bin = 0xffffffffu, run_cnt = 0;
  auto count_tile = [&](uint32_t t) {
    atomicAdd(&s_hist[t & m0], 1u);
    if (bits1 > 0) {
      const unsigned hb = t >> 8;
      if (hb != run_bin) { if (run_cnt) atomicAdd(&s_hist[256 + run_bin], run_cnt); run_bin = hb; run_cnt = 0; }
      ++run_cnt;
    }
  };
  auto flush_hist = [&]() {
    if (!hist) return;
    if (run_cnt) atomicAdd(&s_hist[256 + run_bin], run_cnt);
    __syncthreads();
    if (s_hist[threadIdx.x]) atomicAdd(&hist[threadIdx.x], s_hist[threadIdx.x]);
    if (bits1 > 0 && s_hist[256 + threadIdx.x]) atomicAdd(&hist[256 + threadIdx.x], s_hist[256 + threadIdx.x]);
  };
  const int j = blockIdx.x * kBlock + threadIdx.x;
  const int lane = threadIdx.x & 31;
  // total[1] = live instance count: 0 when the view has no instances OR when K exceeds the capacity (the scan flags the
  // overflow; nothing is emitted, sorted or composited for such a view, so no stale id is ever dereferenced)
  if (total[1] == 0) return;   // block-uniform
  const int N = n_dev ? min(*n_dev, N_host) : N_host;
  bool live = j < N && dkeys[j] != 0xffffffffu;
  int i = 0, cnt = 0;
  long long o = 0;
  float u = 0.f, v = 0.f, ex = 0.f, ey = 0.f;
  float4 r = make_float4(0.f, 0.f, 0.f, 1.f);
  if (live) {
    i = (int)perm[j];
    o = offsets[j];
    cnt = touched[i];
    // the three gathers through the permutation leave together (a key != ~0 means cnt > 0)
    u = uv[(size_t)i * uv_stride]; v = uv[(size_t)i * uv_stride + 1];
    if (tight) { ex = uv[(size_t)i * uv_stride + 2]; ey = uv[(size_t)i * uv_stride + 3]; }
    r = radius[i];
    live = cnt > 0 && o + cnt <= capacity;   // overflow: detected by the host through the total
  }
  if (__any_sync(0xffffffffu, live)) {
    ObbSat s{};
    TileRange tr{0, -1, 0, -1};
    if (live) {
      const TileWindow w = candidate_window(u, v, r.x, ntx, nty);
      s = obb_sat_setup(compute_obb(u, v, r.x, r.y, r.z, r.w));
      tr = sat_aabb_range(s, w);
      // contributing-only lists (view-level path, uv_stride 4: the record is {u, v, extent_x, extent_y})
      if (tight) clip_range_to_extent(tr, u, v, ex, ey);
    }
    const bool big = live && rect_cells(tr) > COOP_CELLS;
    if (live && !big) {
      for (int tx = tr.x0; tx <= tr.x1; ++tx) {
        const SatCol col = sat_col(s, tx);
        for (int ty = tr.y0; ty <= tr.y1; ++ty)
          if (sat_cell(s, col, ty)) {
            const uint32_t t = (uint32_t)(ty * ntx + tx);
            tiles[o] = t;
            vals[o] = (uint32_t)i;
            ++o;
            if (hist) count_tile(t);
          }
      }
    }
    unsigned bm = __ballot_sync(0xffffffffu, big);
    while (bm) {
      const int leader = __ffs(bm) - 1;
      bm &= bm - 1;
      long long ol = __shfl_sync(0xffffffffu, o, leader);
      const uint32_t il = (uint32_t)__shfl_sync(0xffffffffu, i, leader);
      coop_tile_rect(leader, s, tr, [&](bool hit, unsigned ballot, int tx, int ty) {
        if (hit) {
          const long long pos = ol + __popc(ballot & ((1u << lane) - 1u));
          const uint32_t t = (uint32_t)(ty * ntx + tx);
          tiles[pos] = t;
          vals[pos] = il;
          if (hist) count_tile(t);
        }
        ol += __popc(ballot);
      });
    }
  }
  flush_hist();
}

// Emission from what the exact tile count of the per-Gaussian forward left (batch.cu): no SAT, no gathers of the
// screen-space records, one 8-byte record per Gaussian -- the hit mask of a small rectangle, or where a large one's tile ids
// are.  A warp owns 32 consecutive depth ranks; their instances are one contiguous output range.  Small rectangles: each lane
// expands its mask into a per-warp shared-memory staging buffer, the warp writes it out 32 entries per step (coalesced).
// Large rectangles: the warp copies the list.  A large rectangle without a list (the buffer was full) is enumerated with the
// SAT as k_emit does.  Compact mode only (every entry has instances).
__device__ __forceinline__ int rec_count(uint2 r) { return (r.y & 0x80000000u) ? (int)(r.y & 0x7fffffffu) : __popc(r.y); }

// Record mode: a block owns one scan tile of REC_TILE consecutive depth ranks (warp w: REC_IPT sub-chunks of 32), rebuilds their
// output offsets from the scanned block sums + its own prefix over the counts, and emits sub-chunk by sub-chunk.
// resident blocks per SM the register allocation is capped for (r2n sweep of the whole C4 step: 4 / 5 / 6 / 8 = 593.7 / 599.1 /
// 594.5 / 594.0 views/s)
#ifndef GSX_EMIT_MINB
#define GSX_EMIT_MINB 5
#endif
template <typename TileT>   // uint16_t when the tile sort that follows runs on 16-bit keys
__global__ void __launch_bounds__(kBlock, GSX_EMIT_MINB) k_emit_lists(const float *__restrict__ uv, int uv_stride,
                                                       const float4 *__restrict__ radius, const uint32_t *__restrict__ perm,
                                                       const uint2 *__restrict__ srec, const int *__restrict__ block_sums,
                                                       const uint16_t *__restrict__ tlist, int ntx, int nty, int N_host,
                                                       const int *__restrict__ n_dev, const int *__restrict__ total,
                                                       TileT *__restrict__ tiles, uint32_t *__restrict__ vals, int tight,
                                                       unsigned *__restrict__ hist, int bits0, int bits1) {
  static_assert(REC_TILE == kBlock * REC_IPT, "a warp owns REC_IPT sub-chunks of 32 entries");
  __shared__ int s_w[kBlock / 32];
  __shared__ uint16_t s_tile[kBlock / 32][32 * COOP_CELLS];   // per warp: the tiles of a sub-chunk's small rectangles
  __shared__ uint8_t s_own[kBlock / 32][32 * COOP_CELLS];     // ... and the lane each belongs to
  if (total[1] == 0) return;   // empty view or K beyond the capacity: nothing is emitted (block-uniform)
  const int n = min(*n_dev, N_host);
  const int base = blockIdx.x * REC_TILE;
  if (base >= n) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // hist (optional): the digit histograms of the tile sort that follows ([pass][256], pass p = tile-id bits [8p, 8p + 8)),
  // accumulated while the tile ids are in registers, so the sort needs no histogram pass over the 2 K bytes of keys.  Low
  // digit: one shared-memory atomic per entry (neighbouring entries differ).  High digit: the same for long runs of
  // entries (it changes every 256 / n_tiles_x tile rows), so each run of a 32-entry step adds its length once.
  __shared__ unsigned s_hist[2 * 256];
  if (hist) { s_hist[threadIdx.x] = 0u; s_hist[threadIdx.x + 256] = 0u; __syncthreads(); }
  const unsigned m0 = (1u << bits0) - 1u;
  auto count_step = [&](bool ok, unsigned t, int nvalid) {   // called by all 32 lanes; lanes [0, nvalid) hold entries
    if (ok) atomicAdd(&s_hist[t & m0], 1u);
    if (bits1 > 0) {
      const unsigned hb = ok ? (t >> 8) : 0xffffffffu;
      const unsigned prev = __shfl_up_sync(0xffffffffu, hb, 1);
      const bool head = ok && (lane == 0 || prev != hb);
      const unsigned heads = __ballot_sync(0xffffffffu, head);
      if (head) {
        const unsigned after = lane == 31 ? 0u : (heads & ~((2u << lane) - 1u));
        atomicAdd(&s_hist[256 + hb], (unsigned)((after ? __ffs(after) - 1 : nvalid) - lane));
      }
    }
  };
  const int wj0 = base + warp * (32 * REC_IPT);
  int ex[REC_IPT];   // exclusive prefix of the counts inside the warp's 256 entries
  int run = 0;
#pragma unroll
  for (int c = 0; c < REC_IPT; ++c) {
    const int j = wj0 + c * 32 + lane;
    const int cnt = j < n ? rec_count(srec[j]) : 0;
    int incl = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += x;
    }
    ex[c] = run + incl - cnt;
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) s_w[warp] = run;
  __syncthreads();
  int wbase = block_sums[blockIdx.x];
  for (int w = 0; w < warp; ++w) wbase += s_w[w];
  // the (record, id) of sub-chunk c + 1 are requested before sub-chunk c is expanded
  uint2 rec_n = make_uint2(0u, 0u);
  int i_n = 0;
  if (wj0 + lane < n) { rec_n = srec[wj0 + lane]; i_n = (int)perm[wj0 + lane]; }
#pragma unroll 1
  for (int c = 0; c < REC_IPT; ++c) {
    const int j = wj0 + c * 32 + lane;
    const bool live = j < n;
    if (!__any_sync(0xffffffffu, live)) break;
    const uint2 rec = rec_n;
    const int i = i_n;
    if (c + 1 < REC_IPT && j + 32 < n) { rec_n = srec[j + 32]; i_n = (int)perm[j + 32]; }
    int cnt = 0, o = INT_MAX;
    if (live) {
      cnt = rec_count(rec);
      o = wbase + ex[c];
    }
    if (live && (rec.y & 0x80000000u) && rec.x == 0xffffffffu) {   // large rectangle without a list: the SAT, serially (rare)
      const float u = uv[(size_t)i * uv_stride], v = uv[(size_t)i * uv_stride + 1];
      const float4 r = radius[i];
      const ObbSat s = obb_sat_setup(compute_obb(u, v, r.x, r.y, r.z, r.w));
      TileRange tr = sat_aabb_range(s, candidate_window(u, v, r.x, ntx, nty));
      if (tight) clip_range_to_extent(tr, u, v, uv[(size_t)i * uv_stride + 2], uv[(size_t)i * uv_stride + 3]);
      long long w = o;
      for (int tx = tr.x0; tx <= tr.x1; ++tx) {
        const SatCol col = sat_col(s, tx);
        for (int ty = tr.y0; ty <= tr.y1; ++ty)
          if (sat_cell(s, col, ty)) {
            const unsigned t = (unsigned)(ty * ntx + tx);
            tiles[w] = (TileT)t; vals[w] = (uint32_t)i; ++w;
            if (hist) { atomicAdd(&s_hist[t & m0], 1u); if (bits1 > 0) atomicAdd(&s_hist[256 + (t >> 8)], 1u); }
          }
      }
    }
    // small rectangles: every lane expands its hit mask into the warp's staging buffer (tile id + owner lane, at the
    // exclusive prefix of the small counts), then the warp writes the staged entries out 32 per step: consecutive entries of
    // one Gaussian go to consecutive output positions, so the stores coalesce, and no lane searches for its owner
    const bool is_big = (rec.y & 0x80000000u) != 0u;
    const int sc = (live && !is_big) ? cnt : 0;
    int incl = sc;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += x;
    }
    const int sp = incl - sc, stotal = __shfl_sync(0xffffffffu, incl, 31);
    if (sc) {
      const unsigned ny = rec.x >> 16, t00 = rec.x & 0xffffu, colmask = (1u << ny) - 1u;
      int e = sp;
      for (unsigned m = rec.y, cx = 0; m; m >>= ny, ++cx) {   // column by column, as the mask was built (batch.cu)
        unsigned col = m & colmask;
        while (col) {
          const unsigned b = (unsigned)__ffs(col) - 1u;
          col &= col - 1;
          s_tile[warp][e] = (uint16_t)(t00 + b * (unsigned)ntx + cx);
          s_own[warp][e] = (uint8_t)lane;
          ++e;
        }
      }
    }
    __syncwarp();
    for (int e0 = 0; e0 < stotal; e0 += 32) {
      const int e = e0 + lane;
      const bool ok = e < stotal;
      const int own = ok ? s_own[warp][e] : 0;
      const unsigned t = ok ? s_tile[warp][e] : 0u;
      const int oo = __shfl_sync(0xffffffffu, o, own), spo = __shfl_sync(0xffffffffu, sp, own);
      const uint32_t oi = (uint32_t)__shfl_sync(0xffffffffu, i, own);
      if (ok) {
        const int p = oo + e - spo;
        tiles[p] = (TileT)t;
        vals[p] = oi;
      }
      if (hist) count_step(ok, t, min(32, stotal - e0));
    }
    __syncwarp();   // the staging buffer is free for the next sub-chunk
    // large rectangles: the warp copies each list
    unsigned bm = __ballot_sync(0xffffffffu, live && is_big && rec.x != 0xffffffffu);
    while (bm) {
      const int L = __ffs(bm) - 1;
      bm &= bm - 1;
      const int oo = __shfl_sync(0xffffffffu, o, L), n_l = __shfl_sync(0xffffffffu, cnt, L);
      const uint32_t ts = __shfl_sync(0xffffffffu, rec.x, L), oi = (uint32_t)__shfl_sync(0xffffffffu, i, L);
      for (int k0 = 0; k0 < n_l; k0 += 32) {
        const int k = k0 + lane;
        const bool ok = k < n_l;
        const unsigned t = ok ? tlist[ts + (uint32_t)k] : 0u;
        if (ok) {
          tiles[oo + k] = (TileT)t;
          vals[oo + k] = oi;
        }
        if (hist) count_step(ok, t, min(32, n_l - k0));
      }
    }
  }
  if (hist) {
    __syncthreads();
    if (s_hist[threadIdx.x]) atomicAdd(&hist[threadIdx.x], s_hist[threadIdx.x]);
    if (bits1 > 0 && s_hist[256 + threadIdx.x]) atomicAdd(&hist[256 + threadIdx.x], s_hist[256 + threadIdx.x]);
  }
}

// Compacted depth keys arrive in nondeterministic order (atomic append, batch.cu), so Gaussians with EQUAL depth bits
// leave the stable sort in that order; the canonical order breaks depth ties by ascending Gaussian id.
// Runs of equal keys are short (a few per million): the thread at the start of a run insertion-sorts its ids.
__global__ void __launch_bounds__(kBlock) k_fix_ties(const uint32_t *__restrict__ keys, uint32_t *__restrict__ ids,
                                                     int N_host, const int *__restrict__ n_dev) {
  const int n = min(*n_dev, N_host);
  const int j = blockIdx.x * kBlock + threadIdx.x;
  if (j >= n - 1) return;
  const uint32_t k = keys[j];
  if (keys[j + 1] != k || (j > 0 && keys[j - 1] == k)) return;   // not the first element of a run of >= 2
  int e = j + 2;
  while (e < n && keys[e] == k) ++e;
  for (int a = j + 1; a < e; ++a) {
    const uint32_t x = ids[a];
    int b = a - 1;
    while (b >= j && ids[b] > x) { ids[b + 1] = ids[b]; --b; }
    ids[b + 1] = x;
  }
}

// ---- 3-kernel exclusive scan of int32 (N up to 2^31) ------------------------------------------------------
__device__ __forceinline__ int block_reduce_sum(int v, int *s_tmp) {
  for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
  if ((threadIdx.x & 31) == 0) s_tmp[threadIdx.x >> 5] = v;
  __syncthreads();
  int t = 0;
  if (threadIdx.x < 32) {
    t = threadIdx.x < (kBlock / 32) ? s_tmp[threadIdx.x] : 0;
    for (int off = 16; off > 0; off >>= 1) t += __shfl_down_sync(0xffffffffu, t, off);
  }
  return t;   // valid in thread 0
}

// value i of the scanned sequence: in[i], or in[perm[i]] when a permutation is given
__device__ __forceinline__ int scan_value(const int *__restrict__ in, const uint32_t *__restrict__ perm,
                                          const uint32_t *__restrict__ dkeys, long long i) {
  if (!perm) return in[i];
  if (dkeys[i] == 0xffffffffu) return 0;   // no instances: sorted to the end, no gather needed
  return in[perm[i]];
}

__global__ void __launch_bounds__(kBlock) k_scan_reduce(const int *__restrict__ in, const uint32_t *__restrict__ perm,
                                                        const uint32_t *__restrict__ dkeys, int N_host,
                                                        const int *__restrict__ n_dev, int *__restrict__ block_sums) {
  __shared__ int s_tmp[kBlock / 32];
  const int N = n_dev ? min(*n_dev, N_host) : N_host;
  const long long base = (long long)blockIdx.x * SCAN_TILE;
  int v = 0;
#pragma unroll
  for (int k = 0; k < SCAN_IPT; ++k) {
    const long long i = base + k * kBlock + threadIdx.x;
    if (i < N) v += scan_value(in, perm, dkeys, i);
  }
  const int t = block_reduce_sum(v, s_tmp);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = t;
}

// record mode: the tile count of entry j is in the 8-byte record of Gaussian perm[j] (batch.cu); the record is gathered
// once, here, and left in depth order for the emission kernel
__global__ void __launch_bounds__(kBlock) k_scan_reduce_rec(const uint32_t *__restrict__ perm, const uint2 *__restrict__ trec,
                                                            int N_host, const int *__restrict__ n_dev, uint2 *__restrict__ srec,
                                                            int *__restrict__ block_sums) {
  __shared__ int s_tmp[kBlock / 32];
  const int N = min(*n_dev, N_host);
  const long long base = (long long)blockIdx.x * REC_TILE;
  int v = 0;
#pragma unroll
  for (int k = 0; k < REC_IPT; ++k) {
    const long long j = base + k * kBlock + threadIdx.x;
    if (j < N) {
      const uint2 r = trec[perm[j]];
      srec[j] = r;
      v += rec_count(r);
    }
  }
  const int t = block_reduce_sum(v, s_tmp);
  if (threadIdx.x == 0) block_sums[blockIdx.x] = t;
}

// total[0] = K (saturated at INT_MAX), total[1] = live count = K if K <= capacity else 0; an overflow is also recorded in
// *over_flag (pinned host memory mapped into the device, polled by gsx_ctx_overflow without any synchronisation)
__global__ void __launch_bounds__(1024) k_scan_block_sums(int *__restrict__ block_sums, int nb, int *__restrict__ total,
                                                          long long capacity, int *over_flag) {
  __shared__ int s[1024];
  __shared__ long long carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < nb; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const int v = i < nb ? block_sums[i] : 0;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
      const int t = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
      __syncthreads();
      s[threadIdx.x] += t;
      __syncthreads();
    }
    if (i < nb) block_sums[i] = (int)min(carry + s[threadIdx.x] - v, (long long)INT_MAX);
    __syncthreads();
    if (threadIdx.x == 1023) carry += s[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const int K = (int)min(carry, (long long)INT_MAX);
    const bool over = capacity > 0 && carry > capacity;
    total[0] = K;
    total[1] = over ? 0 : K;
    if (over && over_flag && K > *(volatile int *)over_flag) *(volatile int *)over_flag = K;
  }
}

__global__ void __launch_bounds__(kBlock) k_scan_apply(const int *__restrict__ in, const uint32_t *__restrict__ perm,
                                                       const uint32_t *__restrict__ dkeys, int N_host,
                                                       const int *__restrict__ n_dev,
                                                       const int *__restrict__ block_sums, int *__restrict__ out) {
  __shared__ int s_warp[kBlock / 32];
  const int N = n_dev ? min(*n_dev, N_host) : N_host;
  const long long base = (long long)blockIdx.x * SCAN_TILE + (long long)threadIdx.x * SCAN_IPT;
  int v[SCAN_IPT];
  int sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_IPT; ++k) { v[k] = (base + k < N) ? scan_value(in, perm, dkeys, base + k) : 0; sum += v[k]; }
  int incl = sum;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, off);
    if (lane >= off) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int wbase = 0;
  for (int w = 0; w < warp; ++w) wbase += s_warp[w];
  int run = block_sums[blockIdx.x] + wbase + incl - sum;
#pragma unroll
  for (int k = 0; k < SCAN_IPT; ++k) {
    if (base + k < N) out[base + k] = run;
    run += v[k];
  }
}

__global__ void k_copy_u32_to_i32(const uint32_t *__restrict__ in, int *__restrict__ out, long long n_host,
                                  const int *__restrict__ n_dev) {
  const long long n = n_dev ? min((long long)*n_dev, n_host) : n_host;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = (int)in[i];
}

}  // namespace

int tile_key_bits(int n_tiles) {
  int b = 0;
  while ((1 << b) < n_tiles) ++b;
  return 32 + std::max(b, 1);
}

int exclusive_scan_i32(const int *in, const uint32_t *perm, const uint32_t *dkeys, int N, const int *n_dev, int *out,
                       int *block_sums, int *total_dev, cudaStream_t st, long long capacity, int *over_flag) {
  const int nb = std::max(1, ceil_div(N, SCAN_TILE));
  k_scan_reduce<<<nb, kBlock, 0, st>>>(in, perm, dkeys, N, n_dev, block_sums);
  GSX_LAUNCH_CHECK();
  k_scan_block_sums<<<1, 1024, 0, st>>>(block_sums, nb, total_dev, capacity, over_flag);
  GSX_LAUNCH_CHECK();
  k_scan_apply<<<nb, kBlock, 0, st>>>(in, perm, dkeys, N, n_dev, block_sums, out);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}
size_t scan_block_sums_count(int N) { return (size_t)std::max(1, ceil_div(N, std::min(SCAN_TILE, REC_TILE))); }

// Workspace carve-up shared by the per-function F7 entry point and the view-level context.
BinLayout bin_layout(int N, long long capacity) {
  BinLayout L;
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off += align_up(std::max<size_t>(b, 16), 256); return o; };
  L.touched = take(sizeof(int) * (size_t)N);
  L.offsets = take(sizeof(int) * (size_t)N);
  L.block_sums = take(sizeof(int) * scan_block_sums_count(N));
  L.total = take(sizeof(int) * 4);
  L.dk_a = take(sizeof(uint32_t) * (size_t)N);
  L.dk_b = take(sizeof(uint32_t) * (size_t)N);
  L.dv_a = take(sizeof(uint32_t) * (size_t)N);
  L.dv_b = take(sizeof(uint32_t) * (size_t)N);
  L.keys_a = take(sizeof(uint32_t) * (size_t)capacity);
  L.keys_b = take(sizeof(uint32_t) * (size_t)capacity);
  L.vals_a = take(sizeof(uint32_t) * (size_t)capacity);
  L.vals_b = take(sizeof(uint32_t) * (size_t)capacity);
  L.sort_ws_bytes = sort_workspace_bytes_impl(std::max<long long>(capacity, N));
  L.sort_ws = take(L.sort_ws_bytes);
  L.bytes = off;
  return L;
}

// ---- phase 1 (pre-binning): depth sort -> tie fix-up -> scan of the tile counts in depth order.
// compact mode (ckey given): key_a / val_a are the slot's compacted, unordered depth keys + ids (*ccount entries) written
// by the per-Gaussian kernel, so everything here runs over P' entries; otherwise keys are derived from depth[] over N.
int bin_prebin_impl(const float *depth, int depth_stride, const int *touched, int N, bool compact, const int *ccount,
                    PreBin *pb, cudaStream_t st, long long capacity, const uint2 *trec) {
  int in_b = 0, rc;
  pb->perm = nullptr; pb->dkeys = nullptr; pb->count = nullptr; pb->rec_mode = false;
  if (N > 0) {
    if (compact) {
      pb->count = ccount;
    } else {
      k_depth_keys<<<ceil_div(N, kBlock), kBlock, 0, st>>>(depth, depth_stride, touched, N, pb->key_a, pb->val_a);
      GSX_LAUNCH_CHECK();
    }
    rc = sort_pairs_u32_impl(pb->key_a, pb->key_b, pb->val_a, pb->val_b, N, pb->count, 32, pb->sort_ws, pb->sort_ws_bytes, st, &in_b, 0);
    if (rc != GSX_OK) return rc;
    uint32_t *sk = in_b ? pb->key_b : pb->key_a, *sv = in_b ? pb->val_b : pb->val_a;
    if (compact) {
      k_fix_ties<<<ceil_div(N, kBlock), kBlock, 0, st>>>(sk, sv, N, ccount);
      GSX_LAUNCH_CHECK();
    }
    pb->perm = sv; pb->dkeys = sk;
  }
  // record mode: one gather of the 8-byte tile records (left in depth order), block sums, totals; the emission kernel
  // rebuilds the offsets itself.  The direct-binning experiment (GSX_BINNING=direct) needs the offsets array: full scan.
  const char *e = getenv("GSX_BINNING");
  if (compact && trec && pb->srec && N > 0 && !(e && e[0] == 'd')) {
    const int nb = std::max(1, ceil_div(N, REC_TILE));
    k_scan_reduce_rec<<<nb, kBlock, 0, st>>>(pb->perm, trec, N, pb->count, pb->srec, pb->block_sums);
    GSX_LAUNCH_CHECK();
    k_scan_block_sums<<<1, 1024, 0, st>>>(pb->block_sums, nb, pb->total, capacity, pb->over_flag);
    GSX_LAUNCH_CHECK();
    pb->rec_mode = true;
    return GSX_OK;
  }
  return exclusive_scan_i32(touched, pb->perm, pb->dkeys, N, pb->count, pb->offsets, pb->block_sums, pb->total, st, capacity,
                            pb->over_flag);
}

// ---- phase 2: emit in depth order -> tile sort -> ranges.  Leaves the sorted tile ids / Gaussian ids in
// (*sorted_tiles, *sorted_vals) inside the workspace; K is pb.total[0].
// ev (optional): 3 events recorded after emission, after the tile sort and after the ranges (stage timing)
int bin_emit_sort_impl(const float *uv, int uv_stride, const float4 *radius, const int *touched, int ntx, int nty, int N,
                       long long capacity, char *ws, const BinLayout &L, const PreBin &pb, int *ranges,
                       const uint32_t **sorted_tiles, const uint32_t **sorted_vals, cudaStream_t st, cudaEvent_t *ev,
                       int tight, const uint2 *trec, const uint16_t *tlist) {
  uint32_t *ka = (uint32_t *)(ws + L.keys_a), *kb = (uint32_t *)(ws + L.keys_b);
  uint32_t *va = (uint32_t *)(ws + L.vals_a), *vb = (uint32_t *)(ws + L.vals_b);
  int in_b = 0, rc;
  // Direct tile binning (round 2, tilebin.cu, GSX_BINNING=direct): the instance ids go straight to their final place by one
  // stable counting sort on the tile id; its count matrices live in the key buffers the sort path would use.  Bit-identical
  // lists (tests), but measured SLOWER at C4 on B200 (profiles/r2i_*): count pass 0.57 ms + ordered emission 0.83 ms against
  // 0.23 + 0.07 + 0.33 ms for emission + histogram + two radix passes -- the exact SAT enumeration runs twice (152 M + 279 M warp
  // instructions at 17 % occupancy: one CTA of 11 warps per SM holds the 16-bit tile counters), and the scattered 4-byte
  // stores do not merge in L2 (433 MB of DRAM writes + as much read-for-fill for 84 MB of ids).  So emission + radix sort stays
  // the default; the direct path is kept as the measured alternative.
  {
    const char *e = getenv("GSX_BINNING");
    TileBinPlan plan;
    if (e && e[0] == 'd' && N > 0 && tilebin_plan(ntx, nty, N, capacity, L.vals_a - L.keys_a, &plan)) {
      rc = tilebin_run(plan, uv, uv_stride, radius, pb.dkeys, pb.perm, pb.offsets, N, pb.count, pb.total, ntx, nty,
                       (tight && uv_stride == 4) ? 1 : 0, ws + L.keys_a, ranges, va, st, ev);
      if (rc != GSX_OK) return rc;
      *sorted_tiles = nullptr;
      *sorted_vals = va;
      return GSX_OK;
    }
  }
  const int key_bits = tile_key_bits(ntx * nty) - 32;
  // Measured on B200 (profiles/r2c_*): the sort loses its 0.057 ms histogram pass, the emission kernel gains 0.050 ms of
  // shared-memory atomics (20.9 M instances): a wash, so the separate histogram pass stays the default (GSX_FUSED_HIST=1).
  const bool fuse_hist = key_bits <= 16 && getenv("GSX_FUSED_HIST") != nullptr;
  // tile ranges: extracted by the last scatter of the tile sort (round 2; GSX_RANGES=separate: the stand-alone kernel that
  // re-reads the 4 K bytes of sorted tile ids)
  static const bool fused_ranges = [] { const char *e = getenv("GSX_RANGES"); return !(e && e[0] == 's'); }();
  // record mode + fused ranges: nobody needs 32-bit tile keys -- emission and sort run on 16-bit ones (GSX_KEYS16=0: A/B)
  const bool keys16 = pb.rec_mode && !fuse_hist && fused_ranges && key_bits <= 16 && N > 0 &&
                      !(getenv("GSX_KEYS16") && getenv("GSX_KEYS16")[0] == '0');
  // record mode, 16-bit keys: the emission accumulates the tile sort's digit histograms (GSX_EMIT_HIST=0: separate pass)
  const bool emit_hist = keys16 && !(getenv("GSX_EMIT_HIST") && getenv("GSX_EMIT_HIST")[0] == '0');
  if (N > 0) {
    // the tile sort's digit histograms are accumulated by the emission kernel (tile ids of up to 16 bits: 65536 tiles)
    if (fuse_hist) {
      rc = sort_prepare_impl(capacity, key_bits, ws + L.sort_ws, L.sort_ws_bytes, st);
      if (rc != GSX_OK) return rc;
    }
    if (pb.rec_mode && !fuse_hist && keys16) {
      if (emit_hist) {
        rc = sort_prepare_impl(capacity, key_bits, ws + L.sort_ws, L.sort_ws_bytes, st);
        if (rc != GSX_OK) return rc;
      }
      k_emit_lists<uint16_t><<<ceil_div(N, REC_TILE), kBlock, 0, st>>>(uv, uv_stride, radius, pb.perm, pb.srec, pb.block_sums, tlist,
                                                                      ntx, nty, N, pb.count, pb.total, (uint16_t *)ka, va,
                                                                      (tight && uv_stride == 4) ? 1 : 0,
                                                                      emit_hist ? sort_hist_impl(ws + L.sort_ws) : nullptr,
                                                                      std::min(key_bits, 8), std::max(key_bits - 8, 0));
    } else if (pb.rec_mode && !fuse_hist)
      k_emit_lists<uint32_t><<<ceil_div(N, REC_TILE), kBlock, 0, st>>>(uv, uv_stride, radius, pb.perm, pb.srec, pb.block_sums, tlist,
                                                                      ntx, nty, N, pb.count, pb.total, ka, va,
                                                                      (tight && uv_stride == 4) ? 1 : 0, nullptr, 8, 0);
    else
    k_emit<<<ceil_div(N, kBlock), kBlock, 0, st>>>(uv, uv_stride, radius, pb.dkeys, pb.perm, touched, pb.offsets, ntx, nty, N,
                                                   pb.count, pb.total, capacity, ka, va, (tight && uv_stride == 4) ? 1 : 0,
                                                   fuse_hist ? sort_hist_impl(ws + L.sort_ws) : nullptr, std::min(key_bits, 8),
                                                   std::max(key_bits - 8, 0));
    GSX_LAUNCH_CHECK();
  }
  if (ev) GSX_CUDA(cudaEventRecord(ev[0], st));
  if (keys16) {
    rc = sort_tiles_with_ranges_u16_impl((uint16_t *)ka, (uint16_t *)kb, va, vb, capacity, pb.total + 1, key_bits, ntx * nty, ranges,
                                         ws + L.sort_ws, L.sort_ws_bytes, st, &in_b, emit_hist ? 1 : 0);
    if (rc != GSX_OK) return rc;
    *sorted_tiles = nullptr;   // never written (gsx_ctx_build_sorted_keys rebuilds the tile from the ranges)
    *sorted_vals = in_b ? vb : va;
    if (ev) { GSX_CUDA(cudaEventRecord(ev[1], st)); GSX_CUDA(cudaEventRecord(ev[2], st)); }
    return GSX_OK;
  }
  if (fused_ranges) {
    rc = sort_tiles_with_ranges_impl(ka, kb, va, vb, capacity, pb.total + 1, key_bits, ntx * nty, ranges, ws + L.sort_ws,
                                     L.sort_ws_bytes, st, &in_b, (N > 0 && fuse_hist) ? 1 : 0);
    if (rc != GSX_OK) return rc;
    *sorted_tiles = in_b ? kb : ka;
    *sorted_vals = in_b ? vb : va;
    if (ev) { GSX_CUDA(cudaEventRecord(ev[1], st)); GSX_CUDA(cudaEventRecord(ev[2], st)); }
    return GSX_OK;
  }
  rc = sort_pairs_u32_impl(ka, kb, va, vb, capacity, pb.total + 1, key_bits, ws + L.sort_ws, L.sort_ws_bytes, st, &in_b,
                           (N > 0 && fuse_hist) ? 1 : 0);
  if (rc != GSX_OK) return rc;
  *sorted_tiles = in_b ? kb : ka;
  *sorted_vals = in_b ? vb : va;
  if (ev) GSX_CUDA(cudaEventRecord(ev[1], st));
  rc = tile_ranges_u32_impl(*sorted_tiles, capacity, pb.total + 1, ntx * nty, ranges, st);
  if (ev) GSX_CUDA(cudaEventRecord(ev[2], st));
  return rc;
}

// both phases on one stream out of one workspace (per-function F7 path)
int bin_sort_impl(const float *uv, int uv_stride, const float *depth, int depth_stride, const float4 *radius,
                  const int *touched, int ntx, int nty, int N, long long capacity, char *ws, const BinLayout &L, int *ranges,
                  const uint32_t **sorted_tiles, const uint32_t **sorted_vals, cudaStream_t st) {
  PreBin pb;
  pb.key_a = (uint32_t *)(ws + L.dk_a); pb.key_b = (uint32_t *)(ws + L.dk_b);
  pb.val_a = (uint32_t *)(ws + L.dv_a); pb.val_b = (uint32_t *)(ws + L.dv_b);
  pb.offsets = (int *)(ws + L.offsets); pb.block_sums = (int *)(ws + L.block_sums); pb.total = (int *)(ws + L.total);
  pb.sort_ws = ws + L.sort_ws; pb.sort_ws_bytes = L.sort_ws_bytes;
  int rc = bin_prebin_impl(depth, depth_stride, touched, N, false, nullptr, &pb, st, capacity, nullptr);
  if (rc != GSX_OK) return rc;
  return bin_emit_sort_impl(uv, uv_stride, radius, touched, ntx, nty, N, capacity, ws, L, pb, ranges, sorted_tiles, sorted_vals, st,
                            nullptr, 0, nullptr, nullptr);
}

size_t prebin_sort_ws_bytes(int N) { return sort_workspace_bytes_impl(N); }
size_t prebin_block_sums_count(int N) { return scan_block_sums_count(N); }

int count_tiles_launch(const float *uv, int uv_stride, const float4 *radius, const uint8_t *mask, int ntx, int nty, int N, int *touched,
                       unsigned long long *coarse_total, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  k_count<<<ceil_div(N, kBlock), kBlock, 0, st>>>(uv, uv_stride, radius, mask, ntx, nty, N, touched, coarse_total);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int copy_ids_launch(const uint32_t *in, int *out, long long n, const int *n_dev, cudaStream_t st) {
  if (n <= 0) return GSX_OK;
  const int blocks = (int)std::min<long long>(kNumSMs * 8, (n + 255) / 256);
  k_copy_u32_to_i32<<<blocks, 256, 0, st>>>(in, out, n, n_dev);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

// small cached internal workspace for callers that pass none (the reference API has no workspace argument)
// One per device, grown under a mutex.  A grown buffer replaces the old one only after the device is idle (kernels of an
// earlier call may still be queued on it).  Callers that share a device between host threads must pass their own
// workspace (gsx.h): the cached one is handed out without a lease.
static std::mutex g_ws_mu;
static char *g_ws[2][64] = {};
static size_t g_ws_bytes[2][64] = {};
int internal_workspace_slot(int which, size_t bytes, char **out) {   // which: 0 = sorted list (F7), 1 = fused_loss shim
  int dev = 0;
  GSX_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64 || which < 0 || which > 1) { set_error("device index %d out of range", dev); return GSX_E_ARG; }
  std::lock_guard<std::mutex> lk(g_ws_mu);
  if (bytes > g_ws_bytes[which][dev]) {
    if (g_ws[which][dev]) { GSX_CUDA(cudaDeviceSynchronize()); cudaFree(g_ws[which][dev]); }
    g_ws[which][dev] = nullptr; g_ws_bytes[which][dev] = 0;
    GSX_CUDA(cudaMalloc(&g_ws[which][dev], bytes));
    g_ws_bytes[which][dev] = bytes;
  }
  *out = g_ws[which][dev];
  return GSX_OK;
}
int internal_workspace(size_t bytes, char **out) { return internal_workspace_slot(0, bytes, out); }

}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_sorted_list_size(const float *uv, const float *radius, int n_tiles_x, int n_tiles_y, int N,
                         long long *coarse_pairs_host, long long *num_instances_host, void *stream) {
  GSX_REQUIRE_DEV(uv); GSX_REQUIRE_DEV(radius);
  GSX_REQUIRE_ALIGNED(radius, 16);
  cudaStream_t st = (cudaStream_t)stream;
  if (coarse_pairs_host) *coarse_pairs_host = 0;
  if (num_instances_host) *num_instances_host = 0;
  if (N <= 0) return GSX_OK;
  const BinLayout L = bin_layout(N, 1);
  char *ws;
  int rc = internal_workspace(L.bytes + 256, &ws);
  if (rc != GSX_OK) return rc;
  unsigned long long *coarse = (unsigned long long *)(ws + L.bytes);
  GSX_CUDA(cudaMemsetAsync(coarse, 0, sizeof(unsigned long long), st));
  rc = count_tiles_launch(uv, 2, (const float4 *)radius, nullptr, n_tiles_x, n_tiles_y, N, (int *)(ws + L.touched), coarse, st);
  if (rc != GSX_OK) return rc;
  rc = exclusive_scan_i32((int *)(ws + L.touched), nullptr, nullptr, N, nullptr, (int *)(ws + L.offsets), (int *)(ws + L.block_sums),
                          (int *)(ws + L.total), st, 0, nullptr);
  if (rc != GSX_OK) return rc;
  unsigned long long hc = 0;
  int hk = 0;
  GSX_CUDA(cudaMemcpyAsync(&hc, coarse, sizeof(hc), cudaMemcpyDeviceToHost, st));
  GSX_CUDA(cudaMemcpyAsync(&hk, ws + L.total, sizeof(int), cudaMemcpyDeviceToHost, st));
  GSX_CUDA(cudaStreamSynchronize(st));
  if (coarse_pairs_host) *coarse_pairs_host = (long long)hc;
  if (num_instances_host) *num_instances_host = hk;
  return GSX_OK;
}

size_t gsx_sorted_list_workspace_bytes(int N, long long capacity) {
  return bin_layout(std::max(N, 1), std::max<long long>(capacity, 1)).bytes;
}

int gsx_get_sorted_gaussian_list(const float *uv, const float *xyz_c, const float *radius, int n_tiles_x,
                                 int n_tiles_y, int N, long long capacity, int *sorted_gaussians, int *ranges,
                                 int *num_instances_dev, void *workspace, size_t workspace_bytes, void *stream) {
  GSX_REQUIRE_DEV(uv); GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(radius); GSX_REQUIRE_DEV(ranges);
  GSX_REQUIRE_ALIGNED(radius, 16);
  cudaStream_t st = (cudaStream_t)stream;
  const int T = n_tiles_x * n_tiles_y;
  if (N <= 0 || capacity <= 0) {
    GSX_CUDA(cudaMemsetAsync(ranges, 0, sizeof(int) * (size_t)(T + 1), st));
    if (num_instances_dev) GSX_CUDA(cudaMemsetAsync(num_instances_dev, 0, sizeof(int), st));
    return GSX_OK;
  }
  GSX_REQUIRE_DEV(sorted_gaussians);
  const BinLayout L = bin_layout(N, capacity);
  char *ws = (char *)workspace;
  if (!ws) {
    int rc = internal_workspace(L.bytes, &ws);
    if (rc != GSX_OK) return rc;
  } else if (workspace_bytes < L.bytes) {
    set_error("sorted-list workspace %zu < %zu", workspace_bytes, L.bytes);
    return GSX_E_WORKSPACE;
  }
  int rc = count_tiles_launch(uv, 2, (const float4 *)radius, nullptr, n_tiles_x, n_tiles_y, N, (int *)(ws + L.touched),
                              nullptr, st);
  if (rc != GSX_OK) return rc;
  const uint32_t *sk;
  const uint32_t *sv;
  rc = bin_sort_impl(uv, 2, xyz_c + 2, 3, (const float4 *)radius, (const int *)(ws + L.touched), n_tiles_x, n_tiles_y, N,
                     capacity, ws, L, ranges, &sk, &sv, st);
  if (rc != GSX_OK) return rc;
  rc = copy_ids_launch(sv, sorted_gaussians, capacity, (const int *)(ws + L.total) + 1, st);
  if (rc != GSX_OK) return rc;
  if (num_instances_dev)
    GSX_CUDA(cudaMemcpyAsync(num_instances_dev, ws + L.total, sizeof(int), cudaMemcpyDeviceToDevice, st));
  return GSX_OK;
}

}  // extern "C"
