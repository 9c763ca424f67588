// radix_sort.cu -- hand-written onesweep LSD radix sort of (uint64 key, uint32 value) pairs, plus tile-range
// extraction.  Replaces cub::DeviceRadixSort::SortPairs<double,int> over 64 bits + find_tile_boundaries_kernel
// (reference cuda/culling.cu:455-469, :302-343).  Keys are tile<<32 | order-preserving depth bits, so only
// 32 + ceil(log2 T) bits are sorted (6 passes of 8 bits up to 65536 tiles instead of the reference's 8).
//
// Structure (one read of the keys for all histograms, then one read + one write per pass):
//   k_histogram      : per-pass 256-bin digit histograms, shared-memory privatised, one global RED per bin/block
//   k_scan_histogram : exclusive scan of each pass's 256 bins -> global digit bases
//   k_onesweep       : per 4096-key tile: warp-synchronous ballot ranking (stable), decoupled look-back of the
//                      per-digit tile prefixes through a (flag|count) status word, shared-memory reorder so the
//                      global scatter is coalesced per digit run.  Tile ids come from an atomic ticket so a tile
//                      only ever waits on tiles that are already resident (no deadlock).
// HBM-bound: algorithmic bytes = 8n (histogram read) + passes * 24n.  The live count may live on the device
// (n_dev) so the view-level path never syncs with the host.
// The kernels are templated on the key type: the view-level path sorts 32-bit keys twice (depth per Gaussian, then
// tile id per instance, binning.cu) instead of one 64-bit tile|depth key per instance.
#include <limits.h>
#include <stdlib.h>

#include <algorithm>

#include "gsx_common.cuh"

namespace gsx {

namespace {

// GSX_RS_LATE_AGG=1 (build variant): no tile histogram by shared-memory atomics; the aggregate is published after the
// ranking from the per-warp digit counts.  Measured on B200 at C4 (r2k): 16 fewer shared atomics per thread, but successors'
// look-back then waits for the ranking: tile sort 0.356 -> 0.411 ms, depth sort + scan + emission 0.363 -> 0.390 ms.  Stays 0.
#ifndef GSX_RS_LB
#define GSX_RS_LB 8   // predecessors fetched per look-back round (r2o sweep of the C4 step: 4 / 8 / 16 = 599 / 598 / 587 views/s)
#endif
#ifndef GSX_RS_MATCH_ATOMIC
#define GSX_RS_MATCH_ATOMIC 1
#endif
#ifndef GSX_RS_LATE_AGG
#define GSX_RS_LATE_AGG 0
#endif
constexpr int RADIX = 256;
constexpr int MAX_PASSES = 8;
constexpr uint32_t FLAG_AGG = 1u << 30, FLAG_INC = 2u << 30, FLAG_MASK = 3u << 30, VAL_MASK = ~FLAG_MASK;

// Tile shapes of the onesweep kernel (threads x keys per thread).  The per-tile fixed costs -- zeroing and scanning 256
// digit counters per warp, the look-back over 256 status words -- are amortised over the tile, so larger tiles raise the
// bandwidth per pass until registers / shared memory cut the occupancy (CUB's sm_100 tuning for 4-byte keys + 4-byte values
// uses 384-512 threads x ~20 keys, tuning_radix_sort.cuh:92-160).  GSX_RS_CFG=0..3 selects one at run time for A/B runs.
struct RsCfg { int threads, ipt; };
constexpr RsCfg kRsCfg[6] = {{256, 16}, {512, 16}, {384, 20}, {512, 12}, {256, 20}, {256, 24}};
// Measured on B200 against CUB Policy1000 on the C4 keys (profiles/r2b_*sort_vs_cub*): 4-byte keys, 20.9 M pairs: 512 x 16
// is the fastest shape; 2.7 M pairs (one wave of tiles, latency-bound): 512 x 12; 8-byte keys: 256 x 16.
static int rs_cfg(int key_bytes, long long n) {
  static const int forced = [] {
    const char *e = getenv("GSX_RS_CFG");
    const int v = e ? atoi(e) : -1;
    return (v >= 0 && v < 6) ? v : -1;
  }();
  if (forced >= 0) return forced;
  // Round 2, after the ranking moved to shared-memory peer masks (r2m A/B of the whole C4 step, views/s): 256 x 16 for every
  // sort 582.6, 512 x 16 578.8, 384 x 20 578.9, 512 x 12 571.6 (with the ballot ranking the large shapes had won: 512 x 16
  // for the 20.9 M instances, 512 x 12 for the 2.7 M depth keys).
  // 256 x 20 against 256 x 16 / 256 x 24, same A/B: 585.0 / 582.3 / 579.3 (2- and 4-byte keys; 8-byte keys keep 256 x 16)
  (void)n;
  return key_bytes == 8 ? 0 : 4;
}
static int rs_tile(int key_bytes, long long n) { const int c = rs_cfg(key_bytes, n); return kRsCfg[c].threads * kRsCfg[c].ipt; }

__device__ __forceinline__ long long live_count(const int *n_dev, long long n) {
  if (n_dev) { const long long v = *n_dev; return v < n ? v : n; }
  return n;
}

template <typename KeyT>
__global__ void __launch_bounds__(256) k_histogram(const KeyT *__restrict__ keys, long long n_host,
                                                   const int *__restrict__ n_dev, int passes, int key_bits,
                                                   unsigned int *__restrict__ hist /*[passes][256]*/) {
  __shared__ unsigned int sh[MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const long long n = live_count(n_dev, n_host);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const KeyT k = keys[i];
    for (int p = 0; p < passes; ++p) {
      const int bits = min(8, key_bits - 8 * p);
      const unsigned d = (unsigned)(k >> (8 * p)) & ((1u << bits) - 1u);
      atomicAdd(&sh[p * RADIX + d], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

__global__ void __launch_bounds__(RADIX) k_scan_histogram(unsigned int *__restrict__ hist) {
  __shared__ unsigned int s[RADIX];
  unsigned int *h = hist + blockIdx.x * RADIX;
  const unsigned v = h[threadIdx.x];
  s[threadIdx.x] = v;
  __syncthreads();
  for (int off = 1; off < RADIX; off <<= 1) {
    const unsigned t = threadIdx.x >= off ? s[threadIdx.x - off] : 0u;
    __syncthreads();
    s[threadIdx.x] += t;
    __syncthreads();
  }
  h[threadIdx.x] = s[threadIdx.x] - v;
}

__device__ __forceinline__ unsigned ld_volatile(const unsigned *p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_volatile(unsigned *p, unsigned v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// the tile's work after its extent is known; FULL: all THREADS * IPT slots hold keys (every tile but the last), so the bounds
// predicates and the padding digit of the ranking disappear
template <typename KeyT, int THREADS, int IPT, bool FULL>
__device__ __forceinline__ void onesweep_tile(const KeyT *__restrict__ keys_in, KeyT *__restrict__ keys_out,
                                              const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out, int shift,
                                              int bits, const unsigned int *__restrict__ digit_base, unsigned int *__restrict__ status,
                                              int *__restrict__ first_of_key, int key_clamp, int write_keys, unsigned tile,
                                              long long base, int valid, unsigned mask, KeyT *s_keys,
                                              unsigned int (*s_warp)[RADIX + 1], unsigned int (*s_match)[RADIX + 1],
                                              unsigned int *s_excl, int *s_gbase, unsigned int *s_scan) {
  constexpr int WARPS = THREADS / 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  KeyT key[IPT];
  unsigned pos[IPT];
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp * (IPT * 32) + i * 32 + lane;
    key[i] = (FULL || li < valid) ? keys_in[base + li] : ~(KeyT)0;
  }
  // 1. (round 1: a tile histogram by shared-memory atomics came first, so that the aggregate was published before the
  //    ranking.  GSX_RS_LATE_AGG builds: the aggregate is the sum of the per-warp digit counts the ranking leaves anyway.)
#if !GSX_RS_LATE_AGG
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp * (IPT * 32) + i * 32 + lane;
    if (FULL || li < valid) atomicAdd(&s_excl[(unsigned)(key[i] >> shift) & mask], 1u);
  }
  __syncthreads();
  const unsigned total = tid < RADIX ? s_excl[tid] : 0u;   // for digit == tid
  unsigned *my = status + (size_t)tile * RADIX + tid;
  if (tid < RADIX) st_volatile(my, (tile == 0 ? FLAG_INC : FLAG_AGG) | total);
#else
  unsigned total = 0u;
  unsigned *my = status + (size_t)tile * RADIX + tid;
#endif
  // 2. stable rank inside the warp: items in order i, lanes in order; peers found with one ballot per digit bit
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp * (IPT * 32) + i * 32 + lane;
    const unsigned d = (FULL || li < valid) ? ((unsigned)(key[i] >> shift) & mask) : (unsigned)RADIX;
#if GSX_RS_MATCH_ATOMIC
    // peers = lanes holding the same digit, found through shared memory: every lane ORs its lane bit into the warp's word of
    // its digit, the word IS the peer mask; the group's first lane clears it for the next item while it advances the
    // warp's digit counter.  3 shared-memory operations instead of one vote per digit bit, which cost 9 SASS instructions
    // each (shift, mask, compare, VOTE, select, and: 72 of the ~130 instructions per item of an 8-bit pass).
    unsigned *mw = &s_match[warp][d];
    atomicOr(mw, 1u << lane);
    __syncwarp();
    const unsigned peers = *reinterpret_cast<volatile unsigned *>(mw);
    const unsigned lt = peers & ((1u << lane) - 1u);
    const unsigned before = s_warp[warp][d];
    __syncwarp();
    if (lt == 0) { *mw = 0u; s_warp[warp][d] = before + __popc(peers); }
    __syncwarp();
    pos[i] = before + __popc(lt);
  }
#else
    // peers = lanes holding the same digit: one vote per digit bit (+ one for the padding digit RADIX, bit 8).
    // Measured against match.any on B200: tile sort 0.427 -> 0.399 ms per view.
    unsigned peers = 0xffffffffu;
    if (!FULL) {   // a full tile has no padding items
      const bool pad = d >> 8;
      const unsigned vp = __ballot_sync(0xffffffffu, pad);
      peers = pad ? vp : ~vp;
    }
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      if (b < bits) {   // uniform
        const bool bit = (d >> b) & 1u;
        const unsigned vote = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? vote : ~vote;
      }
    }
    const unsigned lt = peers & ((1u << lane) - 1u);
    const unsigned before = s_warp[warp][d];   // (a leader-lane shared atomicAdd + shuffle here measured 38 % slower)
    __syncwarp();
    if (lt == 0) s_warp[warp][d] = before + __popc(peers);
    __syncwarp();
    pos[i] = before + __popc(lt);
  }
#endif
  __syncthreads();
  // per digit: exclusive scan over warps
  if (tid < RADIX) {
    unsigned run = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) { const unsigned t = s_warp[w][tid]; s_warp[w][tid] = run; run += t; }
#if GSX_RS_LATE_AGG
    total = run;
    st_volatile(my, (tile == 0 ? FLAG_INC : FLAG_AGG) | total);
#endif
  }
  // block-exclusive scan of totals over the 256 digits (the first 8 warps)
  unsigned incl = total;
  if (tid < RADIX) {
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned t = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
  }
  __syncthreads();
  if (tid < RADIX) {
    unsigned wbase = 0;
    for (int w = 0; w < warp; ++w) wbase += s_scan[w];
    s_excl[tid] = wbase + incl - total;
    // 3. decoupled look-back for digit == tid (predecessors published their aggregates before ranking)
    unsigned excl = 0;
    if (tile != 0) {
      // windowed: LB predecessors are fetched at once (independent loads in flight) and consumed nearest first up to
      // the first one that is inclusive or not yet published -- with every tile of a small input resident in one wave
      // the walk back is long, and one L2 round trip per hop was the latency of the whole pass
      constexpr int LB = GSX_RS_LB;
      long long t = (long long)tile - 1;
      while (true) {
        unsigned sw[LB];
#pragma unroll
        for (int w = 0; w < LB; ++w)
          sw[w] = (t - w >= 0) ? ld_volatile(status + (size_t)(t - w) * RADIX + tid) : FLAG_INC;   // before tile 0: nothing
        int used = 0;
        bool fin = false;
#pragma unroll
        for (int w = 0; w < LB; ++w) {
          const unsigned f = sw[w] & FLAG_MASK;
          if (!fin && used == w && f != 0) {
            excl += sw[w] & VAL_MASK;
            used = w + 1;
            fin = (f == FLAG_INC);
          }
        }
        if (fin) break;
        t -= used;
      }
      st_volatile(my, FLAG_INC | (excl + total));
    }
    s_gbase[tid] = (int)(digit_base[tid] + excl) - (int)s_excl[tid];
  }
  __syncthreads();
  // 4. local reorder through shared memory, then coalesced scatter
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp * (IPT * 32) + i * 32 + lane;
    if (FULL || li < valid) {   // padding items (digit RADIX) are never written
      const unsigned d = (unsigned)(key[i] >> shift) & mask;
      pos[i] += s_excl[d] + s_warp[warp][d];
      s_keys[pos[i]] = key[i];
    }
  }
  __syncthreads();
  int gidx[IPT];
#pragma unroll
  for (int k = 0; k < IPT; ++k) {
    const int slot = tid + k * THREADS;
    gidx[k] = -1;
    if (FULL || slot < valid) {
      const KeyT kk = s_keys[slot];
      const unsigned d = (unsigned)(kk >> shift) & mask;
      gidx[k] = s_gbase[d] + slot;
      if (write_keys) keys_out[gidx[k]] = kk;   // the last pass of the tile sort leaves only the values (nobody reads its keys)
      if (first_of_key) {   // the keys of a digit run are sorted (earlier passes), so equal keys are adjacent slots
        const bool run_start = slot == 0 || s_keys[slot - 1] != kk;
        if (run_start) atomicMin(first_of_key + (int)min((unsigned long long)kk, (unsigned long long)key_clamp), gidx[k]);
      }
    }
  }
  __syncthreads();
  // values are loaded only now (IPT fewer live registers through the ranking)
  uint32_t *s_vals = reinterpret_cast<uint32_t *>(s_keys);
#pragma unroll
  for (int i = 0; i < IPT; ++i) {
    const int li = warp * (IPT * 32) + i * 32 + lane;
    if (FULL || li < valid) s_vals[pos[i]] = vals_in[base + li];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < IPT; ++k)
    if (gidx[k] >= 0) vals_out[gidx[k]] = s_vals[tid + k * THREADS];
}

template <typename KeyT, int THREADS, int IPT>
__global__ void __launch_bounds__(THREADS, (THREADS <= 256 ? (sizeof(KeyT) == 8 ? 3 : 4) : (sizeof(KeyT) == 8 ? 1 : 2)))
k_onesweep(const KeyT *__restrict__ keys_in, KeyT *__restrict__ keys_out, const uint32_t *__restrict__ vals_in,
           uint32_t *__restrict__ vals_out, long long n_host, const int *__restrict__ n_dev, int shift, int bits,
           const unsigned int *__restrict__ digit_base /*[256] exclusive*/, unsigned int *__restrict__ status,
           unsigned int *__restrict__ ticket, int *__restrict__ first_of_key, int key_clamp, int write_keys) {
  // first_of_key (last pass of the tile sort only, else null): first_of_key[k] = smallest output index holding key k
  // (keys clamped to key_clamp) -- the tile-range extraction fused into the scatter: a run of equal keys
  // is contiguous in the re-ordered tile, so only run starts issue an atomicMin; k_ranges_from_first turns it into ranges
  constexpr int WARPS = THREADS / 32, TILE = THREADS * IPT;
  static_assert(THREADS >= RADIX && THREADS % 32 == 0, "one thread per digit for the digit-parallel steps");
  extern __shared__ __align__(16) unsigned char s_dyn[];
  KeyT *s_keys = reinterpret_cast<KeyT *>(s_dyn);           // [TILE], reused for the values
  __shared__ unsigned int s_warp[WARPS][RADIX + 1];         // per-warp digit counts -> exclusive offsets
  __shared__ unsigned int s_match[GSX_RS_MATCH_ATOMIC ? WARPS : 1][RADIX + 1];   // per-warp peer masks of the item being ranked
  __shared__ unsigned int s_excl[RADIX + 1];                // block-exclusive offset of each digit
  __shared__ int s_gbase[RADIX];                            // global base minus local offset
  __shared__ unsigned int s_scan[RADIX / 32];
  __shared__ unsigned int s_tile;

  const int tid = threadIdx.x;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  for (int i = tid; i < WARPS * (RADIX + 1); i += THREADS) {
    (&s_warp[0][0])[i] = 0;
    if (GSX_RS_MATCH_ATOMIC) (&s_match[0][0])[i] = 0;
  }
  if (tid < RADIX) s_excl[tid] = 0;
  __syncthreads();
  const unsigned tile = s_tile;
  const long long n = live_count(n_dev, n_host);
  const long long base = (long long)tile * TILE;
  if (base >= n) return;
  const int valid = (int)min((long long)TILE, n - base);
  const unsigned mask = (1u << bits) - 1u;
  if (valid == TILE)
    onesweep_tile<KeyT, THREADS, IPT, true>(keys_in, keys_out, vals_in, vals_out, shift, bits, digit_base, status, first_of_key,
                                            key_clamp, write_keys, tile, base, valid, mask, s_keys, s_warp, s_match, s_excl, s_gbase, s_scan);
  else
    onesweep_tile<KeyT, THREADS, IPT, false>(keys_in, keys_out, vals_in, vals_out, shift, bits, digit_base, status, first_of_key,
                                             key_clamp, write_keys, tile, base, valid, mask, s_keys, s_warp, s_match, s_excl, s_gbase, s_scan);
}

// ranges[t] = first i whose tile >= t ; ranges[n_tiles] = n   (reference cuda/culling.cu:302-343; exact integer
// tile ids instead of floor(key / (max_z+1)) -- SURVEY H3).  n == 0 -> all zeros.
template <typename KeyT, int SHIFT>
__global__ void __launch_bounds__(256) k_tile_ranges(const KeyT *__restrict__ keys, long long n_host,
                                                     const int *__restrict__ n_dev, int n_tiles,
                                                     int *__restrict__ ranges) {
  const long long n = live_count(n_dev, n_host);
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n == 0) {
    for (long long t = i0; t <= n_tiles; t += stride) ranges[t] = 0;
    return;
  }
  for (long long i = i0; i < n; i += stride) {
    const int cur = (int)min((unsigned)(keys[i] >> SHIFT), (unsigned)(n_tiles - 1));   // clamp as the reference (:313)
    int prev = -1;
    if (i > 0) prev = (int)min((unsigned)(keys[i - 1] >> SHIFT), (unsigned)(n_tiles - 1));
    for (int t = prev + 1; t <= cur; ++t) ranges[t] = (int)i;
    if (i == n - 1)
      for (int t = cur + 1; t <= n_tiles; ++t) ranges[t] = (int)n;
  }
}

struct SortLayout {
  size_t hist_off, ticket_off, status_off, total;
  int tiles;
};
SortLayout sort_layout(long long n) {
  SortLayout L;
  // sized for the smallest tile shape, so that one workspace serves whichever shape is selected
  constexpr int min_tile = 4096;
  L.tiles = (int)((n + min_tile - 1) / min_tile);
  if (L.tiles < 1) L.tiles = 1;
  size_t off = 0;
  L.hist_off = off; off += align_up(sizeof(unsigned) * MAX_PASSES * RADIX, 256);
  L.ticket_off = off; off += align_up(sizeof(unsigned) * MAX_PASSES, 256);
  L.status_off = off; off += align_up(sizeof(unsigned) * (size_t)MAX_PASSES * L.tiles * RADIX, 256);
  L.total = off;
  return L;
}

template <typename KeyT, int THREADS, int IPT>
static int launch_onesweep(int tiles, const KeyT *kin, KeyT *kout, const uint32_t *vin, uint32_t *vout, long long n,
                           const int *n_dev, int shift, int bits, const unsigned *base, unsigned *status, unsigned *ticket,
                           cudaStream_t st, int *first_of_key, int key_clamp, int write_keys) {
  const size_t dyn = std::max(sizeof(KeyT), sizeof(uint32_t)) * (size_t)THREADS * IPT;   // the key tile is reused for the values
  static bool attr_set = false;   // per instantiation
  if (!attr_set) {
    GSX_CUDA(cudaFuncSetAttribute(k_onesweep<KeyT, THREADS, IPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    attr_set = true;
  }
  k_onesweep<KeyT, THREADS, IPT><<<tiles, THREADS, dyn, st>>>(kin, kout, vin, vout, n, n_dev, shift, bits, base, status, ticket,
                                                              first_of_key, key_clamp, write_keys);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // namespace

// Internal entry (no pointer attribute checks) used by the binning path.  Result lands in keys_b/vals_b when
// the number of passes is odd, else back in keys_a/vals_a; *result_in_b tells which.
// flags bit 0 (SORT_HIST_READY): the caller zeroed the workspace with sort_prepare_impl() and accumulated the per-pass digit
// histograms into sort_hist_impl() itself (the emission kernel does, binning.cu): no memset, no histogram pass here.
template <typename KeyT>
static int sort_pairs_t(KeyT *keys_a, KeyT *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n, const int *n_dev,
                        int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags,
                        KeyT *first_k = nullptr, uint32_t *first_v = nullptr, int *first_of_key = nullptr, int key_clamp = 0,
                        bool drop_last_keys = false) {
  if (key_bits < 1 || key_bits > (int)sizeof(KeyT) * 8) { set_error("key_bits %d out of range", key_bits); return GSX_E_ARG; }
  const int passes = (key_bits + 7) / 8;
  // With an even number of passes the ping-pong a -> b -> a ends in a.  first_k / first_v (a third buffer): the first pass
  // writes there instead, and the remaining odd number of passes ends in b (what CUB's temp storage does for SortPairs).
  const bool via_third = first_k && first_v && !(passes & 1);
  *result_in_b = via_third ? 1 : (passes & 1);
  if (n <= 0) return GSX_OK;
  const SortLayout L = sort_layout(n);
  if (workspace_bytes < L.total) { set_error("sort workspace %zu < %zu", workspace_bytes, L.total); return GSX_E_WORKSPACE; }
  char *ws = (char *)workspace;
  unsigned *hist = (unsigned *)(ws + L.hist_off);
  unsigned *ticket = (unsigned *)(ws + L.ticket_off);
  unsigned *status = (unsigned *)(ws + L.status_off);
  const int cfg = rs_cfg((int)sizeof(KeyT), n);
  const int tile = kRsCfg[cfg].threads * kRsCfg[cfg].ipt;
  const int tiles = (int)((n + tile - 1) / tile);
  if (!(flags & 1)) {
    GSX_CUDA(cudaMemsetAsync(ws, 0, L.status_off + sizeof(unsigned) * (size_t)passes * tiles * RADIX, st));
    const int hblocks = (int)std::min<long long>(kNumSMs * 8, (n + 255) / 256);
    k_histogram<KeyT><<<hblocks, 256, 0, st>>>(keys_a, n, n_dev, passes, key_bits, hist);
    GSX_LAUNCH_CHECK();
  }
  k_scan_histogram<<<passes, RADIX, 0, st>>>(hist);
  GSX_LAUNCH_CHECK();
  KeyT *kin = keys_a, *kout = via_third ? first_k : keys_b;
  uint32_t *vin = vals_a, *vout = via_third ? first_v : vals_b;
  for (int p = 0; p < passes; ++p) {
    const int bits = std::min(8, key_bits - 8 * p);
    unsigned *stp = status + (size_t)p * tiles * RADIX;
    int *fk = p == passes - 1 ? first_of_key : nullptr;
    const int wk = (drop_last_keys && p == passes - 1) ? 0 : 1;
    int rc;
    switch (cfg) {
      case 1: rc = launch_onesweep<KeyT, 512, 16>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
      case 2: rc = launch_onesweep<KeyT, 384, 20>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
      case 3: rc = launch_onesweep<KeyT, 512, 12>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
      case 4: rc = launch_onesweep<KeyT, 256, 20>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
      case 5: rc = launch_onesweep<KeyT, 256, 24>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
      default: rc = launch_onesweep<KeyT, 256, 16>(tiles, kin, kout, vin, vout, n, n_dev, 8 * p, bits, hist + p * RADIX, stp, ticket + p, st, fk, key_clamp, wk); break;
    }
    if (rc != GSX_OK) return rc;
    if (via_third && p == 0) { kin = first_k; vin = first_v; kout = keys_b; vout = vals_b; }
    else if (via_third && p == 1) { kin = keys_b; vin = vals_b; kout = keys_a; vout = vals_a; }
    else { std::swap(kin, kout); std::swap(vin, vout); }
  }
  return GSX_OK;
}

// zeroes histogram, tickets and status words of a sort of up to n keys on key_bits bits (for SORT_HIST_READY callers)
int sort_prepare_impl(long long n, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st) {
  if (n <= 0) return GSX_OK;
  const SortLayout L = sort_layout(n);
  if (workspace_bytes < L.total) { set_error("sort workspace %zu < %zu", workspace_bytes, L.total); return GSX_E_WORKSPACE; }
  const int passes = (key_bits + 7) / 8;
  const int tiles = (int)((n + rs_tile(4, n) - 1) / rs_tile(4, n));   // SORT_HIST_READY callers sort 4-byte keys
  GSX_CUDA(cudaMemsetAsync(workspace, 0, L.status_off + sizeof(unsigned) * (size_t)passes * tiles * RADIX, st));
  return GSX_OK;
}
unsigned *sort_hist_impl(void *workspace) { return (unsigned *)workspace; }   // [passes][256], pass p = key bits [8p, 8p + 8)

int sort_pairs_impl(uint64_t *keys_a, uint64_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                    const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st,
                    int *result_in_b) {
  return sort_pairs_t<uint64_t>(keys_a, keys_b, vals_a, vals_b, n, n_dev, key_bits, workspace, workspace_bytes, st, result_in_b, 0);
}
int sort_pairs_u32_impl(uint32_t *keys_a, uint32_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                        const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st,
                        int *result_in_b, int flags) {
  return sort_pairs_t<uint32_t>(keys_a, keys_b, vals_a, vals_b, n, n_dev, key_bits, workspace, workspace_bytes, st, result_in_b, flags);
}

// ---- tile sort with the range extraction fused into its last scatter: ranges[t] = first i with tile(i) >= t, ranges[T] = n
namespace {
__global__ void __launch_bounds__(1024) k_ranges_init(int *__restrict__ ranges, int n_tiles) {
  for (int t = blockIdx.x * 1024 + threadIdx.x; t <= n_tiles; t += gridDim.x * 1024) ranges[t] = INT_MAX;
}
// suffix minimum over first_of_key (in place): empty tiles take the first index of the next non-empty one, the end is n
__global__ void __launch_bounds__(1024) k_ranges_from_first(int *__restrict__ ranges, int n_tiles, long long n_host,
                                                            const int *__restrict__ n_dev) {
  __shared__ int s_min[1024];
  const int n = (int)live_count(n_dev, n_host);
  const int per = (n_tiles + 1 + 1023) / 1024;               // contiguous chunk of tiles per thread
  const int t0 = threadIdx.x * per, t1 = min(t0 + per, n_tiles + 1);
  int m = INT_MAX;
  for (int t = t1 - 1; t >= t0; --t) m = min(m, t == n_tiles ? n : ranges[t]);
  s_min[threadIdx.x] = m;
  __syncthreads();
  // suffix minimum across the chunks (inclusive), sequential part: 1024 values
  for (int off = 1; off < 1024; off <<= 1) {
    const int o = threadIdx.x + off < 1024 ? s_min[threadIdx.x + off] : INT_MAX;
    __syncthreads();
    s_min[threadIdx.x] = min(s_min[threadIdx.x], o);
    __syncthreads();
  }
  int run = threadIdx.x + 1 < 1024 ? s_min[threadIdx.x + 1] : INT_MAX;   // minimum over all later chunks
  run = min(run, n);
  for (int t = t1 - 1; t >= t0; --t) {
    run = min(run, t == n_tiles ? n : ranges[t]);
    ranges[t] = run;
  }
}
}  // namespace

template <typename KeyT>
static int sort_tiles_with_ranges_t(KeyT *keys_a, KeyT *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                                    const int *n_dev, int key_bits, int n_tiles, int *ranges, void *workspace,
                                    size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags, bool drop_last_keys) {
  k_ranges_init<<<std::max(1, std::min(8, (n_tiles + 1024) / 1024)), 1024, 0, st>>>(ranges, n_tiles);
  GSX_LAUNCH_CHECK();
  int rc = sort_pairs_t<KeyT>(keys_a, keys_b, vals_a, vals_b, n, n_dev, key_bits, workspace, workspace_bytes, st, result_in_b, flags,
                              nullptr, nullptr, ranges, n_tiles - 1, drop_last_keys);
  if (rc != GSX_OK) return rc;
  k_ranges_from_first<<<1, 1024, 0, st>>>(ranges, n_tiles, n, n_dev);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}
int sort_tiles_with_ranges_impl(uint32_t *keys_a, uint32_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                                const int *n_dev, int key_bits, int n_tiles, int *ranges, void *workspace,
                                size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags) {
  return sort_tiles_with_ranges_t<uint32_t>(keys_a, keys_b, vals_a, vals_b, n, n_dev, key_bits, n_tiles, ranges, workspace,
                                            workspace_bytes, st, result_in_b, flags, false);
}
// 16-bit tile ids (up to 65536 tiles): 2 bytes less per key in the emission's write, the histogram's read and both passes'
// reads and the first pass's write; the last pass writes no keys at all (the ranges come out of its scatter, fused)
int sort_tiles_with_ranges_u16_impl(uint16_t *keys_a, uint16_t *keys_b, uint32_t *vals_a, uint32_t *vals_b, long long n,
                                    const int *n_dev, int key_bits, int n_tiles, int *ranges, void *workspace,
                                    size_t workspace_bytes, cudaStream_t st, int *result_in_b, int flags) {
  return sort_tiles_with_ranges_t<uint16_t>(keys_a, keys_b, vals_a, vals_b, n, n_dev, key_bits, n_tiles, ranges, workspace,
                                            workspace_bytes, st, result_in_b, flags, true);
}

static int ranges_grid(long long n, int n_tiles) {
  long long work = std::max<long long>(n, n_tiles + 1);
  return std::max((int)std::min<long long>(kNumSMs * 16, (work + 255) / 256), 1);
}
int tile_ranges_impl(const uint64_t *sorted_keys, long long n, const int *n_dev, int n_tiles, int *ranges,
                     cudaStream_t st) {
  k_tile_ranges<uint64_t, 32><<<ranges_grid(n, n_tiles), 256, 0, st>>>(sorted_keys, n, n_dev, n_tiles, ranges);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}
int tile_ranges_u32_impl(const uint32_t *sorted_tiles, long long n, const int *n_dev, int n_tiles, int *ranges,
                         cudaStream_t st) {
  k_tile_ranges<uint32_t, 0><<<ranges_grid(n, n_tiles), 256, 0, st>>>(sorted_tiles, n, n_dev, n_tiles, ranges);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

size_t sort_workspace_bytes_impl(long long n) { return sort_layout(n).total; }

}  // namespace gsx

using namespace gsx;

template <typename KeyT>
static int sort_pairs_public(KeyT *keys_in, KeyT *keys_out, uint32_t *vals_in, uint32_t *vals_out, long long n, const int *n_dev,
                             int key_bits, void *workspace, size_t workspace_bytes, cudaStream_t st) {
  int in_b = 0;
  const size_t base = align_up(sort_workspace_bytes_impl(n), 256);
  const size_t kb = align_up((size_t)std::max<long long>(n, 1) * sizeof(KeyT), 256), vb = align_up((size_t)std::max<long long>(n, 1) * 4, 256);
  KeyT *tk = nullptr;
  uint32_t *tv = nullptr;
  if (workspace_bytes >= base + kb + vb) { tk = (KeyT *)((char *)workspace + base); tv = (uint32_t *)((char *)workspace + base + kb); }
  int rc = sort_pairs_t<KeyT>(keys_in, keys_out, vals_in, vals_out, n, n_dev, key_bits, workspace, workspace_bytes, st, &in_b, 0, tk, tv);
  if (rc != GSX_OK) return rc;
  if (!in_b && n > 0) {   // even number of passes and no room for the third buffer: copy to the documented output
    GSX_CUDA(cudaMemcpyAsync(keys_out, keys_in, sizeof(KeyT) * n, cudaMemcpyDeviceToDevice, st));
    GSX_CUDA(cudaMemcpyAsync(vals_out, vals_in, sizeof(uint32_t) * n, cudaMemcpyDeviceToDevice, st));
  }
  return GSX_OK;
}

extern "C" {

// internal state + room for a third (key, value) buffer of n 8-byte keys: with it an even number of passes needs no final
// copy into keys_out (smaller workspaces still work: the result is then copied)
size_t gsx_sort_workspace_bytes(long long n) {
  return align_up(sort_workspace_bytes_impl(n), 256) + align_up((size_t)std::max<long long>(n, 1) * 8, 256) +
         align_up((size_t)std::max<long long>(n, 1) * 4, 256);
}

int gsx_sort_pairs_u64(uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in, uint32_t *vals_out, long long n,
                       const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, void *stream) {
  GSX_REQUIRE_DEV(keys_in); GSX_REQUIRE_DEV(keys_out); GSX_REQUIRE_DEV(vals_in); GSX_REQUIRE_DEV(vals_out);
  GSX_REQUIRE_DEV(workspace);
  return sort_pairs_public<uint64_t>(keys_in, keys_out, vals_in, vals_out, n, n_dev, key_bits, workspace, workspace_bytes,
                                     (cudaStream_t)stream);
}

int gsx_sort_pairs_u32(uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in, uint32_t *vals_out, long long n,
                       const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, void *stream) {
  GSX_REQUIRE_DEV(keys_in); GSX_REQUIRE_DEV(keys_out); GSX_REQUIRE_DEV(vals_in); GSX_REQUIRE_DEV(vals_out);
  GSX_REQUIRE_DEV(workspace);
  return sort_pairs_public<uint32_t>(keys_in, keys_out, vals_in, vals_out, n, n_dev, key_bits, workspace, workspace_bytes,
                                     (cudaStream_t)stream);
}

int gsx_tile_ranges(const uint64_t *sorted_keys, long long n, const int *n_dev, int n_tiles, int *ranges,
                    void *stream) {
  GSX_REQUIRE_DEV(ranges);
  if (n > 0) GSX_REQUIRE_DEV(sorted_keys);
  return tile_ranges_impl(sorted_keys, n, n_dev, n_tiles, ranges, (cudaStream_t)stream);
}

}  // extern "C"
