// tilebin.cu -- direct tile binning: the per-tile lists written in their final order by ONE stable counting sort on the
// tile id, instead of "emit instances in depth order, then radix-sort them on the tile id in two passes" (binning.cu).
//
// Replaces, for the second level of the two-level sort, generate_splats_kernel + cub::DeviceRadixSort +
// find_tile_boundaries_kernel of the reference (cuda/culling.cu:255-343, :455-475).  The Gaussians arrive in depth order
// (first level: the 32-bit depth sort of binning.cu).  That list is cut into R contiguous segments, one per warp.  A
// stable counting sort needs, for every (segment, tile), the number of instances the EARLIER segments put into the tile:
//   pass A  k_tilebin_count   every warp enumerates the tiles of its segment's Gaussians (exact 4-axis SAT, in registers)
//                             into a PRIVATE shared-memory histogram of 16-bit counters, one per tile; a CTA leaves the
//                             exclusive prefix over its warps per tile (u16 matrix W[R][T]) and its total (u32 C[G][T])
//   scan    k_tilebin_colscan exclusive prefix of C over the G CTAs per tile + tile totals; k_tilebin_ranges scans the
//                             totals into the tile ranges (the ranges are a by-product: no boundary search, no re-read)
//   pass B  k_tilebin_emit    every warp walks its segment again IN ORDER and writes each instance's Gaussian id straight
//                             to ranges[t] + C[cta][t] + W[warp][t] + (running count of the warp for tile t).
// Order inside a warp: a chunk of 32 Gaussians (lane = Gaussian) is enumerated into per-lane staging slots, the slots are
// then ranked 32 entries at a time in Gaussian-major order -- __match_any_sync on the tile id gives the rank among equal
// tiles of the step, the group's first lane advances the private counter -- so equal (tile) instances keep depth order
// (and ascending id among equal depths): the result is bit-identical to the sort path.  Rectangles of more than COOP_CELLS
// tiles are expanded by the whole warp at their place in the order.
// Traffic per instance: one 4-byte store (the 84 MB of ids of a C4 view stay in L2 while the lists fill) against 8 B
// written + 4 B histogram read + 2 x 16 B of radix passes; the SAT enumeration runs twice instead of once.
// Tiles are processed in bands of whole tile rows (grid.y) when T 16-bit counters per warp would not fit shared memory.
#include <algorithm>

#include "gsx_common.cuh"
#include "gsx_math.cuh"

namespace gsx {

namespace {

constexpr int kSlot = COOP_CELLS;          // staging entries per lane (a "small" rectangle has at most COOP_CELLS cells)
constexpr int LPS = 32 / kSlot;            // lanes (Gaussians) whose slots one ranking step covers
static_assert(kSlot >= 1 && kSlot <= 32 && (32 % kSlot) == 0, "COOP_CELLS must divide 32");
constexpr int kMaxWarps = 16;
constexpr size_t kSmemLimit = 232448 - 1024;   // 227 KB per CTA minus slack for static shared memory

struct Raw {            // what a lane needs of one Gaussian, as loaded
  int i;                // Gaussian id (-1: none)
  float u, v, ex, ey;
  float4 r;
};

// Gaussian id at depth rank j (-1: past the segment or a Gaussian without instances, which sorted to the end)
__device__ __forceinline__ int fetch_id(const uint32_t *__restrict__ dkeys, const uint32_t *__restrict__ perm, int j, int end) {
  return (j < end && dkeys[j] != 0xffffffffu) ? (int)perm[j] : -1;
}
__device__ __forceinline__ Raw fetch_raw(const float *__restrict__ uv, int uv_stride, const float4 *__restrict__ radius, int i) {
  Raw w;
  w.i = i; w.u = w.v = w.ex = w.ey = 0.f; w.r = make_float4(0.f, 0.f, 0.f, 1.f);
  if (i >= 0) {
    if (uv_stride == 4) {
      const float4 q = *reinterpret_cast<const float4 *>(uv + (size_t)i * 4);
      w.u = q.x; w.v = q.y; w.ex = q.z; w.ey = q.w;
    } else {
      const float2 q = *reinterpret_cast<const float2 *>(uv + (size_t)i * 2);
      w.u = q.x; w.v = q.y;
    }
    w.r = radius[i];
  }
  return w;
}

// SAT constants + tile rectangle of a lane's Gaussian, clipped to the band's tile rows [ty_lo, ty_hi)
__device__ __forceinline__ bool setup_item(const Raw &w, int ntx, int nty, int tight, int ty_lo, int ty_hi, ObbSat &s, TileRange &tr) {
  tr = TileRange{0, -1, 0, -1};
  if (w.i < 0) return false;
  const TileWindow win = candidate_window(w.u, w.v, w.r.x, ntx, nty);
  s = obb_sat_setup(compute_obb(w.u, w.v, w.r.x, w.r.y, w.r.z, w.r.w));
  tr = sat_aabb_range(s, win);
  if (tight) clip_range_to_extent(tr, w.u, w.v, w.ex, w.ey);
  tr.y0 = max(tr.y0, ty_lo); tr.y1 = min(tr.y1, ty_hi - 1);
  return rect_cells(tr) > 0;
}

// lanes [lo, hi) as a mask
__device__ __forceinline__ unsigned lane_range(int lo, int hi) {
  if (lo >= hi) return 0u;
  const unsigned up = hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
  return up & ~((1u << lo) - 1u);
}

// Segments balance WORK, not Gaussians: in depth order the near Gaussians cover hundreds of tiles and the far ones one or
// two, so equal-length segments would leave the first warps with most of the instances.  The weight of a Gaussian is its
// instance count + kGaussWeight; its prefix m(j) = offsets[j] + kGaussWeight * j is monotone (offsets = exclusive scan of the
// counts in depth order, binning.cu), so boundary r is the first j with m(j) >= r * M / rows, M = K + kGaussWeight * n.
// A segment thus holds at most M / (rows * kGaussWeight) + 1 Gaussians, which bounds the 16-bit counters (tilebin_plan).
constexpr int kGaussWeight = 16;
__device__ __forceinline__ int segment_start(const int *__restrict__ offsets, int n, long long K, int row, int rows) {
  if (row <= 0) return 0;
  if (row >= rows) return n;
  const long long M = K + (long long)kGaussWeight * n;
  const long long target = (M * row + rows - 1) / rows;
  int lo = 0, hi = n;   // first j in [0, n] with m(j) >= target, m(n) = M
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if ((long long)offsets[mid] + (long long)kGaussWeight * mid >= target) hi = mid; else lo = mid + 1;
  }
  return lo;
}

// ---- pass A ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kMaxWarps * 32) k_tilebin_count(
    const float *__restrict__ uv, int uv_stride, const float4 *__restrict__ radius, const uint32_t *__restrict__ dkeys,
    const uint32_t *__restrict__ perm, const int *__restrict__ offsets, int n_host, const int *__restrict__ n_dev,
    const int *__restrict__ total, int ntx, int nty, int tight, int rows_per_band, int tb, unsigned *__restrict__ Cm,
    unsigned short *__restrict__ Wm, int *__restrict__ seg) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned short *cnt = reinterpret_cast<unsigned short *>(smem_raw);   // [warps][tb]
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = ntx * nty;
  const int band = blockIdx.y, ty_lo = band * rows_per_band, ty_hi = min(nty, ty_lo + rows_per_band);
  const int t0 = ty_lo * ntx, tn = (ty_hi - ty_lo) * ntx;   // this band's tiles: [t0, t0 + tn)
  for (int k = threadIdx.x; k < warps * tb / 2; k += blockDim.x) reinterpret_cast<unsigned *>(cnt)[k] = 0u;
  __syncthreads();
  // total[1] = live instance count: 0 for an empty view or one whose K exceeds the capacity -> nothing is binned
  const int n = total[1] == 0 ? 0 : (n_dev ? min(*n_dev, n_host) : n_host);
  const int rows = gridDim.x * warps, row = blockIdx.x * warps + warp;
  // lanes 0 and 1 search the two ends of the warp's segment; band 0 leaves the starts for pass B
  int jb = lane < 2 ? segment_start(offsets, n, total[0], row + lane, rows) : 0;
  const int j0 = __shfl_sync(0xffffffffu, jb, 0), j1 = __shfl_sync(0xffffffffu, jb, 1);
  if (blockIdx.y == 0 && lane == 0) {
    seg[row] = j0;
    if (row == rows - 1) seg[rows] = j1;
  }
  unsigned short *mine = cnt + (size_t)warp * tb;
  unsigned *mine32 = reinterpret_cast<unsigned *>(mine);
  // two-deep software pipeline: ids of chunk c + 2 and records of chunk c + 1 are in flight under chunk c
  Raw nxt = fetch_raw(uv, uv_stride, radius, fetch_id(dkeys, perm, j0 + lane, j1));
  int id2 = fetch_id(dkeys, perm, j0 + 32 + lane, j1);
  for (int jc = j0; jc < j1; jc += 32) {
    const Raw cur = nxt;
    nxt = fetch_raw(uv, uv_stride, radius, id2);
    id2 = fetch_id(dkeys, perm, jc + 64 + lane, j1);
    ObbSat s{};
    TileRange tr;
    const bool live = setup_item(cur, ntx, nty, tight, ty_lo, ty_hi, s, tr);
    const bool big = live && rect_cells(tr) > COOP_CELLS;
    if (live && !big) {
      for (int tx = tr.x0; tx <= tr.x1; ++tx) {
        const SatCol col = sat_col(s, tx);
        for (int ty = tr.y0; ty <= tr.y1; ++ty)
          if (sat_cell(s, col, ty)) {
            const int t = ty * ntx + tx - t0;
            atomicAdd(&mine32[t >> 1], 1u << ((t & 1) * 16));   // lanes of the warp may meet in a tile; < 65536 per segment
          }
      }
    }
    __syncwarp();   // the plain 16-bit updates below must see the atomics above
    unsigned bm = __ballot_sync(0xffffffffu, big);
    while (bm) {
      const int leader = __ffs(bm) - 1;
      bm &= bm - 1;
      coop_tile_rect(leader, s, tr, [&](bool hit, unsigned, int tx, int ty) {
        if (hit) mine[ty * ntx + tx - t0] += 1;   // the cells of one rectangle are distinct tiles: plain 16-bit update
      });
      __syncwarp();
    }
  }
  __syncthreads();
  // per tile: exclusive prefix over the CTA's warps (u16, coalesced rows of W) and the CTA total (u32 row of C)
  for (int t = threadIdx.x; t < tn; t += blockDim.x) {
    unsigned run = 0;
    for (int w = 0; w < warps; ++w) {
      const unsigned c = cnt[(size_t)w * tb + t];
      Wm[(size_t)(blockIdx.x * warps + w) * T + t0 + t] = (unsigned short)run;
      run += c;
    }
    Cm[(size_t)blockIdx.x * T + t0 + t] = run;
  }
}

// ---- scan -----------------------------------------------------------------------------------------------------------
// block (32 tiles, 8 row groups): C[g][t] <- sum of C[g'][t] over g' < g; tile_tot[t] = column sum
__global__ void __launch_bounds__(256) k_tilebin_colscan(unsigned *__restrict__ Cm, int G, int T, int *__restrict__ tile_tot) {
  __shared__ unsigned part[8][33];
  const int t = blockIdx.x * 32 + threadIdx.x, y = threadIdx.y;
  const int per = (G + 7) / 8, g0 = min(G, y * per), g1 = min(G, g0 + per);
  unsigned sum = 0;
  if (t < T)
    for (int g = g0; g < g1; ++g) sum += Cm[(size_t)g * T + t];
  part[y][threadIdx.x] = sum;
  __syncthreads();
  unsigned run = 0;
  for (int k = 0; k < y; ++k) run += part[k][threadIdx.x];
  if (t < T) {
    for (int g = g0; g < g1; ++g) {
      const unsigned c = Cm[(size_t)g * T + t];
      Cm[(size_t)g * T + t] = run;
      run += c;
    }
    if (y == 7) tile_tot[t] = (int)run;
  }
}

// ranges[t] <- exclusive scan of the tile totals held in ranges[0..T); ranges[T] = K
__global__ void __launch_bounds__(1024) k_tilebin_ranges(int *__restrict__ ranges, int T) {
  __shared__ int s[1024];
  __shared__ int carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b0 = 0; b0 < T; b0 += 1024) {
    const int i = b0 + threadIdx.x;
    const int v = i < T ? ranges[i] : 0;
    s[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
      const int x = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
      __syncthreads();
      s[threadIdx.x] += x;
      __syncthreads();
    }
    if (i < T) ranges[i] = carry + s[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry += s[1023];
    __syncthreads();
  }
  if (threadIdx.x == 0) ranges[T] = carry;
}

// ---- pass B ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kMaxWarps * 32) k_tilebin_emit(
    const float *__restrict__ uv, int uv_stride, const float4 *__restrict__ radius, const uint32_t *__restrict__ dkeys,
    const uint32_t *__restrict__ perm, int n_host, const int *__restrict__ n_dev, const int *__restrict__ total, int ntx, int nty,
    int tight, int rows_per_band, int tb, const unsigned *__restrict__ Cm, const unsigned short *__restrict__ Wm,
    const int *__restrict__ seg, const int *__restrict__ ranges, uint32_t *__restrict__ out_ids) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned *base = reinterpret_cast<unsigned *>(smem_raw);                                   // [tb] first position of this CTA in tile t
  unsigned short *off = reinterpret_cast<unsigned short *>(base + tb) + (size_t)warp * tb;    // [warps][tb] running offset of the warp
  unsigned short *stage = reinterpret_cast<unsigned short *>(base + tb) + (size_t)warps * tb + (size_t)warp * (32 * kSlot);
  const int T = ntx * nty;
  const int band = blockIdx.y, ty_lo = band * rows_per_band, ty_hi = min(nty, ty_lo + rows_per_band);
  const int t0 = ty_lo * ntx, tn = (ty_hi - ty_lo) * ntx;
  const int n = total[1] == 0 ? 0 : (n_dev ? min(*n_dev, n_host) : n_host);   // block-uniform
  if (n == 0) return;
  const int row = blockIdx.x * warps + warp;
  for (int t = threadIdx.x; t < tn; t += blockDim.x) base[t] = (unsigned)ranges[t0 + t] + Cm[(size_t)blockIdx.x * T + t0 + t];
  for (int t = lane; t < tn; t += 32) off[t] = Wm[(size_t)row * T + t0 + t];
  __syncthreads();
  const int j0 = seg[row], j1 = seg[row + 1];
  const unsigned lt = (1u << lane) - 1u;
  // two-deep software pipeline: ids of chunk c + 2 and records of chunk c + 1 are in flight under chunk c
  Raw nxt = fetch_raw(uv, uv_stride, radius, fetch_id(dkeys, perm, j0 + lane, j1));
  int id2 = fetch_id(dkeys, perm, j0 + 32 + lane, j1);
  for (int jc = j0; jc < j1; jc += 32) {
    const Raw cur = nxt;
    nxt = fetch_raw(uv, uv_stride, radius, id2);
    id2 = fetch_id(dkeys, perm, jc + 64 + lane, j1);
    ObbSat s{};
    TileRange tr;
    const bool live = setup_item(cur, ntx, nty, tight, ty_lo, ty_hi, s, tr);
    const bool big = live && rect_cells(tr) > COOP_CELLS;
    int c = 0;   // staged entries of this lane
    if (live && !big) {
      unsigned short *slot = stage + lane * kSlot;
      for (int tx = tr.x0; tx <= tr.x1; ++tx) {
        const SatCol col = sat_col(s, tx);
        for (int ty = tr.y0; ty <= tr.y1; ++ty)
          if (sat_cell(s, col, ty)) slot[c++] = (unsigned short)(ty * ntx + tx - t0);
      }
    }
    __syncwarp();
    const unsigned bigm = __ballot_sync(0xffffffffu, big);
    const unsigned anym = __ballot_sync(0xffffffffu, c > 0);
    // ranks the staged entries of step q (the slots of lanes [q * LPS, (q + 1) * LPS)) owned by lanes in [lo, hi)
    auto rank_step = [&](int q, int lo, int hi) {
      const int owner = q * LPS + lane / kSlot, idx = lane % kSlot;
      const int oc = __shfl_sync(0xffffffffu, c, owner);
      const uint32_t oid = (uint32_t)__shfl_sync(0xffffffffu, cur.i, owner);
      const bool valid = owner >= lo && owner < hi && idx < oc;
      const int t = valid ? stage[owner * kSlot + idx] : 0;
      const unsigned peers = __match_any_sync(0xffffffffu, valid ? t : (0x10000 | lane));   // tiles of a band are < 65536
      const int rank = __popc(peers & lt);
      const unsigned b = valid ? off[t] : 0u;
      __syncwarp();
      if (valid && rank == 0) off[t] = (unsigned short)(b + __popc(peers));
      __syncwarp();
      if (valid) out_ids[base[t] + b + rank] = oid;
    };
    // Gaussian-major walk over the chunk: step q covers LPS lanes; a big Gaussian is expanded at its place in the order
#pragma unroll 1
    for (int q = 0; q < 32 / LPS; ++q) {
      const unsigned grp = lane_range(q * LPS, (q + 1) * LPS);
      if (!((bigm | anym) & grp)) continue;
      int lo = q * LPS;
      unsigned bg = bigm & grp;
      while (bg) {
        const int L = __ffs(bg) - 1;
        bg &= bg - 1;
        if (anym & lane_range(lo, L)) rank_step(q, lo, L);
        lo = L + 1;
        const uint32_t lid = (uint32_t)__shfl_sync(0xffffffffu, cur.i, L);
        coop_tile_rect(L, s, tr, [&](bool hit, unsigned, int tx, int ty) {
          if (hit) {   // the cells of one rectangle are distinct tiles
            const int t = ty * ntx + tx - t0;
            const unsigned b = off[t];
            off[t] = (unsigned short)(b + 1);
            out_ids[base[t] + b] = lid;
          }
        });
        __syncwarp();
      }
      if (anym & lane_range(lo, (q + 1) * LPS)) rank_step(q, lo, (q + 1) * LPS);
    }
    __syncwarp();
  }
}

}  // namespace

// Shape of the two passes for a view of ntx x nty tiles and up to n_max depth-sorted Gaussians; false when the matrices do
// not fit `region_bytes` or fewer than 4 warps per CTA would fit shared memory (the caller then takes the sort path).
bool tilebin_plan(int ntx, int nty, int n_max, long long capacity, size_t region_bytes, TileBinPlan *p) {
  const int T = ntx * nty;
  if (T <= 0 || ntx > 65535) return false;
  // fewest bands that leave >= 8 warps per CTA (every band re-reads the Gaussians); else whatever fits with >= 4
  for (int want = 8; want >= 4; want -= 4) {
    for (int bands = 1; bands <= 8; ++bands) {
      const int rpb = (nty + bands - 1) / bands;
      int tb = rpb * ntx;
      tb += tb & 1;
      if (tb > 65536) continue;
      const size_t per_warp = (size_t)tb * 2 + (size_t)32 * kSlot * 2;
      if ((size_t)tb * 4 + (size_t)want * per_warp > kSmemLimit) continue;
      const int warps = (int)std::min<size_t>(kMaxWarps, (kSmemLimit - (size_t)tb * 4) / per_warp);
      p->bands = (nty + rpb - 1) / rpb; p->rows_per_band = rpb; p->tb = tb; p->warps = warps;
      // 16-bit counters: a CTA's segments hold at most (K / kGaussWeight + n) / G + warps Gaussians, K <= capacity
      const long long need = (capacity / kGaussWeight + n_max + 65000 - 1) / 65000;
      p->ctas = kNumSMs * (int)std::max<long long>(1, (need + kNumSMs - 1) / kNumSMs);
      p->smem_count = (size_t)warps * tb * 2;
      p->smem_emit = (size_t)tb * 4 + (size_t)warps * per_warp;
      p->off_W = align_up((size_t)p->ctas * T * 4, 256);
      p->off_seg = p->off_W + align_up((size_t)p->ctas * warps * T * 2, 256);
      p->bytes = p->off_seg + align_up((size_t)(p->ctas * warps + 1) * 4, 256);
      return p->bytes <= region_bytes;
    }
  }
  return false;
}

// counts -> column scan -> ranges -> ordered emission.  `region`: scratch for the two matrices (plan.bytes).  total: the
// PreBin totals (total[1] = live instance count).  ranges [T + 1] and out_ids [K] are the products.
int tilebin_run(const TileBinPlan &p, const float *uv, int uv_stride, const float4 *radius, const uint32_t *dkeys,
                const uint32_t *perm, const int *offsets, int n_host, const int *n_dev, const int *total, int ntx, int nty,
                int tight, char *region, int *ranges, uint32_t *out_ids, cudaStream_t st, cudaEvent_t *ev) {
  static bool attr_set[64] = {};
  int dev = 0;
  GSX_CUDA(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    GSX_CUDA(cudaFuncSetAttribute(k_tilebin_count, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLimit));
    GSX_CUDA(cudaFuncSetAttribute(k_tilebin_emit, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLimit));
    attr_set[dev] = true;
  }
  const int T = ntx * nty;
  unsigned *Cm = reinterpret_cast<unsigned *>(region);
  unsigned short *Wm = reinterpret_cast<unsigned short *>(region + p.off_W);
  int *seg = reinterpret_cast<int *>(region + p.off_seg);
  const dim3 grid(p.ctas, p.bands), block(p.warps * 32);
  k_tilebin_count<<<grid, block, p.smem_count, st>>>(uv, uv_stride, radius, dkeys, perm, offsets, n_host, n_dev, total, ntx, nty,
                                                    tight, p.rows_per_band, p.tb, Cm, Wm, seg);
  GSX_LAUNCH_CHECK();
  k_tilebin_colscan<<<ceil_div(T, 32), dim3(32, 8), 0, st>>>(Cm, p.ctas, T, ranges);
  GSX_LAUNCH_CHECK();
  k_tilebin_ranges<<<1, 1024, 0, st>>>(ranges, T);
  GSX_LAUNCH_CHECK();
  if (ev) { GSX_CUDA(cudaEventRecord(ev[0], st)); }
  k_tilebin_emit<<<grid, block, p.smem_emit, st>>>(uv, uv_stride, radius, dkeys, perm, n_host, n_dev, total, ntx, nty, tight,
                                                  p.rows_per_band, p.tb, Cm, Wm, seg, ranges, out_ids);
  GSX_LAUNCH_CHECK();
  if (ev) { GSX_CUDA(cudaEventRecord(ev[1], st)); GSX_CUDA(cudaEventRecord(ev[2], st)); }
  return GSX_OK;
}

}  // namespace gsx
