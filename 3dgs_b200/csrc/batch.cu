// batch.cu -- per-Gaussian stages for a BATCH of views in one pass over the parameters.
//
// A training batch renders n views from the same parameters (north star: 8 views at 6 M Gaussians).  Run view by
// view, the per-Gaussian stages read the 236 B/Gaussian parameter block twice per view and read-modify-write the
// 236 B gradient block once per view.  Here:
//   k_batch_fwd<L> : F1-F6 + the exact tile count of F7 for all n views -- parameters (SH rows by TMA bulk copy) are
//                    read ONCE per batch, Sigma and sigmoid(opacity) are computed once per Gaussian, and each view's
//                    compositing records go to that view's slot.
//   k_batch_bwd<L> : F10-F15 for all n views -- parameters read once, each view's packed compositing gradients read
//                    (and re-zeroed) once, gradients summed over the views in registers / shared memory (Sigma's
//                    backward F13 is linear in dSigma, so it runs once on the view-summed dSigma) and written ONCE
//                    per batch; SH gradient rows leave through one cp.async.bulk (or cp.reduce.async.bulk) per block.
// Reference per view: 8 + 7 kernels and 16 thrust compactions (cuda/raster.cu:26-104, cuda/trainer.cu:941-1012).
// Algorithmic bytes per batch at degree 3 (n views, P' = Gaussians that touch a tile in a view):
//   forward  236 N read + 74 N n written (records, radius, depth, count, mask)
//   backward 236 N + 4 N n + 96 P' n (packed gradients read + re-zeroed) read/written + 236 N written.
#include <algorithm>

#include "gsx_common.cuh"
#include "gsx_math.cuh"
#include "gsx_tma.cuh"
#include "gsx_view.cuh"

namespace gsx {

namespace {

constexpr int BB = 128;
#ifndef GSX_BATCH_MINB
#define GSX_BATCH_MINB 4   // backward; the forward kernel uses GSX_BATCH_MINB_FWD
#endif
#ifndef GSX_BATCH_MINB_FWD
#define GSX_BATCH_MINB_FWD 6   // profiles/sweep_batch_minb.sh: forward gains from 6 CTAs/SM, backward loses
#endif

__device__ __forceinline__ float2 alpha_extent_b(float a, float b, float c, float opa) {   // == composite.cu
  if (!(opa >= 0.00392156862f)) return make_float2(-INFINITY, -INFINITY);
  const float det = a * c - b * b;
  if (!(a > 0.0f && c > 0.0f && det > 0.0f) || !(opa <= 1.0f)) return make_float2(INFINITY, INFINITY);
  const float tau2 = 2.0f * __logf(fmaxf(opa * 255.0f, 1.0f)) * 1.001f + 0.02f;
  const float inv = tau2 / det;
  return make_float2(sqrtf(inv * c) * 1.0005f + 0.02f, sqrtf(inv * a) * 1.0005f + 0.02f);
}

// block-cooperative staging of this block's SH rows: one TMA bulk copy when the rows are dense and aligned,
// otherwise a strided cooperative copy.  Block-uniform control flow.
template <int SHF>
__device__ __forceinline__ void stage_rows(float *s_dst, const float *sh, int sh_stride, long long row0, int rows,
                                           bool use_tma, uint64_t *bar) {
  if (use_tma) {
    if (threadIdx.x == 0) {
      const uint32_t bytes = (uint32_t)rows * SHF * 4u;
      mbar_expect_tx(bar, bytes);
      bulk_load(s_dst, sh + row0 * SHF, bytes, bar);
    }
  } else {
    for (int k = threadIdx.x; k < rows * SHF; k += BB) {
      const int r = k / (SHF > 0 ? SHF : 1), c = k - r * SHF;
      s_dst[k] = __ldg(sh + (size_t)(row0 + r) * sh_stride + c);
    }
  }
}

constexpr int VCW = sizeof(ViewConst) / 4;   // words per view (43: consecutive views fall on distinct banks)
static_assert(sizeof(ViewConst) % 4 == 0, "ViewConst is word-sized");
// (view, local Gaussian) slots of a block: BB * nv, sized at launch (a single view needs 1/8 of the 8-view staging,
// which lifts the backward kernel from 3 to 4+ CTAs per SM)

// Work list of a block: the active (view, Gaussian) pairs, compacted.  A thread appends all its pairs with one
// shared-memory atomic; the order is irrelevant for the results (every pair writes to places keyed by (view, Gaussian)).
__device__ __forceinline__ void append_pairs(unsigned bits, int t, unsigned *s_n, uint16_t *s_pair) {
  if (!bits) return;
  unsigned o = atomicAdd(s_n, (unsigned)__popc(bits));
  while (bits) {
    const int v = __ffs(bits) - 1;
    bits &= bits - 1;
    s_pair[o++] = (uint16_t)(v * BB + t);
  }
}

__device__ __forceinline__ void copy_views_to_smem(const BatchConst &bc, float *s_view) {
  const float *src = reinterpret_cast<const float *>(&bc.v[0]);
  for (int k = threadIdx.x; k < bc.nv * VCW; k += BB) s_view[k] = src[k];
}

// Forward.  Phase A (thread = Gaussian): camera transform + cull for every view -> visibility bits; Sigma, sigmoid and
// band 0 once per Gaussian into shared memory.  Phase B (thread = visible pair): Jacobian, conic, radius, extent, exact
// tile count, SH colour -> the view's slot.  Only visible pairs cost anything beyond the cull.
template <int L>
__global__ void __launch_bounds__(BB, GSX_BATCH_MINB_FWD)
k_batch_fwd(const __grid_constant__ BatchConst bc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
            const float *__restrict__ quat, const float *__restrict__ opacity, const float *__restrict__ band0,
            const float *__restrict__ sh, int sh_stride, int chunk_ok) {
  extern __shared__ __align__(16) float s_mem[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ unsigned s_n;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  float *s_sh = s_mem;                               // [BB][SHF]
  float *s_g = s_mem + BB * SHF;                     // [13][BB]: pw 3, sigma 6, opa, band0 3
  float *s_view = s_g + 13 * BB;                     // [nv][VCW]
  uint16_t *s_pair = reinterpret_cast<uint16_t *>(s_view + GSX_MAX_BATCH * VCW);   // [BB * nv]
  const int t = threadIdx.x;
  const int i = blockIdx.x * BB + t;
  const long long row0 = (long long)blockIdx.x * BB;
  const int rows = (int)min((long long)BB, (long long)N - row0);
  const bool use_tma = (L > 0) && chunk_ok && (((rows * SHF * 4) & 15) == 0);
  const int nv = bc.nv;
  if (t == 0) { s_n = 0; if (use_tma) mbar_init(&s_bar, 1); }
  copy_views_to_smem(bc, s_view);
  unsigned vis = 0;
  float3 pw = make_float3(0.f, 0.f, 0.f);
  if (i < N) {
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    for (int v = 0; v < nv; ++v) {
      const ViewConst &vc = bc.v[v];
      const float3 pc = cam_point(vc.view, pw.x, pw.y, pw.z);
      const float2 uv = screen_point(vc.proj, pc, vc.width, vc.height);
      if (keep_gaussian(uv.x, uv.y, pc.z, vc.near_thresh, vc.padding, vc.width, vc.height)) vis |= 1u << v;
    }
  }
  const int any = __syncthreads_or(vis != 0);        // also publishes s_n = 0, the mbarrier and s_view
  if (i < N)
    for (int v = 0; v < nv; ++v)
      if (!((vis >> v) & 1u)) { bc.s[v].mask[i] = 0; bc.s[v].touched[i] = 0; }
  if (!any) return;                                  // whole block culled in every view: only xyz was read
  if (L > 0) stage_rows<SHF>(s_sh, sh, sh_stride, row0, rows, use_tma, &s_bar);
  if (vis) {
    const float4 q = reinterpret_cast<const float4 *>(quat)[i];
    float sig[6];
    sigma_from_qs(q.x, q.y, q.z, q.w, scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], sig);
    s_g[0 * BB + t] = pw.x; s_g[1 * BB + t] = pw.y; s_g[2 * BB + t] = pw.z;
#pragma unroll
    for (int k = 0; k < 6; ++k) s_g[(3 + k) * BB + t] = sig[k];
    s_g[9 * BB + t] = 1.0f / (1.0f + __expf(-opacity[i]));   // render.cu:56, once per Gaussian
    s_g[10 * BB + t] = band0[3 * i]; s_g[11 * BB + t] = band0[3 * i + 1]; s_g[12 * BB + t] = band0[3 * i + 2];
    append_pairs(vis, t, &s_n, s_pair);
  }
  __syncthreads();
  if (L > 0 && use_tma) mbar_wait(&s_bar, 0);
  const int np = (int)s_n;
  const int lane = t & 31;
  for (int jb = t - lane; jb < np; jb += BB) {       // warp-uniform trip count: the tile count below is cooperative
    const int j = jb + lane;
    const bool active = j < np;
    const int slot = active ? s_pair[j] : 0;
    const int v = slot / BB, g = slot - v * BB;
    const int gi = (int)row0 + g;
    const ViewConst &vc = *reinterpret_cast<const ViewConst *>(s_view + v * VCW);
    const SlotPtrs &sp = bc.s[v];
    ObbSat so{};
    TileRange tr{0, -1, 0, -1};
    int cnt = 0;
    float zc = 0.0f;
    if (active) {
      const float px = s_g[0 * BB + g], py = s_g[1 * BB + g], pz = s_g[2 * BB + g];
      float sig[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) sig[k] = s_g[(3 + k) * BB + g];
      const float opa = s_g[9 * BB + g];
      const float b0[3] = {s_g[10 * BB + g], s_g[11 * BB + g], s_g[12 * BB + g]};
      const float3 pc = cam_point(vc.view, px, py, pz);
      const float2 uv = screen_point(vc.proj, pc, vc.width, vc.height);
      float J[6], con[3];
      float4 rad;
      projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
      conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad);
      const float2 ext = alpha_extent_b(con[0], con[1], con[2], opa);
      const TileWindow w = candidate_window(uv.x, uv.y, rad.x, vc.ntx, vc.nty);
      if (w.x1 > w.x0 && w.y1 > w.y0) {
        so = obb_sat_setup(compute_obb(uv.x, uv.y, rad.x, rad.y, rad.z, rad.w));
        tr = sat_aabb_range(so, w);
        if (vc.tight) clip_range_to_extent(tr, uv.x, uv.y, ext.x, ext.y);
      }
      const float3 rgb = sh_to_rgb<L>(px, py, pz, vc.campos, b0, L > 0 ? s_sh + g * SHF : nullptr);
      sp.r0[gi] = make_float4(uv.x, uv.y, ext.x, ext.y);
      sp.r1[gi] = make_float4(con[0], con[1], con[2], opa);
      sp.r2[gi] = make_float4(rgb.x, rgb.y, rgb.z, opa * (1.0f - opa));
      sp.radius[gi] = rad;
      sp.depth[gi] = pc.z;
      sp.mask[gi] = 1;
      zc = pc.z;
    }
    // exact tile count (F7 window + SAT): small rectangles per thread, large ones by the whole warp.  What the enumeration
    // finds is kept for the emission kernel (binning.cu k_emit_lists), which then needs neither the SAT nor the screen-space
    // records: a small rectangle leaves its 16-bit hit mask (cells tx-major, as enumerated) + origin tile + row count in an
    // 8-byte record; a large one reserves its cell count in the batch's tile-list buffer and writes its tile ids there.
    static_assert(COOP_CELLS <= 16, "a small rectangle's hit mask shares the record word with the large-rectangle flag and is staged 16 per lane");
    const bool big = active && rect_cells(tr) > COOP_CELLS;
    unsigned hitmask = 0;
    if (active && !big) {
      int b = 0;
      for (int tx = tr.x0; tx <= tr.x1; ++tx) {
        const SatCol col = sat_col(so, tx);
        for (int ty = tr.y0; ty <= tr.y1; ++ty, ++b) hitmask |= (sat_cell(so, col, ty) ? 1u : 0u) << b;
      }
      cnt = __popc(hitmask);
      if (cnt > 0) sp.trec[gi] = make_uint2((unsigned)(tr.y0 * vc.ntx + tr.x0) | ((unsigned)(tr.y1 - tr.y0 + 1) << 16), hitmask);
    }
    unsigned bm = __ballot_sync(0xffffffffu, big);
    // room for the large rectangles of this warp step: one atomic for all of them (a same-address atomic per rectangle, with
    // its return value on the critical path, cost 8 % of the kernel's stall samples)
    unsigned mybase = 0xffffffffu;   // 0xffffffff: no room in the list buffer -> the emission repeats the SAT for this one
    if (bm && bc.tl_cap) {
      const unsigned need = big ? (unsigned)rect_cells(tr) : 0u;
      unsigned incl = need;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const unsigned x = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += x;
      }
      unsigned at = 0;
      if (lane == 31) at = atomicAdd(bc.tcursor, incl);
      at = __shfl_sync(0xffffffffu, at, 31) + incl - need;
      if (big && (unsigned long long)at + need <= (unsigned long long)bc.tl_cap) mybase = at;
    }
    while (bm) {
      const int leader = __ffs(bm) - 1;
      bm &= bm - 1;
      int c = 0;
      unsigned lbase = mybase;
      lbase = __shfl_sync(0xffffffffu, lbase, leader);
      const int lntx = __shfl_sync(0xffffffffu, vc.ntx, leader);
      coop_tile_rect(leader, so, tr, [&](bool hit, unsigned ballot, int tx, int ty) {
        if (hit && lbase != 0xffffffffu) bc.tlist[lbase + c + __popc(ballot & ((1u << lane) - 1u))] = (uint16_t)(ty * lntx + tx);
        c += __popc(ballot);
      });
      if (lane == leader) {
        cnt = c;
        if (c > 0) sp.trec[gi] = make_uint2(lbase, 0x80000000u | (unsigned)c);
      }
    }
    if (active) sp.touched[gi] = cnt;
    // Gaussians with instances are appended (warp-aggregated, unordered) to the view's compact depth-key list: the
    // depth sort, the scan and the emission of the view then run over P' entries instead of N (binning.cu)
    const bool app = active && cnt > 0;
    const unsigned peers = __match_any_sync(0xffffffffu, app ? v : -1 - lane);
    if (app) {
      const int leader = __ffs(peers) - 1;
      int basep = 0;
      if (lane == leader) basep = atomicAdd(sp.ccount, __popc(peers));
      basep = __shfl_sync(peers, basep, leader);
      const int pos = basep + __popc(peers & ((1u << lane) - 1u));
      sp.ckey[pos] = min(float_sort_bits(zc), 0xfffffffeu);
      sp.cid[pos] = (uint32_t)gi;
    }
  }
}

struct BatchGradOut {
  float *xyz, *scale, *quat, *opacity, *band0, *sh, *uv;
  int sh_stride;
  int overwrite;     // 1: '=' (Gaussians without a gradient get zeros); 0: '+='
  int compat_dsph;   // Q12
  int chunk_ok;      // SH rows of params and grads dense + 16-byte aligned -> bulk async copies
};

// Backward.  Phase A (thread = Gaussian): the views in which the Gaussian touched a tile AND received a non-zero packed
// compositing gradient (all loads issued up front); their 9 gradient values go to shared-memory slots.  Phase B (thread
// = active pair): F11, F12, F14, F15 and the position term of F10 -> d xyz_w (3) and d Sigma (6), in place in the slot.
// Phase C (thread = Gaussian): sums its views, accumulates the SH coefficient gradients g_c * Y_k(view) in registers,
// runs F13 once on the summed d Sigma, re-zeroes the packed gradients it consumed and writes every gradient once.
// COEF = false: the SH coefficient / band-0 gradients are left out (view-parallel multi-GPU mode: they are rebuilt
// from the all-gathered per-view colour gradients by k_sh_from_view_rgb, see below).
template <int L, bool COEF>
__global__ void __launch_bounds__(BB, GSX_BATCH_MINB)
k_batch_bwd(const __grid_constant__ BatchConst bc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
            const float *__restrict__ quat, const float *__restrict__ band0, const float *__restrict__ sh, int sh_stride,
            const BatchGradOut out) {
  extern __shared__ __align__(16) float s_mem[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ unsigned s_n;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  float *s_coef = s_mem;                             // [BB][SHF] coefficients, re-used for the gradient rows
  float *s_g = s_mem + BB * SHF;                     // [12][BB]: pw 3, sigma 6, band0 3
  float *s_view = s_g + 12 * BB;                     // [nv][VCW]
  const int nslot = BB * bc.nv;
  float *s_stage = s_view + GSX_MAX_BATCH * VCW;     // [9][nslot]
  uint16_t *s_pair = reinterpret_cast<uint16_t *>(s_stage + 9 * nslot);
  const int t = threadIdx.x;
  const int i = blockIdx.x * BB + t;
  const long long row0 = (long long)blockIdx.x * BB;
  const int rows = (int)min((long long)BB, (long long)N - row0);
  const bool use_tma = (L > 0) && out.chunk_ok && (((rows * SHF * 4) & 15) == 0);
  const int nv = bc.nv;
  if (t == 0) { s_n = 0; if (use_tma) mbar_init(&s_bar, 1); }
  copy_views_to_smem(bc, s_view);
  // ---- phase A
  unsigned tch = 0;
  if (i < N) {
#pragma unroll
    for (int v = 0; v < GSX_MAX_BATCH; ++v)
      if (v < nv) tch |= (bc.s[v].touched[i] > 0 ? 1u : 0u) << v;
  }
  const int any_t = __syncthreads_or(tch != 0);
  if (!any_t && !out.overwrite) return;
  if (any_t && L > 0) stage_rows<SHF>(s_coef, sh, sh_stride, row0, rows, use_tma, &s_bar);   // overlaps the loads below
#ifndef GSX_BWD_EARLY
#define GSX_BWD_EARLY 1
#endif
#ifndef GSX_BWD_PF
#define GSX_BWD_PF 4
#endif
  // parameters of every touched Gaussian are requested now, together with the packed gradients below, instead of one
  // more dependent round trip after the gradients turned out non-zero (a touched Gaussian whose gradients are all zero
  // -- occluded in every view -- costs 52 B of reads it did not need)
  float3 pw = make_float3(0.f, 0.f, 0.f);
  float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, b0r = 0.f, b0g = 0.f, b0b = 0.f;
  if (GSX_BWD_EARLY && tch) {
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    q = reinterpret_cast<const float4 *>(quat)[i];
    s0 = scale[3 * i]; s1 = scale[3 * i + 1]; s2 = scale[3 * i + 2];
    b0r = band0[3 * i]; b0g = band0[3 * i + 1]; b0b = band0[3 * i + 2];
  }
  unsigned act = 0;
  // GSX_BWD_PF views per trip: their packed gradients are all requested before any is tested (one DRAM round trip per
  // group of views instead of one per view -- 47 % of this kernel's stall samples waited on these dependent loads)
  for (int v0 = 0; v0 < nv; v0 += GSX_BWD_PF) {
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 a[GSX_BWD_PF][3];
    bool tv[GSX_BWD_PF];
#pragma unroll
    for (int u = 0; u < GSX_BWD_PF; ++u) {
      const int v = v0 + u;
      tv[u] = v < nv && ((tch >> v) & 1u);
      a[u][0] = a[u][1] = a[u][2] = z4;
      if (tv[u]) {
        const float4 *ga = bc.s[v].gacc + 3 * (size_t)i;
        a[u][0] = ga[0]; a[u][1] = ga[1]; a[u][2] = ga[2];
      }
    }
#pragma unroll
    for (int u = 0; u < GSX_BWD_PF; ++u) {
      const int v = v0 + u;
      if (!tv[u]) continue;
      const float4 a0 = a[u][0], a1 = a[u][1], a2 = a[u][2];
      const bool nz = (a0.x != 0.f) | (a0.y != 0.f) | (a0.z != 0.f) | (a0.w != 0.f) | (a1.x != 0.f) | (a1.y != 0.f) |
                      (a1.z != 0.f) | (a1.w != 0.f) | (a2.x != 0.f);
      if (!nz) continue;
      act |= 1u << v;
      float *st = s_stage + v * BB + t;
      st[0 * nslot] = a0.x; st[1 * nslot] = a0.y; st[2 * nslot] = a0.z; st[3 * nslot] = a0.w;
      st[4 * nslot] = a1.x; st[5 * nslot] = a1.y; st[6 * nslot] = a1.z; st[7 * nslot] = a1.w; st[8 * nslot] = a2.x;
    }
  }
  if (act) {
    if (!GSX_BWD_EARLY) {
      pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
      q = reinterpret_cast<const float4 *>(quat)[i];
      s0 = scale[3 * i]; s1 = scale[3 * i + 1]; s2 = scale[3 * i + 2];
      b0r = band0[3 * i]; b0g = band0[3 * i + 1]; b0b = band0[3 * i + 2];
    }
    float sig[6];
    sigma_from_qs(q.x, q.y, q.z, q.w, s0, s1, s2, sig);
    s_g[0 * BB + t] = pw.x; s_g[1 * BB + t] = pw.y; s_g[2 * BB + t] = pw.z;
#pragma unroll
    for (int k = 0; k < 6; ++k) s_g[(3 + k) * BB + t] = sig[k];
    s_g[9 * BB + t] = b0r; s_g[10 * BB + t] = b0g; s_g[11 * BB + t] = b0b;
    append_pairs(act, t, &s_n, s_pair);
  }
  __syncthreads();
  if (any_t && L > 0 && use_tma) mbar_wait(&s_bar, 0);
  // ---- phase B
  const int np = (int)s_n;
  for (int j = t; j < np; j += BB) {
    const int slot = s_pair[j];
    const int v = slot / BB, g = slot - v * BB;
    const ViewConst &vc = *reinterpret_cast<const ViewConst *>(s_view + v * VCW);
    float *st = s_stage + slot;
    const float g_r = st[0 * nslot], g_g = st[1 * nslot], g_b = st[2 * nslot];
    const float g_u = st[4 * nslot], g_v = st[5 * nslot];
    const float g_con[3] = {st[6 * nslot], st[7 * nslot], st[8 * nslot]};
    const float px = s_g[0 * BB + g], py = s_g[1 * BB + g], pz = s_g[2 * BB + g];
    float sig[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) sig[k] = s_g[(3 + k) * BB + g];
    const float b0[3] = {s_g[9 * BB + g], s_g[10 * BB + g], s_g[11 * BB + g]};
    const float3 pc = cam_point(vc.view, px, py, pz);
    float J[6], con[3], dS[6], dJ[6];
    float4 rad_unused;
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad_unused);   // same arithmetic as the forward: same bits
    conic_backward(J, sig, vc.view, con, g_con, dS, dJ);                // F11
    float3 gc = make_float3(0.f, 0.f, 0.f), tt;
    if (jacobian_backward(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, dJ, &tt)) { gc.x += tt.x; gc.y += tt.y; gc.z += tt.z; }   // F12
    if (screen_backward(vc.proj, pc.x, pc.y, pc.z, g_u, g_v, vc.width, vc.height, &tt)) { gc.x += tt.x; gc.y += tt.y; gc.z += tt.z; }          // F14
    const float3 gw = cam_point_backward(vc.view, gc.x, gc.y, gc.z);   // F15
    const float3 gsh = sh_dir_backward<L>(px, py, pz, vc.campos, b0, L > 0 ? s_coef + g * SHF : nullptr, g_r, g_g, g_b,
                                          out.compat_dsph != 0);       // F10, position term
    st[0 * nslot] = gsh.x + gw.x; st[1 * nslot] = gsh.y + gw.y; st[2 * nslot] = gsh.z + gw.z;
#pragma unroll
    for (int k = 0; k < 6; ++k) st[(3 + k) * nslot] = dS[k];
  }
  __syncthreads();   // slots hold the pair results; the coefficient rows are free to be overwritten
  // ---- phase C
  float gxyz[3] = {0.f, 0.f, 0.f}, dSt[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, gb0[3] = {0.f, 0.f, 0.f};
  constexpr int NG = (COEF && SHF > 0) ? SHF : 1;
  float gsh[NG];
#pragma unroll
  for (int k = 0; k < NG; ++k) gsh[k] = 0.0f;
  float gopa = 0.0f, gu = 0.0f, gv = 0.0f;
  for (int v = 0; v < nv; ++v) {
    if (!((act >> v) & 1u)) continue;
    float4 *ga = bc.s[v].gacc + 3 * (size_t)i;
    const float4 a0 = ga[0], a1 = ga[1];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    ga[0] = z4; ga[1] = z4; ga[2] = z4;   // the slot's packed gradients are zero again for the next batch
    const float *st = s_stage + v * BB + t;
    gxyz[0] += st[0 * nslot]; gxyz[1] += st[1 * nslot]; gxyz[2] += st[2 * nslot];
#pragma unroll
    for (int k = 0; k < 6; ++k) dSt[k] += st[(3 + k) * nslot];
    gopa += a0.w; gu += a1.x; gv += a1.y;
    if (COEF) {
      float Y[n];
      const ViewConst &vc = *reinterpret_cast<const ViewConst *>(s_view + v * VCW);
      sh_basis_of_point<L>(pw.x, pw.y, pw.z, vc.campos, Y);
      gb0[0] += a0.x * Y[0]; gb0[1] += a0.y * Y[0]; gb0[2] += a0.z * Y[0];
#pragma unroll
      for (int k = 1; k < n; ++k) {
        gsh[3 * (k - 1) + 0] += a0.x * Y[k]; gsh[3 * (k - 1) + 1] += a0.y * Y[k]; gsh[3 * (k - 1) + 2] += a0.z * Y[k];
      }
    }
  }
  float gq[4] = {0.f, 0.f, 0.f, 0.f}, gs[3] = {0.f, 0.f, 0.f};
  if (act) sigma_backward(q.x, q.y, q.z, q.w, s0, s1, s2, dSt, gq, gs);   // F13, once on the view-summed dSigma
  if (i < N && (act || out.overwrite)) {
    if (out.overwrite) {
      out.xyz[3 * i] = gxyz[0]; out.xyz[3 * i + 1] = gxyz[1]; out.xyz[3 * i + 2] = gxyz[2];
      out.scale[3 * i] = gs[0]; out.scale[3 * i + 1] = gs[1]; out.scale[3 * i + 2] = gs[2];
      reinterpret_cast<float4 *>(out.quat)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
      out.opacity[i] = gopa;
      if (COEF) { out.band0[3 * i] = gb0[0]; out.band0[3 * i + 1] = gb0[1]; out.band0[3 * i + 2] = gb0[2]; }
      if (out.uv) { out.uv[2 * i] = gu; out.uv[2 * i + 1] = gv; }
    } else {
      const float ox0 = out.xyz[3 * i], ox1 = out.xyz[3 * i + 1], ox2 = out.xyz[3 * i + 2];
      const float os0 = out.scale[3 * i], os1 = out.scale[3 * i + 1], os2 = out.scale[3 * i + 2];
      float4 *qq = reinterpret_cast<float4 *>(out.quat) + i;
      float4 c = *qq;
      const float oo = out.opacity[i];
      float ob0 = 0.f, ob1 = 0.f, ob2 = 0.f;
      if (COEF) { ob0 = out.band0[3 * i]; ob1 = out.band0[3 * i + 1]; ob2 = out.band0[3 * i + 2]; }
      float ou = 0.f, ov = 0.f;
      if (out.uv) { ou = out.uv[2 * i]; ov = out.uv[2 * i + 1]; }
      out.xyz[3 * i] = ox0 + gxyz[0]; out.xyz[3 * i + 1] = ox1 + gxyz[1]; out.xyz[3 * i + 2] = ox2 + gxyz[2];
      out.scale[3 * i] = os0 + gs[0]; out.scale[3 * i + 1] = os1 + gs[1]; out.scale[3 * i + 2] = os2 + gs[2];
      c.x += gq[0]; c.y += gq[1]; c.z += gq[2]; c.w += gq[3];
      *qq = c;
      out.opacity[i] = oo + gopa;
      if (COEF) { out.band0[3 * i] = ob0 + gb0[0]; out.band0[3 * i + 1] = ob1 + gb0[1]; out.band0[3 * i + 2] = ob2 + gb0[2]; }
      if (out.uv) { out.uv[2 * i] = ou + gu; out.uv[2 * i + 1] = ov + gv; }
    }
  }
  if (L > 0 && COEF) {
    float *grow = s_coef + t * SHF;
#pragma unroll
    for (int k = 0; k < SHF; ++k) grow[k] = gsh[k];
    if (use_tma) {
      fence_proxy_async_smem();
      __syncthreads();
      if (t == 0) {
        float *dst = out.sh + row0 * SHF;
        const uint32_t bytes = (uint32_t)rows * SHF * 4u;
        if (out.overwrite) bulk_store(dst, s_coef, bytes);
        else bulk_reduce_add_f32(dst, s_coef, bytes);
        bulk_commit();
        bulk_wait_read_all();   // shared memory must outlive the bulk store
      }
    } else {
      __syncthreads();
      for (int k = t; k < rows * SHF; k += BB) {
        const int r = k / (SHF > 0 ? SHF : 1), c = k - r * SHF;
        float *o = out.sh + (size_t)(row0 + r) * out.sh_stride + c;
        *o = out.overwrite ? s_coef[k] : *o + s_coef[k];
      }
    }
  }
}

// View-parallel multi-GPU mode.  d rgb of one view, dense [N][3] (zeros where the Gaussian got no gradient): the rank-1
// factor of that view's SH coefficient gradient, d sh[k][c] = d rgb_c * Y_k(dir_view)
// (cuda/spherical_harmonics_backward.cu:100-104).  12 B/Gaussian/view cross NVLink instead of 192 B/Gaussian.
__global__ void __launch_bounds__(256) k_extract_view_rgb(const float4 *__restrict__ gacc, const int *__restrict__ touched,
                                                          int N, float *__restrict__ drgb) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= N) return;
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  if (touched[i] > 0) a = gacc[3 * (size_t)i];
  drgb[3 * (size_t)i] = a.x; drgb[3 * (size_t)i + 1] = a.y; drgb[3 * (size_t)i + 2] = a.z;
}

// Fused extraction + all-gather: the same d rgb rows leave the SM straight into EVERY rank's gather buffer -- one
// multimem.st per 16 bytes through the NVSwitch multicast address (the switch replicates; SASS STG.E.128.STRONG.SYS on a
// multicast mapping), or one plain st.global per peer pointer (P2P over NVLink) when no multicast mapping exists.
// The transfer overlaps the compositing of the rank's next views / the per-Gaussian backward; no NCCL call, no staging.
struct PeerPtrs { float *p[8]; int n; };
__device__ __forceinline__ void multimem_st_v4(float *mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__global__ void __launch_bounds__(256) k_extract_view_rgb_peers(const float4 *__restrict__ gacc,
                                                                const int *__restrict__ touched,
                                                                const float *__restrict__ src_dense /*or null*/, int N,
                                                                float *mc /*multicast address of the view's [N][3] or null*/,
                                                                const PeerPtrs peers) {
  // one ADDRESS-aligned 16-byte group of the flat [3N] array per thread (a view's block may start at any 4-byte offset)
  const float *base = mc ? mc : peers.p[0];
  const int mis = (int)((reinterpret_cast<uintptr_t>(base) >> 2) & 3);   // identical on every rank (symmetric offsets)
  const long long q = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long total = 3LL * N;
  const long long first = 4 * q - mis;
  if (first >= total) return;
  float v[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const long long idx = first + e;
    float x = 0.0f;
    if (idx >= 0 && idx < total) {
      if (src_dense) {   // push of an already extracted dense block (runs on a side stream, gsx_push_view_rgb_peers)
        x = src_dense[idx];
      } else {
        const int i = (int)(idx / 3), c = (int)(idx - 3LL * i);
        if (touched[i] > 0) {
          const float4 a = gacc[3 * (size_t)i];
          x = c == 0 ? a.x : (c == 1 ? a.y : a.z);
        }
      }
    }
    v[e] = x;
  }
  if (first >= 0 && first + 3 < total) {
    const float4 o = make_float4(v[0], v[1], v[2], v[3]);
    if (mc) multimem_st_v4(mc + first, o);
    else
      for (int r = 0; r < peers.n; ++r) *reinterpret_cast<float4 *>(peers.p[r] + first) = o;
  } else {   // ragged head / tail: scalar stores to every peer
    for (int e = 0; e < 4; ++e) {
      const long long idx = first + e;
      if (idx < 0 || idx >= total) continue;
      if (mc) asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc + idx), "f"(v[e]) : "memory");
      else
        for (int r = 0; r < peers.n; ++r) peers.p[r][idx] = v[e];
    }
  }
}

// In-switch allreduce (NVLS): rank r sums its 1/world slice of the buffer over ALL ranks with multimem.ld_reduce (SASS
// LDGMC.ADD: the NVSwitch adds the ranks' copies) and broadcasts the sums with multimem.st.  Each byte crosses a GPU's
// links once in each direction.  Callers barrier across ranks before (inputs written) and after (outputs complete).
__global__ void __launch_bounds__(256) k_allreduce_multimem(float *mc, long long n4_begin, long long n4_end) {
  // few CTAs, four independent 16-byte reductions in flight per thread: the kernel shares the GPU with the HBM-bound SH
  // rebuild, so it should hold the links busy without occupying every SM
  const long long stride = (long long)gridDim.x * 256;
  long long q = n4_begin + (long long)blockIdx.x * 256 + threadIdx.x;
  for (; q + 3 * stride < n4_end; q += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                   : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w)
                   : "l"(mc + 4 * (q + u * stride))
                   : "memory");
#pragma unroll
    for (int u = 0; u < 4; ++u) multimem_st_v4(mc + 4 * (q + u * stride), v[u]);
  }
  for (; q < n4_end; q += stride) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(mc + 4 * q)
                 : "memory");
    multimem_st_v4(mc + 4 * q, v);
  }
}

struct ViewCams { float campos[GSX_MAX_VIEWS_TOTAL][3]; int nv; };

// SH coefficient + band-0 gradients of a whole batch from the per-view colour gradients of ALL its views (gathered
// from every rank): out[k][c] = sum_v drgb[v][i][c] * Y_k(dir_v(i)).  One pass: reads 12 B per (view, Gaussian), writes
// the 180 + 12 B gradient row once (rows leave through one bulk async store per block).
template <int L>
__global__ void __launch_bounds__(BB) k_sh_from_view_rgb(const __grid_constant__ ViewCams vc, int N,
                                                          const float *__restrict__ xyz, const float *__restrict__ drgb,
                                                          size_t view_stride, float *__restrict__ sh_grad, int sh_stride,
                                                          float *__restrict__ band0_grad, int overwrite, int chunk_ok) {
  extern __shared__ __align__(16) float s_rows[];
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int t = threadIdx.x;
  const int i = blockIdx.x * BB + t;
  const long long row0 = (long long)blockIdx.x * BB;
  const int rows = (int)min((long long)BB, (long long)N - row0);
  const bool use_tma = (L > 0) && chunk_ok && (((rows * SHF * 4) & 15) == 0);
  float gsh[SHF > 0 ? SHF : 1], gb0[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < SHF; ++k) gsh[k] = 0.0f;
  if (i < N) {
    const float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
    for (int v = 0; v < vc.nv; ++v) {
      const float *g = drgb + (size_t)v * view_stride + 3 * (size_t)i;
      const float gr = g[0], gg = g[1], gb = g[2];
      if (gr == 0.f && gg == 0.f && gb == 0.f) continue;
      float Y[n];
      sh_basis_of_point<L>(px, py, pz, make_float3(vc.campos[v][0], vc.campos[v][1], vc.campos[v][2]), Y);
      gb0[0] += gr * Y[0]; gb0[1] += gg * Y[0]; gb0[2] += gb * Y[0];
#pragma unroll
      for (int k = 1; k < n; ++k) {
        gsh[3 * (k - 1) + 0] += gr * Y[k]; gsh[3 * (k - 1) + 1] += gg * Y[k]; gsh[3 * (k - 1) + 2] += gb * Y[k];
      }
    }
    if (overwrite) { band0_grad[3 * i] = gb0[0]; band0_grad[3 * i + 1] = gb0[1]; band0_grad[3 * i + 2] = gb0[2]; }
    else { band0_grad[3 * i] += gb0[0]; band0_grad[3 * i + 1] += gb0[1]; band0_grad[3 * i + 2] += gb0[2]; }
  }
  if (L > 0) {
    float *grow = s_rows + t * SHF;
#pragma unroll
    for (int k = 0; k < SHF; ++k) grow[k] = gsh[k];
    if (use_tma) {
      fence_proxy_async_smem();
      __syncthreads();
      if (t == 0) {
        float *dst = sh_grad + row0 * SHF;
        const uint32_t bytes = (uint32_t)rows * SHF * 4u;
        if (overwrite) bulk_store(dst, s_rows, bytes);
        else bulk_reduce_add_f32(dst, s_rows, bytes);
        bulk_commit();
        bulk_wait_read_all();
      }
    } else {
      __syncthreads();
      for (int k = t; k < rows * SHF; k += BB) {
        const int r = k / (SHF > 0 ? SHF : 1), c = k - r * SHF;
        float *o = sh_grad + (size_t)(row0 + r) * sh_stride + c;
        *o = overwrite ? s_rows[k] : *o + s_rows[k];
      }
    }
  }
}

int sh_floats(int l_max) { return ((l_max + 1) * (l_max + 1) - 1) * 3; }

}  // namespace

int batch_fwd_launch(const BatchConst &bc, const gsx_params_t &p, int N, int l_max, cudaStream_t st) {
  if (N <= 0 || bc.nv <= 0) return GSX_OK;
  const int shf = sh_floats(l_max);
  const int ok = l_max > 0 && p.sh_stride == shf && (((uintptr_t)p.sh) & 15) == 0;
  const size_t sm = ((size_t)BB * shf + 13 * BB + GSX_MAX_BATCH * VCW) * sizeof(float) + (size_t)BB * bc.nv * sizeof(uint16_t);
  const int grid = ceil_div(N, BB);
#define GSX_BF(L)                                                                                                     \
  do {                                                                                                                \
    GSX_CUDA(cudaFuncSetAttribute(k_batch_fwd<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));              \
    k_batch_fwd<L><<<grid, BB, sm, st>>>(bc, N, p.xyz, p.scale, p.quaternion, p.opacity, p.band0, p.sh, p.sh_stride, ok); \
  } while (0)
  switch (l_max) {
    case 0: GSX_BF(0); break;
    case 1: GSX_BF(1); break;
    case 2: GSX_BF(2); break;
    case 3: GSX_BF(3); break;
    default: set_error("l_max %d out of range 0..3", l_max); return GSX_E_ARG;
  }
#undef GSX_BF
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int batch_bwd_launch(const BatchConst &bc, const gsx_params_t &p, int N, int l_max, const gsx_grads_t &g, int flags,
                     cudaStream_t st) {
  if (N <= 0 || bc.nv <= 0) return GSX_OK;
  const int shf = sh_floats(l_max);
  BatchGradOut o;
  o.xyz = g.xyz; o.scale = g.scale; o.quat = g.quaternion; o.opacity = g.opacity; o.band0 = g.band0; o.sh = g.sh;
  o.uv = g.uv;
  o.sh_stride = p.sh_stride;
  o.overwrite = (flags & GSX_GRAD_OVERWRITE) ? 1 : 0;
  o.compat_dsph = (flags & GSX_SH_DSPH_CORRECT) ? 0 : 1;
  o.chunk_ok = l_max > 0 && p.sh_stride == shf && (((uintptr_t)p.sh) & 15) == 0 && (((uintptr_t)g.sh) & 15) == 0;
  const size_t sm = ((size_t)BB * shf + 12 * BB + GSX_MAX_BATCH * VCW + (size_t)9 * BB * bc.nv) * sizeof(float) + (size_t)BB * bc.nv * sizeof(uint16_t);
  const int grid = ceil_div(N, BB);
  const bool coef = !(flags & GSX_GRAD_NO_SH);
#define GSX_BB(L)                                                                                                   \
  do {                                                                                                              \
    if (coef) {                                                                                                     \
      GSX_CUDA(cudaFuncSetAttribute(k_batch_bwd<L, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));    \
      k_batch_bwd<L, true><<<grid, BB, sm, st>>>(bc, N, p.xyz, p.scale, p.quaternion, p.band0, p.sh, p.sh_stride, o); \
    } else {                                                                                                        \
      GSX_CUDA(cudaFuncSetAttribute(k_batch_bwd<L, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));   \
      k_batch_bwd<L, false><<<grid, BB, sm, st>>>(bc, N, p.xyz, p.scale, p.quaternion, p.band0, p.sh, p.sh_stride, o); \
    }                                                                                                               \
  } while (0)
  switch (l_max) {
    case 0: GSX_BB(0); break;
    case 1: GSX_BB(1); break;
    case 2: GSX_BB(2); break;
    case 3: GSX_BB(3); break;
    default: set_error("l_max %d out of range 0..3", l_max); return GSX_E_ARG;
  }
#undef GSX_BB
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int extract_view_rgb_launch(const float4 *gacc, const int *touched, int N, float *drgb, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  k_extract_view_rgb<<<ceil_div(N, 256), 256, 0, st>>>(gacc, touched, N, drgb);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int extract_view_rgb_peers_launch(const float4 *gacc, const int *touched, const float *src_dense, int N, float *mc,
                                  float *const *peers, int n_peers, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  if (!mc && (n_peers < 1 || n_peers > 8)) { set_error("n_peers %d out of range 1..8", n_peers); return GSX_E_ARG; }
  PeerPtrs pp;
  pp.n = mc ? 0 : n_peers;
  for (int r = 0; r < 8; ++r) pp.p[r] = (!mc && r < n_peers) ? peers[r] : nullptr;
  const long long groups = (3LL * N + 3 + 3) / 4;   // + up to 3 floats of head misalignment
  k_extract_view_rgb_peers<<<ceil_div(groups, 256), 256, 0, st>>>(gacc, touched, src_dense, N, mc, pp);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int allreduce_multimem_launch(float *mc, size_t n_floats, int rank, int world, cudaStream_t st) {
  if (!mc || world < 1 || rank < 0 || rank >= world || (n_floats & 3)) { set_error("bad multimem allreduce arguments"); return GSX_E_ARG; }
  const long long n4 = (long long)(n_floats / 4);
  const long long per = (n4 + world - 1) / world;
  const long long b = std::min<long long>(n4, per * rank), e = std::min<long long>(n4, per * (rank + 1));
  if (e <= b) return GSX_OK;
  const int blocks = (int)std::min<long long>(48, (e - b + 255) / 256);
  k_allreduce_multimem<<<blocks, 256, 0, st>>>(mc, b, e);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int sh_from_view_rgb_launch(const float *xyz, int N, int l_max, const float *campos_host, int n_views, const float *drgb,
                            size_t view_stride, float *sh_grad, int sh_stride, float *band0_grad, int flags,
                            cudaStream_t st) {
  if (N <= 0 || n_views <= 0) return GSX_OK;
  if (n_views > GSX_MAX_VIEWS_TOTAL) { set_error("n_views %d > %d", n_views, GSX_MAX_VIEWS_TOTAL); return GSX_E_ARG; }
  ViewCams vc;
  vc.nv = n_views;
  for (int v = 0; v < n_views; ++v)
    for (int k = 0; k < 3; ++k) vc.campos[v][k] = campos_host[3 * v + k];
  const int shf = sh_floats(l_max);
  const int ok = l_max > 0 && sh_stride == shf && (((uintptr_t)sh_grad) & 15) == 0;
  const int overwrite = (flags & GSX_GRAD_OVERWRITE) ? 1 : 0;
  const size_t sm = (size_t)BB * shf * sizeof(float);
  const int grid = ceil_div(N, BB);
#define GSX_SV(L) \
  k_sh_from_view_rgb<L><<<grid, BB, sm, st>>>(vc, N, xyz, drgb, view_stride, sh_grad, sh_stride, band0_grad, overwrite, ok)
  switch (l_max) {
    case 0: GSX_SV(0); break;
    case 1: GSX_SV(1); break;
    case 2: GSX_SV(2); break;
    case 3: GSX_SV(3); break;
    default: set_error("l_max %d out of range 0..3", l_max); return GSX_E_ARG;
  }
#undef GSX_SV
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // namespace gsx
