// composite3.cu -- view-level tile compositing, third generation (F8 forward, reference cuda/render.cu:6-108; F9 backward,
// reference cuda/render_backward.cu:11-223).  Same mapping as k_composite_fwd2 / k_composite_bwd2 (composite.cu): one
// 128-thread CTA per 16x16 tile, a warp owns an 8x8 pixel block, a lane the two pixels (column lane&7, rows lane>>3 and
// +4), per-warp culling of the staged entries by their alpha >= 1/255 extent box, early termination by ballot.
// What changed (round 2; ncu of the v2 kernels: 610 M warp instructions per backward launch for 4.7 G algorithmic
// FMA-equivalents, issue-bound at 80 % -- 173 SASS instructions per contributing warp-hit, a shared-memory float
// atomicAdd that compiles to an ATOMS.CAST.SPIN loop, and a synchronous __ldg -> st.shared -> barrier staging):
//   * the two pixels of a lane are evaluated with PACKED fp32x2 arithmetic (fma.rn.f32x2 / mul / add -> SASS FFMA2,
//     FMUL2, FADD2 on sm_100a): one issue slot per pair instead of two for everything that is not MUFU / compare / select;
//   * the per-entry invariants of the exponent are hoisted out of the pixel loop (reference render.cu:64-66 does the same
//     with basic/linear/quad): log2(e) * power = dy * (hc * dy + B) + A with A, B, hc per (lane, entry), so the exponent of a
//     pixel pair is two FFMA2 and exp() is one bare MUFU.EX2 per pixel;
//   * backward: the colour accumulated behind an entry only ever enters through its dot product with the pixel's image
//     gradient, so ONE running scalar s = sum_c acc_c * g_c replaces the three running colours (6 packed operations per
//     entry instead of 12); dL/dpower uses the un-clamped alpha directly (G * opa, render_backward.cu:154);
//   * batches of 128 records are staged with cp.async (LDGSTS, 16-byte gathers straight into shared memory, no registers)
//     into a DOUBLE-BUFFERED ring: the records of batch b+1 and the ids of batch b+2 are in flight while batch b is
//     evaluated (north_star: "per-tile Gaussian batches staged into shared memory with TMA/cp.async");
//   * backward: no shared-memory accumulation at all.  The 9 per-entry sums are final values (the linear map from the
//     moments to the uv / conic / opacity gradients runs on the lane's two-pixel partial sums), reduced across the warp by
//     the 12-shuffle transposing butterfly, and leave as ONE red.global.add.f32 instruction whose 9 active lanes hit the 9
//     consecutive floats of the Gaussian's packed gradient row (36 bytes, 1-2 L2 sectors): no CAS loop, no per-batch
//     zeroing, no flush pass, no CTA barrier on the reduction path.
// Bound: FP32 issue (no tensor cores: the path has no dense contraction).  Algorithmic work per SURVEY 8d.
#include <stdlib.h>

#include "gsx_common.cuh"

// r2n sweep of the whole C4 step (profiles/sweep_c4_minb.sh, views/s): backward cap 5 / 6 / 7 / 8 / 9 CTAs per SM = 575 / 585 / 591 /
// 583 / 578 (78 -> 71 registers at 7; shared memory still holds 6 CTAs, the gain is the tighter schedule); with the forward
// at 14 (below): 595
#ifndef GSX_C4_BWD_MINB
#define GSX_C4_BWD_MINB 7
#endif
#ifndef GSX_C4_DEFAULT
#define GSX_C4_DEFAULT 1   // B200, C4 (profiles/r2d_*): 453 M -> 356 M warp instructions, 0.518 -> 0.484 ms per view
#endif
#ifndef GSX_C3_DEFAULT_CTA
#define GSX_C3_DEFAULT_CTA 128
#endif
// resident CTAs per SM the register allocation is capped for (measured on B200, profiles/r2c_*: backward 8 -> 10 CTAs
// (59 -> 48 registers, 4 bytes spilled) 0.540 -> 0.518 ms, 9 CTAs 0.554; forward 8 -> 12 CTAs (50 -> 40 registers) 0.262 -> 0.260)
#ifndef GSX_C3_BWD_MINB
#define GSX_C3_BWD_MINB 10
#endif
// r2n: forward cap 10 / 12 / 14 / 16 = 577 / 585 / 588 / 588 views/s (32 registers + 8 bytes spilled from 14 on)
#ifndef GSX_C3_FWD_MINB
#define GSX_C3_FWD_MINB 14
#endif

namespace gsx {
namespace {

constexpr float ALPHA_MIN = 0.00392156862f;   // render.cu:74, render_backward.cu:123
constexpr float LOG2E = 1.4426950408889634f;

__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 bc(float x) { return make_float2(x, x); }
__device__ __forceinline__ float ex2(float x) {   // one MUFU.EX2; results below 2^-126 flush to zero (far under 1/255)
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp(float x) {   // 1 - alpha in [0.01, 1]: one MUFU.RCP, no slow path
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void cp_async16(void *smem, const void *g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int NT>
struct Ring {   // two stages of NT records (one per thread of the CTA)
  float4 r0[2][NT], r1[2][NT], r2[2][NT];
};
// NT = 128: one CTA per 16x16 tile, four warps share the staged batches (CTA barriers around each batch).
// NT = 32 (GSX_COMPOSITE_CTA=32, kept for A/B): one single-warp CTA per 8x8 pixel block: the warp stages its own batches of
// 32 records (the tile's list is then fetched four times, from L2), needs no CTA barrier, terminates and retires on its
// own.  Measured on B200 (profiles/r2c_*): the barrier stalls disappear (1.35 -> 0 cycles per issued instruction) but the
// kernel executes 5 % more instructions at the same 3.1 instructions / cycle / SM: backward 0.540 -> 0.555 ms, forward
// 0.262 -> 0.270 ms.  The tile-per-CTA shape stays the default: the kernels are bound by issue slots, not by barriers.
template <int NT>
__device__ __forceinline__ void cta_sync() { if (NT == 32) __syncwarp(); else __syncthreads(); }
template <int NT>
__device__ __forceinline__ bool cta_all(bool p) { return NT == 32 ? p : (__syncthreads_and(p) != 0); }

template <int NT>
__device__ __forceinline__ void stage(Ring<NT> &s, int buf, int tid, int g, const float4 *__restrict__ r0,
                                      const float4 *__restrict__ r1, const float4 *__restrict__ r2) {
  if (g >= 0) {
    cp_async16(&s.r0[buf][tid], r0 + g);
    cp_async16(&s.r1[buf][tid], r1 + g);
    cp_async16(&s.r2[buf][tid], r2 + g);
  }
  cp_commit();   // every thread commits a (possibly empty) group per batch: wait_group counts stay uniform
}

// per (lane, entry) invariants of log2(e) * power = dy * (hc * dy + B) + A
struct Quad { float A, B, hc; };
__device__ __forceinline__ Quad quad_of(float dx, float a, float b, float c) {
  const float dxs = dx * (-0.5f * LOG2E);
  Quad q;
  q.A = (a * dx) * dxs;
  q.B = b * (dxs + dxs);
  q.hc = c * (-0.5f * LOG2E);
  return q;
}

// ---------------------------------------------------------------------------------------------------------------- forward
template <int NT>
__global__ void __launch_bounds__(NT, (NT == 32 ? 32 : GSX_C3_FWD_MINB))
k_composite_fwd3(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const float4 *__restrict__ r2,
                 const int *__restrict__ sorted, const int *__restrict__ ranges, const int *__restrict__ order, int ntx, int W, int H, float bg,
                 int *__restrict__ splats_per_pixel, float *__restrict__ weight_per_pixel, float *__restrict__ image) {
  constexpr int CT = NT;
  __shared__ Ring<NT> s;
  const int tslot = NT == 32 ? blockIdx.x >> 2 : blockIdx.x;
  const int tile = order ? order[tslot] : tslot;   // longest lists first (k_tile_order): the tail of the grid is short tiles
  const int tid = threadIdx.x, lane = tid & 31, warp = NT == 32 ? (blockIdx.x & 3) : (tid >> 5);
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 8;
  const int px = X0 + (lane & 7), py0 = Y0 + (lane >> 3);
  const float fpx = (float)px;
  const float2 fpy = make_float2((float)py0, (float)(py0 + 4));
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 7);
  const int start = ranges[tile];
  const int total = ranges[tile + 1] - start;
  const bool in0 = px < W && py0 < H, in1 = px < W && py0 + 4 < H;

  float2 T = bc(1.0f), cr = bc(0.0f), cg = bc(0.0f), cb = bc(0.0f);
  bool done0 = !in0, done1 = !in1;
  int count0 = total, count1 = total;
  bool warp_done = __ballot_sync(0xffffffffu, !(done0 && done1)) == 0;

  const int nbatch = (total + CT - 1) / CT;
  int id_next = tid < total ? sorted[start + tid] : -1;
  if (nbatch > 0) stage(s, 0, tid, id_next, r0, r1, r2);
  id_next = CT + tid < total ? sorted[start + CT + tid] : -1;
  for (int b = 0; b < nbatch; ++b) {
    const int buf = b & 1, base = b * CT;
    // records of batch b+1 and ids of batch b+2 go in flight before batch b is touched (the other stage was last read in
    // iteration b-1, which ended with a CTA barrier)
    stage(s, buf ^ 1, tid, b + 1 < nbatch ? id_next : -1, r0, r1, r2);
    id_next = base + 2 * CT + tid < total ? sorted[start + base + 2 * CT + tid] : -1;
    cp_wait<1>();
    cta_sync<NT>();
    const int nb = min(CT, total - base);
    if (!warp_done) {
      for (int c = 0; c < nb; c += 32) {
        const int j = c + lane;
        bool hit = false;
        if (j < nb) {
          const float4 q = s.r0[buf][j];
          hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
        }
        unsigned m = __ballot_sync(0xffffffffu, hit);
        while (m) {
          const int e = c + __ffs(m) - 1;
          m &= m - 1;
          const float2 uv = *reinterpret_cast<const float2 *>(&s.r0[buf][e]);
          const float4 q1 = s.r1[buf][e];
          const Quad q = quad_of(uv.x - fpx, q1.x, q1.y, q1.z);
          const float2 dy = add2(bc(uv.y), make_float2(-fpy.x, -fpy.y));
          const float2 p2 = fma2(dy, fma2(bc(q.hc), dy, bc(q.B)), bc(q.A));
          const float a0 = fminf(0.99f, q1.w * ex2(fminf(0.0f, p2.x)));
          const float a1 = fminf(0.99f, q1.w * ex2(fminf(0.0f, p2.y)));
          const bool t0 = a0 > ALPHA_MIN && !done0, t1 = a1 > ALPHA_MIN && !done1;
          if (!__any_sync(0xffffffffu, t0 || t1)) continue;
          // branch-free blend: a pixel that does not take the entry runs it with alpha = 0 (T * 1, colour + 0: exact no-ops)
          const float4 q2 = s.r2[buf][e];
          const float2 al = make_float2(t0 ? a0 : 0.0f, t1 ? a1 : 0.0f);
          const float2 w = mul2(al, T);
          T = mul2(T, add2(bc(1.0f), make_float2(-al.x, -al.y)));   // T (1 - alpha), render.cu:75, same two roundings
          cr = fma2(bc(q2.x), w, cr); cg = fma2(bc(q2.y), w, cg); cb = fma2(bc(q2.z), w, cb);
          if (t0 && T.x < 0.0001f) { done0 = true; count0 = base + e + 1; }   // render.cu:76: the saturating splat is counted
          if (t1 && T.y < 0.0001f) { done1 = true; count1 = base + e + 1; }
        }
        if (__ballot_sync(0xffffffffu, !(done0 && done1)) == 0) { warp_done = true; break; }
      }
    }
    if (cta_all<NT>(warp_done)) break;
  }
  cp_wait<0>();   // nothing may still be landing in shared memory when the CTA retires
  if (in0) {
    const size_t q = (size_t)py0 * W + px;
    splats_per_pixel[q] = count0; weight_per_pixel[q] = T.x;
    image[3 * q + 0] = cr.x + T.x * bg; image[3 * q + 1] = cg.x + T.x * bg; image[3 * q + 2] = cb.x + T.x * bg;
  }
  if (in1) {
    const size_t q = (size_t)(py0 + 4) * W + px;
    splats_per_pixel[q] = count1; weight_per_pixel[q] = T.y;
    image[3 * q + 0] = cr.y + T.y * bg; image[3 * q + 1] = cg.y + T.y * bg; image[3 * q + 2] = cb.y + T.y * bg;
  }
}

// ---------------------------------------------------------------------------------------------------------------- backward
// STATS: counts warp-hits (extent-box survivors inside the warp's range) and contributing warp-hits instead of writing
// gradients (diagnostics for the roofline arithmetic, gsx_ctx_count_hits).
template <bool STATS, int NT>
__global__ void __launch_bounds__(NT, (NT == 32 ? 32 : GSX_C3_BWD_MINB))
k_composite_bwd3(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const float4 *__restrict__ r2,
                 const int *__restrict__ sorted, const int *__restrict__ ranges, const int *__restrict__ order, int ntx, int W, int H, float bg,
                 const int *__restrict__ splats_per_pixel, const float *__restrict__ weight_per_pixel,
                 const float *__restrict__ grad_image, float *__restrict__ packed, unsigned long long *__restrict__ stats) {
  constexpr int CT = NT;
  __shared__ Ring<NT> s;
  __shared__ int s_ids[2][CT];
  __shared__ int s_max;
  const int tslot = NT == 32 ? blockIdx.x >> 2 : blockIdx.x;
  const int tile = order ? order[tslot] : tslot;   // longest lists first (k_tile_order): the tail of the grid is short tiles
  const int tid = threadIdx.x, lane = tid & 31, warp = NT == 32 ? (blockIdx.x & 3) : (tid >> 5);
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 8;
  const int px = X0 + (lane & 7), py0 = Y0 + (lane >> 3);
  const float fpx = (float)px;
  const float2 nfpy = make_float2(-(float)py0, -(float)(py0 + 4));
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 7);
  const int start = ranges[tile];
  const float half_w = 0.5f * (float)W, half_h = 0.5f * (float)H;

  // per-pixel state; a pixel outside the image has count 0 and never takes an entry
  int count0 = 0, count1 = 0;
  float2 T = bc(0.0f), gx = bc(0.0f), gy = bc(0.0f), gz = bc(0.0f);
  if (px < W && py0 < H) {
    const size_t q = (size_t)py0 * W + px;
    count0 = splats_per_pixel[q]; T.x = weight_per_pixel[q];
    gx.x = grad_image[3 * q]; gy.x = grad_image[3 * q + 1]; gz.x = grad_image[3 * q + 2];
  }
  if (px < W && py0 + 4 < H) {
    const size_t q = (size_t)(py0 + 4) * W + px;
    count1 = splats_per_pixel[q]; T.y = weight_per_pixel[q];
    gx.y = grad_image[3 * q]; gy.y = grad_image[3 * q + 1]; gz.y = grad_image[3 * q + 2];
  }
  float2 tb;   // -T_final * (bg gx + bg gy + bg gz): the background term of d alpha, render_backward.cu:149-152
  {
    float b0 = 0.0f, b1 = 0.0f;
    b0 += bg * gx.x; b0 += bg * gy.x; b0 += bg * gz.x;
    b1 += bg * gx.y; b1 += bg * gy.y; b1 += bg * gz.y;
    tb = make_float2(-T.x * b0, -T.y * b1);
  }
  float2 sacc = bc(0.0f);   // sum_c (colour accumulated behind the current entry)_c * g_c
  const int warp_max = __reduce_max_sync(0xffffffffu, max(count0, count1));
  int cta_max = warp_max;   // render_backward.cu:74 (start index = max count - 1); a lone warp starts at its own maximum
  if (NT != 32) {
    if (tid == 0) s_max = 0;
    __syncthreads();
    if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
    __syncthreads();
    cta_max = s_max;
  }

  // after the butterfly, lane (lane & 3) == 0 holds value k = 4 b4 + 2 b3 + b2 of {r, g, b, u, v, ca, cb, cc} and lane 2 the
  // ninth (opacity); their places in the packed row {r,g,b,opa}{u,v,ca,cb}{cc}
  const int kv = ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
  const bool red_lane = (lane & 3) == 0 || lane == 2;
  const int red_off = lane == 2 ? 3 : (kv < 3 ? kv : kv + 1);
  unsigned long long n_hit = 0, n_contrib = 0;

  // batch j (back to front): entries [lo_j, top_j), top_j = cta_max - j * CT
  const int nbatch = (cta_max + CT - 1) / CT;
  auto id_of = [&](int j) {
    const int top = cta_max - j * CT, lo = max(0, top - CT);
    return (j < nbatch && tid < top - lo) ? sorted[start + lo + tid] : -1;
  };
  int id_next = id_of(0);
  if (nbatch > 0) { if (id_next >= 0) s_ids[0][tid] = id_next; stage(s, 0, tid, id_next, r0, r1, r2); }
  id_next = id_of(1);
  for (int b = 0; b < nbatch; ++b) {
    const int buf = b & 1;
    const int top = cta_max - b * CT, lo = max(0, top - CT);
    if (b + 1 < nbatch && id_next >= 0) s_ids[buf ^ 1][tid] = id_next;
    stage(s, buf ^ 1, tid, b + 1 < nbatch ? id_next : -1, r0, r1, r2);
    id_next = id_of(b + 2);
    cp_wait<1>();
    cta_sync<NT>();
    const int hi = min(top, warp_max);
    for (int chi = hi; chi > lo; chi -= 32) {
      const int j = chi - 1 - lane;   // lane 0 = farthest entry first (back to front)
      bool hit = false;
      if (j >= lo) {
        const float4 q = s.r0[buf][j - lo];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      if (STATS) n_hit += __popc(m);
      while (m) {
        const int idx = chi - 1 - (__ffs(m) - 1);
        m &= m - 1;
        const int e = idx - lo;
        const float2 uv = *reinterpret_cast<const float2 *>(&s.r0[buf][e]);
        const float4 q1 = s.r1[buf][e];
        const float dx = uv.x - fpx;
        const Quad q = quad_of(dx, q1.x, q1.y, q1.z);
        const float2 dy = add2(bc(uv.y), nfpy);
        const float2 p2 = fma2(dy, fma2(bc(q.hc), dy, bc(q.B)), bc(q.A));
        const float2 au = mul2(bc(q1.w), make_float2(ex2(fminf(0.0f, p2.x)), ex2(fminf(0.0f, p2.y))));   // opa * G
        // valid: alpha = min(0.99, opa G) >= 1/255 (render_backward.cu:123: '>=', the forward used '>') and the pixel had
        // not terminated before this entry
        const bool v0 = au.x >= ALPHA_MIN && idx < count0, v1 = au.y >= ALPHA_MIN && idx < count1;
        if (!__any_sync(0xffffffffu, v0 || v1)) continue;
        if (STATS) { ++n_contrib; continue; }
        // Branch-free body: a pixel that does not take the entry runs with alpha = 0 and G opa = 0, which leaves its state
        // untouched (T * rcp(1) = T, s * 1 + 0) and contributes exact zeros.
        const float4 q2 = s.r2[buf][e];
        const float2 aum = make_float2(v0 ? au.x : 0.0f, v1 ? au.y : 0.0f);
        const float2 al = make_float2(fminf(0.99f, aum.x), fminf(0.99f, aum.y));
        const float2 om = add2(bc(1.0f), make_float2(-al.x, -al.y));
        const float2 iom = make_float2(rcp(om.x), rcp(om.y));          // render_backward.cu:132 divides; 1 - alpha >= 0.01
        T = mul2(T, iom);                                              // transmittance in front of this entry
        const float2 w = mul2(al, T);
        const float2 cgd = fma2(bc(q2.z), gz, fma2(bc(q2.y), gy, mul2(bc(q2.x), gx)));   // colour . g
        // d alpha = T (colour - behind) . g - T_final / (1 - alpha) * bg . g     (render_backward.cu:140-152)
        const float2 ga = fma2(tb, iom, mul2(add2(cgd, make_float2(-sacc.x, -sacc.y)), T));
        sacc = fma2(om, sacc, mul2(al, cgd));                          // behind <- alpha colour + (1 - alpha) behind
        const float2 gp = mul2(aum, ga);                               // d L / d power = G opa d alpha
        const float2 wr = mul2(w, gx), wg = mul2(w, gy), wb = mul2(w, gz);
        const float2 gpy = mul2(gp, dy), gpyy = mul2(gpy, dy);
        // the lane's two-pixel sums of the moments of dL/dpower, mapped to the entry's gradients (render_backward.cu:
        // 154-166, 186-197): d u = (-a S gp dx - b S gp dy) W/2 (Q5: NDC-scaled), d a = -1/2 S gp dx^2, d b = -S gp dx dy, ...
        const float s0 = gp.x + gp.y, s4 = gpy.x + gpy.y, s7 = gpyy.x + gpyy.y;
        const float s3 = dx * s0;
        float v[8];
        v[0] = wr.x + wr.y; v[1] = wg.x + wg.y; v[2] = wb.x + wb.y;
        v[3] = (q1.x * s3 + q1.y * s4) * -half_w;
        v[4] = (q1.z * s4 + q1.y * s3) * -half_h;
        v[5] = (dx * s3) * -0.5f; v[6] = -(dx * s4); v[7] = s7 * -0.5f;
        float vo = s0 * (1.0f - q1.w);                                 // d opacity logit: sum of G d alpha opa (1 - opa)
        // transposing butterfly: 9 values over 32 lanes in 12 shuffles
        float a4[4], a2[2], a1;
        {
          const bool up = lane & 16;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float send = up ? v[k] : v[k + 4], keep = up ? v[k + 4] : v[k];
            a4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
          }
          vo += __shfl_xor_sync(0xffffffffu, vo, 16);
        }
        {
          const bool up = lane & 8;
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const float send = up ? a4[k] : a4[k + 2], keep = up ? a4[k + 2] : a4[k];
            a2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
          }
          vo += __shfl_xor_sync(0xffffffffu, vo, 8);
        }
        {
          const bool up = lane & 4;
          const float send = up ? a2[0] : a2[1], keep = up ? a2[1] : a2[0];
          a1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
          vo += __shfl_xor_sync(0xffffffffu, vo, 4);
        }
        {
          const bool up = lane & 2;   // lanes with bit 1 set carry the ninth value from here on
          const float send = up ? a1 : vo, keep = up ? vo : a1;
          a1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        }
        a1 += __shfl_xor_sync(0xffffffffu, a1, 1);
        // Q7 (render_backward.cu:170): the reference drops the whole entry when every lane's opacity partial is exactly
        // zero -- which is what happens when opa (1 - opa) underflows to 0 (q2.w).  One RED instruction, 9 active lanes.
        if (red_lane && q2.w != 0.0f) atomicAdd(packed + (size_t)s_ids[buf][e] * 12 + red_off, a1);
      }
    }
    cta_sync<NT>();   // the stage is free for batch b+2
  }
  cp_wait<0>();
  if (STATS) {
    if (lane == 0) { atomicAdd(stats, n_hit); atomicAdd(stats + 1, n_contrib); }
  }
}


// ------------------------------------------------------------------------------------------ backward, two-phase reduction
// k_composite_bwd3 spends 55 of its 127 instructions per contributing warp-hit on turning 64 per-pixel partials into 9
// per-entry sums: the moment products, a 12-shuffle / 16-select / 15-add transposing butterfly and the linear map to the
// gradients.  Everything the 9 sums need from a pixel is TWO scalars, w = alpha T (colour gradient) and gp = dL/dpower
// (geometry gradients); the rest are per-pixel constants (image gradient, pixel coordinates).  So:
//   phase 1 (lane = pixel pair, per contributing hit): evaluate the entry as before, store the lane's (w, gp) pairs and the
//            entry's 8 parameters into a per-warp shared-memory buffer of HB hits -- no cross-lane traffic at all;
//   phase 2 (lane = (buffered hit h, pixel-row pair g), once per HB hits): the lane runs over the 8 columns of its two
//            pixel rows accumulating sum w g_c (3), sum gp, sum gp cx, sum gp cx^2 with packed FMAs against per-pixel
//            constants (cx = column, compile-time; cy = row, a per-lane constant, so the y-moments follow from the
//            x-moments by three multiplies); two shuffle steps merge the four row pairs; the moments about the block origin
//            are shifted to the entry's centre (dx = (u - X0) - cx), mapped to the gradients (render_backward.cu:154-197)
//            and leave as two red.global.add.v4.f32 + one scalar RED per hit, 8 hits per instruction.
// Instructions per contributing warp-hit: ~65 + ~22 instead of 127; shared memory 5.3 KB per warp.
constexpr int HB = 8;             // buffered hits per flush (lanes: 8 hits x 4 row pairs)
constexpr int HROW = 66;          // floats per buffered hit row: 64 + 2 -> the 8 hits of a 16-lane phase hit distinct banks
struct WarpBuf {
  float w[HB][HROW], gp[HB][HROW];   // [hit][2 * source lane + pixel]
  float4 meta[HB][2];                // {u, v, conic a, b} {conic c, opa, opa (1 - opa), Gaussian id (int bits)}
  float g[32][6];                    // per source lane: gx.x, gx.y, gy.x, gy.y, gz.x, gz.y (image gradient of its two pixels)
};

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// phase 2: reduce the nbuf buffered hits of this warp and add them to the packed gradient rows
__device__ __forceinline__ void flush_hits(WarpBuf &wb, int nbuf, int lane, float X0f, float Y0f, float half_w, float half_h,
                                           float *__restrict__ packed) {
  __syncwarp();
  const int h = lane & 7, g = lane >> 3;             // buffered hit, row pair (pixel rows g and g + 4 of the 8x8 block)
  const float2 cy = make_float2((float)g, (float)(g + 4));
  // two independent accumulator sets (even / odd columns): the six chains are otherwise 8 dependent FMAs long each
  float2 Ra[2] = {bc(0.f), bc(0.f)}, Ga[2] = {bc(0.f), bc(0.f)}, Ba[2] = {bc(0.f), bc(0.f)};
  float2 S0a[2] = {bc(0.f), bc(0.f)}, Sxa[2] = {bc(0.f), bc(0.f)}, Sxxa[2] = {bc(0.f), bc(0.f)};
#pragma unroll
  for (int c = 0; c < 8; ++c) {                      // source lane j = 8 g + c: column c, rows g and g + 4
    const int j = 8 * g + c, a = c & 1;
    const float2 w2 = *reinterpret_cast<const float2 *>(&wb.w[h][2 * j]);
    const float2 p2 = *reinterpret_cast<const float2 *>(&wb.gp[h][2 * j]);
    const float2 gx = *reinterpret_cast<const float2 *>(&wb.g[j][0]);
    const float2 gy = *reinterpret_cast<const float2 *>(&wb.g[j][2]);
    const float2 gz = *reinterpret_cast<const float2 *>(&wb.g[j][4]);
    Ra[a] = fma2(w2, gx, Ra[a]); Ga[a] = fma2(w2, gy, Ga[a]); Ba[a] = fma2(w2, gz, Ba[a]);
    S0a[a] = add2(S0a[a], p2);
    if (c > 0) { Sxa[a] = fma2(p2, bc((float)c), Sxa[a]); Sxxa[a] = fma2(p2, bc((float)(c * c)), Sxxa[a]); }
  }
  const float2 R = add2(Ra[0], Ra[1]), G = add2(Ga[0], Ga[1]), B = add2(Ba[0], Ba[1]);
  const float2 S0 = add2(S0a[0], S0a[1]), Sx = add2(Sxa[0], Sxa[1]), Sxx = add2(Sxxa[0], Sxxa[1]);
  // the lane's two rows -> scalars (cy is constant along a row: the y-moments follow from S0 and Sx)
  float v[9];
  v[0] = R.x + R.y; v[1] = G.x + G.y; v[2] = B.x + B.y;
  v[3] = S0.x + S0.y;                                 // S0
  v[4] = Sx.x + Sx.y;                                 // Sx
  v[5] = Sxx.x + Sxx.y;                               // Sxx
  v[6] = cy.x * S0.x + cy.y * S0.y;                   // Sy
  v[7] = cy.x * Sx.x + cy.y * Sx.y;                   // Sxy
  v[8] = (cy.x * cy.x) * S0.x + (cy.y * cy.y) * S0.y; // Syy
#pragma unroll
  for (int k = 0; k < 9; ++k) {                       // sum over the four row pairs (lanes h, h + 8, h + 16, h + 24)
    v[k] += __shfl_xor_sync(0xffffffffu, v[k], 8);
    v[k] += __shfl_xor_sync(0xffffffffu, v[k], 16);
  }
  if (g == 0 && h < nbuf) {
    const float4 m0 = wb.meta[h][0], m1 = wb.meta[h][1];
    if (m1.z != 0.0f) {   // Q7 (render_backward.cu:170): opa (1 - opa) underflowed -> the reference drops the entry
      const float A = m0.x - X0f, Bc = m0.y - Y0f;    // entry centre relative to the block origin: dx = A - cx, dy = Bc - cy
      const float S0_ = v[3], Sx_ = v[4], Sxx_ = v[5], Sy_ = v[6], Sxy_ = v[7], Syy_ = v[8];
      const float gdx = A * S0_ - Sx_, gdy = Bc * S0_ - Sy_;                       // sum gp dx, sum gp dy
      const float gdxx = A * (A * S0_ - 2.0f * Sx_) + Sxx_;                         // sum gp dx^2
      const float gdyy = Bc * (Bc * S0_ - 2.0f * Sy_) + Syy_;                       // sum gp dy^2
      const float gdxy = A * (Bc * S0_ - Sy_) - Bc * Sx_ + Sxy_;                    // sum gp dx dy
      float *o = packed + (size_t)__float_as_int(m1.w) * 12;
      red_add_v4(o, v[0], v[1], v[2], S0_ * (1.0f - m1.y));
      red_add_v4(o + 4, (m0.z * gdx + m0.w * gdy) * -half_w, (m1.x * gdy + m0.w * gdx) * -half_h, -0.5f * gdxx, -gdxy);
      atomicAdd(o + 8, -0.5f * gdyy);
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(128, GSX_C4_BWD_MINB)
k_composite_bwd4(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const float4 *__restrict__ r2,
                 const int *__restrict__ sorted, const int *__restrict__ ranges, const int *__restrict__ order, int ntx, int W, int H, float bg,
                 const int *__restrict__ splats_per_pixel, const float *__restrict__ weight_per_pixel,
                 const float *__restrict__ grad_image, float *__restrict__ packed) {
  constexpr int CT = 128;
  __shared__ Ring<CT> s;
  __shared__ int s_ids[2][CT];
  __shared__ int s_max;
  __shared__ __align__(16) WarpBuf s_wb[4];
  const int tile = order ? order[blockIdx.x] : blockIdx.x;   // longest lists first (k_tile_order)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 8;
  const int px = X0 + (lane & 7), py0 = Y0 + (lane >> 3);
  const float fpx = (float)px;
  const float2 nfpy = make_float2(-(float)py0, -(float)(py0 + 4));
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 7);
  const int start = ranges[tile];
  const float half_w = 0.5f * (float)W, half_h = 0.5f * (float)H;
  WarpBuf &wb = s_wb[warp];

  int count0 = 0, count1 = 0;
  float2 T = bc(0.0f), gx = bc(0.0f), gy = bc(0.0f), gz = bc(0.0f);
  if (px < W && py0 < H) {
    const size_t q = (size_t)py0 * W + px;
    count0 = splats_per_pixel[q]; T.x = weight_per_pixel[q];
    gx.x = grad_image[3 * q]; gy.x = grad_image[3 * q + 1]; gz.x = grad_image[3 * q + 2];
  }
  if (px < W && py0 + 4 < H) {
    const size_t q = (size_t)(py0 + 4) * W + px;
    count1 = splats_per_pixel[q]; T.y = weight_per_pixel[q];
    gx.y = grad_image[3 * q]; gy.y = grad_image[3 * q + 1]; gz.y = grad_image[3 * q + 2];
  }
  wb.g[lane][0] = gx.x; wb.g[lane][1] = gx.y; wb.g[lane][2] = gy.x; wb.g[lane][3] = gy.y; wb.g[lane][4] = gz.x; wb.g[lane][5] = gz.y;
  float2 tb;   // -T_final * (bg gx + bg gy + bg gz), render_backward.cu:149-152
  {
    float b0 = 0.0f, b1 = 0.0f;
    b0 += bg * gx.x; b0 += bg * gy.x; b0 += bg * gz.x;
    b1 += bg * gx.y; b1 += bg * gy.y; b1 += bg * gz.y;
    tb = make_float2(-T.x * b0, -T.y * b1);
  }
  float2 sacc = bc(0.0f);
  const int warp_max = __reduce_max_sync(0xffffffffu, max(count0, count1));
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
  __syncthreads();
  const int cta_max = s_max;   // render_backward.cu:74
  int nbuf = 0;                // buffered contributing hits of this warp (warp-uniform)

  const int nbatch = (cta_max + CT - 1) / CT;
  auto id_of = [&](int j) {
    const int top = cta_max - j * CT, lo = max(0, top - CT);
    return (j < nbatch && tid < top - lo) ? sorted[start + lo + tid] : -1;
  };
  int id_next = id_of(0);
  if (nbatch > 0) { if (id_next >= 0) s_ids[0][tid] = id_next; stage(s, 0, tid, id_next, r0, r1, r2); }
  id_next = id_of(1);
  for (int b = 0; b < nbatch; ++b) {
    const int buf = b & 1;
    const int top = cta_max - b * CT, lo = max(0, top - CT);
    if (b + 1 < nbatch && id_next >= 0) s_ids[buf ^ 1][tid] = id_next;
    stage(s, buf ^ 1, tid, b + 1 < nbatch ? id_next : -1, r0, r1, r2);
    id_next = id_of(b + 2);
    cp_wait<1>();
    __syncthreads();
    const int hi = min(top, warp_max);
    for (int chi = hi; chi > lo; chi -= 32) {
      const int j = chi - 1 - lane;   // lane 0 = farthest entry first (back to front)
      bool hit = false;
      if (j >= lo) {
        const float4 q = s.r0[buf][j - lo];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int idx = chi - 1 - (__ffs(m) - 1);
        m &= m - 1;
        const int e = idx - lo;
        const float2 uv = *reinterpret_cast<const float2 *>(&s.r0[buf][e]);
        const float4 q1 = s.r1[buf][e];
        const float dx = uv.x - fpx;
        const Quad q = quad_of(dx, q1.x, q1.y, q1.z);
        const float2 dy = add2(bc(uv.y), nfpy);
        const float2 p2 = fma2(dy, fma2(bc(q.hc), dy, bc(q.B)), bc(q.A));
        const float2 au = mul2(bc(q1.w), make_float2(ex2(fminf(0.0f, p2.x)), ex2(fminf(0.0f, p2.y))));   // opa * G
        const bool v0 = au.x >= ALPHA_MIN && idx < count0, v1 = au.y >= ALPHA_MIN && idx < count1;
        if (!__any_sync(0xffffffffu, v0 || v1)) continue;
        const float4 q2 = s.r2[buf][e];
        const float2 aum = make_float2(v0 ? au.x : 0.0f, v1 ? au.y : 0.0f);
        const float2 al = make_float2(fminf(0.99f, aum.x), fminf(0.99f, aum.y));
        const float2 om = add2(bc(1.0f), make_float2(-al.x, -al.y));
        const float2 iom = make_float2(rcp(om.x), rcp(om.y));
        T = mul2(T, iom);
        const float2 w = mul2(al, T);
        const float2 cgd = fma2(bc(q2.z), gz, fma2(bc(q2.y), gy, mul2(bc(q2.x), gx)));
        const float2 ga = fma2(tb, iom, mul2(add2(cgd, make_float2(-sacc.x, -sacc.y)), T));
        sacc = fma2(om, sacc, mul2(al, cgd));
        const float2 gp = mul2(aum, ga);
        // phase 1 ends here: park the lane's (w, gp) pairs and the entry's parameters
        *reinterpret_cast<float2 *>(&wb.w[nbuf][2 * lane]) = w;
        *reinterpret_cast<float2 *>(&wb.gp[nbuf][2 * lane]) = gp;
        if (lane < 2) wb.meta[nbuf][lane] = lane == 0 ? make_float4(uv.x, uv.y, q1.x, q1.y)
                                                      : make_float4(q1.z, q1.w, q2.w, __int_as_float(s_ids[buf][e]));
        if (++nbuf == HB) { flush_hits(wb, HB, lane, (float)X0, (float)Y0, half_w, half_h, packed); nbuf = 0; }
      }
    }
    __syncthreads();   // the stage is free for batch b+2
  }
  cp_wait<0>();
  if (nbuf > 0) flush_hits(wb, nbuf, lane, (float)X0, (float)Y0, half_w, half_h, packed);
}

}  // namespace

// GSX_COMPOSITE_CTA=32 selects the single-warp CTAs (one per 8x8 pixel block), 128 the tile-per-CTA kernels
static int cta_threads() {
  static const int v = [] { const char *e = getenv("GSX_COMPOSITE_CTA"); const int x = e ? atoi(e) : GSX_C3_DEFAULT_CTA; return x == 32 ? 32 : 128; }();
  return v;
}

// GSX_COMPOSITE_BWD=butterfly selects k_composite_bwd3 (per-hit shuffle reduction), two_phase k_composite_bwd4
static bool bwd_two_phase() {
  static const int v = [] {
    const char *e = getenv("GSX_COMPOSITE_BWD");
    if (!e) return GSX_C4_DEFAULT;
    return e[0] == 't' ? 1 : 0;
  }();
  return v != 0;
}

int composite_fwd3_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, float bg, int *spp, float *wpp, float *image, cudaStream_t st, const int *order) {
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  if (cta_threads() == 32)
    k_composite_fwd3<32><<<ntx * nty * 4, 32, 0, st>>>(r0, r1, r2, sorted, ranges, order, ntx, W, H, bg, spp, wpp, image);
  else
    k_composite_fwd3<128><<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, order, ntx, W, H, bg, spp, wpp, image);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int composite_bwd3_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, float bg, const int *spp, const float *wpp, const float *grad_image, float *packed_grads,
                          cudaStream_t st, const int *order) {
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  if (bwd_two_phase()) {
    k_composite_bwd4<<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, order, ntx, W, H, bg, spp, wpp, grad_image, packed_grads);
    GSX_LAUNCH_CHECK();
    return GSX_OK;
  }
  if (cta_threads() == 32)
    k_composite_bwd3<false, 32><<<ntx * nty * 4, 32, 0, st>>>(r0, r1, r2, sorted, ranges, order, ntx, W, H, bg, spp, wpp, grad_image,
                                                            packed_grads, nullptr);
  else
    k_composite_bwd3<false, 128><<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, order, ntx, W, H, bg, spp, wpp, grad_image,
                                                            packed_grads, nullptr);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

// order[k] = the tile with the k-th longest list, approximately: a one-block counting sort of the T list lengths into 256
// length classes, longest class first.  The compositing grids take their tiles in this order, so the CTAs that are still
// running when the grid drains are short ones (measured on the forward at C4: SMs active 92 % of the launch in tile order).
// Opt-in (GSX_TILE_ORDER=1): measured, it gains nothing (api.cu).
namespace {
__global__ void __launch_bounds__(1024) k_tile_order(const int *__restrict__ ranges, int T, int *__restrict__ order) {
  __shared__ int s_cnt[256];
  __shared__ int s_max;
  if (threadIdx.x < 256) s_cnt[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_max = 1;
  __syncthreads();
  int m = 1;
  for (int t = threadIdx.x; t < T; t += 1024) m = max(m, ranges[t + 1] - ranges[t]);
  for (int off = 16; off > 0; off >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0) atomicMax(&s_max, m);
  __syncthreads();
  const float scale = 255.0f / (float)s_max;
  auto cls = [&](int len) { return 255 - min(255, (int)((float)len * scale)); };   // class 0 = longest
  for (int t = threadIdx.x; t < T; t += 1024) atomicAdd(&s_cnt[cls(ranges[t + 1] - ranges[t])], 1);
  __syncthreads();
  if (threadIdx.x < 32) {   // exclusive scan of the 256 class counts (8 per lane)
    int v[8], sum = 0;
    for (int k = 0; k < 8; ++k) { v[k] = s_cnt[threadIdx.x * 8 + k]; sum += v[k]; }
    int incl = sum;
    for (int off = 1; off < 32; off <<= 1) {
      const int x = __shfl_up_sync(0xffffffffu, incl, off);
      if ((int)threadIdx.x >= off) incl += x;
    }
    int run = incl - sum;
    for (int k = 0; k < 8; ++k) { s_cnt[threadIdx.x * 8 + k] = run; run += v[k]; }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += 1024) order[atomicAdd(&s_cnt[cls(ranges[t + 1] - ranges[t])], 1)] = t;
}
}  // namespace
int tile_order_launch(const int *ranges, int T, int *order, cudaStream_t st) {
  k_tile_order<<<1, 1024, 0, st>>>(ranges, T, order);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

// out_dev[2] = {warp-hits, contributing warp-hits} of the backward over the current lists (diagnostics)
int composite_hits_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, const int *spp, const float *wpp, const float *grad_image, unsigned long long *out_dev,
                          cudaStream_t st) {
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  k_composite_bwd3<true, 128><<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, nullptr, ntx, W, H, 0.0f, spp, wpp, grad_image, nullptr,
                                                         out_dev);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // namespace gsx
