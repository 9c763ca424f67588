// composite.cu -- tile compositing forward (F8, reference cuda/render.cu:6-108) and backward (F9, reference
// cuda/render_backward.cu:11-223).
//
// B200 design (the reference runs one warp per 16x16 tile with every lane re-loading each Gaussian from global):
//   * one 256-thread CTA per 16x16 tile, one pixel per thread; warp w owns an 8x4 pixel block;
//   * the tile's sorted list is consumed in batches of 256 entries staged once into shared memory as three float4
//     records {u,v,ex,ey} {a,b,c,sigmoid(opacity)} {r,g,b,opa(1-opa)} with coalesced id loads + 16-byte gathers;
//   * warp-level culling: each lane tests one staged entry's alpha>1/255 extent box (ex,ey: conservative half
//     widths of the level set opa*exp(power) = 1/255) against the warp's 8x4 block, a ballot yields the surviving
//     entries, and only those are evaluated.  Skipped entries cannot change T, colour or the saturation index, so
//     results are identical to evaluating everything;
//   * early termination per warp by ballot and per CTA by __syncthreads_and;
//   * backward: 9 per-pixel partials are reduced across the warp with a transposing butterfly (9 + 5 shuffles
//     instead of 45), merged across the tile's 8 warps in shared memory, and flushed with ONE set of global
//     reductions per (tile, Gaussian): two red.global.add.v4.f32 + one scalar red into a packed [N][12] buffer
//     (view-level path) or 9 scalar atomics into the reference's separate arrays (per-function path).
// Bound: FP32/issue (no tensor cores: not a dense contraction).  Algorithmic work per SURVEY 8d.
#include <stdlib.h>

#include "gsx_common.cuh"

// minimum resident CTAs per SM of the view-level compositing kernels (register cap); overridable for A/B builds
#ifndef GSX_CBWD_MINB
#define GSX_CBWD_MINB 8
#endif
#ifndef GSX_CFWD_MINB
#define GSX_CFWD_MINB 1
#endif
namespace gsx {

constexpr float ALPHA_MIN = 0.00392156862f;   // render.cu:74, render_backward.cu:123
constexpr int CB = 256;                        // threads per CTA == staged entries per batch

// Conservative half-extents (pixels) of {alpha >= 1/255}: q(d) = a dx^2 + 2b dx dy + c dy^2 <= 2 ln(255 opa).
// (-inf,-inf): can never reach the threshold.  (inf,inf): conic not positive definite -> never cull.
__device__ __forceinline__ float2 alpha_extent(float a, float b, float c, float opa) {
  if (!(opa >= ALPHA_MIN)) return make_float2(-INFINITY, -INFINITY);
  const float det = a * c - b * b;
  if (!(a > 0.0f && c > 0.0f && det > 0.0f) || !(opa <= 1.0f)) return make_float2(INFINITY, INFINITY);
  const float tau2 = 2.0f * __logf(fmaxf(opa * 255.0f, 1.0f)) * 1.001f + 0.02f;
  const float inv = tau2 / det;
  return make_float2(sqrtf(inv * c) * 1.0005f + 0.02f, sqrtf(inv * a) * 1.0005f + 0.02f);
}

__device__ __forceinline__ float sigmoid_fast(float x) { return 1.0f / (1.0f + __expf(-x)); }   // render.cu:56

// exp(x) for x <= 0 as one MUFU.EX2 (render.cu:73 uses __expf = ex2.approx(x * log2 e) plus a denormal-range rescale;
// results below 2^-126 flush to zero here, far under the 1/255 alpha threshold, so every kept alpha is identical)
__device__ __forceinline__ float exp_neg(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x * 1.4426950408889634f));
  return y;
}
// 1/x for x in [0.01, 1] (1 - alpha, alpha <= 0.99): one MUFU.RCP, no slow path
__device__ __forceinline__ float rcp_fast(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct GaussSrc {          // where a Gaussian's compositing record comes from
  // packed (view-level path): three float4 arrays indexed by global id
  const float4 *r0, *r1, *r2;
  // separate (per-function path, reference layouts)
  const float *uv, *opacity, *conic, *rgb;
};

template <bool PACKED>
__device__ __forceinline__ void stage_entry(const GaussSrc &s, int g, float4 *o0, float4 *o1, float4 *o2) {
  if (PACKED) {
    *o0 = __ldg(s.r0 + g); *o1 = __ldg(s.r1 + g); *o2 = __ldg(s.r2 + g);
  } else {
    const float u = s.uv[2 * g], v = s.uv[2 * g + 1];
    const float a = s.conic[3 * g], b = s.conic[3 * g + 1], c = s.conic[3 * g + 2];
    const float opa = sigmoid_fast(s.opacity[g]);
    const float2 e = alpha_extent(a, b, c, opa);
    *o0 = make_float4(u, v, e.x, e.y);
    *o1 = make_float4(a, b, c, opa);
    *o2 = make_float4(s.rgb[3 * g], s.rgb[3 * g + 1], s.rgb[3 * g + 2], opa * (1.0f - opa));
  }
}

template <bool PACKED>
__global__ void __launch_bounds__(CB)
k_composite_fwd(GaussSrc src, const int *__restrict__ sorted, const int *__restrict__ ranges, int ntx, int W, int H,
                float bg, int *__restrict__ splats_per_pixel, float *__restrict__ weight_per_pixel,
                float *__restrict__ image) {
  __shared__ float4 s_r0[CB], s_r1[CB], s_r2[CB];
  const int tile = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 4;
  const int px = X0 + (lane & 7), py = Y0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float fpx = (float)px, fpy = (float)py;
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 3);
  const int start = ranges[tile];
  const int total = ranges[tile + 1] - start;

  float T = 1.0f, cr = 0.0f, cg = 0.0f, cb = 0.0f;
  bool done = !inside;
  int count = total;
  bool warp_done = __ballot_sync(0xffffffffu, !done) == 0;

  for (int base = 0; base < total; base += CB) {
    if (__syncthreads_and(warp_done)) break;
    const int nb = min(CB, total - base);
    if (tid < nb) stage_entry<PACKED>(src, sorted[start + base + tid], &s_r0[tid], &s_r1[tid], &s_r2[tid]);
    __syncthreads();
    if (warp_done) continue;
    for (int c = 0; c < nb; c += 32) {
      const int j = c + lane;
      bool hit = false;
      if (j < nb) {
        const float4 q = s_r0[j];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int e = c + __ffs(m) - 1;
        m &= m - 1;
        const float4 q0 = s_r0[e], q1 = s_r1[e];
        const float dx = q0.x - fpx, dy = q0.y - fpy;
        const float power = fminf(0.0f, -0.5f * (q1.x * dx * dx + q1.z * dy * dy) - q1.y * dx * dy);
        const float alpha = fminf(0.99f, q1.w * exp_neg(power));
        if (alpha > ALPHA_MIN && !done) {
          const float4 q2 = s_r2[e];
          const float test_T = T * (1.0f - alpha);
          const float w = alpha * T;
          cr += q2.x * w; cg += q2.y * w; cb += q2.z * w;
          T = test_T;
          if (test_T < 0.0001f) { done = true; count = base + e + 1; }
        }
      }
      if (__ballot_sync(0xffffffffu, !done) == 0) { warp_done = true; break; }
    }
  }
  if (inside) {
    const size_t p = (size_t)py * W + px;
    splats_per_pixel[p] = count;
    weight_per_pixel[p] = T;
    image[3 * p + 0] = cr + T * bg;
    image[3 * p + 1] = cg + T * bg;
    image[3 * p + 2] = cb + T * bg;
  }
}

// View-level forward, two pixels per lane (a warp owns an 8x8 block, a 128-thread CTA one tile): the per-entry overhead of
// the hit loop (next-hit extraction, two shared-memory record loads, dx and its products) is shared by the two pixels.
// Per-pixel arithmetic identical to k_composite_fwd.
__global__ void __launch_bounds__(128, GSX_CFWD_MINB)
k_composite_fwd2(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const float4 *__restrict__ r2,
                 const int *__restrict__ sorted, const int *__restrict__ ranges, int ntx, int W, int H, float bg,
                 int *__restrict__ splats_per_pixel, float *__restrict__ weight_per_pixel, float *__restrict__ image) {
  constexpr int CF = 128;
  __shared__ float4 s_r0[CF], s_r1[CF], s_r2[CF];
  const int tile = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 8;
  const int px = X0 + (lane & 7);
  const float fpx = (float)px;
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 7);
  const int start = ranges[tile];
  const int total = ranges[tile + 1] - start;

  float fpy[2], T[2], cr[2], cg[2], cb[2];
  bool done[2], inside[2];
  int count[2];
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    const int py = Y0 + (lane >> 3) + 4 * p;
    fpy[p] = (float)py;
    inside[p] = px < W && py < H;
    done[p] = !inside[p];
    T[p] = 1.0f; cr[p] = cg[p] = cb[p] = 0.0f;
    count[p] = total;
  }
  bool warp_done = __ballot_sync(0xffffffffu, !(done[0] && done[1])) == 0;

  for (int base = 0; base < total; base += CF) {
    if (__syncthreads_and(warp_done)) break;
    const int nb = min(CF, total - base);
    if (tid < nb) {
      const int g = sorted[start + base + tid];
      s_r0[tid] = __ldg(r0 + g); s_r1[tid] = __ldg(r1 + g); s_r2[tid] = __ldg(r2 + g);
    }
    __syncthreads();
    if (warp_done) continue;
    for (int c = 0; c < nb; c += 32) {
      const int j = c + lane;
      bool hit = false;
      if (j < nb) {
        const float4 q = s_r0[j];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int e = c + __ffs(m) - 1;
        m &= m - 1;
        const float4 q0 = s_r0[e], q1 = s_r1[e];
        const float dx = q0.x - fpx;
        float alpha[2];
        bool take = false;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          const float dy = q0.y - fpy[p];
          const float power = fminf(0.0f, -0.5f * (q1.x * dx * dx + q1.z * dy * dy) - q1.y * dx * dy);
          alpha[p] = fminf(0.99f, q1.w * exp_neg(power));
          take |= (alpha[p] > ALPHA_MIN && !done[p]);
        }
        if (take) {
          const float4 q2 = s_r2[e];
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            if (alpha[p] > ALPHA_MIN && !done[p]) {
              const float test_T = T[p] * (1.0f - alpha[p]);
              const float w = alpha[p] * T[p];
              cr[p] += q2.x * w; cg[p] += q2.y * w; cb[p] += q2.z * w;
              T[p] = test_T;
              if (test_T < 0.0001f) { done[p] = true; count[p] = base + e + 1; }
            }
          }
        }
      }
      if (__ballot_sync(0xffffffffu, !(done[0] && done[1])) == 0) { warp_done = true; break; }
    }
  }
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    if (inside[p]) {
      const size_t q = (size_t)(Y0 + (lane >> 3) + 4 * p) * W + px;
      splats_per_pixel[q] = count[p];
      weight_per_pixel[q] = T[p];
      image[3 * q + 0] = cr[p] + T[p] * bg;
      image[3 * q + 1] = cg[p] + T[p] * bg;
      image[3 * q + 2] = cb[p] + T[p] * bg;
    }
  }
}

struct GradDst {
  float *packed;                                   // [N][12]: {r,g,b,opa} {u,v,ca,cb} {cc,-,-,-}
  float *grad_rgb, *grad_opacity, *grad_uv, *grad_conic;   // reference layouts
};

__device__ __forceinline__ void red_add_v4(float *addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

template <bool PACKED>
__global__ void __launch_bounds__(CB)
k_composite_bwd(GaussSrc src, const int *__restrict__ sorted, const int *__restrict__ ranges, int ntx, int W, int H,
                float bg, const int *__restrict__ splats_per_pixel, const float *__restrict__ weight_per_pixel,
                const float *__restrict__ grad_image, GradDst dst) {
  __shared__ float4 s_r0[CB], s_r1[CB], s_r2[CB];
  __shared__ float s_acc[CB * 9];
  __shared__ int s_ids[CB];
  __shared__ int s_max;
  const int tile = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * 4;
  const int px = X0 + (lane & 7), py = Y0 + (lane >> 3);
  const bool inside = px < W && py < H;
  const float fpx = (float)px, fpy = (float)py;
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + 3);
  const int start = ranges[tile];
  const float half_w = 0.5f * (float)W, half_h = 0.5f * (float)H;

  int count = 0;
  float T = 0.0f, Tf = 0.0f, gx = 0.0f, gy = 0.0f, gz = 0.0f;
  if (inside) {
    const size_t p = (size_t)py * W + px;
    count = splats_per_pixel[p];
    T = Tf = weight_per_pixel[p];
    gx = grad_image[3 * p]; gy = grad_image[3 * p + 1]; gz = grad_image[3 * p + 2];
  }
  float bgdot = 0.0f;   // render_backward.cu:149-152
  bgdot += bg * gx; bgdot += bg * gy; bgdot += bg * gz;
  float ar = 0.0f, ag = 0.0f, ab = 0.0f;   // colour accumulated behind the current entry
  const int warp_max = __reduce_max_sync(0xffffffffu, count);
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
  __syncthreads();
  const int cta_max = s_max;   // render_backward.cu:74 (start index = max count - 1)

  for (int top = cta_max; top > 0; top -= CB) {
    const int lo = max(0, top - CB);
    const int nb = top - lo;
    __syncthreads();
    if (tid < nb) {
      const int g = sorted[start + lo + tid];
      s_ids[tid] = g;
      stage_entry<PACKED>(src, g, &s_r0[tid], &s_r1[tid], &s_r2[tid]);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) s_acc[tid + k * CB] = 0.0f;
    __syncthreads();
    const int hi = min(top, warp_max);
    for (int chi = hi; chi > lo; chi -= 32) {
      const int j = chi - 1 - lane;   // lane 0 = farthest entry first (back to front)
      bool hit = false;
      if (j >= lo) {
        const float4 q = s_r0[j - lo];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int idx = chi - 1 - (__ffs(m) - 1);
        m &= m - 1;
        const int e = idx - lo;
        const float4 q0 = s_r0[e], q1 = s_r1[e];
        const float dx = q0.x - fpx, dy = q0.y - fpy;
        const float power = fminf(0.0f, -0.5f * (q1.x * dx * dx + q1.z * dy * dy) - q1.y * dx * dy);
        const float gg = exp_neg(power);
        const float alpha = fminf(0.99f, q1.w * gg);
        const bool valid = inside && (alpha >= ALPHA_MIN) && (idx < count);
        if (!__any_sync(0xffffffffu, valid)) continue;
        // Branch-free body: a lane that does not take the entry runs it with alpha = 0 and g = 0, which leaves its
        // pixel state untouched and contributes exact zeros.  Per lane only the moments of dL/dpower are formed
        // (gp, gp dx, gp dy, gp dx^2, gp dx dy, gp dy^2); the per-entry linear combinations that turn them into the
        // uv / conic / opacity gradients (render_backward.cu:154-166) are applied once per entry at the flush.
        const float4 q2 = s_r2[e];
        const float al = valid ? alpha : 0.0f;
        const float gm = valid ? gg : 0.0f;
        const float om = 1.0f - al;
        const float inv_om = rcp_fast(om);   // render_backward.cu:132 divides; 1 - alpha >= 0.01
        T = valid ? T * inv_om : T;
        const float w = al * T;
        float ga = 0.0f;
        ga += (q2.x - ar) * gx; ga += (q2.y - ag) * gy; ga += (q2.z - ab) * gz;
        ga *= T;
        ga += (-Tf * inv_om) * bgdot;
        ar = al * q2.x + om * ar; ag = al * q2.y + om * ag; ab = al * q2.z + om * ab;
        const float gp = gm * (ga * q1.w);   // d L / d power
        const float gpx = gp * dx, gpy = gp * dy;
        float v[8];
        v[0] = w * gx; v[1] = w * gy; v[2] = w * gz;
        v[3] = gpx; v[4] = gpy; v[5] = gpx * dx; v[6] = gpx * dy; v[7] = gpy * dy;
        float vo = gp;
        // transposing butterfly: 9 values over 32 lanes in 12 shuffles (the ninth rides along for three steps and
        // then shares the last two shuffles with the other eight)
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
        // lane with (lane&3)==0 holds value k = 4*b4 + 2*b3 + b2 ; lane 2 publishes the ninth
        const int k = (lane == 2) ? 8 : ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
        if ((lane & 3) == 0 || lane == 2) atomicAdd(&s_acc[e * 9 + k], a1);
      }
    }
    __syncthreads();
    if (tid < nb) {
      float m[9];
      bool any = false;
#pragma unroll
      for (int k = 0; k < 9; ++k) { m[k] = s_acc[tid * 9 + k]; any |= (m[k] != 0.0f); }
      // Q7 (render_backward.cu:170): the reference drops the whole entry when every lane's opacity partial is
      // exactly zero -- which is what happens when opa*(1-opa) underflows to 0.
      const float4 q1 = s_r1[tid];
      const float w2 = s_r2[tid].w;
      if (any && w2 != 0.0f) {
        float a[9];
        a[0] = m[0]; a[1] = m[1]; a[2] = m[2];
        a[3] = (-q1.x * m[3] - q1.y * m[4]) * half_w;   // Q5: NDC-scaled
        a[4] = (-q1.z * m[4] - q1.y * m[3]) * half_h;
        a[5] = -0.5f * m[5]; a[6] = -m[6]; a[7] = -0.5f * m[7];
        a[8] = m[8] * (w2 / q1.w);                      // sum of g * grad_alpha * opa * (1 - opa)
        const int g = s_ids[tid];
        if (PACKED) {
          float *o = dst.packed + (size_t)g * 12;
          red_add_v4(o, a[0], a[1], a[2], a[8]);
          red_add_v4(o + 4, a[3], a[4], a[5], a[6]);
          atomicAdd(o + 8, a[7]);
        } else {
          atomicAdd(dst.grad_rgb + 3 * g + 0, a[0]); atomicAdd(dst.grad_rgb + 3 * g + 1, a[1]);
          atomicAdd(dst.grad_rgb + 3 * g + 2, a[2]);
          atomicAdd(dst.grad_opacity + g, a[8]);
          atomicAdd(dst.grad_uv + 2 * g + 0, a[3]); atomicAdd(dst.grad_uv + 2 * g + 1, a[4]);
          atomicAdd(dst.grad_conic + 3 * g + 0, a[5]); atomicAdd(dst.grad_conic + 3 * g + 1, a[6]);
          atomicAdd(dst.grad_conic + 3 * g + 2, a[7]);
        }
      }
    }
  }
}

// View-level backward, two pixels per lane: a warp owns an 8x8 block (lane -> column lane&7, rows lane>>3 and +4), a
// 128-thread CTA one tile.  The per-pixel arithmetic is k_composite_bwd's; what changes is the cost of the cross-lane
// reduction (12 shuffles + selects + the shared-memory publish per contributing warp-hit, more than half of that kernel's
// instructions): the two pixels of a lane are summed in registers first, so it is paid once per 64 pixels.
template <int PPL>
__global__ void __launch_bounds__(256 / PPL, GSX_CBWD_MINB)
k_composite_bwd2(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const float4 *__restrict__ r2,
                 const int *__restrict__ sorted, const int *__restrict__ ranges, int ntx, int W, int H, float bg,
                 const int *__restrict__ splats_per_pixel, const float *__restrict__ weight_per_pixel,
                 const float *__restrict__ grad_image, float *__restrict__ packed) {
  constexpr int CB2 = 256 / PPL;   // threads per CTA == staged entries per batch
  constexpr int WROWS = 4 * PPL;   // rows of a warp's block
  __shared__ float4 s_r0[CB2], s_r1[CB2], s_r2[CB2];
  __shared__ float s_acc[CB2 * 9];
  __shared__ int s_ids[CB2];
  __shared__ int s_max;
  const int tile = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int X0 = (tile % ntx) * 16 + (warp & 1) * 8, Y0 = (tile / ntx) * 16 + (warp >> 1) * WROWS;
  const int px = X0 + (lane & 7);
  const float fpx = (float)px;
  const float bx0 = (float)X0, bx1 = (float)(X0 + 7), by0 = (float)Y0, by1 = (float)(Y0 + WROWS - 1);
  const int start = ranges[tile];
  const float half_w = 0.5f * (float)W, half_h = 0.5f * (float)H;

  int count[PPL];
  float fpy[PPL], T[PPL], Tf[PPL], gx[PPL], gy[PPL], gz[PPL], bgdot[PPL], ar[PPL], ag[PPL], ab[PPL];
  bool inside[PPL];
#pragma unroll
  for (int p = 0; p < PPL; ++p) {
    const int py = Y0 + (lane >> 3) + 4 * p;
    fpy[p] = (float)py;
    inside[p] = px < W && py < H;
    count[p] = 0; T[p] = Tf[p] = 0.0f; gx[p] = gy[p] = gz[p] = 0.0f;
    if (inside[p]) {
      const size_t q = (size_t)py * W + px;
      count[p] = splats_per_pixel[q];
      T[p] = Tf[p] = weight_per_pixel[q];
      gx[p] = grad_image[3 * q]; gy[p] = grad_image[3 * q + 1]; gz[p] = grad_image[3 * q + 2];
    }
    float b = 0.0f;   // render_backward.cu:149-152
    b += bg * gx[p]; b += bg * gy[p]; b += bg * gz[p];
    bgdot[p] = b;
    ar[p] = ag[p] = ab[p] = 0.0f;   // colour accumulated behind the current entry
  }
  int cmax = count[0];
#pragma unroll
  for (int p = 1; p < PPL; ++p) cmax = max(cmax, count[p]);
  const int warp_max = __reduce_max_sync(0xffffffffu, cmax);
  if (tid == 0) s_max = 0;
  __syncthreads();
  if (lane == 0 && warp_max > 0) atomicMax(&s_max, warp_max);
  __syncthreads();
  const int cta_max = s_max;   // render_backward.cu:74 (start index = max count - 1)

  for (int top = cta_max; top > 0; top -= CB2) {
    const int lo = max(0, top - CB2);
    const int nb = top - lo;
    __syncthreads();
    if (tid < nb) {
      const int g = sorted[start + lo + tid];
      s_ids[tid] = g;
      s_r0[tid] = __ldg(r0 + g); s_r1[tid] = __ldg(r1 + g); s_r2[tid] = __ldg(r2 + g);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) s_acc[tid + k * CB2] = 0.0f;
    __syncthreads();
    const int hi = min(top, warp_max);
    for (int chi = hi; chi > lo; chi -= 32) {
      const int j = chi - 1 - lane;   // lane 0 = farthest entry first (back to front)
      bool hit = false;
      if (j >= lo) {
        const float4 q = s_r0[j - lo];
        hit = !(q.x + q.z < bx0 || q.x - q.z > bx1 || q.y + q.w < by0 || q.y - q.w > by1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int idx = chi - 1 - (__ffs(m) - 1);
        m &= m - 1;
        const int e = idx - lo;
        const float4 q0 = s_r0[e], q1 = s_r1[e];
        const float dx = q0.x - fpx;
        float dy[PPL], gg[PPL], alpha[PPL];
        bool valid[PPL], anyv = false;
#pragma unroll
        for (int p = 0; p < PPL; ++p) {
          dy[p] = q0.y - fpy[p];
          const float power = fminf(0.0f, -0.5f * (q1.x * dx * dx + q1.z * dy[p] * dy[p]) - q1.y * dx * dy[p]);
          gg[p] = exp_neg(power);
          alpha[p] = fminf(0.99f, q1.w * gg[p]);
          valid[p] = inside[p] && (alpha[p] >= ALPHA_MIN) && (idx < count[p]);
          anyv |= valid[p];
        }
        if (!__any_sync(0xffffffffu, anyv)) continue;
        const float4 q2 = s_r2[e];
        float v[8], vo = 0.0f;
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = 0.0f;
#pragma unroll
        for (int p = 0; p < PPL; ++p) {   // branch-free: a pixel that does not take the entry runs with alpha = 0, g = 0
          const float al = valid[p] ? alpha[p] : 0.0f;
          const float gm = valid[p] ? gg[p] : 0.0f;
          const float om = 1.0f - al;
          const float inv_om = rcp_fast(om);   // render_backward.cu:132 divides; 1 - alpha >= 0.01
          T[p] = valid[p] ? T[p] * inv_om : T[p];
          const float w = al * T[p];
          float ga = 0.0f;
          ga += (q2.x - ar[p]) * gx[p]; ga += (q2.y - ag[p]) * gy[p]; ga += (q2.z - ab[p]) * gz[p];
          ga *= T[p];
          ga += (-Tf[p] * inv_om) * bgdot[p];
          ar[p] = al * q2.x + om * ar[p]; ag[p] = al * q2.y + om * ag[p]; ab[p] = al * q2.z + om * ab[p];
          const float gp = gm * (ga * q1.w);   // d L / d power
          const float gpx = gp * dx, gpy = gp * dy[p];
          v[0] += w * gx[p]; v[1] += w * gy[p]; v[2] += w * gz[p];
          v[3] += gpx; v[4] += gpy; v[5] += gpx * dx; v[6] += gpx * dy[p]; v[7] += gpy * dy[p];
          vo += gp;
        }
        // transposing butterfly: 9 values over 32 lanes in 12 shuffles (as k_composite_bwd)
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
          const bool up = lane & 2;
          const float send = up ? a1 : vo, keep = up ? vo : a1;
          a1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
        }
        a1 += __shfl_xor_sync(0xffffffffu, a1, 1);
        const int k = (lane == 2) ? 8 : ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1);
        if ((lane & 3) == 0 || lane == 2) atomicAdd(&s_acc[e * 9 + k], a1);
      }
    }
    __syncthreads();
    if (tid < nb) {
      float m[9];
      bool any = false;
#pragma unroll
      for (int k = 0; k < 9; ++k) { m[k] = s_acc[tid * 9 + k]; any |= (m[k] != 0.0f); }
      const float4 q1 = s_r1[tid];
      const float w2 = s_r2[tid].w;
      if (any && w2 != 0.0f) {   // Q7 (render_backward.cu:170), see k_composite_bwd
        float *o = packed + (size_t)s_ids[tid] * 12;
        red_add_v4(o, m[0], m[1], m[2], m[8] * (w2 / q1.w));
        red_add_v4(o + 4, (-q1.x * m[3] - q1.y * m[4]) * half_w, (-q1.z * m[4] - q1.y * m[3]) * half_h, -0.5f * m[5], -m[6]);
        atomicAdd(o + 8, -0.5f * m[7]);
      }
    }
  }
}

// Work counters for the roofline arithmetic (bench.py only): S = sum of splats_per_pixel (pixel-entries the
// reference evaluates), S+ = those with alpha >= 1/255.  One thread per pixel, no staging: not a hot path.
__global__ void __launch_bounds__(CB)
k_count_work(const float4 *__restrict__ r0, const float4 *__restrict__ r1, const int *__restrict__ sorted,
             const int *__restrict__ ranges, int ntx, int W, int H, const int *__restrict__ spp,
             unsigned long long *__restrict__ out) {
  const int tile = blockIdx.x, tid = threadIdx.x;
  const int px = (tile % ntx) * 16 + (tid & 15), py = (tile / ntx) * 16 + (tid >> 4);
  unsigned long long s = 0, sp = 0;
  if (px < W && py < H) {
    const int start = ranges[tile];
    const int count = spp[(size_t)py * W + px];
    s = (unsigned long long)count;
    for (int i = 0; i < count; ++i) {
      const int g = sorted[start + i];
      const float4 q0 = __ldg(r0 + g), q1 = __ldg(r1 + g);
      const float dx = q0.x - (float)px, dy = q0.y - (float)py;
      const float power = fminf(0.0f, -0.5f * (q1.x * dx * dx + q1.z * dy * dy) - q1.y * dx * dy);
      sp += (fminf(0.99f, q1.w * __expf(power)) >= ALPHA_MIN) ? 1ull : 0ull;
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, off);
    sp += __shfl_down_sync(0xffffffffu, sp, off);
  }
  if ((tid & 31) == 0) { atomicAdd(out, s); atomicAdd(out + 1, sp); }
}

int count_work_launch(const float4 *r0, const float4 *r1, const int *sorted, const int *ranges, int W, int H,
                      const int *spp, unsigned long long *out_dev, cudaStream_t st) {
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  k_count_work<<<ntx * nty, CB, 0, st>>>(r0, r1, sorted, ranges, ntx, W, H, spp, out_dev);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

// composite3.cu: the round-2 kernels (packed fp32x2 arithmetic, cp.async ring, direct per-warp reductions).  Default;
// GSX_COMPOSITE=v2 in the environment selects the round-1 kernels of this file for A/B measurements.
int composite_fwd3_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, float bg, int *spp, float *wpp, float *image, cudaStream_t st, const int *order);
int composite_bwd3_launch(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges, int W,
                          int H, float bg, const int *spp, const float *wpp, const float *grad_image, float *packed_grads,
                          cudaStream_t st, const int *order);
static bool use_v2() {
  static const int v = [] { const char *e = getenv("GSX_COMPOSITE"); return (e && e[0] == 'v' && e[1] == '2') ? 1 : 0; }();
  return v != 0;
}

int composite_fwd_packed(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges,
                         int W, int H, float bg, int *spp, float *wpp, float *image, cudaStream_t st, const int *order) {
  if (!use_v2()) return composite_fwd3_launch(r0, r1, r2, sorted, ranges, W, H, bg, spp, wpp, image, st, order);
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  k_composite_fwd2<<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, ntx, W, H, bg, spp, wpp, image);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int composite_bwd_packed(const float4 *r0, const float4 *r1, const float4 *r2, const int *sorted, const int *ranges,
                         int W, int H, float bg, const int *spp, const float *wpp, const float *grad_image,
                         float *packed_grads, cudaStream_t st, const int *order) {
  if (!use_v2()) return composite_bwd3_launch(r0, r1, r2, sorted, ranges, W, H, bg, spp, wpp, grad_image, packed_grads, st, order);
  const int ntx = (W + 15) / 16, nty = (H + 15) / 16;
  k_composite_bwd2<2><<<ntx * nty, 128, 0, st>>>(r0, r1, r2, sorted, ranges, ntx, W, H, bg, spp, wpp, grad_image, packed_grads);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_render_image(const float *uv, const float *opacity, const float *conic, const float *rgb, float background,
                     const int *sorted_splats, const int *splat_range_by_tile, int width, int height,
                     int *splats_per_pixel, float *weight_per_pixel, float *image, void *stream) {
  GSX_REQUIRE_DEV(uv); GSX_REQUIRE_DEV(opacity); GSX_REQUIRE_DEV(conic); GSX_REQUIRE_DEV(rgb);
  GSX_REQUIRE_DEV(sorted_splats); GSX_REQUIRE_DEV(splat_range_by_tile); GSX_REQUIRE_DEV(splats_per_pixel);
  GSX_REQUIRE_DEV(weight_per_pixel); GSX_REQUIRE_DEV(image);
  if (width <= 0 || height <= 0) return GSX_OK;
  const int ntx = (width + 15) / 16, nty = (height + 15) / 16;
  GaussSrc s{};
  s.uv = uv; s.opacity = opacity; s.conic = conic; s.rgb = rgb;
  k_composite_fwd<false><<<ntx * nty, CB, 0, (cudaStream_t)stream>>>(s, sorted_splats, splat_range_by_tile, ntx, width,
                                                                     height, background, splats_per_pixel,
                                                                     weight_per_pixel, image);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_render_image_backward(const float *uv, const float *opacity, const float *conic, const float *rgb,
                              float background, const int *sorted_splats, const int *splat_range_by_tile,
                              const int *splats_per_pixel, const float *weight_per_pixel, const float *grad_image,
                              int width, int height, int N, float *grad_rgb, float *grad_opacity, float *grad_uv,
                              float *grad_conic, void *stream) {
  (void)N;
  GSX_REQUIRE_DEV(uv); GSX_REQUIRE_DEV(opacity); GSX_REQUIRE_DEV(conic); GSX_REQUIRE_DEV(rgb);
  GSX_REQUIRE_DEV(sorted_splats); GSX_REQUIRE_DEV(splat_range_by_tile); GSX_REQUIRE_DEV(splats_per_pixel);
  GSX_REQUIRE_DEV(weight_per_pixel); GSX_REQUIRE_DEV(grad_image); GSX_REQUIRE_DEV(grad_rgb);
  GSX_REQUIRE_DEV(grad_opacity); GSX_REQUIRE_DEV(grad_uv); GSX_REQUIRE_DEV(grad_conic);
  if (width <= 0 || height <= 0) return GSX_OK;
  const int ntx = (width + 15) / 16, nty = (height + 15) / 16;
  GaussSrc s{};
  s.uv = uv; s.opacity = opacity; s.conic = conic; s.rgb = rgb;
  GradDst d{};
  d.grad_rgb = grad_rgb; d.grad_opacity = grad_opacity; d.grad_uv = grad_uv; d.grad_conic = grad_conic;
  k_composite_bwd<false><<<ntx * nty, CB, 0, (cudaStream_t)stream>>>(s, sorted_splats, splat_range_by_tile, ntx, width,
                                                                     height, background, splats_per_pixel,
                                                                     weight_per_pixel, grad_image, d);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // extern "C"
