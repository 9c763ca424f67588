// optim.cu -- SURVEY 8f rank 2: the step right after the splatting path.  Adam over the GLOBAL-index gradient
// buffers the path produces, every parameter group and the densification statistic in ONE launch.
//
// Reference: cuda/optimizer.cu:6-44 (adam_kernel / adam_step, one launch per parameter array on arrays compacted by
// the cull mask) driven by TrainerImpl::optimizer_step, cuda/trainer.cu:1027-1158: 16-19 compactions, 5-6 adam_step
// launches, 16-19 scatters, then the accumulators uv_grad_accum += |grad_uv| and grad_accum_dur += 1 for the visible
// Gaussians (trainer.cu:1138-1157, PositionalGradientNorm trainer.cu:1018-1025).  Here the mask is read in place: a
// Gaussian whose mask byte is 0 keeps its parameters AND its moments untouched, exactly what compaction + scatter do.
//
// HBM-bound: 28 B per updated float (param, grad, m, v read; param, m, v written) + 1 mask byte and 24 B of
// statistics per Gaussian.  float4 streams (ld/st.global.cs: nothing is reused), GSX_ADAM_BATCH float4s of each stream
// requested per thread before the first use, grid-stride over 148 x 8 CTAs per group.
#include "gsx_common.cuh"

namespace gsx {
namespace {

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

struct AdamGroupDev {
  float *p;
  const float *g;
  float *m, *v;
  long long n;        // N * width floats
  int width;          // floats per Gaussian (the S of adam_step)
  float lr;
  int vec;            // all four pointers 16-byte aligned
  unsigned div_mul;   // element index -> Gaussian index without a division: (e * div_mul) >> div_shift, exact for e < 2^31
  int div_shift;      //   (0 = n too large for it: plain 64-bit division)
};

struct AdamArgs {
  AdamGroupDev grp[GSX_ADAM_MAX_GROUPS];
  int n_groups;
  const uint8_t *mask;   // [N] or null (every Gaussian is updated)
  int N;
  float b1, b2, eps, bias1, bias2;
  const float2 *uv_grad;   // [N][2] or null (no statistics)
  float *uv_accum;         // [N]
  int *dur;                // [N]
};

// The reference's update (cuda/optimizer.cu:14-28): a NaN gradient counts as 0.  Roundings are spelled out -- one fused
// multiply-add per moment, as nvcc contracts the reference's expressions, IEEE division and square root -- so the float4
// and the scalar path give the same bits.  The guards keep exact zeros (moments of Gaussians no view has touched yet,
// gradients of culled ones) off the slow path of the IEEE division / square root sequences: x/y and sqrt(x) ARE x for
// x = +-0 and y > 0, so results are unchanged; without them a realistic gradient buffer (half of it zeros) runs at 66 %
// of the bandwidth of a dense one.
__device__ __forceinline__ float div_pos(float x, float y) { return (x == 0.0f && y > 0.0f) ? x : __fdiv_rn(x, y); }
__device__ __forceinline__ void adam_elem(float &p, float g, float &m, float &v, const float lr, const float b1,
                                          const float b2, const float eps, const float bias1, const float bias2) {
  if (isnan(g)) g = 0.0f;
  m = __fmaf_rn(b1, m, __fmul_rn(1.0f - b1, g));
  v = __fmaf_rn(b2, v, __fmul_rn(__fmul_rn(1.0f - b2, g), g));
  const float m_hat = div_pos(m, bias1);
  const float v_hat = div_pos(v, bias2);
  const float root = v_hat == 0.0f ? v_hat : __fsqrt_rn(v_hat);
  const float step = div_pos(__fmul_rn(-lr, m_hat), __fadd_rn(root, eps));
  p = __fadd_rn(p, step);
}

constexpr int kAdamBlock = 256;
// float4s of each of the four streams a thread requests before it touches the first one: the kernel is a pure stream
// (no reuse), so the bytes in flight per SM are what sets the achieved bandwidth
#ifndef GSX_ADAM_BATCH
#define GSX_ADAM_BATCH 1
#endif
constexpr int kAdamBatch = GSX_ADAM_BATCH;
#ifndef GSX_ADAM_MINB
#define GSX_ADAM_MINB 4
#endif

__device__ __forceinline__ long long adam_row_of(const AdamGroupDev &G, long long e) {
  return G.div_shift ? (long long)(((unsigned long long)(unsigned)e * G.div_mul) >> G.div_shift) : e / G.width;
}

// which of the four floats of float4 q belong to Gaussians the mask keeps
__device__ __forceinline__ unsigned adam_keep_bits(const AdamGroupDev &G, const uint8_t *__restrict__ mask, long long q) {
  const long long e = q * 4;
  long long i = adam_row_of(G, e);
  int r = (int)(e - i * G.width);
  unsigned keep = 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    keep |= (mask[i] ? 1u : 0u) << c;
    if (++r == G.width && c < 3) { r = 0; ++i; }
  }
  return keep;
}

__global__ void __launch_bounds__(kAdamBlock, GSX_ADAM_MINB) k_adam_fused(const AdamArgs a) {
  const int gi = blockIdx.y;
  const long long tid = (long long)blockIdx.x * kAdamBlock + threadIdx.x;
  const long long nthr = (long long)gridDim.x * kAdamBlock;

  if (gi == a.n_groups) {   // densification statistics of the visible Gaussians
    for (long long i = tid; i < a.N; i += nthr) {
      if (a.mask && !a.mask[i]) continue;
      const float2 d = a.uv_grad[i];
      a.uv_accum[i] += sqrtf(d.x * d.x + d.y * d.y);
      a.dur[i] += 1;
    }
    return;
  }

  const AdamGroupDev G = a.grp[gi];
  const long long nq = G.vec ? (G.n >> 2) : 0;
  float4 *const P4 = reinterpret_cast<float4 *>(G.p);
  const float4 *const G4 = reinterpret_cast<const float4 *>(G.g);
  float4 *const M4 = reinterpret_cast<float4 *>(G.m);
  float4 *const V4 = reinterpret_cast<float4 *>(G.v);
  for (long long q0 = tid; q0 < nq; q0 += nthr * kAdamBatch) {
    unsigned keep[kAdamBatch];
    float4 p[kAdamBatch], g[kAdamBatch], m[kAdamBatch], v[kAdamBatch];
#pragma unroll
    for (int u = 0; u < kAdamBatch; ++u) {
      const long long q = q0 + u * nthr;
      keep[u] = q < nq ? (a.mask ? adam_keep_bits(G, a.mask, q) : 0xFu) : 0u;
    }
#pragma unroll
    for (int u = 0; u < kAdamBatch; ++u) {   // every request of the batch leaves before the first use
      const long long q = q0 + u * nthr;
      if (keep[u]) {   // a float4 whose Gaussians are all culled costs no HBM traffic
        p[u] = __ldcs(P4 + q);
        g[u] = __ldcs(G4 + q);
        m[u] = __ldcs(M4 + q);
        v[u] = __ldcs(V4 + q);
      }
    }
#pragma unroll
    for (int u = 0; u < kAdamBatch; ++u) {
      if (!keep[u]) continue;
      const long long q = q0 + u * nthr;
      if (keep[u] & 1u) adam_elem(p[u].x, g[u].x, m[u].x, v[u].x, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
      if (keep[u] & 2u) adam_elem(p[u].y, g[u].y, m[u].y, v[u].y, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
      if (keep[u] & 4u) adam_elem(p[u].z, g[u].z, m[u].z, v[u].z, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
      if (keep[u] & 8u) adam_elem(p[u].w, g[u].w, m[u].w, v[u].w, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
      __stcs(P4 + q, p[u]);
      __stcs(M4 + q, m[u]);
      __stcs(V4 + q, v[u]);
    }
  }
  // scalar remainder (everything when the arrays are not 16-byte aligned)
  for (long long e = nq * 4 + tid; e < G.n; e += nthr) {
    if (a.mask && !a.mask[e / G.width]) continue;
    float p = G.p[e], m = G.m[e], v = G.v[e];
    adam_elem(p, G.g[e], m, v, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
    G.p[e] = p;
    G.m[e] = m;
    G.v[e] = v;
  }
}

int launch_adam(const AdamArgs &a, cudaStream_t stream) {
  long long longest = a.uv_grad ? a.N : 0;
  for (int g = 0; g < a.n_groups; ++g) {
    const long long work = a.grp[g].vec ? (a.grp[g].n / 4 + kAdamBatch - 1) / kAdamBatch + 3 : a.grp[g].n;
    if (work > longest) longest = work;
  }
  if (longest <= 0) return GSX_OK;
  long long gx = (longest + kAdamBlock - 1) / kAdamBlock;
  if (gx > kNumSMs * 8) gx = kNumSMs * 8;
  const dim3 grid((unsigned)gx, (unsigned)(a.n_groups + (a.uv_grad ? 1 : 0)));
  k_adam_fused<<<grid, kAdamBlock, 0, stream>>>(a);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

AdamGroupDev make_group(float *p, const float *g, float *m, float *v, long long n, int width, float lr) {
  AdamGroupDev d;
  d.p = p; d.g = g; d.m = m; d.v = v; d.n = n; d.width = width; d.lr = lr;
  d.vec = aligned16(p) && aligned16(g) && aligned16(m) && aligned16(v);
  d.div_mul = 0; d.div_shift = 0;
  if (n < (1LL << 31) && width <= (1 << 16)) {   // e / width == (e * ceil(2^(31+l) / width)) >> (31+l) for e < 2^31, l = ceil(log2 width)
    int l = 0;
    while ((1LL << l) < width) ++l;
    d.div_shift = 31 + l;
    d.div_mul = (unsigned)(((1ULL << d.div_shift) + (unsigned)width - 1) / (unsigned)width);
  }
  return d;
}

}  // namespace
}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_adam_step(float *params, const float *param_grads, float *exp_avg, float *exp_avg_sq, float lr, float b1,
                  float b2, float eps, float bias1, float bias2, long long n_floats, void *stream) {
  GSX_REQUIRE_DEV(params); GSX_REQUIRE_DEV(param_grads); GSX_REQUIRE_DEV(exp_avg); GSX_REQUIRE_DEV(exp_avg_sq);
  if (n_floats <= 0) return GSX_OK;
  AdamArgs a;
  memset(&a, 0, sizeof(a));
  a.grp[0] = make_group(params, param_grads, exp_avg, exp_avg_sq, n_floats, 1, lr);
  a.n_groups = 1;
  a.N = 0;
  a.b1 = b1; a.b2 = b2; a.eps = eps; a.bias1 = bias1; a.bias2 = bias2;
  return launch_adam(a, (cudaStream_t)stream);
}

int gsx_adam_step_fused(const gsx_adam_group_t *groups_host, int n_groups, const uint8_t *mask, int N, float b1,
                        float b2, float eps, float bias1, float bias2, const float *uv_grad, float *uv_grad_accum,
                        int *grad_accum_dur, void *stream) {
  if (!groups_host && n_groups > 0) { set_error("groups_host is null"); return GSX_E_NULL; }
  if (n_groups < 0 || n_groups > GSX_ADAM_MAX_GROUPS) { set_error("n_groups %d outside 0..%d", n_groups, GSX_ADAM_MAX_GROUPS); return GSX_E_ARG; }
  if (N <= 0) return GSX_OK;
  if (mask) GSX_REQUIRE_DEV(mask);
  AdamArgs a;
  memset(&a, 0, sizeof(a));
  for (int g = 0; g < n_groups; ++g) {
    const gsx_adam_group_t &h = groups_host[g];
    if (h.width <= 0) { set_error("group %d: width %d", g, h.width); return GSX_E_ARG; }
    GSX_REQUIRE_DEV(h.param); GSX_REQUIRE_DEV(h.grad); GSX_REQUIRE_DEV(h.exp_avg); GSX_REQUIRE_DEV(h.exp_avg_sq);
    a.grp[g] = make_group(h.param, h.grad, h.exp_avg, h.exp_avg_sq, (long long)N * h.width, h.width, h.lr);
  }
  a.n_groups = n_groups;
  a.mask = mask;
  a.N = N;
  a.b1 = b1; a.b2 = b2; a.eps = eps; a.bias1 = bias1; a.bias2 = bias2;
  if (uv_grad) {
    GSX_REQUIRE_DEV(uv_grad); GSX_REQUIRE_DEV(uv_grad_accum); GSX_REQUIRE_DEV(grad_accum_dur);
    a.uv_grad = reinterpret_cast<const float2 *>(uv_grad);
    a.uv_accum = uv_grad_accum;
    a.dur = grad_accum_dur;
  }
  return launch_adam(a, (cudaStream_t)stream);
}

}  // extern "C"
