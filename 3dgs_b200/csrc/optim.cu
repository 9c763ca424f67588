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
// statistics per Gaussian.  float4 streams, grid-stride over 148 x 8 CTAs per group.
#include "gsx_common.cuh"

namespace gsx {
namespace {

struct AdamGroupDev {
  float *p;
  const float *g;
  float *m, *v;
  long long n;   // N * width floats
  int width;     // floats per Gaussian (the S of adam_step)
  float lr;
  int vec;       // all four pointers 16-byte aligned
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

// The reference's update, expression for expression (cuda/optimizer.cu:14-28): a NaN gradient counts as 0.
__device__ __forceinline__ void adam_elem(float &p, float g, float &m, float &v, const float lr, const float b1,
                                          const float b2, const float eps, const float bias1, const float bias2) {
  if (isnan(g)) g = 0.0f;
  m = b1 * m + (1.0f - b1) * g;
  v = b2 * v + (1.0f - b2) * g * g;
  const float m_hat = m / bias1;
  const float v_hat = v / bias2;
  const float step = -lr * m_hat / (sqrtf(v_hat) + eps);
  p += step;
}

constexpr int kAdamBlock = 256;

__global__ void __launch_bounds__(kAdamBlock) k_adam_fused(const AdamArgs a) {
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
  for (long long q = tid; q < nq; q += nthr) {
    unsigned keep = 0xFu;
    if (a.mask) {
      const long long e = q * 4;
      long long i = e / G.width;
      int r = (int)(e - i * G.width);
      keep = 0;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        keep |= (a.mask[i] ? 1u : 0u) << c;
        if (++r == G.width && c < 3) { r = 0; ++i; }
      }
      if (!keep) continue;
    }
    float4 p = reinterpret_cast<float4 *>(G.p)[q];
    const float4 g = reinterpret_cast<const float4 *>(G.g)[q];
    float4 m = reinterpret_cast<float4 *>(G.m)[q];
    float4 v = reinterpret_cast<float4 *>(G.v)[q];
    if (keep & 1u) adam_elem(p.x, g.x, m.x, v.x, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
    if (keep & 2u) adam_elem(p.y, g.y, m.y, v.y, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
    if (keep & 4u) adam_elem(p.z, g.z, m.z, v.z, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
    if (keep & 8u) adam_elem(p.w, g.w, m.w, v.w, G.lr, a.b1, a.b2, a.eps, a.bias1, a.bias2);
    reinterpret_cast<float4 *>(G.p)[q] = p;
    reinterpret_cast<float4 *>(G.m)[q] = m;
    reinterpret_cast<float4 *>(G.v)[q] = v;
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
    const long long work = a.grp[g].vec ? (a.grp[g].n + 3) / 4 : a.grp[g].n;
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

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

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
  a.grp[0] = {params, param_grads, exp_avg, exp_avg_sq, n_floats, 1, lr,
              aligned16(params) && aligned16(param_grads) && aligned16(exp_avg) && aligned16(exp_avg_sq)};
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
    a.grp[g] = {h.param, h.grad, h.exp_avg, h.exp_avg_sq, (long long)N * h.width, h.width, h.lr,
                aligned16(h.param) && aligned16(h.grad) && aligned16(h.exp_avg) && aligned16(h.exp_avg_sq)};
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
