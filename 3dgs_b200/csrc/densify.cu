// densify.cu -- SURVEY 8f rank 4 (second half): adaptive density control on the path's arrays -- clone, split, prune.
//
// Reference: clone_gaussians / split_gaussians (include/gsplat_cuda/adaptive_density.cuh, cuda/adaptive_density.cu:8-206)
// and their caller TrainerImpl::adaptive_density_step (cuda/trainer.cu:416-775): ~10 thrust transforms / counts / scans
// with a host round trip each, 12 temporary vectors for the new Gaussians, 18 thrust::copy_if compactions of the
// parameters and Adam moments into temporaries, 12 fills, 18 + 12 copies back, and a curand state per thread re-seeded
// with time(NULL) at every call.
//
// Here (no tensor cores: byte-moving and a few transcendental per Gaussian, HBM-bound):
//   gsx_densify_plan  : ONE pass over the per-Gaussian statistics decides prune / clone / split (same predicates, same
//                       float expressions as trainer.cu:416-512), then a 3-kernel scan turns the decisions into
//                       destination rows; the three counts come back in one 12-byte copy (the step's only host sync,
//                       needed for the capacity check the reference does too).
//   gsx_densify_apply : ONE launch moves every array -- 6 parameter arrays and their 12 Adam moment arrays, blockIdx.y =
//                       array -- from the old set to the new one (out of place, coalesced reads): survivors keep their
//                       order at the front, clones are appended after them (moments zero), split children after the
//                       clones; a second small launch over the split parents writes the children's positions and scales.
//   Counter-based RNG  : Philox4x32-10 keyed by an explicit seed, counter = (parent index, child) -> the samples do not
//                       depend on the launch shape, need no state array, and a run can be reproduced; the reference-named
//                       C++ shim (compat.cu) seeds with time(NULL) like the reference.
//   gsx_clone_gaussians / gsx_split_gaussians: the reference's two functions with their exact argument meaning (mask +
//                       exclusive-scan write ids in, separate output buffers), for drop-in callers and the reference's
//                       own tests/adaptive_density_test.cpp.
#include <algorithm>

#include "gsx_common.cuh"

namespace gsx {
namespace {

constexpr int DB = 256;
constexpr int D_IPT = 8;
constexpr int D_TILE = DB * D_IPT;

// ---- Philox4x32-10 (Salmon et al., SC'11): counter-based, no state -----------------------------------------------------
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1,
                                              uint32_t out[4]) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
// 24 random bits -> (0, 1), exactly representable
__host__ __device__ inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f) + (0.5f / 16777216.0f); }
// three standard normals for child j of parent i (Box-Muller on one Philox block; the fourth normal is unused)
__device__ inline void normal3(unsigned long long seed, uint32_t i, uint32_t j, float n[3]) {
  uint32_t r[4];
  philox4x32_10(i, j, 0x3D65u, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), r);
  const float ra = sqrtf(-2.0f * logf(u01(r[0]))), ta = 6.283185307179586f * u01(r[1]);
  const float rb = sqrtf(-2.0f * logf(u01(r[2]))), tb = 6.283185307179586f * u01(r[3]);
  n[0] = ra * cosf(ta); n[1] = ra * sinf(ta); n[2] = rb * cosf(tb);
}

// rotation of the normalised quaternion (w, x, y, z), adaptive_density.cu:95-125
__device__ inline void quat_rot(float qw, float qx, float qy, float qz, float R[9]) {
  const float inv = rsqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
  const float w = qw * inv, x = qx * inv, y = qy * inv, z = qz * inv;
  const float x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
  R[0] = 1.0f - 2.0f * (y2 + z2); R[1] = 2.0f * (xy - wz); R[2] = 2.0f * (xz + wy);
  R[3] = 2.0f * (xy + wz); R[4] = 1.0f - 2.0f * (x2 + z2); R[5] = 2.0f * (yz - wx);
  R[6] = 2.0f * (xz - wy); R[7] = 2.0f * (yz + wx); R[8] = 1.0f - 2.0f * (x2 + y2);
}

// children of one split parent: centre + R (n * exp(scale)), log(exp(scale) / factor)   (adaptive_density.cu:127-151)
__device__ inline void split_children(unsigned long long seed, uint32_t i, const float xyz[3], const float scale[3],
                                      const float q[4], float factor, float cx[2][3], float cs[3]) {
  const float es[3] = {expf(scale[0]), expf(scale[1]), expf(scale[2])};
  float R[9];
  quat_rot(q[0], q[1], q[2], q[3], R);
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    float n[3];
    normal3(seed, i, (uint32_t)j, n);
    const float vx = n[0] * es[0], vy = n[1] * es[1], vz = n[2] * es[2];
    cx[j][0] = xyz[0] + (vx * R[0] + vy * R[1] + vz * R[2]);
    cx[j][1] = xyz[1] + (vx * R[3] + vy * R[4] + vz * R[5]);
    cx[j][2] = xyz[2] + (vx * R[6] + vy * R[7] + vz * R[8]);
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) cs[k] = logf(es[k] / factor);
}

// ---- reference-shaped entry points ---------------------------------------------------------------------------------------
// one thread per output float of a row: rows of the selected Gaussians go to row write_ids[i] * copies (+ j)
__global__ void __launch_bounds__(DB) k_copy_rows_masked(const float *__restrict__ src, const uint8_t *__restrict__ mask,
                                                         const int *__restrict__ write_ids, long long total, int width,
                                                         int copies, float *__restrict__ dst) {
  for (long long e = (long long)blockIdx.x * DB + threadIdx.x; e < total; e += (long long)gridDim.x * DB) {
    const long long i = e / width;
    if (!mask[i]) continue;
    const int k = (int)(e - i * width);
    const float v = src[e];
    const long long row = (long long)write_ids[i] * copies;
    for (int j = 0; j < copies; ++j) dst[(row + j) * width + k] = v;
  }
}

__global__ void __launch_bounds__(DB) k_split_geometry(int N, float factor, unsigned long long seed,
                                                       const uint8_t *__restrict__ mask, const int *__restrict__ write_ids,
                                                       const float *__restrict__ xyz_in, const float *__restrict__ scale_in,
                                                       const float *__restrict__ quat_in, float *__restrict__ xyz_out,
                                                       float *__restrict__ scale_out) {
  const int i = blockIdx.x * DB + threadIdx.x;
  if (i >= N || !mask[i]) return;
  const float xyz[3] = {xyz_in[3 * (size_t)i], xyz_in[3 * (size_t)i + 1], xyz_in[3 * (size_t)i + 2]};
  const float sc[3] = {scale_in[3 * (size_t)i], scale_in[3 * (size_t)i + 1], scale_in[3 * (size_t)i + 2]};
  const float q[4] = {quat_in[4 * (size_t)i], quat_in[4 * (size_t)i + 1], quat_in[4 * (size_t)i + 2], quat_in[4 * (size_t)i + 3]};
  float cx[2][3], cs[3];
  split_children(seed, (uint32_t)i, xyz, sc, q, factor, cx, cs);
  const size_t row = (size_t)write_ids[i] * 2;
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) { xyz_out[(row + j) * 3 + k] = cx[j][k]; scale_out[(row + j) * 3 + k] = cs[k]; }
}

int copy_rows_masked(const float *src, const uint8_t *mask, const int *wid, int N, int width, int copies, float *dst,
                     cudaStream_t st) {
  if (N <= 0 || width <= 0) return GSX_OK;
  const long long total = (long long)N * width;
  const int blocks = (int)std::min<long long>(kNumSMs * 16, (total + DB - 1) / DB);
  k_copy_rows_masked<<<blocks, DB, 0, st>>>(src, mask, wid, total, width, copies, dst);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

// ---- fused step --------------------------------------------------------------------------------------------------------
enum : uint8_t { ACT_KEEP = 0, ACT_PRUNE = 1, ACT_CLONE = 2, ACT_SPLIT = 3 };

struct PlanCfg {
  float grad_threshold, op_threshold_logit, max_scale, clone_scale_threshold;
};

// trainer.cu:416-504 -- the same predicates on the same float expressions (expf of the log-scales, logit threshold)
__device__ __forceinline__ uint8_t decide(const PlanCfg c, float opacity_logit, float s1l, float s2l, float s3l, float accum,
                                          int dur) {
  const float avg = dur == 0 ? 0.0f : accum / (float)dur;                 // ComputeAvgGrad
  const float max_s = fmaxf(expf(s1l), fmaxf(expf(s2l), expf(s3l)));      // ComputeScaleMax / IdentifyPrune
  bool prune;
  if (opacity_logit < c.op_threshold_logit) prune = true;
  else if (avg > c.grad_threshold && (max_s / 1.6f) <= c.max_scale) prune = false;
  else prune = max_s > c.max_scale;
  if (prune) return ACT_PRUNE;
  if (avg > c.grad_threshold) return max_s <= c.clone_scale_threshold ? ACT_CLONE : ACT_SPLIT;   // IdentifyClone / Split
  return ACT_KEEP;
}

struct Cnt3 { int keep, clone, split; };   // keep = survives at the front (everything but pruned and split)
__device__ __forceinline__ Cnt3 cnt_of(uint8_t a) { return Cnt3{(a == ACT_KEEP || a == ACT_CLONE) ? 1 : 0, a == ACT_CLONE ? 1 : 0, a == ACT_SPLIT ? 1 : 0}; }

__global__ void __launch_bounds__(DB) k_plan_decide(const PlanCfg c, int N, const float *__restrict__ opacity,
                                                    const float *__restrict__ scale, const float *__restrict__ accum,
                                                    const int *__restrict__ dur, uint8_t *__restrict__ action,
                                                    int *__restrict__ block_sums /*[3][nb]*/, int nb) {
  __shared__ int s_tmp[3][DB / 32];
  const long long base = (long long)blockIdx.x * D_TILE;
  int k = 0, cl = 0, sp = 0;
#pragma unroll
  for (int u = 0; u < D_IPT; ++u) {
    const long long i = base + u * DB + threadIdx.x;
    if (i < N) {
      const uint8_t a = decide(c, opacity[i], scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], accum[i], dur[i]);
      action[i] = a;
      const Cnt3 t = cnt_of(a);
      k += t.keep; cl += t.clone; sp += t.split;
    }
  }
  for (int off = 16; off > 0; off >>= 1) {
    k += __shfl_down_sync(0xffffffffu, k, off); cl += __shfl_down_sync(0xffffffffu, cl, off); sp += __shfl_down_sync(0xffffffffu, sp, off);
  }
  if ((threadIdx.x & 31) == 0) { s_tmp[0][threadIdx.x >> 5] = k; s_tmp[1][threadIdx.x >> 5] = cl; s_tmp[2][threadIdx.x >> 5] = sp; }
  __syncthreads();
  if (threadIdx.x < 3) {
    int t = 0;
    for (int w = 0; w < DB / 32; ++w) t += s_tmp[threadIdx.x][w];
    block_sums[threadIdx.x * nb + blockIdx.x] = t;
  }
}

// exclusive scan of the three block-sum rows (one block; nb is a few thousand), totals -> counts[0..2]
__global__ void __launch_bounds__(1024) k_plan_scan_blocks(int *__restrict__ block_sums, int nb, int *__restrict__ counts) {
  __shared__ int s[1024];
  __shared__ int carry;
  for (int row = 0; row < 3; ++row) {
    int *b = block_sums + (size_t)row * nb;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nb; b0 += 1024) {
      const int i = b0 + threadIdx.x;
      const int v = i < nb ? b[i] : 0;
      s[threadIdx.x] = v;
      __syncthreads();
      for (int off = 1; off < 1024; off <<= 1) {
        const int t = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
        __syncthreads();
        s[threadIdx.x] += t;
        __syncthreads();
      }
      if (i < nb) b[i] = carry + s[threadIdx.x] - v;
      __syncthreads();
      if (threadIdx.x == 1023) carry += s[1023];
      __syncthreads();
    }
    if (threadIdx.x == 0) counts[row] = carry;
    __syncthreads();
  }
}

// destination rows: dst_keep[i] = new index of a survivor (or -1), dst_new[i] = first appended row of a clone / split
// parent relative to the end of the survivors' block of its kind (or -1)
__global__ void __launch_bounds__(DB) k_plan_rows(int N, const uint8_t *__restrict__ action, const int *__restrict__ block_sums,
                                                  int nb, int *__restrict__ dst_keep, int *__restrict__ dst_new) {
  __shared__ int s_w[3][DB / 32];
  const long long base = (long long)blockIdx.x * D_TILE + (long long)threadIdx.x * D_IPT;
  uint8_t a[D_IPT];
  int k = 0, cl = 0, sp = 0;
#pragma unroll
  for (int u = 0; u < D_IPT; ++u) {
    a[u] = base + u < N ? action[base + u] : (uint8_t)ACT_PRUNE;
    const Cnt3 t = cnt_of(a[u]);
    k += t.keep; cl += t.clone; sp += t.split;
  }
  int ik = k, ic = cl, is = sp;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int tk = __shfl_up_sync(0xffffffffu, ik, off), tc = __shfl_up_sync(0xffffffffu, ic, off), ts = __shfl_up_sync(0xffffffffu, is, off);
    if (lane >= off) { ik += tk; ic += tc; is += ts; }
  }
  if (lane == 31) { s_w[0][warp] = ik; s_w[1][warp] = ic; s_w[2][warp] = is; }
  __syncthreads();
  int bk = block_sums[blockIdx.x], bcl = block_sums[nb + blockIdx.x], bs = block_sums[2 * nb + blockIdx.x];
  for (int w = 0; w < warp; ++w) { bk += s_w[0][w]; bcl += s_w[1][w]; bs += s_w[2][w]; }
  int rk = bk + ik - k, rc = bcl + ic - cl, rs = bs + is - sp;
#pragma unroll
  for (int u = 0; u < D_IPT; ++u) {
    if (base + u >= N) break;
    const Cnt3 t = cnt_of(a[u]);
    dst_keep[base + u] = t.keep ? rk : -1;
    dst_new[base + u] = t.clone ? rc : (t.split ? rs : -1);
    rk += t.keep; rc += t.clone; rs += t.split;
  }
}

constexpr int MAX_ARR = 18;
struct MoveArr { const float *src; float *dst; int width; int kind; };   // kind: 0 parameter, 1 Adam moment, 2 xyz / scale
struct MoveArgs { MoveArr a[MAX_ARR]; int n; };

// One launch for every array (blockIdx.y = array): reads are contiguous; survivors land contiguously in order, clones /
// split children in the appended blocks.  Moments of appended Gaussians are zero (trainer.cu:694-746); the positions and
// scales of split children are written by k_apply_split_geometry, so kind 2 arrays skip them here.
__global__ void __launch_bounds__(DB) k_apply_move(const __grid_constant__ MoveArgs A, long long N,
                                                   const uint8_t *__restrict__ action, const int *__restrict__ dst_keep,
                                                   const int *__restrict__ dst_new, int n_keep, int n_clone) {
  const MoveArr ar = A.a[blockIdx.y];
  const int w = ar.width;
  const long long total = N * w;
  for (long long e = (long long)blockIdx.x * DB + threadIdx.x; e < total; e += (long long)gridDim.x * DB) {
    const long long i = e / w;
    const int k = (int)(e - i * w);
    const uint8_t a = action[i];
    if (a == ACT_PRUNE) continue;
    const float v = ar.src[e];
    if (a != ACT_SPLIT) ar.dst[(long long)dst_keep[i] * w + k] = v;
    if (a == ACT_CLONE) ar.dst[((long long)n_keep + dst_new[i]) * w + k] = ar.kind == 1 ? 0.0f : v;
    if (a == ACT_SPLIT && ar.kind != 2) {
      const long long row = (long long)n_keep + n_clone + 2LL * dst_new[i];
      const float o = ar.kind == 1 ? 0.0f : v;
      ar.dst[row * w + k] = o;
      ar.dst[(row + 1) * w + k] = o;
    }
  }
}

__global__ void __launch_bounds__(DB) k_apply_split_geometry(int N, float factor, unsigned long long seed,
                                                             const uint8_t *__restrict__ action, const int *__restrict__ dst_new,
                                                             int n_keep, int n_clone, const float *__restrict__ xyz_in,
                                                             const float *__restrict__ scale_in, const float *__restrict__ quat_in,
                                                             float *__restrict__ xyz_out, float *__restrict__ scale_out) {
  const int i = blockIdx.x * DB + threadIdx.x;
  if (i >= N || action[i] != ACT_SPLIT) return;
  const float xyz[3] = {xyz_in[3 * (size_t)i], xyz_in[3 * (size_t)i + 1], xyz_in[3 * (size_t)i + 2]};
  const float sc[3] = {scale_in[3 * (size_t)i], scale_in[3 * (size_t)i + 1], scale_in[3 * (size_t)i + 2]};
  const float q[4] = {quat_in[4 * (size_t)i], quat_in[4 * (size_t)i + 1], quat_in[4 * (size_t)i + 2], quat_in[4 * (size_t)i + 3]};
  float cx[2][3], cs[3];
  split_children(seed, (uint32_t)i, xyz, sc, q, factor, cx, cs);
  const size_t row = (size_t)n_keep + n_clone + 2 * (size_t)dst_new[i];
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) { xyz_out[(row + j) * 3 + k] = cx[j][k]; scale_out[(row + j) * 3 + k] = cs[k]; }
}

struct DensLayout { size_t action, dst_keep, dst_new, block_sums, counts, bytes; int nb; };
DensLayout dens_layout(int N) {
  DensLayout L;
  L.nb = std::max(1, ceil_div(N, D_TILE));
  size_t off = 0;
  auto take = [&](size_t b) { size_t o = off; off += align_up(std::max<size_t>(b, 16), 256); return o; };
  L.action = take((size_t)N);
  L.dst_keep = take(sizeof(int) * (size_t)N);
  L.dst_new = take(sizeof(int) * (size_t)N);
  L.block_sums = take(sizeof(int) * 3 * (size_t)L.nb);
  L.counts = take(sizeof(int) * 4);
  L.bytes = off;
  return L;
}

}  // namespace
}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_clone_gaussians(int N, int num_sh_coef, const uint8_t *mask, const int *write_ids, const float *xyz_in,
                        const float *rgb_in, const float *op_in, const float *scale_in, const float *quat_in,
                        const float *sh_in, float *xyz_out, float *rgb_out, float *op_out, float *scale_out,
                        float *quat_out, float *sh_out, void *stream) {
  GSX_REQUIRE_DEV(mask); GSX_REQUIRE_DEV(write_ids);   // what the reference asserts (adaptive_density.cu:161-162)
  if (N <= 0) return GSX_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const float *src[6] = {xyz_in, rgb_in, op_in, scale_in, quat_in, sh_in};
  float *dst[6] = {xyz_out, rgb_out, op_out, scale_out, quat_out, sh_out};
  const int width[6] = {3, 3, 1, 3, 4, 3 * std::max(num_sh_coef, 0)};
  for (int a = 0; a < 6; ++a) {
    if (width[a] == 0) continue;
    if (!src[a] || !dst[a]) { set_error("null array %d", a); return GSX_E_NULL; }
    int rc = copy_rows_masked(src[a], mask, write_ids, N, width[a], 1, dst[a], st);
    if (rc != GSX_OK) return rc;
  }
  return GSX_OK;
}

int gsx_split_gaussians(int N, float scale_factor, int num_sh_coef, const uint8_t *mask, const int *write_ids,
                        const float *xyz_in, const float *rgb_in, const float *op_in, const float *scale_in,
                        const float *quat_in, const float *sh_in, float *xyz_out, float *rgb_out, float *op_out,
                        float *scale_out, float *quat_out, float *sh_out, unsigned long long seed, void *stream) {
  GSX_REQUIRE_DEV(mask); GSX_REQUIRE_DEV(write_ids);
  if (N <= 0) return GSX_OK;
  if (!(scale_factor > 0.0f)) { set_error("scale_factor must be positive"); return GSX_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  const float *src[4] = {rgb_in, op_in, quat_in, sh_in};
  float *dst[4] = {rgb_out, op_out, quat_out, sh_out};
  const int width[4] = {3, 1, 4, 3 * std::max(num_sh_coef, 0)};
  for (int a = 0; a < 4; ++a) {
    if (width[a] == 0) continue;
    if (!src[a] || !dst[a]) { set_error("null array %d", a); return GSX_E_NULL; }
    int rc = copy_rows_masked(src[a], mask, write_ids, N, width[a], 2, dst[a], st);
    if (rc != GSX_OK) return rc;
  }
  if (!xyz_in || !scale_in || !quat_in || !xyz_out || !scale_out) { set_error("null geometry array"); return GSX_E_NULL; }
  k_split_geometry<<<ceil_div(N, DB), DB, 0, st>>>(N, scale_factor, seed, mask, write_ids, xyz_in, scale_in, quat_in, xyz_out,
                                                   scale_out);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

size_t gsx_densify_workspace_bytes(int N) { return dens_layout(std::max(N, 1)).bytes; }

int gsx_densify_plan(const gsx_density_config_t *cfg, int N, const float *opacity, const float *scale,
                     const float *uv_grad_accum, const int *grad_accum_dur, void *workspace, size_t workspace_bytes,
                     gsx_density_counts_t *counts_host, void *stream) {
  if (!cfg || !counts_host) { set_error("null config / counts"); return GSX_E_NULL; }
  counts_host->n_in = N; counts_host->n_prune = counts_host->n_clone = counts_host->n_split = 0; counts_host->n_out = N;
  if (N <= 0) return GSX_OK;
  GSX_REQUIRE_DEV(opacity); GSX_REQUIRE_DEV(scale); GSX_REQUIRE_DEV(uv_grad_accum); GSX_REQUIRE_DEV(grad_accum_dur);
  GSX_REQUIRE_DEV(workspace);
  const DensLayout L = dens_layout(N);
  if (workspace_bytes < L.bytes) { set_error("densify workspace %zu < %zu", workspace_bytes, L.bytes); return GSX_E_WORKSPACE; }
  if (!(cfg->delete_opacity_threshold > 0.0f && cfg->delete_opacity_threshold < 1.0f)) { set_error("delete_opacity_threshold must be in (0, 1)"); return GSX_E_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  char *ws = (char *)workspace;
  PlanCfg pc;
  pc.grad_threshold = cfg->uv_grad_threshold;
  pc.op_threshold_logit = logf(cfg->delete_opacity_threshold) - logf(1.0f - cfg->delete_opacity_threshold);   // trainer.cu:544
  pc.max_scale = cfg->scene_extent * 0.1f;                 // trainer.cu:539
  pc.clone_scale_threshold = cfg->scene_extent * 0.01f;    // trainer.cu:540
  uint8_t *action = (uint8_t *)(ws + L.action);
  int *bs = (int *)(ws + L.block_sums), *counts = (int *)(ws + L.counts);
  k_plan_decide<<<L.nb, DB, 0, st>>>(pc, N, opacity, scale, uv_grad_accum, grad_accum_dur, action, bs, L.nb);
  GSX_LAUNCH_CHECK();
  k_plan_scan_blocks<<<1, 1024, 0, st>>>(bs, L.nb, counts);
  GSX_LAUNCH_CHECK();
  k_plan_rows<<<L.nb, DB, 0, st>>>(N, action, bs, L.nb, (int *)(ws + L.dst_keep), (int *)(ws + L.dst_new));
  GSX_LAUNCH_CHECK();
  int h[3] = {0, 0, 0};
  GSX_CUDA(cudaMemcpyAsync(h, counts, sizeof(h), cudaMemcpyDeviceToHost, st));
  GSX_CUDA(cudaStreamSynchronize(st));
  counts_host->n_clone = h[1]; counts_host->n_split = h[2];
  counts_host->n_prune = N - h[0] - h[2];                  // everything that neither survives nor is split
  counts_host->n_out = h[0] + h[1] + 2 * h[2];
  return GSX_OK;
}

int gsx_densify_apply(const gsx_density_config_t *cfg, const gsx_density_counts_t *counts, const gsx_gaussian_set_t *in,
                      const gsx_gaussian_set_t *out, const void *workspace, size_t workspace_bytes, unsigned long long seed,
                      void *stream) {
  if (!cfg || !counts || !in || !out) { set_error("null argument"); return GSX_E_NULL; }
  const int N = counts->n_in;
  if (N <= 0) return GSX_OK;
  if (cfg->max_gaussians > 0 && counts->n_out > cfg->max_gaussians) {   // trainer.cu:579-585: the reference skips the step
    set_error("densification would exceed max_gaussians (%d > %d): skipped", counts->n_out, cfg->max_gaussians);
    return GSX_E_CAPACITY;
  }
  GSX_REQUIRE_DEV(workspace);
  const DensLayout L = dens_layout(N);
  if (workspace_bytes < L.bytes) { set_error("densify workspace %zu < %zu", workspace_bytes, L.bytes); return GSX_E_WORKSPACE; }
  if (!(cfg->split_scale_factor > 0.0f)) { set_error("split_scale_factor must be positive"); return GSX_E_ARG; }
  const char *ws = (const char *)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  const int n_keep = counts->n_out - counts->n_clone - 2 * counts->n_split;
  const int shw = in->sh_width;
  if (shw < 0 || shw > in->sh_stride || in->sh_stride != out->sh_stride) { set_error("bad sh_width / sh_stride"); return GSX_E_ARG; }
  const float *ps[6] = {in->xyz, in->scale, in->quaternion, in->opacity, in->band0, in->sh};
  float *pd[6] = {out->xyz, out->scale, out->quaternion, out->opacity, out->band0, out->sh};
  const int width[6] = {3, 3, 4, 1, 3, in->sh_stride};
  MoveArgs A;
  A.n = 0;
  for (int a = 0; a < 6; ++a) {
    if (width[a] == 0 || (a == 5 && shw == 0)) continue;
    if (!ps[a] || !pd[a]) { set_error("null parameter array %d", a); return GSX_E_NULL; }
    if (ps[a] == pd[a]) { set_error("densify is out of place: array %d aliases its output", a); return GSX_E_ARG; }
    A.a[A.n++] = MoveArr{ps[a], pd[a], width[a], (a == 0 || a == 1) ? 2 : 0};
    const float *ms[2] = {in->exp_avg[a], in->exp_avg_sq[a]};
    float *md[2] = {out->exp_avg[a], out->exp_avg_sq[a]};
    for (int m = 0; m < 2; ++m) {
      if (!ms[m] && !md[m]) continue;                      // moments are optional (a caller without an optimiser)
      if (!ms[m] || !md[m] || ms[m] == md[m]) { set_error("moment array %d/%d null or aliased", a, m); return GSX_E_ARG; }
      A.a[A.n++] = MoveArr{ms[m], md[m], width[a], 1};
    }
  }
  const uint8_t *action = (const uint8_t *)(ws + L.action);
  const int *dst_keep = (const int *)(ws + L.dst_keep), *dst_new = (const int *)(ws + L.dst_new);
  dim3 grid((unsigned)std::min<long long>(kNumSMs * 8, ((long long)N * 4 + DB - 1) / DB), (unsigned)A.n);
  k_apply_move<<<grid, DB, 0, st>>>(A, (long long)N, action, dst_keep, dst_new, n_keep, counts->n_clone);
  GSX_LAUNCH_CHECK();
  if (counts->n_split > 0) {
    k_apply_split_geometry<<<ceil_div(N, DB), DB, 0, st>>>(N, cfg->split_scale_factor, seed, action, dst_new, n_keep,
                                                           counts->n_clone, in->xyz, in->scale, in->quaternion, out->xyz,
                                                           out->scale);
    GSX_LAUNCH_CHECK();
  }
  return GSX_OK;
}

}  // extern "C"
