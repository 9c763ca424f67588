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
#include "gsx_common.cuh"
#include "gsx_math.cuh"
#include "gsx_tma.cuh"
#include "gsx_view.cuh"

namespace gsx {

namespace {

constexpr int BB = 128;
#ifndef GSX_BATCH_MINB
#define GSX_BATCH_MINB 4
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

template <int L>
__global__ void __launch_bounds__(BB, GSX_BATCH_MINB)
k_batch_fwd(const __grid_constant__ BatchConst bc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
            const float *__restrict__ quat, const float *__restrict__ opacity, const float *__restrict__ band0,
            const float *__restrict__ sh, int sh_stride, int chunk_ok) {
  extern __shared__ __align__(16) float s_sh[];
  __shared__ __align__(8) uint64_t s_bar;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int i = blockIdx.x * BB + threadIdx.x;
  const long long row0 = (long long)blockIdx.x * BB;
  const int rows = (int)min((long long)BB, (long long)N - row0);
  const bool use_tma = (L > 0) && chunk_ok && (((rows * SHF * 4) & 15) == 0);
  const int nv = bc.nv;
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
  if (use_tma && threadIdx.x == 0) mbar_init(&s_bar, 1);
  const int any = __syncthreads_or(vis != 0);
  if (!any) {   // whole block culled in every view: no parameter traffic beyond xyz
    if (i < N)
      for (int v = 0; v < nv; ++v) { bc.s[v].mask[i] = 0; bc.s[v].touched[i] = 0; }
    return;
  }
  if (L > 0) stage_rows<SHF>(s_sh, sh, sh_stride, row0, rows, use_tma, &s_bar);
  float sig[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, b0[3] = {0.f, 0.f, 0.f};
  float opa = 0.f;
  if (vis) {
    const float4 q = reinterpret_cast<const float4 *>(quat)[i];
    sigma_from_qs(q.x, q.y, q.z, q.w, scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], sig);
    opa = 1.0f / (1.0f + __expf(-opacity[i]));   // render.cu:56, once per Gaussian
    b0[0] = band0[3 * i]; b0[1] = band0[3 * i + 1]; b0[2] = band0[3 * i + 2];
  }
  if (L > 0) {
    if (use_tma) mbar_wait(&s_bar, 0);
    else __syncthreads();
  }
  if (i >= N) return;
  const float *shp = L > 0 ? s_sh + threadIdx.x * SHF : nullptr;
  for (int v = 0; v < nv; ++v) {
    const SlotPtrs &sp = bc.s[v];
    if (!((vis >> v) & 1u)) { sp.mask[i] = 0; sp.touched[i] = 0; continue; }
    const ViewConst &vc = bc.v[v];
    const float3 pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    const float2 uv = screen_point(vc.proj, pc, vc.width, vc.height);
    float J[6], con[3];
    float4 rad;
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad);
    const float2 ext = alpha_extent_b(con[0], con[1], con[2], opa);
    int cnt = 0;
    const TileWindow w = candidate_window(uv.x, uv.y, rad.x, vc.ntx, vc.nty);
    if (w.x1 > w.x0 && w.y1 > w.y0) {
      const ObbSat so = obb_sat_setup(compute_obb(uv.x, uv.y, rad.x, rad.y, rad.z, rad.w));
      const TileRange tr = sat_aabb_range(so, w);
      for (int tx = tr.x0; tx <= tr.x1; ++tx)
        for (int ty = tr.y0; ty <= tr.y1; ++ty) cnt += sat_tile_rot(so, tx, ty) ? 1 : 0;
    }
    const float3 rgb = sh_to_rgb<L>(pw.x, pw.y, pw.z, vc.campos, b0, shp);
    sp.r0[i] = make_float4(uv.x, uv.y, ext.x, ext.y);
    sp.r1[i] = make_float4(con[0], con[1], con[2], opa);
    sp.r2[i] = make_float4(rgb.x, rgb.y, rgb.z, opa * (1.0f - opa));
    sp.radius[i] = rad;
    sp.depth[i] = pc.z;
    sp.touched[i] = cnt;
    sp.mask[i] = 1;
  }
}

struct BatchGradOut {
  float *xyz, *scale, *quat, *opacity, *band0, *sh, *uv;
  int sh_stride;
  int overwrite;     // 1: '=' (Gaussians without a gradient get zeros); 0: '+='
  int compat_dsph;   // Q12
  int chunk_ok;      // SH rows of params and grads dense + 16-byte aligned -> bulk async copies
};

template <int L>
__global__ void __launch_bounds__(BB, GSX_BATCH_MINB)
k_batch_bwd(const __grid_constant__ BatchConst bc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
            const float *__restrict__ quat, const float *__restrict__ band0, const float *__restrict__ sh, int sh_stride,
            const BatchGradOut out) {
  extern __shared__ __align__(16) float s_mem[];
  __shared__ __align__(8) uint64_t s_bar;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  float *s_coef = s_mem, *s_grad = s_mem + BB * SHF;
  const int i = blockIdx.x * BB + threadIdx.x;
  const long long row0 = (long long)blockIdx.x * BB;
  const int rows = (int)min((long long)BB, (long long)N - row0);
  const bool use_tma = (L > 0) && out.chunk_ok && (((rows * SHF * 4) & 15) == 0);
  const int nv = bc.nv;
  unsigned vis = 0;
  if (i < N)
    for (int v = 0; v < nv; ++v) vis |= (bc.s[v].touched[i] > 0 ? 1u : 0u) << v;
  if (use_tma && threadIdx.x == 0) mbar_init(&s_bar, 1);
  const int any = __syncthreads_or(vis != 0);
  if (!any && !out.overwrite) return;
  if (any && L > 0) stage_rows<SHF>(s_coef, sh, sh_stride, row0, rows, use_tma, &s_bar);
  float *grow = s_grad + threadIdx.x * SHF;
  if (L > 0) {
#pragma unroll
    for (int k = 0; k < SHF; ++k) grow[k] = 0.0f;
  }
  float3 gxyz = make_float3(0.f, 0.f, 0.f), pw = gxyz;
  float gs[3] = {0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f}, gb0[3] = {0.f, 0.f, 0.f};
  float dSt[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float gopa = 0.0f, gu = 0.0f, gv = 0.0f;
  float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, sig[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, b0[3] = {0.f, 0.f, 0.f};
  if (vis) {
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    q = reinterpret_cast<const float4 *>(quat)[i];
    s0 = scale[3 * i]; s1 = scale[3 * i + 1]; s2 = scale[3 * i + 2];
    b0[0] = band0[3 * i]; b0[1] = band0[3 * i + 1]; b0[2] = band0[3 * i + 2];
    sigma_from_qs(q.x, q.y, q.z, q.w, s0, s1, s2, sig);
  }
  if (any && L > 0) {
    if (use_tma) mbar_wait(&s_bar, 0);
    else __syncthreads();
  }
  const float *crow = s_coef + threadIdx.x * SHF;
  for (int v = 0; v < nv; ++v) {
    if (!((vis >> v) & 1u)) continue;
    const ViewConst &vc = bc.v[v];
    float4 *ga = bc.s[v].gacc + 3 * (size_t)i;
    const float4 a0 = ga[0], a1 = ga[1], a2 = ga[2];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    ga[0] = z4; ga[1] = z4; ga[2] = z4;   // the slot's packed gradients are zero again for the next batch
    const float g_con[3] = {a1.z, a1.w, a2.x};
    const float3 pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    float J[6], con[3], dS[6], dJ[6];
    float4 rad_unused;
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad_unused);   // same arithmetic as the forward: same bits
    conic_backward(J, sig, vc.view, con, g_con, dS, dJ);                // F11
    float3 gc = make_float3(0.f, 0.f, 0.f), t;
    if (jacobian_backward(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, dJ, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }   // F12
    if (screen_backward(vc.proj, pc.x, pc.y, pc.z, a1.x, a1.y, vc.width, vc.height, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }        // F14
    const float3 gw = cam_point_backward(vc.view, gc.x, gc.y, gc.z);   // F15
    const float3 gsh = sh_backward<L>(pw.x, pw.y, pw.z, vc.campos, b0, crow, a0.x, a0.y, a0.z, out.compat_dsph != 0, grow,
                                      gb0, true);                      // F10
    gxyz.x += gsh.x + gw.x; gxyz.y += gsh.y + gw.y; gxyz.z += gsh.z + gw.z;
#pragma unroll
    for (int k = 0; k < 6; ++k) dSt[k] += dS[k];
    gopa += a0.w; gu += a1.x; gv += a1.y;
  }
  if (vis) sigma_backward(q.x, q.y, q.z, q.w, s0, s1, s2, dSt, gq, gs);   // F13, once on the view-summed dSigma
  if (i < N && (vis || out.overwrite)) {
    if (out.overwrite) {
      out.xyz[3 * i] = gxyz.x; out.xyz[3 * i + 1] = gxyz.y; out.xyz[3 * i + 2] = gxyz.z;
      out.scale[3 * i] = gs[0]; out.scale[3 * i + 1] = gs[1]; out.scale[3 * i + 2] = gs[2];
      reinterpret_cast<float4 *>(out.quat)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
      out.opacity[i] = gopa;
      out.band0[3 * i] = gb0[0]; out.band0[3 * i + 1] = gb0[1]; out.band0[3 * i + 2] = gb0[2];
      if (out.uv) { out.uv[2 * i] = gu; out.uv[2 * i + 1] = gv; }
    } else {
      const float ox0 = out.xyz[3 * i], ox1 = out.xyz[3 * i + 1], ox2 = out.xyz[3 * i + 2];
      const float os0 = out.scale[3 * i], os1 = out.scale[3 * i + 1], os2 = out.scale[3 * i + 2];
      float4 *qq = reinterpret_cast<float4 *>(out.quat) + i;
      float4 c = *qq;
      const float oo = out.opacity[i];
      const float ob0 = out.band0[3 * i], ob1 = out.band0[3 * i + 1], ob2 = out.band0[3 * i + 2];
      float ou = 0.f, ov = 0.f;
      if (out.uv) { ou = out.uv[2 * i]; ov = out.uv[2 * i + 1]; }
      out.xyz[3 * i] = ox0 + gxyz.x; out.xyz[3 * i + 1] = ox1 + gxyz.y; out.xyz[3 * i + 2] = ox2 + gxyz.z;
      out.scale[3 * i] = os0 + gs[0]; out.scale[3 * i + 1] = os1 + gs[1]; out.scale[3 * i + 2] = os2 + gs[2];
      c.x += gq[0]; c.y += gq[1]; c.z += gq[2]; c.w += gq[3];
      *qq = c;
      out.opacity[i] = oo + gopa;
      out.band0[3 * i] = ob0 + gb0[0]; out.band0[3 * i + 1] = ob1 + gb0[1]; out.band0[3 * i + 2] = ob2 + gb0[2];
      if (out.uv) { out.uv[2 * i] = ou + gu; out.uv[2 * i + 1] = ov + gv; }
    }
  }
  if (L > 0) {
    if (use_tma) {
      fence_proxy_async_smem();
      __syncthreads();
      if (threadIdx.x == 0) {
        float *dst = out.sh + row0 * SHF;
        const uint32_t bytes = (uint32_t)rows * SHF * 4u;
        if (out.overwrite) bulk_store(dst, s_grad, bytes);
        else bulk_reduce_add_f32(dst, s_grad, bytes);
        bulk_commit();
        bulk_wait_read_all();   // shared memory must outlive the bulk store
      }
    } else {
      __syncthreads();
      for (int k = threadIdx.x; k < rows * SHF; k += BB) {
        const int r = k / (SHF > 0 ? SHF : 1), c = k - r * SHF;
        float *o = out.sh + (size_t)(row0 + r) * out.sh_stride + c;
        *o = out.overwrite ? s_grad[k] : *o + s_grad[k];
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
  const size_t sm = (size_t)BB * shf * sizeof(float);
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
  const size_t sm = (size_t)2 * BB * shf * sizeof(float);
  const int grid = ceil_div(N, BB);
#define GSX_BB(L)                                                                                                   \
  do {                                                                                                              \
    GSX_CUDA(cudaFuncSetAttribute(k_batch_bwd<L>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));            \
    k_batch_bwd<L><<<grid, BB, sm, st>>>(bc, N, p.xyz, p.scale, p.quaternion, p.band0, p.sh, p.sh_stride, o);         \
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

}  // namespace gsx
