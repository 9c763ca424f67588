// preprocess.cu -- fused per-Gaussian stages of the view-level path.
//
// Forward  = F1 camera transform + F2 projection + F3 cull + F4 sigma + F5 Jacobian/conic/radius + F6 SH->rgb +
//            the exact tile count of F7, in ONE pass over the Gaussians (the reference runs 8 kernels, 8 thrust
//            compactions and round-trips xyz_c, uv, sigma, J, dir and Y through HBM; cuda/raster.cu:26-104).
// Backward = F10 SH + F11 conic + F12 Jacobian + F13 sigma + F14 projection + F15 camera transform in one pass
//            (reference: 7 kernels + 8 compactions, cuda/trainer.cu:941-1012), reading the packed [N][12]
//            compositing gradients and writing GLOBAL-index parameter gradients.
// Both are HBM-bound streaming kernels: SH coefficients (180 B/Gaussian at degree 3) are moved through shared
// memory with coalesced 16-byte accesses; everything else is one read/write per Gaussian.
// Algorithmic bytes: forward 12N + 272P, backward 496P at degree 3.
#include "gsx_common.cuh"
#include "gsx_math.cuh"
#include "gsx_tma.cuh"
#include "gsx_view.cuh"

namespace gsx {

__device__ __forceinline__ float2 alpha_extent_pre(float a, float b, float c, float opa) {   // == composite.cu
  if (!(opa >= 0.00392156862f)) return make_float2(-INFINITY, -INFINITY);
  const float det = a * c - b * b;
  if (!(a > 0.0f && c > 0.0f && det > 0.0f) || !(opa <= 1.0f)) return make_float2(INFINITY, INFINITY);
  const float tau2 = 2.0f * __logf(fmaxf(opa * 255.0f, 1.0f)) * 1.001f + 0.02f;
  const float inv = tau2 / det;
  return make_float2(sqrtf(inv * c) * 1.0005f + 0.02f, sqrtf(inv * a) * 1.0005f + 0.02f);
}

constexpr int PB = 256;
#ifndef GSX_PRE_MINB
#define GSX_PRE_MINB 3
#endif

// cooperative, coalesced copy of this block's SH rows into shared memory (rows are contiguous when the row
// stride equals the number of floats used)
__device__ __forceinline__ void load_rows_to_smem(const float *__restrict__ g, float *s, long long row0, int rows,
                                                  int row_floats) {
  const long long f0 = row0 * row_floats;
  const int total = rows * row_floats;
  const float *src = g + f0;
  if ((((uintptr_t)src) & 15) == 0) {
    const int n4 = total >> 2;
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
    float4 *d4 = reinterpret_cast<float4 *>(s);
    for (int i = threadIdx.x; i < n4; i += PB) d4[i] = __ldg(s4 + i);
    for (int i = (n4 << 2) + threadIdx.x; i < total; i += PB) s[i] = __ldg(src + i);
  } else {
    for (int i = threadIdx.x; i < total; i += PB) s[i] = __ldg(src + i);
  }
}

template <int L>
__global__ void __launch_bounds__(PB, GSX_PRE_MINB)
k_preprocess_fwd(ViewConst vc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
                 const float *__restrict__ quat, const float *__restrict__ opacity, const float *__restrict__ band0,
                 const float *__restrict__ sh, int sh_stride, int chunk_ok, float4 *__restrict__ r0,
                 float4 *__restrict__ r1, float4 *__restrict__ r2, float4 *__restrict__ radius,
                 float *__restrict__ depth, uint8_t *__restrict__ mask, int *__restrict__ touched) {
  extern __shared__ __align__(16) float s_sh[];
  __shared__ __align__(8) uint64_t s_bar;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int i = blockIdx.x * PB + threadIdx.x;
  const long long row0 = (long long)blockIdx.x * PB;
  const int rows = (int)min((long long)PB, (long long)N - row0);
  const uint32_t chunk_bytes = (uint32_t)rows * SHF * 4u;
  // SH rows of this block are one contiguous span: stream it with ONE bulk async copy (TMA) while the threads do
  // the geometry; block-uniform decision
  const bool use_tma = (L > 0) && chunk_ok && ((chunk_bytes & 15u) == 0);
  bool keep = false;
  float3 pw = make_float3(0.f, 0.f, 0.f), pc = pw;
  float2 uv = make_float2(0.f, 0.f);
  if (i < N) {
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    uv = screen_point(vc.proj, pc, vc.width, vc.height);
    keep = keep_gaussian(uv.x, uv.y, pc.z, vc.near_thresh, vc.padding, vc.width, vc.height);
  }
  if (use_tma) {
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    const int any = __syncthreads_or(keep ? 1 : 0);
    if (!any) {                       // whole block culled: no SH traffic at all
      if (i < N) { mask[i] = 0; touched[i] = 0; }
      return;
    }
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, chunk_bytes);
      bulk_load(s_sh, sh + row0 * SHF, chunk_bytes, &s_bar);
    }
  }
  float4 rad = make_float4(0.f, 0.f, 0.f, 1.f);
  float con[3] = {0.f, 0.f, 0.f};
  float opa = 0.f;
  float2 ext = make_float2(0.f, 0.f);
  int cnt = 0;
  if (keep) {
    const float4 q = reinterpret_cast<const float4 *>(quat)[i];
    float sig[6], J[6];
    sigma_from_qs(q.x, q.y, q.z, q.w, scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], sig);
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad);
    opa = 1.0f / (1.0f + __expf(-opacity[i]));   // render.cu:56, evaluated once per Gaussian
    ext = alpha_extent_pre(con[0], con[1], con[2], opa);
    // exact tile count (F7 window + SAT)
    const TileWindow w = candidate_window(uv.x, uv.y, rad.x, vc.ntx, vc.nty);
    if (w.x1 > w.x0 && w.y1 > w.y0) {
      const ObbSat so = obb_sat_setup(compute_obb(uv.x, uv.y, rad.x, rad.y, rad.z, rad.w));
      const TileRange tr = sat_aabb_range(so, w);
      for (int tx = tr.x0; tx <= tr.x1; ++tx)
        for (int ty = tr.y0; ty <= tr.y1; ++ty) cnt += sat_tile_rot(so, tx, ty) ? 1 : 0;
    }
    r0[i] = make_float4(uv.x, uv.y, ext.x, ext.y);
    radius[i] = rad;
    depth[i] = pc.z;
    touched[i] = cnt;
    mask[i] = 1;
  } else if (i < N) {
    mask[i] = 0; touched[i] = 0;
  }
  // thread 0 always waits so the shared buffer outlives the copy
  if (use_tma && (keep || threadIdx.x == 0)) mbar_wait(&s_bar, 0);
  if (!keep) return;
  const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
  const float *shp = nullptr;
  if (L > 0) shp = use_tma ? (s_sh + threadIdx.x * SHF) : (sh + (size_t)i * sh_stride);
  const float3 rgb = sh_to_rgb<L>(pw.x, pw.y, pw.z, vc.campos, b0, shp);
  r1[i] = make_float4(con[0], con[1], con[2], opa);
  r2[i] = make_float4(rgb.x, rgb.y, rgb.z, opa * (1.0f - opa));
}

struct GradOut {
  float *xyz, *scale, *quat, *opacity, *band0, *sh, *uv;
  int sh_stride;
  int overwrite;   // 1: '=' semantics for every array (culled Gaussians get zeros); 0: '+=' (multi-view batches)
  int compat_dsph;
  int chunk_ok;    // SH rows (params and grads) are dense and 16-byte aligned -> bulk async copies
};

template <int L>
__global__ void __launch_bounds__(PB, GSX_PRE_MINB)
k_preprocess_bwd(ViewConst vc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
                 const float *__restrict__ quat, const float *__restrict__ band0, const float *sh,
                 int sh_stride, const float4 *__restrict__ r1, const uint8_t *__restrict__ mask,
                 float4 *__restrict__ gacc /*[N][3] float4, zeroed after use*/, GradOut out) {
  extern __shared__ __align__(16) float s_sh[];
  __shared__ __align__(8) uint64_t s_bar;
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int i = blockIdx.x * PB + threadIdx.x;
  const bool keep = (i < N) && mask[i];
  const long long row0 = (long long)blockIdx.x * PB;
  const int rows = (int)min((long long)PB, (long long)N - row0);
  const uint32_t chunk_bytes = (uint32_t)rows * SHF * 4u;
  const bool use_tma = (L > 0) && out.chunk_ok && ((chunk_bytes & 15u) == 0);
  if (use_tma) {
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    const int any = __syncthreads_or(keep ? 1 : 0);
    if (!any) {
      if (out.overwrite) {   // whole block culled: zero its gradient rows
        float *dst = out.sh + row0 * SHF;
        for (int k = threadIdx.x; k < rows * SHF; k += PB) dst[k] = 0.0f;
        if (i < N) {
          out.xyz[3 * i] = 0.f; out.xyz[3 * i + 1] = 0.f; out.xyz[3 * i + 2] = 0.f;
          out.scale[3 * i] = 0.f; out.scale[3 * i + 1] = 0.f; out.scale[3 * i + 2] = 0.f;
          reinterpret_cast<float4 *>(out.quat)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
          out.opacity[i] = 0.f;
          out.band0[3 * i] = 0.f; out.band0[3 * i + 1] = 0.f; out.band0[3 * i + 2] = 0.f;
          if (out.uv) { out.uv[2 * i] = 0.f; out.uv[2 * i + 1] = 0.f; }
        }
      }
      return;
    }
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, chunk_bytes);
      bulk_load(s_sh, sh + row0 * SHF, chunk_bytes, &s_bar);
    }
  }
  float3 gxyz = make_float3(0.f, 0.f, 0.f), pw = gxyz;
  float gs[3] = {0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f}, gb0[3] = {0.f, 0.f, 0.f}, g_rgb[3] = {0.f, 0.f, 0.f};
  float gopa = 0.0f, gu = 0.0f, gv = 0.0f;
  if (keep) {
    const float4 a0 = gacc[3 * (size_t)i], a1 = gacc[3 * (size_t)i + 1], a2 = gacc[3 * (size_t)i + 2];
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    const float4 q = reinterpret_cast<const float4 *>(quat)[i];
    const float s0 = scale[3 * i], s1 = scale[3 * i + 1], s2 = scale[3 * i + 2];
    const float4 c4 = r1[i];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    gacc[3 * (size_t)i] = z4; gacc[3 * (size_t)i + 1] = z4; gacc[3 * (size_t)i + 2] = z4;
    g_rgb[0] = a0.x; g_rgb[1] = a0.y; g_rgb[2] = a0.z;
    gopa = a0.w;
    gu = a1.x; gv = a1.y;
    const float g_con[3] = {a1.z, a1.w, a2.x};
    const float3 pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    float sig[6], J[6];
    sigma_from_qs(q.x, q.y, q.z, q.w, s0, s1, s2, sig);
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    const float con[3] = {c4.x, c4.y, c4.z};
    // F11 (trainer.cu:986)
    float dS[6], dJ[6];
    conic_backward(J, sig, vc.view, con, g_con, dS, dJ);
    // F12 (trainer.cu:996) then F14 (trainer.cu:1005) accumulate into the xyz_c gradient in that order
    float3 gc = make_float3(0.f, 0.f, 0.f), t;
    if (jacobian_backward(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, dJ, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }
    // F13 (trainer.cu:1000)
    sigma_backward(q.x, q.y, q.z, q.w, s0, s1, s2, dS, gq, gs);
    if (screen_backward(vc.proj, pc.x, pc.y, pc.z, gu, gv, vc.width, vc.height, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }
    // F15 (trainer.cu:1009); the SH term of d xyz_w (F10, trainer.cu:979) is added below in the reference's order
    gxyz = cam_point_backward(vc.view, gc.x, gc.y, gc.z);
  }
  if (use_tma) {
    mbar_wait(&s_bar, 0);
    float *row = s_sh + threadIdx.x * SHF;
    if (keep) {
      const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
      // coefficients are read from the row and replaced in place by their gradients
      const float3 gsh = sh_backward<L>(pw.x, pw.y, pw.z, vc.campos, b0, row, g_rgb[0], g_rgb[1], g_rgb[2],
                                        out.compat_dsph != 0, row, gb0, false);
      gxyz = make_float3(gsh.x + gxyz.x, gsh.y + gxyz.y, gsh.z + gxyz.z);
    } else if (i < N) {
#pragma unroll
      for (int k = 0; k < SHF; ++k) row[k] = 0.0f;
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x == 0) {
      float *dst = out.sh + row0 * SHF;
      if (out.overwrite) bulk_store(dst, s_sh, chunk_bytes);
      else bulk_reduce_add_f32(dst, s_sh, chunk_bytes);
      bulk_commit();
    }
  } else if (keep) {
    float shg[SHF > 0 ? SHF : 1];
    const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
    const float3 gsh = sh_backward<L>(pw.x, pw.y, pw.z, vc.campos, b0, L > 0 ? sh + (size_t)i * sh_stride : nullptr,
                                      g_rgb[0], g_rgb[1], g_rgb[2], out.compat_dsph != 0, shg, gb0, false);
    gxyz = make_float3(gsh.x + gxyz.x, gsh.y + gxyz.y, gsh.z + gxyz.z);
    if (L > 0) {
      float *o = out.sh + (size_t)i * out.sh_stride;
#pragma unroll
      for (int k = 0; k < SHF; ++k) o[k] = out.overwrite ? shg[k] : o[k] + shg[k];
    }
  } else if (L > 0 && i < N && out.overwrite) {
    float *o = out.sh + (size_t)i * out.sh_stride;
#pragma unroll
    for (int k = 0; k < SHF; ++k) o[k] = 0.0f;
  }
  if (i < N && (keep || out.overwrite)) {
    if (out.overwrite) {
      out.xyz[3 * i] = gxyz.x; out.xyz[3 * i + 1] = gxyz.y; out.xyz[3 * i + 2] = gxyz.z;
      out.scale[3 * i] = gs[0]; out.scale[3 * i + 1] = gs[1]; out.scale[3 * i + 2] = gs[2];
      reinterpret_cast<float4 *>(out.quat)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
      out.opacity[i] = gopa;
      out.band0[3 * i] = gb0[0]; out.band0[3 * i + 1] = gb0[1]; out.band0[3 * i + 2] = gb0[2];
      if (out.uv) { out.uv[2 * i] = gu; out.uv[2 * i + 1] = gv; }
    } else {
      // load every old value first, then store (independent round trips)
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
  if (use_tma && threadIdx.x == 0) bulk_wait_read_all();   // shared memory must outlive the bulk store
}

static int sh_chunk_ok(int l_max, int stride, const void *a, const void *b) {
  const int shf = ((l_max + 1) * (l_max + 1) - 1) * 3;
  return l_max > 0 && stride == shf && (((uintptr_t)a) & 15) == 0 && (((uintptr_t)b) & 15) == 0;
}
static size_t sh_smem_bytes(int l_max, int ok) {
  const int shf = ((l_max + 1) * (l_max + 1) - 1) * 3;
  return ok ? (size_t)PB * shf * sizeof(float) : 0;
}

int preprocess_fwd_launch(const gsx_view_t &v, const gsx_params_t &p, int N, float4 *r0, float4 *r1, float4 *r2,
                          float4 *radius, float *depth, uint8_t *mask, int *touched, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  const ViewConst vc = make_view_const(v);
  const int grid = ceil_div(N, PB);
  const int ok = sh_chunk_ok(v.l_max, p.sh_stride, p.sh, nullptr);
  const size_t sm = sh_smem_bytes(v.l_max, ok);
#define GSX_FWD(L)                                                                                                   \
  k_preprocess_fwd<L><<<grid, PB, sm, st>>>(vc, N, p.xyz, p.scale, p.quaternion, p.opacity, p.band0, p.sh, p.sh_stride, \
                                            ok, r0, r1, r2, radius, depth, mask, touched)
  switch (v.l_max) {
    case 0: GSX_FWD(0); break;
    case 1: GSX_FWD(1); break;
    case 2: GSX_FWD(2); break;
    case 3: GSX_FWD(3); break;
    default: set_error("l_max %d out of range 0..3", v.l_max); return GSX_E_ARG;
  }
#undef GSX_FWD
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int preprocess_bwd_launch(const gsx_view_t &v, const gsx_params_t &p, int N, const float4 *r1, const uint8_t *mask,
                          float4 *gacc, const gsx_grads_t &g, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  const ViewConst vc = make_view_const(v);
  GradOut o;
  o.xyz = g.xyz; o.scale = g.scale; o.quat = g.quaternion; o.opacity = g.opacity; o.band0 = g.band0; o.sh = g.sh;
  o.uv = g.uv;
  o.sh_stride = p.sh_stride;
  o.overwrite = (v.flags & GSX_GRAD_OVERWRITE) ? 1 : 0;
  o.compat_dsph = (v.flags & GSX_SH_DSPH_CORRECT) ? 0 : 1;
  const int grid = ceil_div(N, PB);
  o.chunk_ok = sh_chunk_ok(v.l_max, p.sh_stride, p.sh, g.sh);
  const size_t sm = sh_smem_bytes(v.l_max, o.chunk_ok);
#define GSX_BWD(L)                                                                                                  \
  k_preprocess_bwd<L><<<grid, PB, sm, st>>>(vc, N, p.xyz, p.scale, p.quaternion, p.band0, p.sh, p.sh_stride, r1, mask, \
                                            gacc, o)
  switch (v.l_max) {
    case 0: GSX_BWD(0); break;
    case 1: GSX_BWD(1); break;
    case 2: GSX_BWD(2); break;
    case 3: GSX_BWD(3); break;
    default: set_error("l_max %d out of range 0..3", v.l_max); return GSX_E_ARG;
  }
#undef GSX_BWD
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // namespace gsx
