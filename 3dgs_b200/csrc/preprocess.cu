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

namespace gsx {

__device__ __forceinline__ float2 alpha_extent_pre(float a, float b, float c, float opa) {   // == composite.cu
  if (!(opa >= 0.00392156862f)) return make_float2(-INFINITY, -INFINITY);
  const float det = a * c - b * b;
  if (!(a > 0.0f && c > 0.0f && det > 0.0f) || !(opa <= 1.0f)) return make_float2(INFINITY, INFINITY);
  const float tau2 = 2.0f * __logf(fmaxf(opa * 255.0f, 1.0f)) * 1.001f + 0.02f;
  const float inv = tau2 / det;
  return make_float2(sqrtf(inv * c) * 1.0005f + 0.02f, sqrtf(inv * a) * 1.0005f + 0.02f);
}

struct ViewConst {     // passed by value (kernel parameter space, no device copy of the camera needed)
  float view[12];
  float proj[16];
  float3 campos;
  float fx, fy, tan_fovx, tan_fovy;
  int width, height, ntx, nty;
  float near_thresh;
  int padding;
  float mh_dist;
};

constexpr int PB = 256;

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
__global__ void __launch_bounds__(PB)
k_preprocess_fwd(ViewConst vc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
                 const float *__restrict__ quat, const float *__restrict__ opacity, const float *__restrict__ band0,
                 const float *__restrict__ sh, int sh_stride, float4 *__restrict__ r0, float4 *__restrict__ r1,
                 float4 *__restrict__ r2, float4 *__restrict__ radius, float *__restrict__ depth,
                 uint8_t *__restrict__ mask, int *__restrict__ touched) {
  extern __shared__ __align__(16) float s_sh[];
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int i = blockIdx.x * PB + threadIdx.x;
  bool keep = false;
  float3 pw = make_float3(0.f, 0.f, 0.f), pc = pw;
  float2 uv = make_float2(0.f, 0.f);
  if (i < N) {
    pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    uv = screen_point(vc.proj, pc, vc.width, vc.height);
    keep = keep_gaussian(uv.x, uv.y, pc.z, vc.near_thresh, vc.padding, vc.width, vc.height);
  }
  const bool chunked = (L > 0) && (sh_stride == SHF);
  if (L > 0) {
    const int any = __syncthreads_or(keep ? 1 : 0);
    if (!any) {
      if (i < N) { mask[i] = 0; touched[i] = 0; }
      return;
    }
    if (chunked) {
      const long long row0 = (long long)blockIdx.x * PB;
      const int rows = (int)min((long long)PB, (long long)N - row0);
      load_rows_to_smem(sh, s_sh, row0, rows, SHF);
      __syncthreads();
    }
  }
  if (i >= N) return;
  if (!keep) { mask[i] = 0; touched[i] = 0; return; }

  const float4 q = reinterpret_cast<const float4 *>(quat)[i];
  float sig[6], J[6], con[3];
  sigma_from_qs(q.x, q.y, q.z, q.w, scale[3 * i], scale[3 * i + 1], scale[3 * i + 2], sig);
  projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
  float4 rad;
  conic_and_radius(J, sig, vc.view, vc.mh_dist, con, &rad);
  const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
  const float *shp = nullptr;
  if (L > 0) shp = chunked ? (s_sh + threadIdx.x * SHF) : (sh + (size_t)i * sh_stride);
  const float3 rgb = sh_to_rgb<L>(pw.x, pw.y, pw.z, vc.campos, b0, shp);
  const float opa = 1.0f / (1.0f + __expf(-opacity[i]));   // render.cu:56, evaluated once per Gaussian
  const float2 ext = alpha_extent_pre(con[0], con[1], con[2], opa);

  // exact tile count (F7 window + SAT)
  int cnt = 0;
  {
    const TileWindow w = candidate_window(uv.x, uv.y, rad.x, vc.ntx, vc.nty);
    if (w.x1 > w.x0 && w.y1 > w.y0) {
      const ObbSat s = obb_sat_setup(compute_obb(uv.x, uv.y, rad.x, rad.y, rad.z, rad.w));
      for (int tx = w.x0; tx < w.x1; ++tx)
        for (int ty = w.y0; ty < w.y1; ++ty) cnt += sat_tile(s, tx, ty) ? 1 : 0;
    }
  }
  r0[i] = make_float4(uv.x, uv.y, ext.x, ext.y);
  r1[i] = make_float4(con[0], con[1], con[2], opa);
  r2[i] = make_float4(rgb.x, rgb.y, rgb.z, opa * (1.0f - opa));
  radius[i] = rad;
  depth[i] = pc.z;
  mask[i] = 1;
  touched[i] = cnt;
}

struct GradOut {
  float *xyz, *scale, *quat, *opacity, *band0, *sh, *uv;
  int sh_stride;
  int overwrite;   // 1: '=' semantics for every array (culled Gaussians get zeros); 0: '+=' (multi-view batches)
  int compat_dsph;
};

template <int L>
__global__ void __launch_bounds__(PB)
k_preprocess_bwd(ViewConst vc, int N, const float *__restrict__ xyz, const float *__restrict__ scale,
                 const float *__restrict__ quat, const float *__restrict__ band0, const float *__restrict__ sh,
                 int sh_stride, const float4 *__restrict__ r1, const uint8_t *__restrict__ mask,
                 float4 *__restrict__ gacc /*[N][3] float4, zeroed after use*/, GradOut out) {
  extern __shared__ __align__(16) float s_sh[];
  constexpr int n = (L + 1) * (L + 1);
  constexpr int SHF = (n - 1) * 3;
  const int i = blockIdx.x * PB + threadIdx.x;
  const bool keep = (i < N) && mask[i];
  const bool chunked = (L > 0) && (sh_stride == SHF) && (out.sh_stride == SHF);
  const long long row0 = (long long)blockIdx.x * PB;
  const int rows = (int)min((long long)PB, (long long)N - row0);
  if (L > 0 && chunked) {
    const int any = __syncthreads_or(keep ? 1 : 0);
    if (!any) {
      if (out.overwrite) {   // whole block culled: zero its SH gradient rows, coalesced
        float *dst = out.sh + row0 * SHF;
        for (int k = threadIdx.x; k < rows * SHF; k += PB) dst[k] = 0.0f;
      }
    } else {
      load_rows_to_smem(sh, s_sh, row0, rows, SHF);
    }
    __syncthreads();
    if (!any) {
      if (i < N && out.overwrite) {
        out.xyz[3 * i] = 0.f; out.xyz[3 * i + 1] = 0.f; out.xyz[3 * i + 2] = 0.f;
        out.scale[3 * i] = 0.f; out.scale[3 * i + 1] = 0.f; out.scale[3 * i + 2] = 0.f;
        reinterpret_cast<float4 *>(out.quat)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        out.opacity[i] = 0.f;
        out.band0[3 * i] = 0.f; out.band0[3 * i + 1] = 0.f; out.band0[3 * i + 2] = 0.f;
        if (out.uv) { out.uv[2 * i] = 0.f; out.uv[2 * i + 1] = 0.f; }
      }
      return;
    }
  }
  float shg[SHF > 0 ? SHF : 1];
#pragma unroll
  for (int k = 0; k < SHF; ++k) shg[k] = 0.0f;
  float3 gxyz = make_float3(0.f, 0.f, 0.f);
  float gs[3] = {0.f, 0.f, 0.f}, gq[4] = {0.f, 0.f, 0.f, 0.f}, gb0[3] = {0.f, 0.f, 0.f};
  float gopa = 0.0f, gu = 0.0f, gv = 0.0f;
  if (keep) {
    const float4 a0 = gacc[3 * (size_t)i], a1 = gacc[3 * (size_t)i + 1], a2 = gacc[3 * (size_t)i + 2];
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    gacc[3 * (size_t)i] = z4; gacc[3 * (size_t)i + 1] = z4; gacc[3 * (size_t)i + 2] = z4;
    const float g_rgb[3] = {a0.x, a0.y, a0.z};
    gopa = a0.w;
    gu = a1.x; gv = a1.y;
    const float g_con[3] = {a1.z, a1.w, a2.x};
    const float3 pw = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    const float3 pc = cam_point(vc.view, pw.x, pw.y, pw.z);
    const float4 q = reinterpret_cast<const float4 *>(quat)[i];
    const float s0 = scale[3 * i], s1 = scale[3 * i + 1], s2 = scale[3 * i + 2];
    float sig[6], J[6];
    sigma_from_qs(q.x, q.y, q.z, q.w, s0, s1, s2, sig);
    projection_jacobian(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, J);
    const float4 c4 = r1[i];
    const float con[3] = {c4.x, c4.y, c4.z};
    // F10 (trainer.cu:979)
    const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
    const float *shp = nullptr;
    if (L > 0) shp = chunked ? (s_sh + threadIdx.x * SHF) : (sh + (size_t)i * sh_stride);
    gxyz = sh_backward<L>(pw.x, pw.y, pw.z, vc.campos, b0, shp, g_rgb[0], g_rgb[1], g_rgb[2], out.compat_dsph != 0, shg,
                          gb0, false);
    // F11 (trainer.cu:986)
    float dS[6], dJ[6];
    conic_backward(J, sig, vc.view, con, g_con, dS, dJ);
    // F12 (trainer.cu:996) then F14 (trainer.cu:1005) accumulate into xyz_c grad in that order
    float3 gc = make_float3(0.f, 0.f, 0.f), t;
    if (jacobian_backward(pc.x, pc.y, pc.z, vc.fx, vc.fy, vc.tan_fovx, vc.tan_fovy, dJ, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }
    // F13 (trainer.cu:1000)
    sigma_backward(q.x, q.y, q.z, q.w, s0, s1, s2, dS, gq, gs);
    if (screen_backward(vc.proj, pc.x, pc.y, pc.z, gu, gv, vc.width, vc.height, &t)) { gc.x += t.x; gc.y += t.y; gc.z += t.z; }
    // F15 (trainer.cu:1009)
    t = cam_point_backward(vc.view, gc.x, gc.y, gc.z);
    gxyz.x += t.x; gxyz.y += t.y; gxyz.z += t.z;
  }
  if (L > 0 && chunked) {
    __syncthreads();   // everyone finished reading the coefficients
    if (i < N) {
#pragma unroll
      for (int k = 0; k < SHF; ++k) s_sh[threadIdx.x * SHF + k] = shg[k];
    }
    __syncthreads();
    float *dst = out.sh + row0 * SHF;
    const int total = rows * SHF;
    if (out.overwrite) {
      for (int k = threadIdx.x; k < total; k += PB) dst[k] = s_sh[k];
    } else {
      for (int k = threadIdx.x; k < total; k += PB) { const float v = s_sh[k]; if (v != 0.0f) dst[k] += v; }
    }
  }
  if (i >= N) return;
  if (!keep && !out.overwrite) return;
  if (L > 0 && !chunked) {
    float *o = out.sh + (size_t)i * out.sh_stride;
    if (out.overwrite) {
#pragma unroll
      for (int k = 0; k < SHF; ++k) o[k] = shg[k];
    } else {
#pragma unroll
      for (int k = 0; k < SHF; ++k) o[k] += shg[k];
    }
  }
  if (out.overwrite) {
    out.xyz[3 * i] = gxyz.x; out.xyz[3 * i + 1] = gxyz.y; out.xyz[3 * i + 2] = gxyz.z;
    out.scale[3 * i] = gs[0]; out.scale[3 * i + 1] = gs[1]; out.scale[3 * i + 2] = gs[2];
    reinterpret_cast<float4 *>(out.quat)[i] = make_float4(gq[0], gq[1], gq[2], gq[3]);
    out.opacity[i] = gopa;
    out.band0[3 * i] = gb0[0]; out.band0[3 * i + 1] = gb0[1]; out.band0[3 * i + 2] = gb0[2];
    if (out.uv) { out.uv[2 * i] = gu; out.uv[2 * i + 1] = gv; }
  } else {
    out.xyz[3 * i] += gxyz.x; out.xyz[3 * i + 1] += gxyz.y; out.xyz[3 * i + 2] += gxyz.z;
    out.scale[3 * i] += gs[0]; out.scale[3 * i + 1] += gs[1]; out.scale[3 * i + 2] += gs[2];
    float4 *qq = reinterpret_cast<float4 *>(out.quat) + i;
    float4 c = *qq;
    c.x += gq[0]; c.y += gq[1]; c.z += gq[2]; c.w += gq[3];
    *qq = c;
    out.opacity[i] += gopa;
    out.band0[3 * i] += gb0[0]; out.band0[3 * i + 1] += gb0[1]; out.band0[3 * i + 2] += gb0[2];
    if (out.uv) { out.uv[2 * i] += gu; out.uv[2 * i + 1] += gv; }
  }
}

static ViewConst make_view_const(const gsx_view_t &v) {
  ViewConst c;
  for (int k = 0; k < 12; ++k) c.view[k] = v.view[k];
  for (int k = 0; k < 16; ++k) c.proj[k] = v.proj[k];
  c.campos = make_float3(v.campos[0], v.campos[1], v.campos[2]);
  c.fx = v.focal_x; c.fy = v.focal_y; c.tan_fovx = v.tan_fovx; c.tan_fovy = v.tan_fovy;
  c.width = v.width; c.height = v.height;
  c.ntx = (v.width + 15) / 16; c.nty = (v.height + 15) / 16;
  c.near_thresh = v.near_thresh; c.padding = v.cull_padding; c.mh_dist = v.mh_dist;
  return c;
}

static size_t sh_smem_bytes(int l_max, int stride) {
  const int shf = ((l_max + 1) * (l_max + 1) - 1) * 3;
  return (l_max > 0 && stride == shf) ? (size_t)PB * shf * sizeof(float) : 0;
}

int preprocess_fwd_launch(const gsx_view_t &v, const gsx_params_t &p, int N, float4 *r0, float4 *r1, float4 *r2,
                          float4 *radius, float *depth, uint8_t *mask, int *touched, cudaStream_t st) {
  if (N <= 0) return GSX_OK;
  const ViewConst vc = make_view_const(v);
  const int grid = ceil_div(N, PB);
  const size_t sm = sh_smem_bytes(v.l_max, p.sh_stride);
#define GSX_FWD(L)                                                                                                   \
  k_preprocess_fwd<L><<<grid, PB, sm, st>>>(vc, N, p.xyz, p.scale, p.quaternion, p.opacity, p.band0, p.sh, p.sh_stride, \
                                            r0, r1, r2, radius, depth, mask, touched)
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
  const size_t sm = sh_smem_bytes(v.l_max, p.sh_stride);
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
