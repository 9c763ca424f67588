// pointwise.cu -- per-function mirrors of the reference's per-Gaussian kernels (F1-F6, F10-F15), exported
// through the C ABI so the reference's gtest fixtures and trainer call sites link unchanged.  The fused
// view-level path (batch.cu) uses the same device math from gsx_math.cuh.
#include "gsx_common.cuh"
#include "gsx_math.cuh"

namespace gsx {
namespace {

constexpr int kBlock = 256;

__global__ void __launch_bounds__(kBlock) k_cam_points(const float *__restrict__ xyz_w, const float *__restrict__ view,
                                                       int N, float *__restrict__ xyz_c) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float v[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) v[k] = __ldg(view + k);
  const float3 c = cam_point(v, xyz_w[3 * i], xyz_w[3 * i + 1], xyz_w[3 * i + 2]);
  xyz_c[3 * i] = c.x; xyz_c[3 * i + 1] = c.y; xyz_c[3 * i + 2] = c.z;
}

__global__ void __launch_bounds__(kBlock) k_project(const float *__restrict__ xyz, const float *__restrict__ proj, int N,
                                                    int width, int height, float *__restrict__ uv) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float p[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) p[k] = __ldg(proj + k);
  const float2 o = screen_point(p, make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), width, height);
  reinterpret_cast<float2 *>(uv)[i] = o;
}

__global__ void __launch_bounds__(kBlock) k_cull(const float *__restrict__ uv, const float *__restrict__ xyz, int N,
                                                 float near_thresh, int padding, int width, int height,
                                                 uint8_t *__restrict__ mask) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  mask[i] = keep_gaussian(uv[2 * i], uv[2 * i + 1], xyz[3 * i + 2], near_thresh, padding, width, height) ? 1 : 0;
}

__global__ void __launch_bounds__(kBlock) k_sigma(const float *__restrict__ q, const float *__restrict__ s, int N,
                                                  float *__restrict__ sigma) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  const float4 qq = reinterpret_cast<const float4 *>(q)[i];
  float sg[6];
  sigma_from_qs(qq.x, qq.y, qq.z, qq.w, s[3 * i], s[3 * i + 1], s[3 * i + 2], sg);
#pragma unroll
  for (int k = 0; k < 6; ++k) sigma[6 * i + k] = sg[k];
}

__global__ void __launch_bounds__(kBlock) k_conic(const float *__restrict__ xyz, const float *__restrict__ view,
                                                  const float *__restrict__ sigma, float fx, float fy, float tan_fovx,
                                                  float tan_fovy, float mh_dist, int N, float *__restrict__ J,
                                                  float *__restrict__ conic, float4 *__restrict__ radius) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float v[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) v[k] = __ldg(view + k);
  float Jl[6], S[6], c[3];
  projection_jacobian(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], fx, fy, tan_fovx, tan_fovy, Jl);
#pragma unroll
  for (int k = 0; k < 6; ++k) { S[k] = sigma[6 * i + k]; J[6 * i + k] = Jl[k]; }
  float4 r;
  conic_and_radius(Jl, S, v, mh_dist, c, &r);
  conic[3 * i] = c[0]; conic[3 * i + 1] = c[1]; conic[3 * i + 2] = c[2];
  radius[i] = r;
}

template <int L>
__global__ void __launch_bounds__(kBlock) k_sh_fwd(const float *__restrict__ xyz, const float *__restrict__ sh,
                                                   const float *__restrict__ band0, float3 campos, int N,
                                                   float *__restrict__ rgb) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  constexpr int n = (L + 1) * (L + 1);
  const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
  const float3 c = sh_to_rgb<L>(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], campos, b0,
                                L > 0 ? sh + (size_t)i * (n - 1) * 3 : nullptr);
  rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
}

template <int L>
__global__ void __launch_bounds__(kBlock) k_sh_bwd(const float *__restrict__ xyz, const float *__restrict__ band0,
                                                   const float *__restrict__ sh, float3 campos,
                                                   const float *__restrict__ rgb_grad, int N, bool compat,
                                                   float *__restrict__ sh_grad, float *__restrict__ b0_grad,
                                                   float *__restrict__ xyz_grad) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  constexpr int n = (L + 1) * (L + 1);
  const float b0[3] = {band0[3 * i], band0[3 * i + 1], band0[3 * i + 2]};
  const float3 g = sh_backward<L>(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], campos, b0,
                                  L > 0 ? sh + (size_t)i * (n - 1) * 3 : nullptr, rgb_grad[3 * i], rgb_grad[3 * i + 1],
                                  rgb_grad[3 * i + 2], compat, L > 0 ? sh_grad + (size_t)i * (n - 1) * 3 : nullptr,
                                  b0_grad + 3 * i, false);
  xyz_grad[3 * i] += g.x; xyz_grad[3 * i + 1] += g.y; xyz_grad[3 * i + 2] += g.z;
}

__global__ void __launch_bounds__(kBlock) k_conic_bwd(const float *__restrict__ J, const float *__restrict__ sigma,
                                                      const float *__restrict__ view, const float *__restrict__ conic,
                                                      const float *__restrict__ gc, int N, float *__restrict__ J_grad,
                                                      float *__restrict__ sigma_grad) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float v[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) v[k] = __ldg(view + k);
  float Jl[6], S[6], dS[6], dJ[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) { Jl[k] = J[6 * i + k]; S[k] = sigma[6 * i + k]; }
  const float c[3] = {conic[3 * i], conic[3 * i + 1], conic[3 * i + 2]};
  const float g[3] = {gc[3 * i], gc[3 * i + 1], gc[3 * i + 2]};
  conic_backward(Jl, S, v, c, g, dS, dJ);
#pragma unroll
  for (int k = 0; k < 6; ++k) { sigma_grad[6 * i + k] += dS[k]; J_grad[6 * i + k] += dJ[k]; }
}

__global__ void __launch_bounds__(kBlock) k_jac_bwd(const float *__restrict__ xyz, float fx, float fy, float tan_fovx,
                                                    float tan_fovy, const float *__restrict__ J_grad, int N,
                                                    float *__restrict__ xyz_grad) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float dJ[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) dJ[k] = J_grad[6 * i + k];
  float3 o;
  if (!jacobian_backward(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], fx, fy, tan_fovx, tan_fovy, dJ, &o)) return;
  xyz_grad[3 * i] += o.x; xyz_grad[3 * i + 1] += o.y; xyz_grad[3 * i + 2] += o.z;
}

__global__ void __launch_bounds__(kBlock) k_sigma_bwd(const float *__restrict__ q, const float *__restrict__ s,
                                                      const float *__restrict__ dSigma, int N, float *__restrict__ dQ,
                                                      float *__restrict__ dS) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  const float4 qq = reinterpret_cast<const float4 *>(q)[i];
  float g[6], oq[4], os[3];
#pragma unroll
  for (int k = 0; k < 6; ++k) g[k] = dSigma[6 * i + k];
  sigma_backward(qq.x, qq.y, qq.z, qq.w, s[3 * i], s[3 * i + 1], s[3 * i + 2], g, oq, os);
  reinterpret_cast<float4 *>(dQ)[i] = make_float4(oq[0], oq[1], oq[2], oq[3]);
  dS[3 * i] = os[0]; dS[3 * i + 1] = os[1]; dS[3 * i + 2] = os[2];
}

__global__ void __launch_bounds__(kBlock) k_project_bwd(const float *__restrict__ xyz, const float *__restrict__ proj,
                                                        const float *__restrict__ uv_grad, int N, int width, int height,
                                                        float *__restrict__ xyz_grad) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float p[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) p[k] = __ldg(proj + k);
  float3 o;
  if (!screen_backward(p, xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], uv_grad[2 * i], uv_grad[2 * i + 1], width, height,
                       &o))
    return;
  xyz_grad[3 * i] += o.x; xyz_grad[3 * i + 1] += o.y; xyz_grad[3 * i + 2] += o.z;
}

__global__ void __launch_bounds__(kBlock) k_cam_points_bwd(const float *__restrict__ view,
                                                           const float *__restrict__ g, int N,
                                                           float *__restrict__ out) {
  const int i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= N) return;
  float v[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) v[k] = __ldg(view + k);
  const float3 o = cam_point_backward(v, g[3 * i], g[3 * i + 1], g[3 * i + 2]);
  out[3 * i] += o.x; out[3 * i + 1] += o.y; out[3 * i + 2] += o.z;
}

inline dim3 grid_for(int N) { return dim3((unsigned)ceil_div(N, kBlock)); }

}  // namespace
}  // namespace gsx

using namespace gsx;

extern "C" {

int gsx_compute_camera_space_points(const float *xyz_w, const float *view, int N, float *xyz_c, void *stream) {
  GSX_REQUIRE_DEV(xyz_w); GSX_REQUIRE_DEV(view); GSX_REQUIRE_DEV(xyz_c);
  if (N <= 0) return GSX_OK;
  k_cam_points<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(xyz_w, view, N, xyz_c);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_project_to_screen(const float *xyz_c, const float *proj, int N, int width, int height, float *uv,
                          void *stream) {
  GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(proj); GSX_REQUIRE_DEV(uv);
  GSX_REQUIRE_ALIGNED(uv, 8);
  if (N <= 0) return GSX_OK;
  k_project<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(xyz_c, proj, N, width, height, uv);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_cull_gaussians(const float *uv, const float *xyz_c, int N, float near_thresh, int padding, int width,
                       int height, uint8_t *mask, void *stream) {
  GSX_REQUIRE_DEV(uv); GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(mask);
  if (N <= 0) return GSX_OK;
  k_cull<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(uv, xyz_c, N, near_thresh, padding, width, height, mask);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_sigma(const float *quaternion, const float *scale, int N, float *sigma, void *stream) {
  GSX_REQUIRE_DEV(quaternion); GSX_REQUIRE_DEV(scale); GSX_REQUIRE_DEV(sigma);
  GSX_REQUIRE_ALIGNED(quaternion, 16);
  if (N <= 0) return GSX_OK;
  k_sigma<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(quaternion, scale, N, sigma);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_conic(const float *xyz_c, const float *view, const float *sigma, float focal_x, float focal_y,
                      float tan_fovx, float tan_fovy, float mh_dist, int N, float *J, float *conic, float *radius,
                      void *stream) {
  GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(view); GSX_REQUIRE_DEV(sigma); GSX_REQUIRE_DEV(J); GSX_REQUIRE_DEV(conic);
  GSX_REQUIRE_DEV(radius);
  GSX_REQUIRE_ALIGNED(radius, 16);
  if (N <= 0) return GSX_OK;
  k_conic<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(xyz_c, view, sigma, focal_x, focal_y, tan_fovx, tan_fovy,
                                                            mh_dist, N, J, conic, reinterpret_cast<float4 *>(radius));
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_precompute_spherical_harmonics(const float *xyz_w, const float *sh, const float *band0, float cx, float cy,
                                       float cz, int l_max, int N, float *rgb, void *stream) {
  GSX_REQUIRE_DEV(xyz_w); GSX_REQUIRE_DEV(band0); GSX_REQUIRE_DEV(rgb);
  if (l_max > 0) GSX_REQUIRE_DEV(sh);   // spherical_harmonics.cu:68-69
  if (l_max < 0 || l_max > 3) { set_error("l_max %d out of range 0..3", l_max); return GSX_E_ARG; }
  if (N <= 0) return GSX_OK;
  const float3 cp = make_float3(cx, cy, cz);
  cudaStream_t st = (cudaStream_t)stream;
  switch (l_max) {
    case 0: k_sh_fwd<0><<<grid_for(N), kBlock, 0, st>>>(xyz_w, sh, band0, cp, N, rgb); break;
    case 1: k_sh_fwd<1><<<grid_for(N), kBlock, 0, st>>>(xyz_w, sh, band0, cp, N, rgb); break;
    case 2: k_sh_fwd<2><<<grid_for(N), kBlock, 0, st>>>(xyz_w, sh, band0, cp, N, rgb); break;
    default: k_sh_fwd<3><<<grid_for(N), kBlock, 0, st>>>(xyz_w, sh, band0, cp, N, rgb); break;
  }
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_precompute_spherical_harmonics_backward(const float *xyz_w, const float *band0, const float *sh, float cx,
                                                float cy, float cz, const float *rgb_grad, int l_max, int N,
                                                int flags, float *sh_grad, float *band0_grad, float *xyz_grad,
                                                void *stream) {
  GSX_REQUIRE_DEV(xyz_w); GSX_REQUIRE_DEV(band0); GSX_REQUIRE_DEV(rgb_grad); GSX_REQUIRE_DEV(band0_grad);
  GSX_REQUIRE_DEV(xyz_grad);
  if (l_max > 0) { GSX_REQUIRE_DEV(sh); GSX_REQUIRE_DEV(sh_grad); }
  if (l_max < 0 || l_max > 3) { set_error("l_max %d out of range 0..3", l_max); return GSX_E_ARG; }
  if (N <= 0) return GSX_OK;
  const float3 cp = make_float3(cx, cy, cz);
  const bool compat = !(flags & GSX_SH_DSPH_CORRECT);
  cudaStream_t st = (cudaStream_t)stream;
  switch (l_max) {
    case 0: k_sh_bwd<0><<<grid_for(N), kBlock, 0, st>>>(xyz_w, band0, sh, cp, rgb_grad, N, compat, sh_grad, band0_grad, xyz_grad); break;
    case 1: k_sh_bwd<1><<<grid_for(N), kBlock, 0, st>>>(xyz_w, band0, sh, cp, rgb_grad, N, compat, sh_grad, band0_grad, xyz_grad); break;
    case 2: k_sh_bwd<2><<<grid_for(N), kBlock, 0, st>>>(xyz_w, band0, sh, cp, rgb_grad, N, compat, sh_grad, band0_grad, xyz_grad); break;
    default: k_sh_bwd<3><<<grid_for(N), kBlock, 0, st>>>(xyz_w, band0, sh, cp, rgb_grad, N, compat, sh_grad, band0_grad, xyz_grad); break;
  }
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_conic_backward(const float *J, const float *sigma, const float *view, const float *conic,
                               const float *conic_grad, int N, float *J_grad, float *sigma_grad, void *stream) {
  GSX_REQUIRE_DEV(J); GSX_REQUIRE_DEV(sigma); GSX_REQUIRE_DEV(view); GSX_REQUIRE_DEV(conic); GSX_REQUIRE_DEV(conic_grad);
  GSX_REQUIRE_DEV(J_grad); GSX_REQUIRE_DEV(sigma_grad);
  if (N <= 0) return GSX_OK;
  k_conic_bwd<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(J, sigma, view, conic, conic_grad, N, J_grad, sigma_grad);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_projection_jacobian_backward(const float *xyz_c, float focal_x, float focal_y, float tan_fovx,
                                             float tan_fovy, const float *J_grad, int N, float *xyz_c_grad,
                                             void *stream) {
  GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(J_grad); GSX_REQUIRE_DEV(xyz_c_grad);
  if (N <= 0) return GSX_OK;
  k_jac_bwd<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(xyz_c, focal_x, focal_y, tan_fovx, tan_fovy, J_grad, N,
                                                              xyz_c_grad);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_sigma_backward(const float *quaternion, const float *scale, const float *sigma_grad, int N,
                               float *quat_grad, float *scale_grad, void *stream) {
  GSX_REQUIRE_DEV(quaternion); GSX_REQUIRE_DEV(scale); GSX_REQUIRE_DEV(sigma_grad); GSX_REQUIRE_DEV(quat_grad);
  GSX_REQUIRE_DEV(scale_grad);
  GSX_REQUIRE_ALIGNED(quaternion, 16); GSX_REQUIRE_ALIGNED(quat_grad, 16);
  if (N <= 0) return GSX_OK;
  k_sigma_bwd<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(quaternion, scale, sigma_grad, N, quat_grad, scale_grad);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_project_to_screen_backward(const float *xyz_c, const float *proj, const float *uv_grad, int N, int width,
                                   int height, float *xyz_c_grad, void *stream) {
  GSX_REQUIRE_DEV(xyz_c); GSX_REQUIRE_DEV(proj); GSX_REQUIRE_DEV(uv_grad); GSX_REQUIRE_DEV(xyz_c_grad);
  if (N <= 0) return GSX_OK;
  k_project_bwd<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(xyz_c, proj, uv_grad, N, width, height, xyz_c_grad);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

int gsx_compute_camera_space_points_backward(const float *xyz_w, const float *view, const float *xyz_c_grad, int N,
                                             float *xyz_w_grad, void *stream) {
  (void)xyz_w;  // unused by the reference kernel too (projection_backward.cu:95-137)
  GSX_REQUIRE_DEV(view); GSX_REQUIRE_DEV(xyz_c_grad); GSX_REQUIRE_DEV(xyz_w_grad);
  if (N <= 0) return GSX_OK;
  k_cam_points_bwd<<<grid_for(N), kBlock, 0, (cudaStream_t)stream>>>(view, xyz_c_grad, N, xyz_w_grad);
  GSX_LAUNCH_CHECK();
  return GSX_OK;
}

}  // extern "C"
