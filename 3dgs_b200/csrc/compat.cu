// compat.cu -- the 15 reference-named C++ entry points (include/gsplat_cuda/cuda_forward.cuh, cuda_backward.cuh)
// and the two steps around the path: adam_step (optimizer.cuh), fused_loss (cuda_forward.cuh:144)
// on top of the gsx C ABI, with the reference's error convention: any failure prints and exit(EXIT_FAILURE)
// (reference cuda/checks.cuh:17-38).
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include "../../include/gsplat_cuda/cuda_backward.cuh"
#include "../../include/gsplat_cuda/cuda_forward.cuh"
#include "../../include/gsplat_cuda/adaptive_density.cuh"
#include "../../include/gsplat_cuda/optimizer.cuh"
#include "../../include/gsx.h"

namespace gsx { int internal_workspace_slot(int which, size_t bytes, char **out); }   // binning.cu

static void gsx_or_die(int rc, const char *fn) {
  if (rc != GSX_OK) {
    fprintf(stderr, "Assertion failed in %s: %s (gsx status %d)\n", fn, gsx_last_error(), rc);
    exit(EXIT_FAILURE);
  }
}
#define DIE_UNLESS(call) gsx_or_die((call), __func__)

void compute_conic(float *const xyz, const float *view, float *const sigma, const float focal_x, const float focal_y,
                   const float tan_fovx, const float tan_fovy, const float mh_dist, const int N, float *J, float *conic,
                   float4 *radius, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_conic(xyz, view, sigma, focal_x, focal_y, tan_fovx, tan_fovy, mh_dist, N, J, conic,
                               reinterpret_cast<float *>(radius), stream));
}
void compute_sigma(float *const quaternion, float *const scale, const int N, float *sigma, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_sigma(quaternion, scale, N, sigma, stream));
}
void compute_camera_space_points(float *const xyz_w, const float *view, const int N, float *xyz_c, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_camera_space_points(xyz_w, view, N, xyz_c, stream));
}
void project_to_screen(float *const xyz, const float *proj, const int N, const int width, const int height, float *uv,
                       cudaStream_t stream) {
  DIE_UNLESS(gsx_project_to_screen(xyz, proj, N, width, height, uv, stream));
}
void cull_gaussians(float *const uv, float *const xyz, const int N, const float near_thresh, const int padding,
                    const int width, const int height, bool *mask, cudaStream_t stream) {
  static_assert(sizeof(bool) == 1, "bool mask is one byte");
  DIE_UNLESS(gsx_cull_gaussians(uv, xyz, N, near_thresh, padding, width, height, reinterpret_cast<uint8_t *>(mask), stream));
}
// Two-phase protocol of the reference (cuda/culling.cu:398-407): phase 1 reports the number of coarse candidate
// pairs (an element count despite the parameter's name); phase 2 receives a buffer of that many ints.
void get_sorted_gaussian_list(const float *uv, const float *xyz, const float4 *radius, const int n_tiles_x,
                              const int n_tiles_y, const int N, size_t &sorted_gaussian_bytes, int *sorted_gaussians,
                              int *splat_start_end_idx_by_tile_idx, cudaStream_t stream) {
  if (sorted_gaussians == nullptr) {
    long long coarse = 0, k = 0;
    DIE_UNLESS(gsx_sorted_list_size(uv, reinterpret_cast<const float *>(radius), n_tiles_x, n_tiles_y, N, &coarse, &k,
                                    stream));
    sorted_gaussian_bytes = (size_t)coarse;
    return;
  }
  DIE_UNLESS(gsx_get_sorted_gaussian_list(uv, xyz, reinterpret_cast<const float *>(radius), n_tiles_x, n_tiles_y, N,
                                          (long long)sorted_gaussian_bytes, sorted_gaussians,
                                          splat_start_end_idx_by_tile_idx, nullptr, nullptr, 0, stream));
}
void precompute_spherical_harmonics(const float *xyz, const float *sh_coefficients, const float *sh_coeffs_band_0,
                                    const float3 campos, const int l_max, const int N, float *rgb, cudaStream_t stream) {
  DIE_UNLESS(gsx_precompute_spherical_harmonics(xyz, sh_coefficients, sh_coeffs_band_0, campos.x, campos.y, campos.z,
                                                l_max, N, rgb, stream));
}
void render_image(const float *uv, const float *opacity, const float *conic, const float *rgb,
                  const float background_opacity, const int *sorted_splats, const int *splat_range_by_tile,
                  const int image_width, const int image_height, int *splats_per_pixel, float *weight_per_pixel,
                  float *image, cudaStream_t stream) {
  DIE_UNLESS(gsx_render_image(uv, opacity, conic, rgb, background_opacity, sorted_splats, splat_range_by_tile,
                              image_width, image_height, splats_per_pixel, weight_per_pixel, image, stream));
}

void project_to_screen_backward(const float *const xyz_c, const float *const proj, const float *const uv_grad_out,
                                const int N, const int width, const int height, float *xyz_c_grad_in,
                                cudaStream_t stream) {
  DIE_UNLESS(gsx_project_to_screen_backward(xyz_c, proj, uv_grad_out, N, width, height, xyz_c_grad_in, stream));
}
void compute_camera_space_points_backward(const float *const xyz_w, const float *const view,
                                          const float *const xyz_c_grad_out, const int N, float *xyz_w_grad_in,
                                          cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_camera_space_points_backward(xyz_w, view, xyz_c_grad_out, N, xyz_w_grad_in, stream));
}
void compute_projection_jacobian_backward(const float *const xyz_c, const float focal_x, const float focal_y,
                                          const float tan_fovx, const float tan_fovy, const float *const J_grad_out,
                                          const int N, float *xyz_c_grad_in, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_projection_jacobian_backward(xyz_c, focal_x, focal_y, tan_fovx, tan_fovy, J_grad_out, N,
                                                      xyz_c_grad_in, stream));
}
void compute_conic_backward(const float *const J, const float *const sigma, const float *const view,
                            const float *const conic, const float *const conic_grad_out, const int N, float *J_grad_in,
                            float *sigma_grad_in, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_conic_backward(J, sigma, view, conic, conic_grad_out, N, J_grad_in, sigma_grad_in, stream));
}
void compute_sigma_backward(const float *const quaternion, const float *const scale, const float *const sigma_grad_out,
                            const int N, float *quaternion_grad_in, float *scale_grad_in, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_sigma_backward(quaternion, scale, sigma_grad_out, N, quaternion_grad_in, scale_grad_in, stream));
}
void precompute_spherical_harmonics_backward(const float *const xyz_c, const float *const rgb_vals,
                                             const float *const sh_coeffs, const float3 campos,
                                             const float *const rgb_grad_out, const int l_max, const int N,
                                             float *sh_grad_in, float *sh_grad_band_0_in, float *xyz_c_grad_in,
                                             cudaStream_t stream) {
  DIE_UNLESS(gsx_precompute_spherical_harmonics_backward(xyz_c, rgb_vals, sh_coeffs, campos.x, campos.y, campos.z,
                                                         rgb_grad_out, l_max, N, 0, sh_grad_in, sh_grad_band_0_in,
                                                         xyz_c_grad_in, stream));
}
void render_image_backward(const float *const uvs, const float *const opacity, const float *const conic,
                           const float *const rgb, const float background_opacity, const int *const sorted_splats,
                           const int *const splat_range_by_tile, const int *const num_splats_per_pixel,
                           const float *const final_weight_per_pixel, const float *const grad_image,
                           const int image_width, const int image_height, float *grad_rgb, float *grad_opacity,
                           float *grad_uv, float *grad_conic, cudaStream_t stream) {
  DIE_UNLESS(gsx_render_image_backward(uvs, opacity, conic, rgb, background_opacity, sorted_splats, splat_range_by_tile,
                                       num_splats_per_pixel, final_weight_per_pixel, grad_image, image_width,
                                       image_height, 0, grad_rgb, grad_opacity, grad_uv, grad_conic, stream));
}

// SURVEY 8f rank 2 (reference cuda/optimizer.cu:30-44)
void adam_step(float *params, float *const param_grads, float *exp_avg, float *exp_avg_sq, const float lr,
               const float b1, const float b2, const float eps, const float bias1, const float bias2, const int N,
               const int S, cudaStream_t stream) {
  DIE_UNLESS(gsx_adam_step(params, param_grads, exp_avg, exp_avg_sq, lr, b1, b2, eps, bias1, bias2, (long long)N * S,
                           stream));
}

// SURVEY 8f rank 3 (reference cuda/loss.cu:440-471): returns the mean loss, i.e. synchronises like the reference does.
// The reference allocates its temporaries per call; the shim keeps one grow-only workspace per device.
float fused_loss(const float *predicted_data, const float *gt_data, int rows, int cols, const float ssim_weight,
                 float *image_grad, cudaStream_t stream) {
  const size_t need = gsx_fused_loss_workspace_bytes(rows, cols);
  char *ws = nullptr;   // per-device, mutex-guarded, grow-only (binning.cu internal_workspace_slot)
  if (gsx::internal_workspace_slot(1, need, &ws) != GSX_OK) {
    fprintf(stderr, "Assertion failed in fused_loss: cannot allocate %zu workspace bytes\n", need);
    exit(EXIT_FAILURE);
  }
  float loss = 0.0f;
  DIE_UNLESS(gsx_fused_loss(predicted_data, gt_data, rows, cols, ssim_weight, image_grad, ws, need, nullptr, &loss,
                            stream));
  return loss;
}

// SURVEY 8f rank 4 (reference cuda/culling.cu:345-359)
void compute_morton_codes(const int N, const float *d_xyz, const float x_max, const float y_max, const float z_max,
                          const float x_min, const float y_min, const float z_min, uint64_t *codes, cudaStream_t stream) {
  DIE_UNLESS(gsx_compute_morton_codes(N, d_xyz, x_max, y_max, z_max, x_min, y_min, z_min, 0, codes, stream));
}

// SURVEY 8f rank 4, second half (reference cuda/adaptive_density.cu:155-206).  The reference seeds its curand states with
// time(NULL) (:200); the shim does the same for the counter-based generator (a call counter is mixed in so that two calls
// within one second do not repeat the stream).
void clone_gaussians(const int N, const int num_sh_coef, const bool *mask, const int *write_ids, const float *xyz_in,
                     const float *rgb_in, const float *op_in, const float *scale_in, const float *quat_in,
                     const float *sh_in, float *xyz_out, float *rgb_out, float *op_out, float *scale_out,
                     float *quat_out, float *sh_out, cudaStream_t stream) {
  DIE_UNLESS(gsx_clone_gaussians(N, num_sh_coef, reinterpret_cast<const uint8_t *>(mask), write_ids, xyz_in, rgb_in, op_in,
                                 scale_in, quat_in, sh_in, xyz_out, rgb_out, op_out, scale_out, quat_out, sh_out, stream));
}
void split_gaussians(const int N, const float scale_factor, const int num_sh_coef, const bool *mask,
                     const int *write_ids, const float *xyz_in, const float *rgb_in, const float *op_in,
                     const float *scale_in, const float *quat_in, const float *sh_in, float *xyz_out, float *rgb_out,
                     float *op_out, float *scale_out, float *quat_out, float *sh_out, cudaStream_t stream) {
  static unsigned long long calls = 0;
  const unsigned long long seed = ((unsigned long long)time(NULL) << 20) ^ (++calls);
  DIE_UNLESS(gsx_split_gaussians(N, scale_factor, num_sh_coef, reinterpret_cast<const uint8_t *>(mask), write_ids, xyz_in,
                                 rgb_in, op_in, scale_in, quat_in, sh_in, xyz_out, rgb_out, op_out, scale_out, quat_out,
                                 sh_out, seed, stream));
}
