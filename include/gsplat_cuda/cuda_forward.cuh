// Drop-in declarations for the forward half of the reference's gsplat_cuda library, implemented by libgsx.so
// (3dgs_b200/csrc/compat.cu on top of the C ABI in include/gsx.h).  Signatures match
// /root/reference/include/gsplat_cuda/cuda_forward.cuh:26-131 exactly so tests/cuda_forward_test.cpp,
// cuda/raster.cu and cuda/trainer.cu compile and link unchanged.  Error convention as the reference
// (cuda/checks.cuh): a null / non-device pointer prints to stderr and exits.
// Also provided: fused_loss (the step before the backward; reference cuda_forward.cuh:144).
// and compute_morton_codes (reference cuda_forward.cuh:186).
// Not provided (outside the hot path): compute_psnr, accumulate_gradients (never defined by the reference).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

inline constexpr int TILE_SIZE_FWD = 16;

// F5: J, conic (a,b,c) and float4 {r_major, r_minor, sin, cos} per Gaussian
void compute_conic(float *const xyz, const float *view, float *const sigma, const float focal_x, const float focal_y,
                   const float tan_fovx, const float tan_fovy, const float mh_dist, const int N, float *J, float *conic,
                   float4 *radius, cudaStream_t stream = 0);
// F4: sigma (xx,xy,xz,yy,yz,zz) from quaternion (w,x,y,z) and log-scale
void compute_sigma(float *const quaternion, float *const scale, const int N, float *sigma, cudaStream_t stream = 0);
// F1
void compute_camera_space_points(float *const xyz_w, const float *view, const int N, float *xyz_c,
                                 cudaStream_t stream = 0);
// F2
void project_to_screen(float *const xyz, const float *proj, const int N, const int width, const int height, float *uv,
                       cudaStream_t stream = 0);
// F3: mask true = keep
void cull_gaussians(float *const uv, float *const xyz, const int N, const float near_thresh, const int padding,
                    const int width, const int height, bool *mask, cudaStream_t stream = 0);
// F7: two-phase protocol -- sorted_gaussians == nullptr stores the candidate-pair count in sorted_gaussian_bytes
void get_sorted_gaussian_list(const float *uv, const float *xyz, const float4 *radius, const int n_tiles_x,
                              const int n_tiles_y, const int N, size_t &sorted_gaussian_bytes, int *sorted_gaussians,
                              int *splat_start_end_idx_by_tile_idx, cudaStream_t stream = 0);
// F6: xyz is the world position
void precompute_spherical_harmonics(const float *xyz, const float *sh_coefficients, const float *sh_coeffs_band_0,
                                    const float3 campos, const int l_max, const int N, float *rgb,
                                    cudaStream_t stream = 0);
// F8
void render_image(const float *uv, const float *opacity, const float *conic, const float *rgb,
                  const float background_opacity, const int *sorted_splats, const int *splat_range_by_tile,
                  const int image_width, const int image_height, int *splats_per_pixel, float *weight_per_pixel,
                  float *image, cudaStream_t stream = 0);
// mean L1 + SSIM loss (returned: synchronises) and d loss / d predicted into image_grad
float fused_loss(const float *predicted_data, const float *gt_data, int rows, int cols, const float ssim_weight,
                 float *image_grad, cudaStream_t stream = 0);
// 63-bit Morton codes (21 bits per axis) of xyz[N][3] inside the given bounding box
void compute_morton_codes(const int N, const float *d_xyz, const float x_max, const float y_max, const float z_max,
                          const float x_min, const float y_min, const float z_min, uint64_t *codes,
                          cudaStream_t stream = 0);
