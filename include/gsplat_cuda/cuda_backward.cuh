// Drop-in declarations for the backward half of the reference's gsplat_cuda library, implemented by libgsx.so.
// Signatures match /root/reference/include/gsplat_cuda/cuda_backward.cuh:21-123.  Accumulation semantics as the
// reference: F9, F11, F12, F14, F15 and the xyz output of F10 add into caller-zeroed buffers; F13 and the
// coefficient outputs of F10 overwrite.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

inline constexpr int TILE_SIZE_BWD = 16;

// F14
void project_to_screen_backward(const float *const xyz_c, const float *const proj, const float *const uv_grad_out,
                                const int N, const int width, const int height, float *xyz_c_grad_in,
                                cudaStream_t stream = 0);
// F15
void compute_camera_space_points_backward(const float *const xyz_w, const float *const view,
                                          const float *const xyz_c_grad_out, const int N, float *xyz_w_grad_in,
                                          cudaStream_t stream = 0);
// F12
void compute_projection_jacobian_backward(const float *const xyz_c, const float focal_x, const float focal_y,
                                          const float tan_fovx, const float tan_fovy, const float *const J_grad_out,
                                          const int N, float *xyz_c_grad_in, cudaStream_t stream = 0);
// F11
void compute_conic_backward(const float *const J, const float *const sigma, const float *const view,
                            const float *const conic, const float *const conic_grad_out, const int N, float *J_grad_in,
                            float *sigma_grad_in, cudaStream_t stream = 0);
// F13
void compute_sigma_backward(const float *const quaternion, const float *const scale, const float *const sigma_grad_out,
                            const int N, float *quaternion_grad_in, float *scale_grad_in, cudaStream_t stream = 0);
// F10
void precompute_spherical_harmonics_backward(const float *const xyz_c, const float *const rgb_vals,
                                             const float *const sh_coeffs, const float3 campos,
                                             const float *const rgb_grad_out, const int l_max, const int N,
                                             float *sh_grad_in, float *sh_grad_band_0_in, float *xyz_c_grad_in,
                                             cudaStream_t stream = 0);
// F9
void render_image_backward(const float *const uvs, const float *const opacity, const float *const conic,
                           const float *const rgb, const float background_opacity, const int *const sorted_splats,
                           const int *const splat_range_by_tile, const int *const num_splats_per_pixel,
                           const float *const final_weight_per_pixel, const float *const grad_image,
                           const int image_width, const int image_height, float *grad_rgb, float *grad_opacity,
                           float *grad_uv, float *grad_conic, cudaStream_t stream = 0);
