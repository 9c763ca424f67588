// Drop-in declarations for the reference's densification entry points, implemented by libgsx.so
// (3dgs_b200/csrc/compat.cu on gsx_clone_gaussians / gsx_split_gaussians).  Same names, parameter types and order as
// /root/reference/include/gsplat_cuda/adaptive_density.cuh:28-31,54-58, so cuda/trainer.cu:616-652 and
// tests/adaptive_density_test.cpp link unchanged.  The fused step -- decisions, scans and one move of every parameter and
// optimiser array -- is gsx_densify_plan / gsx_densify_apply in include/gsx.h.
#pragma once
#include <cuda_runtime.h>

// Every Gaussian i with mask[i] set is copied to row write_ids[i] of the output buffers (write_ids = exclusive scan of the
// mask, computed by the caller); rows of unselected Gaussians are not touched.  sh arrays hold num_sh_coef * 3 floats per row.
void clone_gaussians(const int N, const int num_sh_coef, const bool *mask, const int *write_ids, const float *xyz_in,
                     const float *rgb_in, const float *op_in, const float *scale_in, const float *quat_in,
                     const float *sh_in, float *xyz_out, float *rgb_out, float *op_out, float *scale_out,
                     float *quat_out, float *sh_out, cudaStream_t stream = 0);

// Every selected Gaussian produces two children at rows 2 * write_ids[i] and 2 * write_ids[i] + 1: centre sampled from the
// parent's own distribution, log-scale lowered by log(scale_factor), everything else copied.  Like the reference, the
// random stream is seeded from the wall clock; gsx_split_gaussians takes the seed as an argument.
void split_gaussians(const int N, const float scale_factor, const int num_sh_coef, const bool *mask,
                     const int *write_ids, const float *xyz_in, const float *rgb_in, const float *op_in,
                     const float *scale_in, const float *quat_in, const float *sh_in, float *xyz_out, float *rgb_out,
                     float *op_out, float *scale_out, float *quat_out, float *sh_out, cudaStream_t stream = 0);
