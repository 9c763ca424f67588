// Drop-in declaration for the reference's Adam entry point, implemented by libgsx.so.
// Signature matches /root/reference/include/gsplat_cuda/optimizer.cuh:27-29; N*S floats are updated in place.
// The fused form over all parameter groups and the cull mask is gsx_adam_step_fused (include/gsx.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

inline constexpr float B1 = 0.9f;
inline constexpr float B2 = 0.999f;
inline constexpr float EPS = 1e-8f;

void adam_step(float *params, float *const param_grads, float *exp_avg, float *exp_avg_sq, const float lr,
               const float b1, const float b2, const float eps, const float bias1, const float bias2, const int N,
               const int S, cudaStream_t stream = 0);
