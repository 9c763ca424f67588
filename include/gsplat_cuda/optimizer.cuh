// Drop-in declaration for the reference's Adam entry point, implemented by libgsx.so
// (3dgs_b200/csrc/compat.cu on gsx_adam_step).  Same name, parameter types and order as
// /root/reference/include/gsplat_cuda/optimizer.cuh:27-29, so cuda/trainer.cu and tests/optimizer_test.cpp link unchanged.
// The fused form -- every parameter group, the cull mask and the densification accumulators in one launch -- is
// gsx_adam_step_fused in include/gsx.h.
#pragma once
#include <cuda_runtime.h>

// Adam hyper-parameters the reference trainer passes (cuda/trainer.cu:1052-1124)
inline constexpr float B1 = 0.9f, B2 = 0.999f, EPS = 1e-8f;

void adam_step(float *params,             // [N*S] updated in place: p -= lr * (m / bias1) / (sqrt(v / bias2) + eps)
               float *const param_grads,  // [N*S] read only; a NaN gradient counts as 0
               float *exp_avg,            // [N*S] first moment m, in place
               float *exp_avg_sq,         // [N*S] second moment v, in place
               const float lr, const float b1, const float b2, const float eps,
               const float bias1,         // 1 - b1^t, computed by the caller
               const float bias2,         // 1 - b2^t
               const int N, const int S,  // N rows of S floats
               cudaStream_t stream = 0);
