// profiles/ubench.cu -- issue-rate micro-benchmarks behind the compositing kernels' roofline model (B200, sm_100a):
// FFMA vs packed FFMA2 (fma.rn.f32x2), their mixes with integer ALU / MUFU / SHFL / LDS work, in warp-instructions per SM
// per clock and in TFLOP/s.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o profiles/bin/ubench profiles/ubench.cu
// Run on the box: profiles/bin/ubench > gpurun_out/ubench.json
#include <cuda_runtime.h>
#include <stdio.h>

#define ITERS 2048
template <int MODE>
__global__ void __launch_bounds__(256) k(float *out, int iters) {
  float2 a[4], m = make_float2(0.999f, 0.998f), c = make_float2(1e-4f * blockIdx.x, 2e-4f);
  unsigned x0 = threadIdx.x, x1 = threadIdx.x * 3u, x2 = blockIdx.x, x3 = 7u;
  __shared__ float4 sm[256];
  sm[threadIdx.x] = make_float4(1.f, 2.f, 3.f, 4.f);
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  float4 acc4 = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (MODE == 0) {   // 8 FFMA
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }
      } else if (MODE == 1) {   // 4 FFMA2 (same flops as MODE 0)
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = __ffma2_rn(a[i], m, c);
      } else if (MODE == 2) {   // 8 FFMA + 8 integer ALU
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }
        x0 = (x0 ^ x1) + 1u; x1 = (x1 & x2) | x3; x2 = (x2 ^ x3) + 3u; x3 = (x3 | x0) ^ x1;
        x0 = (x0 ^ x2) + 5u; x1 = (x1 & x3) | x0; x2 = (x2 ^ x1) + 7u; x3 = (x3 | x2) ^ x0;
      } else if (MODE == 3) {   // 4 FFMA2 + 8 integer ALU
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = __ffma2_rn(a[i], m, c);
        x0 = (x0 ^ x1) + 1u; x1 = (x1 & x2) | x3; x2 = (x2 ^ x3) + 3u; x3 = (x3 | x0) ^ x1;
        x0 = (x0 ^ x2) + 5u; x1 = (x1 & x3) | x0; x2 = (x2 ^ x1) + 7u; x3 = (x3 | x2) ^ x0;
      } else if (MODE == 4) {   // 8 MUFU.EX2
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i].x));
          asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i].y));
        }
      } else if (MODE == 5) {   // 8 SHFL
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i].x = __shfl_xor_sync(0xffffffffu, a[i].x, 16); a[i].y = __shfl_xor_sync(0xffffffffu, a[i].y, 8); }
      } else if (MODE == 6) {   // 8 LDS.128 (uniform address: broadcast)
#pragma unroll
        for (int i = 0; i < 8; ++i) { const float4 v = sm[(it + u + i) & 255]; acc4.x += v.x; acc4.y += v.y; acc4.z += v.z; acc4.w += v.w; }
      } else if (MODE == 7) {   // 4 FFMA2 + 2 MUFU + 2 SHFL: the compositing mix
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = __ffma2_rn(a[i], m, c);
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(c.x));
        asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(c.y));
        m.x = __shfl_xor_sync(0xffffffffu, m.x, 16); m.y = __shfl_xor_sync(0xffffffffu, m.y, 8);
      } else if (MODE == 8) {   // 8 FSEL-like selects (ALU pipe) + 4 FFMA2
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = __ffma2_rn(a[i], m, c);
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i].x = (x0 & (1u << i)) ? a[i].x : a[i].y; a[i].y = (x1 & (1u << i)) ? a[i].y : c.x; }
      }
    }
  }
  float s = acc4.x + acc4.y + acc4.z + acc4.w + (float)(x0 + x1 + x2 + x3);
#pragma unroll
  for (int i = 0; i < 4; ++i) s += a[i].x + a[i].y;
  s += m.x + m.y + c.x + c.y;
  if (s == 123.456f) out[0] = s;
}

template <int MODE>
static double run(float *d, int blocks, float *ms_out) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<blocks, 256>>>(d, 64);
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0);
    k<MODE><<<blocks, 256>>>(d, ITERS);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  *ms_out = best;
  return best;
}

int main() {
  float *d; cudaMalloc(&d, 16);
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  const int sms = p.multiProcessorCount, blocks = sms * 8;
  // per-iteration-unit instruction counts of each mode: {fp32 lanes-FMA per thread, warp instructions per thread}
  const char *names[9] = {"8 FFMA", "4 FFMA2", "8 FFMA + 8 IALU", "4 FFMA2 + 8 IALU", "8 MUFU.EX2", "8 SHFL", "8 LDS.128 broadcast (+32 FADD)",
                          "4 FFMA2 + 2 MUFU + 2 SHFL", "4 FFMA2 + 8 SEL"};
  const double fma[9] = {8, 8, 8, 8, 0, 0, 0, 8, 8};
  const double inst[9] = {8, 4, 16, 12, 8, 8, 40, 8, 12};
  float ms[9];
  run<0>(d, blocks, &ms[0]); run<1>(d, blocks, &ms[1]); run<2>(d, blocks, &ms[2]); run<3>(d, blocks, &ms[3]); run<4>(d, blocks, &ms[4]);
  run<5>(d, blocks, &ms[5]); run<6>(d, blocks, &ms[6]); run<7>(d, blocks, &ms[7]); run<8>(d, blocks, &ms[8]);
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_mhz_max\": %d, \"cases\": [", p.name, sms, clk_khz / 1000);
  for (int i = 0; i < 9; ++i) {
    const double units = 8.0 * ITERS * 256.0 * blocks;   // unrolled bodies executed, all threads
    const double sec = ms[i] * 1e-3;
    const double tflops = 2.0 * fma[i] * units / sec / 1e12;
    const double winst_per_sm_clk = inst[i] * units / 32.0 / sec / sms / (clk_khz * 1e3);
    printf("%s{\"case\": \"%s\", \"ms\": %.4f, \"fp32_tflops\": %.2f, \"warp_inst_per_sm_per_clk_at_max_clock\": %.3f}", i ? ", " : "", names[i], ms[i],
           tflops, winst_per_sm_clk);
  }
  printf("]}\n");
  return 0;
}
