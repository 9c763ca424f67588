// loss.cu -- SURVEY 8f rank 3: the step right before the backward of the splatting path.  L1 + SSIM loss of the rendered
// image against the ground truth and its gradient d loss / d image (the grad_image gsx_backward consumes).
//
// Reference: fused_loss, cuda/loss.cu:58-471 (declared include/gsplat_cuda/cuda_forward.cuh:144): SSIM with an 11x11
// Gaussian window (sigma 1.5), replicate padding for the image statistics, zero padding for the back-convolution of the
// three derivative maps, loss = mean over H*W*3 of (1-w)|p-g| + w(1-ssim), L1 subgradient -1 at p == g.  The reference
// allocates three H*W*3 temporaries and a thrust vector per call and reads the loss back synchronously; here the caller
// passes a workspace, the mean loss is left on the device (and copied to the host only if asked).
//
// Two kernels (the back-convolution needs the derivative maps of neighbouring tiles).  32x16-pixel tile per 256-thread
// CTA; the interleaved RGB tile + halo is read once, coalesced, into planar shared memory; separable convolution with
// the symmetric taps paired like the reference, so the two agree to rounding; derivative maps are stored PLANAR
// ([map][channel][H][W]) so the second kernel's loads coalesce.  Bytes per pixel: 24 read + 36 written, then 36 + 24
// read + 12 written = 132 B (274 MB at 1920x1080, mostly L2-resident between the two kernels).
#include <mutex>

#include "gsx_common.cuh"

namespace gsx {
namespace {

// 11-tap Gaussian, sigma 1.5, as float32 (the reference's table, cuda/loss.cu:14-17)
__constant__ float c_gauss[11] = {0.001028380123898387f,  0.0075987582094967365f, 0.036000773310661316f,
                                  0.10936068743467331f,   0.21300552785396576f,   0.26601171493530273f,
                                  0.21300552785396576f,   0.10936068743467331f,   0.036000773310661316f,
                                  0.0075987582094967365f, 0.001028380123898387f};

constexpr int kHalo = 5;
constexpr int kTX = 32, kTY = 16;                 // pixels per CTA
constexpr int kBX = 32, kBY = 8;                  // threads per CTA: a thread owns rows ty and ty + 8 of column tx
constexpr int kSX = kTX + 2 * kHalo;              // 42
constexpr int kSY = kTY + 2 * kHalo;              // 26
constexpr int kSP = kSX + 1;                      // padded row pitch of the staged tile
constexpr int kThreads = kBX * kBY;
// GSX_LOSS_REG=1 (build variant): sliding windows in registers -- a thread produces 4 adjacent columns of the horizontal
// pass from 14 staged values and 2 adjacent rows of the vertical pass from 12, instead of 11 shared-memory loads per
// output.  Tried because the capture shows 63 % / 55 % of the shared-memory wavefront peak
// (profiles/r1_v16_ncu_full_summary.txt); measured on the same box it changes nothing (0.260 vs 0.261 ms at 1920x1080,
// profiles/r1_loss_ab_variants.jsonl, both parity-green): the kernels are bound by instruction issue as a whole (74 % SM
// throughput: index arithmetic of the staging loops, 7 IEEE divisions per pixel and channel), not by that pipe.  Default
// stays 0.
#ifndef GSX_LOSS_REG
#define GSX_LOSS_REG 0
#endif
constexpr int kHP = kTX + 1;                      // row pitch of the horizontal-pass buffer (conflict-free 4-column stores)
[[maybe_unused]] constexpr int kCG = kTX / 4;     // 4-column groups per tile row (GSX_LOSS_REG)
constexpr float kC1 = (0.01f * 1.0f) * (0.01f * 1.0f);   // cuda/loss.cu:19-25
constexpr float kC2 = (0.03f * 1.0f) * (0.03f * 1.0f);

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// symmetric 11-tap sum over a strided line: pairs (l + r) * w for d = 1..5, centre last (reference order)
template <typename F>
__device__ __forceinline__ float tap11(F at) {
  float s = 0.0f;
#pragma unroll
  for (int d = 1; d <= kHalo; ++d) s += (at(-d) + at(d)) * c_gauss[kHalo - d];
  s += at(0) * c_gauss[kHalo];
  return s;
}

__global__ void __launch_bounds__(kThreads) k_loss_fwd(int H, int W, float ssim_weight, const float *__restrict__ pred,
                                                       const float *__restrict__ gt, float *__restrict__ maps,
                                                       float *__restrict__ loss_sum) {
  __shared__ float sP[3][kSY][kSP], sG[3][kSY][kSP];   // planar staged tile + halo: prediction, ground truth
  __shared__ float sH[5][kSY][kHP];                    // horizontal pass of one channel: E[x], E[x^2], E[y], E[y^2], E[xy]
  __shared__ float sRed[kThreads / 32];
  const int tid = threadIdx.y * kBX + threadIdx.x;
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY;
  const size_t plane = (size_t)H * W;

  // interleaved RGB rows of the tile + halo, replicate padding (cuda/loss.cu:41-46)
  for (int e = tid; e < kSY * kSX * 3; e += kThreads) {
    const int row = e / (kSX * 3), k = e - row * (kSX * 3);
    const int px = k / 3, c = k - px * 3;
    const int gy = clampi(y0 + row - kHalo, 0, H - 1), gx = clampi(x0 + px - kHalo, 0, W - 1);
    const size_t src = ((size_t)gy * W + gx) * 3 + c;
    sP[c][row][px] = pred[src];
    sG[c][row][px] = gt[src];
  }
  __syncthreads();

  float loss = 0.0f;
  for (int c = 0; c < 3; ++c) {
#if GSX_LOSS_REG
    for (int e = tid; e < kSY * kCG; e += kThreads) {
      const int row = e / kCG, col0 = (e - row * kCG) * 4;
      float pv[4 + 2 * kHalo], gv[4 + 2 * kHalo];
#pragma unroll
      for (int k = 0; k < 4 + 2 * kHalo; ++k) {
        pv[k] = sP[c][row][col0 + k];
        gv[k] = sG[c][row][col0 + k];
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        sH[0][row][col0 + j] = tap11([&](int d) { return pv[j + kHalo + d]; });
        sH[1][row][col0 + j] = tap11([&](int d) { return pv[j + kHalo + d] * pv[j + kHalo + d]; });
        sH[2][row][col0 + j] = tap11([&](int d) { return gv[j + kHalo + d]; });
        sH[3][row][col0 + j] = tap11([&](int d) { return gv[j + kHalo + d] * gv[j + kHalo + d]; });
        sH[4][row][col0 + j] = tap11([&](int d) { return pv[j + kHalo + d] * gv[j + kHalo + d]; });
      }
    }
#else
    for (int e = tid; e < kSY * kTX; e += kThreads) {
      const int row = e / kTX, col = e - row * kTX;
      const float *p = &sP[c][row][col + kHalo], *g = &sG[c][row][col + kHalo];
      sH[0][row][col] = tap11([&](int d) { return p[d]; });
      sH[1][row][col] = tap11([&](int d) { return p[d] * p[d]; });
      sH[2][row][col] = tap11([&](int d) { return g[d]; });
      sH[3][row][col] = tap11([&](int d) { return g[d] * g[d]; });
      sH[4][row][col] = tap11([&](int d) { return p[d] * g[d]; });
    }
#endif
    __syncthreads();
#if GSX_LOSS_REG
    float o2[2][5];
#pragma unroll
    for (int s = 0; s < 5; ++s) {
      float v[2 + 2 * kHalo];
#pragma unroll
      for (int k = 0; k < 2 + 2 * kHalo; ++k) v[k] = sH[s][2 * threadIdx.y + k][threadIdx.x];
#pragma unroll
      for (int h = 0; h < 2; ++h) o2[h][s] = tap11([&](int d) { return v[h + kHalo + d]; });
    }
#endif
#pragma unroll
    for (int half = 0; half < 2; ++half) {
#if GSX_LOSS_REG
      const int ly = 2 * threadIdx.y + half, lx = threadIdx.x;
#else
      const int ly = threadIdx.y + half * kBY, lx = threadIdx.x;
#endif
      const int y = y0 + ly, x = x0 + lx;
      if (y < H && x < W) {
        float o[5];
#if GSX_LOSS_REG
#pragma unroll
        for (int s = 0; s < 5; ++s) o[s] = o2[half][s];
#else
#pragma unroll
        for (int s = 0; s < 5; ++s) o[s] = tap11([&](int d) { return sH[s][ly + kHalo + d][lx]; });
#endif
        // SSIM of the window centred on the pixel and its partial derivatives (cuda/loss.cu:214-252)
        const float mu1 = o[0], mu2 = o[2];
        const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2;
        const float sigma1_sq = o[1] - mu1_sq, sigma2_sq = o[3] - mu2_sq, sigma12 = o[4] - mu1 * mu2;
        const float A = mu1_sq + mu2_sq + kC1, B = sigma1_sq + sigma2_sq + kC2;
        const float C_ = 2.f * mu1 * mu2 + kC1, D_ = 2.f * sigma12 + kC2;
        const float ssim = (C_ * D_) / (A * B);
        const float pv = sP[c][ly + kHalo][lx + kHalo], gv = sG[c][ly + kHalo][lx + kHalo];
        loss += (1.0f - ssim_weight) * fabsf(pv - gv) + ssim_weight * (1.0f - ssim);
        const float d_mu1 = ((mu2 * 2.f * D_) / (A * B) - (mu2 * 2.f * C_) / (A * B) - (mu1 * 2.f * C_ * D_) / (A * A * B) +
                             (mu1 * 2.f * C_ * D_) / (A * B * B));
        const float d_sigma1_sq = (-C_ * D_) / (A * B * B);
        const float d_sigma12 = (2.f * C_) / (A * B);
        const size_t at = (size_t)c * plane + (size_t)y * W + x;
        maps[at] = -ssim_weight * d_mu1;
        maps[3 * plane + at] = -ssim_weight * d_sigma1_sq;
        maps[6 * plane + at] = -ssim_weight * d_sigma12;
      }
    }
    __syncthreads();
  }

  // CTA sum of the per-pixel losses: shuffles, then one atomic per CTA
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss += __shfl_xor_sync(0xffffffffu, loss, o);
  if ((tid & 31) == 0) sRed[tid >> 5] = loss;
  __syncthreads();
  if (tid == 0) {
    float t = 0.0f;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) t += sRed[w];
    atomicAdd(loss_sum, t);
  }
}

__global__ void __launch_bounds__(kThreads) k_loss_bwd(int H, int W, float ssim_weight, const float *__restrict__ pred,
                                                       const float *__restrict__ gt, const float *__restrict__ maps,
                                                       float *__restrict__ image_grad, const float *__restrict__ loss_sum,
                                                       float *__restrict__ loss_mean) {
  __shared__ float sD[3][kSY][kSP];   // the three derivative maps of one channel, zero outside the image
  __shared__ float sH[3][kSY][kHP];
  const int tid = threadIdx.y * kBX + threadIdx.x;
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY;
  const size_t plane = (size_t)H * W;
  const float grad_scale = 1.0f / (float)(H * W * 3);
  if (loss_mean && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *loss_mean = *loss_sum / (float)(H * W * 3);

  for (int c = 0; c < 3; ++c) {
    for (int e = tid; e < 3 * kSY * kSX; e += kThreads) {
      const int m = e / (kSY * kSX), k = e - m * (kSY * kSX);
      const int row = k / kSX, px = k - row * kSX;
      const int gy = y0 + row - kHalo, gx = x0 + px - kHalo;
      const bool in = gy >= 0 && gy < H && gx >= 0 && gx < W;   // zero padding (cuda/loss.cu:48-52)
      sD[m][row][px] = in ? maps[(size_t)(3 * m + c) * plane + (size_t)gy * W + gx] : 0.0f;
    }
    __syncthreads();
#if GSX_LOSS_REG
    for (int e = tid; e < 3 * kSY * kCG; e += kThreads) {
      const int m = e / (kSY * kCG), k2 = e - m * (kSY * kCG);
      const int row = k2 / kCG, col0 = (k2 - row * kCG) * 4;
      float v[4 + 2 * kHalo];
#pragma unroll
      for (int k = 0; k < 4 + 2 * kHalo; ++k) v[k] = sD[m][row][col0 + k];
#pragma unroll
      for (int j = 0; j < 4; ++j) sH[m][row][col0 + j] = tap11([&](int d) { return v[j + kHalo + d]; });
    }
#else
    for (int e = tid; e < kSY * kTX; e += kThreads) {
      const int row = e / kTX, col = e - row * kTX;
#pragma unroll
      for (int m = 0; m < 3; ++m) {
        const float *s = &sD[m][row][col + kHalo];
        sH[m][row][col] = tap11([&](int d) { return s[d]; });
      }
    }
#endif
    __syncthreads();
#if GSX_LOSS_REG
    float sum2[2][3];
#pragma unroll
    for (int m = 0; m < 3; ++m) {
      float v[2 + 2 * kHalo];
#pragma unroll
      for (int k = 0; k < 2 + 2 * kHalo; ++k) v[k] = sH[m][2 * threadIdx.y + k][threadIdx.x];
#pragma unroll
      for (int h = 0; h < 2; ++h) sum2[h][m] = tap11([&](int d) { return v[h + kHalo + d]; });
    }
#endif
#pragma unroll
    for (int half = 0; half < 2; ++half) {
#if GSX_LOSS_REG
      const int ly = 2 * threadIdx.y + half, lx = threadIdx.x;
#else
      const int ly = threadIdx.y + half * kBY, lx = threadIdx.x;
#endif
      const int y = y0 + ly, x = x0 + lx;
      if (y < H && x < W) {
        float sum[3];
#if GSX_LOSS_REG
#pragma unroll
        for (int m = 0; m < 3; ++m) sum[m] = sum2[half][m];
#else
#pragma unroll
        for (int m = 0; m < 3; ++m) sum[m] = tap11([&](int d) { return sH[m][ly + kHalo + d][lx]; });
#endif
        const size_t at = ((size_t)y * W + x) * 3 + c;
        const float p1 = pred[at], p2 = gt[at];
        const float ssim_part = sum[0] + (2.f * p1) * sum[1] + p2 * sum[2];          // cuda/loss.cu:418
        const float l1_part = (1.0f - ssim_weight) * ((p1 > p2) ? 1.0f : -1.0f);      // cuda/loss.cu:421
        image_grad[at] = (ssim_part + l1_part) * grad_scale;
      }
    }
    __syncthreads();
  }
}

// ---- scratch of gsx_fused_loss_static: file-scope device memory, so that entry point never allocates --------------------
// Same carve-up as the caller workspace of gsx_fused_loss: float [0] = loss sum, float [1] = mean when the caller passes no
// loss_dev, +256 bytes: the nine planar derivative maps.  Every GPU that loads libgsx.so holds these 36 MiB + 256 B until the
// process exits.  In HOST code the name below is only a stand-in for the symbol, not a device address: static_scratch()
// asks the runtime for the address (once per device, with that device current) and the kernels are handed that pointer.
constexpr size_t kStaticScratchBytes = 256 + (size_t)9 * GSX_FUSED_LOSS_STATIC_MAX_PIXELS * sizeof(float);
static __device__ __align__(256) unsigned char g_loss_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void *g_static_ptr[kMaxDevices] = {};   // device address of g_loss_static_scratch on device d (each device has its own copy)

int static_scratch(char **out) {
  int dev = -1;
  GSX_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { set_error("device ordinal %d outside [0, %d)", dev, kMaxDevices); return GSX_E_ARG; }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    void *p = nullptr;
    GSX_CUDA(cudaGetSymbolAddress(&p, g_loss_static_scratch));
    if (!p) { set_error("cudaGetSymbolAddress returned a null address on device %d", dev); return GSX_E_CUDA; }
    g_static_ptr[dev] = p;
  }
  *out = static_cast<char *>(g_static_ptr[dev]);
  return GSX_OK;
}

}  // namespace
}  // namespace gsx

using namespace gsx;

extern "C" {

size_t gsx_fused_loss_workspace_bytes(int rows, int cols) {
  if (rows <= 0 || cols <= 0) return 256;
  return 256 + (size_t)9 * rows * cols * sizeof(float);   // [0]: loss sum; +256: three derivative maps x three channels
}

int gsx_fused_loss(const float *predicted, const float *gt, int rows, int cols, float ssim_weight, float *image_grad,
                   void *workspace, size_t workspace_bytes, float *loss_dev, float *loss_host, void *stream) {
  GSX_REQUIRE_DEV(predicted); GSX_REQUIRE_DEV(gt); GSX_REQUIRE_DEV(image_grad); GSX_REQUIRE_DEV(workspace);
  if (loss_dev) GSX_REQUIRE_DEV(loss_dev);
  if (rows <= 0 || cols <= 0) { set_error("image %d x %d", cols, rows); return GSX_E_ARG; }
  if ((long long)rows * cols * 3 >= (1LL << 31)) { set_error("image %d x %d too large", cols, rows); return GSX_E_ARG; }
  if (workspace_bytes < gsx_fused_loss_workspace_bytes(rows, cols)) {
    set_error("workspace %zu < %zu bytes", workspace_bytes, gsx_fused_loss_workspace_bytes(rows, cols));
    return GSX_E_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float *sum = reinterpret_cast<float *>(workspace);
  float *mean = loss_dev ? loss_dev : sum + 1;
  float *maps = reinterpret_cast<float *>(reinterpret_cast<char *>(workspace) + 256);
  GSX_CUDA(cudaMemsetAsync(sum, 0, sizeof(float), st));
  const dim3 block(kBX, kBY), grid(ceil_div(cols, kTX), ceil_div(rows, kTY));
  k_loss_fwd<<<grid, block, 0, st>>>(rows, cols, ssim_weight, predicted, gt, maps, sum);
  GSX_LAUNCH_CHECK();
  k_loss_bwd<<<grid, block, 0, st>>>(rows, cols, ssim_weight, predicted, gt, maps, image_grad, sum, mean);
  GSX_LAUNCH_CHECK();
  if (loss_host) {   // the reference's behaviour: the call returns the loss (one synchronisation)
    GSX_CUDA(cudaMemcpyAsync(loss_host, mean, sizeof(float), cudaMemcpyDeviceToHost, st));
    GSX_CUDA(cudaStreamSynchronize(st));
  }
  return GSX_OK;
}

// The same two kernels over the library's static device scratch: no cudaMalloc / cudaFree / allocating call on any path.
int gsx_fused_loss_static(const float *predicted, const float *gt, int rows, int cols, float ssim_weight, float *image_grad,
                          float *loss_dev, float *loss_host, void *stream) {
  GSX_REQUIRE_DEV(predicted); GSX_REQUIRE_DEV(gt); GSX_REQUIRE_DEV(image_grad);
  if (loss_dev) GSX_REQUIRE_DEV(loss_dev);
  if (rows <= 0 || cols <= 0) { set_error("image %d x %d", cols, rows); return GSX_E_ARG; }
  if ((long long)rows * cols > (long long)GSX_FUSED_LOSS_STATIC_MAX_PIXELS) {
    set_error("gsx_fused_loss_static: image %d x %d = %lld pixels exceeds the static scratch's maximum of %d pixels "
              "(use gsx_fused_loss with a caller workspace)", cols, rows, (long long)rows * cols,
              (int)GSX_FUSED_LOSS_STATIC_MAX_PIXELS);
    return GSX_E_WORKSPACE;
  }
  char *scratch = nullptr;
  if (int rc = static_scratch(&scratch)) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float *sum = reinterpret_cast<float *>(scratch);
  float *mean = loss_dev ? loss_dev : sum + 1;
  float *maps = reinterpret_cast<float *>(scratch + 256);
  // the array keeps the previous call's contents: the loss sum must start at zero (the maps are fully overwritten for
  // every pixel inside rows x cols before the second kernel reads them, and it reads nothing outside)
  GSX_CUDA(cudaMemsetAsync(sum, 0, sizeof(float), st));
  const dim3 block(kBX, kBY), grid(ceil_div(cols, kTX), ceil_div(rows, kTY));
  k_loss_fwd<<<grid, block, 0, st>>>(rows, cols, ssim_weight, predicted, gt, maps, sum);
  GSX_LAUNCH_CHECK();
  k_loss_bwd<<<grid, block, 0, st>>>(rows, cols, ssim_weight, predicted, gt, maps, image_grad, sum, mean);
  GSX_LAUNCH_CHECK();
  if (loss_host) {
    GSX_CUDA(cudaMemcpyAsync(loss_host, mean, sizeof(float), cudaMemcpyDeviceToHost, st));
    GSX_CUDA(cudaStreamSynchronize(st));
  }
  return GSX_OK;
}

}  // extern "C"
