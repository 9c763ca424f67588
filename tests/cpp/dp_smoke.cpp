// tests/cpp/dp_smoke.cpp -- a C++ caller of the multi-GPU step that uses NOTHING but the C ABI of libgsx.so (include/gsx.h)
// and the CUDA runtime: what a maintainer of the reference's trainer (cuda/trainer.cu:1350-1360) would write.  One process
// per rank:   dp_smoke <rank> <world> <id_file> [N] [views]
// Rank 0 creates the NCCL id and publishes it through <id_file>; every rank renders its views of a small synthetic scene
// through gsx_dp_step and checks the batch-summed gradients against the same batch run on its own GPU alone
// (gsx_batch_* over all views): relative error <= 1e-3 per array.  Prints DP_SMOKE_OK on success.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <string>
#include <vector>

#include "gsx.h"

#define CK(call)                                                                                   \
  do {                                                                                             \
    int rc_ = (call);                                                                              \
    if (rc_ != 0) { fprintf(stderr, "[rank %d] %s -> %d: %s\n", g_rank, #call, rc_, gsx_last_error()); exit(2); } \
  } while (0)
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) { fprintf(stderr, "[rank %d] %s -> %s\n", g_rank, #call, cudaGetErrorString(e_)); exit(3); } \
  } while (0)
static int g_rank = 0;

static uint64_t s_rng = 0x3D65;
static double urand() { s_rng = s_rng * 6364136223846793005ull + 1442695040888963407ull; return (double)(s_rng >> 11) / 9007199254740992.0; }
static double nrand() { return sqrt(-2.0 * log(urand() + 1e-300)) * cos(6.283185307179586 * urand()); }

template <typename T> static T *dev(const std::vector<T> &h) {
  T *d = nullptr;
  CU(cudaMalloc(&d, h.size() * sizeof(T)));
  CU(cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return d;
}

// camera on the radius-4 orbit looking at the origin (3dgs_b200/scenes.py orbit_camera; projection as trainer.cu:1299-1320)
static gsx_view_t make_view(int W, int H, double az_deg, int l_max) {
  gsx_view_t v;
  memset(&v, 0, sizeof(v));
  const double a = az_deg * M_PI / 180.0, c[3] = {4 * sin(a), 0.0, -4 * cos(a)};
  double f[3] = {-c[0], -c[1], -c[2]}, n = sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  for (double &x : f) x /= n;
  double r[3] = {1.0 * f[2] - 0.0 * f[1], 0.0 * f[0] - 0.0 * f[2], 0.0 * f[1] - 1.0 * f[0]};   // up x f, up = (0, 1, 0)
  n = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  for (double &x : r) x /= n;
  const double u[3] = {f[1] * r[2] - f[2] * r[1], f[2] * r[0] - f[0] * r[2], f[0] * r[1] - f[1] * r[0]};
  const double *R[3] = {r, u, f};
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) v.view[4 * i + j] = (float)R[i][j];
    v.view[4 * i + 3] = (float)-(R[i][0] * c[0] + R[i][1] * c[1] + R[i][2] * c[2]);
  }
  v.view[15] = 1.0f;
  const double fx = 1.2 * W, fy = 1.2 * W, zn = 0.01, zf = 100.0;
  const double tx = W / (2 * fx), ty = H / (2 * fy);
  v.proj[0] = (float)(2 * zn / (2 * tx * zn)); v.proj[5] = (float)(2 * zn / (2 * ty * zn));
  v.proj[10] = (float)(zf / (zf - zn)); v.proj[11] = (float)(-(zf * zn) / (zf - zn)); v.proj[14] = 1.0f;
  for (int k = 0; k < 3; ++k) v.campos[k] = (float)c[k];
  v.focal_x = (float)fx; v.focal_y = (float)fy; v.tan_fovx = (float)tx; v.tan_fovy = (float)ty;
  v.width = W; v.height = H; v.near_thresh = 0.3f; v.cull_padding = 100; v.mh_dist = 3.0f; v.background = 0.1f;
  v.l_max = l_max; v.flags = 0;
  return v;
}

static double rel_err(const float *d_a, const float *d_b, size_t n) {
  std::vector<float> a(n), b(n);
  CU(cudaMemcpy(a.data(), d_a, n * sizeof(float), cudaMemcpyDeviceToHost));
  CU(cudaMemcpy(b.data(), d_b, n * sizeof(float), cudaMemcpyDeviceToHost));
  double num = 0, den = 0;
  for (size_t i = 0; i < n; ++i) { num += ((double)a[i] - b[i]) * ((double)a[i] - b[i]); den += (double)b[i] * b[i]; }
  if (!(den > 0)) return num > 0 ? 1e30 : 0.0;
  return sqrt(num / den);
}

int main(int argc, char **argv) {
  if (argc < 4) { fprintf(stderr, "usage: dp_smoke <rank> <world> <id_file> [N] [views]\n"); return 1; }
  const int rank = g_rank = atoi(argv[1]), world = atoi(argv[2]);
  const std::string id_file = argv[3];
  const int N = argc > 4 ? atoi(argv[4]) : 20000, NV = argc > 5 ? atoi(argv[5]) : 4;
  const int W = 208, H = 144, l_max = 3, SHF = 45;
  CU(cudaSetDevice(0));   // the launcher gives every rank its own CUDA_VISIBLE_DEVICES
  // ---- scene (identical on every rank: same seed)
  std::vector<float> xyz(3 * N), scale(3 * N), quat(4 * N), opa(N), b0(3 * N), sh((size_t)SHF * N);
  for (int i = 0; i < N; ++i) {
    for (int k = 0; k < 3; ++k) { xyz[3 * i + k] = (float)(1.2 * nrand()); scale[3 * i + k] = (float)(log(0.03) + 0.5 * nrand()); b0[3 * i + k] = (float)(3 * urand() - 1.5); }
    double q[4], n = 0;
    for (double &x : q) { x = nrand(); n += x * x; }
    for (int k = 0; k < 4; ++k) quat[4 * i + k] = (float)(q[k] / sqrt(n));
    opa[i] = (float)(2 * nrand());
    for (int k = 0; k < SHF; ++k) sh[(size_t)SHF * i + k] = (float)(0.15 * nrand());
  }
  gsx_params_t p;
  p.xyz = dev(xyz); p.scale = dev(scale); p.quaternion = dev(quat); p.opacity = dev(opa); p.band0 = dev(b0); p.sh = dev(sh);
  p.sh_stride = SHF;
  std::vector<gsx_view_t> views;
  for (int v = 0; v < NV; ++v) views.push_back(make_view(W, H, 360.0 * v / NV + 7.0, l_max));
  std::vector<float *> gi(NV);
  for (int v = 0; v < NV; ++v) {
    std::vector<float> g((size_t)W * H * 3);
    for (float &x : g) x = (float)(nrand() * 1e-3);
    gi[v] = dev(g);
  }
  gsx_ctx_t *ctx = nullptr;
  CK(gsx_ctx_create(&ctx, N, W, H, 16LL * N));
  // ---- the whole batch on this GPU alone: the answer every rank must end up with
  gsx_grads_t ref;
  const size_t widths[6] = {3, 3, 4, 1, 3, (size_t)SHF};
  float **rp[6] = {&ref.xyz, &ref.scale, &ref.quaternion, &ref.opacity, &ref.band0, &ref.sh};
  for (int a = 0; a < 6; ++a) CU(cudaMalloc(rp[a], widths[a] * N * sizeof(float)));
  ref.uv = nullptr;
  CK(gsx_batch_reserve(ctx, NV));
  CK(gsx_batch_forward_params(ctx, views.data(), NV, &p, N, nullptr));
  for (int v = 0; v < NV; ++v) { CK(gsx_batch_forward_view(ctx, v, nullptr)); CK(gsx_batch_backward_view(ctx, v, gi[v], nullptr)); }
  CK(gsx_batch_backward_params(ctx, &p, N, &ref, GSX_GRAD_OVERWRITE, nullptr));
  CU(cudaDeviceSynchronize());
  gsx_stats_t st;
  CK(gsx_ctx_stats(ctx, &st, nullptr));
  // ---- NCCL id: rank 0 publishes it through a file (a C++ trainer would use MPI_Bcast or a socket)
  unsigned char id[GSX_DP_ID_BYTES];
  if (rank == 0) {
    CK(gsx_dp_unique_id(id));
    const std::string tmp = id_file + ".tmp";
    FILE *f = fopen(tmp.c_str(), "wb");
    if (!f || fwrite(id, 1, sizeof(id), f) != sizeof(id)) { fprintf(stderr, "cannot write %s\n", tmp.c_str()); return 4; }
    fclose(f);
    rename(tmp.c_str(), id_file.c_str());
  } else {
    FILE *f = nullptr;
    for (int t = 0; t < 600 && !(f = fopen(id_file.c_str(), "rb")); ++t) usleep(100000);
    if (!f || fread(id, 1, sizeof(id), f) != sizeof(id)) { fprintf(stderr, "[rank %d] no NCCL id in %s\n", rank, id_file.c_str()); return 4; }
    fclose(f);
  }
  gsx_dp_t *dp = nullptr;
  CK(gsx_dp_init(&dp, ctx, id, rank, world, N, NV, l_max, SHF));
  const int nv_local = NV / world;
  std::vector<const float *> mine(nv_local);
  for (int k = 0; k < nv_local; ++k) mine[k] = gi[rank * nv_local + k];
  for (int it = 0; it < 2; ++it) CK(gsx_dp_step(dp, views.data(), &p, mine.data(), nullptr, nullptr));
  CU(cudaDeviceSynchronize());
  gsx_grads_t got;
  CK(gsx_dp_grads(dp, &got));
  const float *gp[6] = {got.xyz, got.scale, got.quaternion, got.opacity, got.band0, got.sh};
  const char *names[6] = {"xyz", "scale", "quaternion", "opacity", "band0", "sh"};
  double worst = 0;
  for (int a = 0; a < 6; ++a) {
    const double e = rel_err(gp[a], *rp[a], widths[a] * N);
    worst = e > worst ? e : worst;
    if (!(e <= 1e-3)) { fprintf(stderr, "[rank %d] d %s: relative error %.3e > 1e-3\n", rank, names[a], e); return 5; }
  }
  // ---- the host-buffer entry point must give the same gradients and hand the images back
  std::vector<float *> h_gi(nv_local), h_img(nv_local);
  const size_t img = (size_t)W * H * 3 * sizeof(float);
  for (int k = 0; k < nv_local; ++k) {
    CU(cudaMallocHost(&h_gi[k], img)); CU(cudaMallocHost(&h_img[k], img));
    CU(cudaMemcpy(h_gi[k], mine[k], img, cudaMemcpyDeviceToHost));
    memset(h_img[k], 0, img);
  }
  std::vector<float> before((size_t)3 * N);
  CU(cudaMemcpy(before.data(), got.xyz, before.size() * sizeof(float), cudaMemcpyDeviceToHost));
  CK(gsx_dp_render_host(dp, views.data(), &p, (const float *const *)h_gi.data(), h_img.data(), nullptr, nullptr));
  const double e2 = rel_err(got.xyz, *rp[0], (size_t)3 * N);
  double img_sum = 0;
  for (size_t i = 0; i < (size_t)W * H * 3; ++i) img_sum += h_img[nv_local - 1][i];
  if (!(e2 <= 1e-3) || !(img_sum > 0.05 * W * H * 3)) { fprintf(stderr, "[rank %d] render_host: d xyz %.3e, image sum %.3f\n", rank, e2, img_sum); return 6; }
  printf("DP_SMOKE_OK rank %d/%d N %d views %d P %d K %lld worst_rel_err %.2e render_host_rel_err %.2e\n", rank, world, N, NV, st.P,
         st.K, worst, e2);
  CK(gsx_dp_destroy(dp));
  CK(gsx_ctx_destroy(ctx));
  return 0;
}
