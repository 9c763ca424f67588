// dp.cu -- the view-parallel multi-GPU step behind the C ABI (SURVEY 8b last row / 8e; north_star: "host code stays C++",
// "NCCL is called from C++").  One process (or host thread) per GPU, parameters replicated, the views of a batch sharded
// over the ranks, the per-Gaussian parameter gradients summed over all views.  The reference has no multi-GPU path (SURVEY
// 5: no NCCL / MPI anywhere); the caller this serves is its training loop, cuda/trainer.cu:1350-1360
// (rasterize_image -> backward_pass -> optimizer_step), widened to a batch of views.
//
// Exchange (same decomposition as 3dgs_b200/dp.py, which is now a ctypes caller of these entry points): a view's SH
// coefficient gradient is rank 1, d sh[k][c] = d rgb_c * Y_k(dir_view) (cuda/spherical_harmonics_backward.cu:100-104), so
//   * each local view's dense d rgb [N][3] is all-gathered (ncclAllGather, one per local view, issued on a high-priority
//     communication stream as soon as that view's compositing backward is done -- under the next views' compute),
//   * only the 44 B/Gaussian geometry gradients (xyz, scale, quaternion, opacity; one flat buffer) are allreduced
//     (ncclAllReduce, under the SH rebuild),
//   * every rank rebuilds the batch-summed band-0 + SH gradients from all views' d rgb (gsx_sh_grads_from_view_rgb),
//   * optionally the Adam step runs as the epilogue of the exchange on every rank (identical inputs -> replicas stay
//     bit-identical), so no second pass over the gradients is launched by the caller.
// NCCL is bound at run time (dlopen of libnccl.so.2 -- the copy already mapped into the process when there is one, e.g.
// torch's), so libgsx.so loads and the single-GPU path runs on machines without NCCL.
#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "gsx_common.cuh"

namespace gsx {
namespace {

struct NcclApi {
  void *h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int *) = nullptr;
  // optional (user-buffer registration: lets NCCL run its NVLS / zero-copy algorithms on our buffers)
  ncclResult_t (*MemAlloc)(void **, size_t) = nullptr;
  ncclResult_t (*MemFree)(void *) = nullptr;
  ncclResult_t (*CommRegister)(const ncclComm_t, void *, size_t, void **) = nullptr;
  ncclResult_t (*CommDeregister)(const ncclComm_t, void *) = nullptr;
};
NcclApi g_nccl;

int load_nccl() {
  if (g_nccl.h) return GSX_OK;
  const char *names[4] = {getenv("GSX_NCCL_LIB"), "libnccl.so.2", "libnccl.so", nullptr};
  void *h = nullptr;
  for (int i = 0; i < 3 && !h; ++i)
    if (names[i] && names[i][0]) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
  if (!h) { set_error("cannot load NCCL (libnccl.so.2): %s", dlerror()); return GSX_E_ARG; }
#define GSX_SYM(field, name, required)                                                             \
  *(void **)(&g_nccl.field) = dlsym(h, name);                                                       \
  if (required && !g_nccl.field) { set_error("NCCL symbol %s missing", name); dlclose(h); return GSX_E_ARG; }
  GSX_SYM(GetUniqueId, "ncclGetUniqueId", true)
  GSX_SYM(CommInitRank, "ncclCommInitRank", true)
  GSX_SYM(CommDestroy, "ncclCommDestroy", true)
  GSX_SYM(AllGather, "ncclAllGather", true)
  GSX_SYM(AllReduce, "ncclAllReduce", true)
  GSX_SYM(GetErrorString, "ncclGetErrorString", true)
  GSX_SYM(GetVersion, "ncclGetVersion", false)
  GSX_SYM(MemAlloc, "ncclMemAlloc", false)
  GSX_SYM(MemFree, "ncclMemFree", false)
  GSX_SYM(CommRegister, "ncclCommRegister", false)
  GSX_SYM(CommDeregister, "ncclCommDeregister", false)
#undef GSX_SYM
  g_nccl.h = h;
  return GSX_OK;
}

#define GSX_NCCL(call)                                                                                         \
  do {                                                                                                         \
    ncclResult_t _r = (call);                                                                                  \
    if (_r != ncclSuccess) {                                                                                   \
      ::gsx::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(_r));                \
      return GSX_E_CUDA;                                                                                       \
    }                                                                                                          \
  } while (0)

}  // namespace
}  // namespace gsx

using namespace gsx;

struct gsx_dp {
  gsx_ctx_t *ctx = nullptr;
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1, N = 0, n_views = 0, vpr = 0, l_max = 0, sh_stride = 0;
  cudaStream_t cs = nullptr;                       // communication stream (high priority)
  cudaStream_t cs_in = nullptr, cs_out = nullptr;  // host pipeline of gsx_dp_render_host
  cudaEvent_t ev_view = nullptr, ev_gathered = nullptr, ev_geom = nullptr, ev_reduced = nullptr, ev_go = nullptr;
  cudaEvent_t ev_img = nullptr, ev_dl[2] = {nullptr, nullptr}, ev_in[GSX_MAX_VIEWS] = {};
  cudaEvent_t tm[8] = {};                          // phase timing (gsx_dp_timing)
  bool timing = false;
  // buffers: [drgb: n_views x N x 3 | geom: xyz 3N, scale 3N, quaternion 4N, opacity N (16-byte aligned segments)]
  float *base = nullptr;
  bool base_nccl = false;
  void *reg = nullptr;
  size_t n_drgb = 0, n_geom = 0, off[4] = {0, 0, 0, 0};
  float *band0_grad = nullptr, *sh_grad = nullptr;
  float *grad_stage[GSX_MAX_VIEWS] = {};
  float *img_stage[2] = {nullptr, nullptr};
  size_t img_bytes = 0;
  bool dl_pending[2] = {false, false};             // a download from image stage i is (or may be) in flight
  std::vector<float> campos_slots;                 // camera positions in gather-buffer slot order
};

static inline int slot_of(const gsx_dp *d, int rank, int k) { return k * d->world + rank; }   // interleaved: the ranks' k-th views are contiguous

extern "C" {

int gsx_dp_unique_id(void *id128_out) {
  if (!id128_out) return GSX_E_NULL;
  int rc = load_nccl();
  if (rc != GSX_OK) return rc;
  static_assert(sizeof(ncclUniqueId) == GSX_DP_ID_BYTES, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  GSX_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128_out, &id, sizeof(id));
  return GSX_OK;
}

int gsx_dp_destroy(gsx_dp_t *d) {
  if (!d) return GSX_OK;
  cudaDeviceSynchronize();
  if (d->reg && g_nccl.CommDeregister) g_nccl.CommDeregister(d->comm, d->reg);
  if (d->comm) g_nccl.CommDestroy(d->comm);
  if (d->base) { if (d->base_nccl) g_nccl.MemFree(d->base); else cudaFree(d->base); }
  cudaFree(d->band0_grad); cudaFree(d->sh_grad);
  for (int k = 0; k < GSX_MAX_VIEWS; ++k) { cudaFree(d->grad_stage[k]); if (d->ev_in[k]) cudaEventDestroy(d->ev_in[k]); }
  cudaFree(d->img_stage[0]); cudaFree(d->img_stage[1]);
  cudaEvent_t evs[8] = {d->ev_view, d->ev_gathered, d->ev_geom, d->ev_reduced, d->ev_go, d->ev_img, d->ev_dl[0], d->ev_dl[1]};
  for (cudaEvent_t e : evs) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : d->tm) if (e) cudaEventDestroy(e);
  if (d->cs) cudaStreamDestroy(d->cs);
  if (d->cs_in) cudaStreamDestroy(d->cs_in);
  if (d->cs_out) cudaStreamDestroy(d->cs_out);
  delete d;
  return GSX_OK;
}

int gsx_dp_init(gsx_dp_t **out, gsx_ctx_t *ctx, const void *id128, int rank, int world, int N, int n_views_total, int l_max,
                int sh_stride) {
  if (!out || !ctx || !id128) { set_error("null argument"); return GSX_E_NULL; }
  if (world < 1 || rank < 0 || rank >= world || N <= 0 || n_views_total < 1 || n_views_total > GSX_MAX_VIEWS_TOTAL ||
      n_views_total % world || n_views_total / world > GSX_MAX_VIEWS || l_max < 0 || l_max > 3) {
    set_error("bad gsx_dp_init arguments (world %d rank %d N %d views %d l_max %d): the batch must shard evenly, <= %d views per rank",
              world, rank, N, n_views_total, l_max, GSX_MAX_VIEWS);
    return GSX_E_ARG;
  }
  const int shf = ((l_max + 1) * (l_max + 1) - 1) * 3;
  if (l_max > 0 && sh_stride < shf) { set_error("sh_stride %d < %d", sh_stride, shf); return GSX_E_ARG; }
  int rc = load_nccl();
  if (rc != GSX_OK) return rc;
  gsx_dp *d = new gsx_dp();
  d->ctx = ctx; d->rank = rank; d->world = world; d->N = N; d->n_views = n_views_total; d->vpr = n_views_total / world;
  d->l_max = l_max; d->sh_stride = sh_stride;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclResult_t r = g_nccl.CommInitRank(&d->comm, world, id, rank);
  if (r != ncclSuccess) { set_error("ncclCommInitRank -> %s", g_nccl.GetErrorString(r)); delete d; return GSX_E_CUDA; }
  rc = gsx_batch_reserve(ctx, d->vpr);
  if (rc != GSX_OK) { gsx_dp_destroy(d); return rc; }
  int lo = 0, hi = 0;
  auto fail = [&](const char *what) { set_error("gsx_dp_init: %s failed: %s", what, cudaGetErrorString(cudaGetLastError())); gsx_dp_destroy(d); return GSX_E_CUDA; };
  if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) return fail("priority range");
  // high priority: the collectives' CTAs are scheduled ahead of the (many, short) CTAs of the kernels they overlap
  if (cudaStreamCreateWithPriority(&d->cs, cudaStreamNonBlocking, hi) != cudaSuccess) return fail("stream");
  cudaEvent_t *evs[8] = {&d->ev_view, &d->ev_gathered, &d->ev_geom, &d->ev_reduced, &d->ev_go, &d->ev_img, &d->ev_dl[0], &d->ev_dl[1]};
  for (cudaEvent_t *e : evs) if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess) return fail("event");
  for (cudaEvent_t &e : d->tm) if (cudaEventCreate(&e) != cudaSuccess) return fail("event");
  d->n_drgb = align_up((size_t)n_views_total * N * 3, 4);
  const int w[4] = {3, 3, 4, 1};
  size_t o = 0;
  for (int a = 0; a < 4; ++a) { d->off[a] = o; o += align_up((size_t)N * w[a], 4); }
  d->n_geom = o;
  const size_t bytes = (d->n_drgb + d->n_geom) * sizeof(float);
  // NCCL-allocated + registered buffers let NCCL use its NVLS / zero-copy paths on them; plain cudaMalloc otherwise
  if (world > 1 && g_nccl.MemAlloc && g_nccl.CommRegister && !getenv("GSX_DP_NO_REGISTER")) {
    void *p = nullptr;
    if (g_nccl.MemAlloc(&p, bytes) == ncclSuccess && p) {
      d->base = (float *)p; d->base_nccl = true;
      if (g_nccl.CommRegister(d->comm, p, bytes, &d->reg) != ncclSuccess) d->reg = nullptr;
    }
  }
  if (!d->base && cudaMalloc(&d->base, bytes) != cudaSuccess) return fail("cudaMalloc(exchange buffers)");
  if (cudaMemset(d->base, 0, bytes) != cudaSuccess) return fail("memset");
  if (cudaMalloc(&d->band0_grad, (size_t)N * 3 * sizeof(float)) != cudaSuccess) return fail("cudaMalloc(band0 grad)");
  if (cudaMalloc(&d->sh_grad, std::max<size_t>((size_t)N * sh_stride, 4) * sizeof(float)) != cudaSuccess) return fail("cudaMalloc(sh grad)");
  cudaMemset(d->band0_grad, 0, (size_t)N * 3 * sizeof(float));
  cudaMemset(d->sh_grad, 0, std::max<size_t>((size_t)N * sh_stride, 4) * sizeof(float));
  d->campos_slots.assign((size_t)3 * n_views_total, 0.0f);
  *out = d;
  return GSX_OK;
}

// device pointers of the batch-summed gradients this context produces (geometry gradients live in the allreduce buffer)
int gsx_dp_grads(gsx_dp_t *d, gsx_grads_t *g) {
  if (!d || !g) return GSX_E_NULL;
  float *geom = d->base + d->n_drgb;
  g->xyz = geom + d->off[0]; g->scale = geom + d->off[1]; g->quaternion = geom + d->off[2]; g->opacity = geom + d->off[3];
  g->band0 = d->band0_grad; g->sh = d->sh_grad; g->uv = nullptr;
  return GSX_OK;
}

int gsx_dp_enable_timing(gsx_dp_t *d, int on) { if (!d) return GSX_E_NULL; d->timing = on != 0; return GSX_OK; }
// ms_out[6] = views phase, gather tail (after the last view), per-Gaussian backward, allreduce, SH rebuild, whole step
int gsx_dp_timing(gsx_dp_t *d, float *ms) {
  if (!d || !ms) return GSX_E_NULL;
  GSX_CUDA(cudaDeviceSynchronize());
  const int a[6] = {0, 1, 1, 4, 6, 0}, b[6] = {1, 2, 3, 5, 7, 7};
  for (int i = 0; i < 6; ++i)
    if (cudaEventElapsedTime(&ms[i], d->tm[a[i]], d->tm[b[i]]) != cudaSuccess) { cudaGetLastError(); ms[i] = -1.0f; }
  return GSX_OK;
}

// plain sum-allreduce of a caller buffer on the caller's stream (the un-optimised baseline of SURVEY 8e: the whole
// 236 B/Gaussian gradient block in one collective)
int gsx_dp_allreduce(gsx_dp_t *d, float *buf, size_t n_floats, void *stream) {
  if (!d || !buf) return GSX_E_NULL;
  if (d->world == 1 || n_floats == 0) return GSX_OK;
  GSX_NCCL(g_nccl.AllReduce(buf, buf, n_floats, ncclFloat, ncclSum, d->comm, (cudaStream_t)stream));
  return GSX_OK;
}

static int dp_step_impl(gsx_dp_t *d, const gsx_view_t *views_all, const gsx_params_t *p, const float *const *grad_images_dev,
                        const float *const *grad_images_host, float *const *images_host, const gsx_adam_epilogue_t *adam,
                        cudaStream_t st, bool sync) {
  NvtxRange nvtx_("gsx_dp_step");
  gsx_ctx_t *c = d->ctx;
  const int N = d->N, nv = d->vpr, W = d->world;
  const bool host = grad_images_host != nullptr;
  long long need = 0;
  if (gsx_ctx_overflow(c, &need) != GSX_OK) return GSX_E_CAPACITY;   // raised by an earlier step: that step lost views
  for (int v = 0; v < d->n_views; ++v) {                             // camera positions in slot order (every rank needs all)
    const int s = slot_of(d, v / nv, v % nv);
    for (int k = 0; k < 3; ++k) d->campos_slots[3 * s + k] = views_all[v].campos[k];
  }
  const gsx_view_t *mine = views_all + (size_t)d->rank * nv;
  int rc;
  if (host) {   // uploads of the loss gradients start at once on their own stream, under the per-Gaussian forward
    if (!d->cs_in) {
      GSX_CUDA(cudaStreamCreateWithFlags(&d->cs_in, cudaStreamNonBlocking));
      GSX_CUDA(cudaStreamCreateWithFlags(&d->cs_out, cudaStreamNonBlocking));
    }
    size_t bytes = 0;
    for (int k = 0; k < nv; ++k) bytes = std::max(bytes, (size_t)mine[k].width * mine[k].height * 3 * sizeof(float));
    if (bytes > d->img_bytes) {
      for (int k = 0; k < GSX_MAX_VIEWS; ++k) { cudaFree(d->grad_stage[k]); d->grad_stage[k] = nullptr; }
      cudaFree(d->img_stage[0]); cudaFree(d->img_stage[1]);
      GSX_CUDA(cudaMalloc(&d->img_stage[0], bytes)); GSX_CUDA(cudaMalloc(&d->img_stage[1], bytes));
      d->img_bytes = bytes;
    }
    GSX_CUDA(cudaEventRecord(d->ev_go, st));
    GSX_CUDA(cudaStreamWaitEvent(d->cs_in, d->ev_go, 0));
    for (int k = 0; k < nv; ++k) {
      if (!grad_images_host[k]) { set_error("null gradient image %d", k); return GSX_E_NULL; }
      if (!d->grad_stage[k]) GSX_CUDA(cudaMalloc(&d->grad_stage[k], d->img_bytes));
      if (!d->ev_in[k]) GSX_CUDA(cudaEventCreateWithFlags(&d->ev_in[k], cudaEventDisableTiming));
      GSX_CUDA(cudaMemcpyAsync(d->grad_stage[k], grad_images_host[k], (size_t)mine[k].width * mine[k].height * 3 * sizeof(float),
                               cudaMemcpyHostToDevice, d->cs_in));
      GSX_CUDA(cudaEventRecord(d->ev_in[k], d->cs_in));
    }
  }
  if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[0], st));
  rc = gsx_batch_forward_params(c, mine, nv, p, N, st);
  if (rc != GSX_OK) return rc;
  gsx_state_t state;
  for (int k = 0; k < nv; ++k) {
    if (host && d->dl_pending[k & 1]) { GSX_CUDA(cudaStreamWaitEvent(st, d->ev_dl[k & 1], 0)); d->dl_pending[k & 1] = false; }   // image stage k & 1 is free again (a download of this or of the previous step)
    rc = gsx_batch_forward_view(c, k, st);
    if (rc != GSX_OK) return rc;
    if (host && images_host && images_host[k]) {   // view k's image goes back to the host while its backward runs
      const size_t bytes = (size_t)mine[k].width * mine[k].height * 3 * sizeof(float);
      gsx_ctx_state(c, &state);
      GSX_CUDA(cudaMemcpyAsync(d->img_stage[k & 1], state.image, bytes, cudaMemcpyDeviceToDevice, st));
      GSX_CUDA(cudaEventRecord(d->ev_img, st));
      GSX_CUDA(cudaStreamWaitEvent(d->cs_out, d->ev_img, 0));
      GSX_CUDA(cudaMemcpyAsync(images_host[k], d->img_stage[k & 1], bytes, cudaMemcpyDeviceToHost, d->cs_out));
      GSX_CUDA(cudaEventRecord(d->ev_dl[k & 1], d->cs_out));
      d->dl_pending[k & 1] = true;
    }
    const float *gi = host ? d->grad_stage[k] : grad_images_dev[k];
    if (host) GSX_CUDA(cudaStreamWaitEvent(st, d->ev_in[k], 0));
    rc = gsx_batch_backward_view(c, k, gi, st);
    if (rc != GSX_OK) return rc;
    float *blk = d->base + (size_t)k * W * N * 3;              // [world][N][3]: the ranks' k-th views
    rc = gsx_batch_extract_view_rgb_grad(c, k, blk + (size_t)d->rank * N * 3, st);
    if (rc != GSX_OK) return rc;
    if (W > 1) {   // all-gather of the ranks' k-th views on the communication stream, under the following views' compute
      GSX_CUDA(cudaEventRecord(d->ev_view, st));
      GSX_CUDA(cudaStreamWaitEvent(d->cs, d->ev_view, 0));
      GSX_NCCL(g_nccl.AllGather(blk + (size_t)d->rank * N * 3, blk, (size_t)N * 3, ncclFloat, d->comm, d->cs));
    }
  }
  if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[1], st));
  if (W > 1) {
    if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[2], d->cs));
    GSX_CUDA(cudaEventRecord(d->ev_gathered, d->cs));
  }
  gsx_grads_t g;
  gsx_dp_grads(d, &g);
  rc = gsx_batch_backward_params(c, p, N, &g, GSX_GRAD_OVERWRITE | GSX_GRAD_NO_SH, st);   // geometry gradients -> flat buffer
  if (rc != GSX_OK) return rc;
  if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[3], st));
  if (W > 1) {   // allreduce under the SH rebuild
    GSX_CUDA(cudaEventRecord(d->ev_geom, st));
    GSX_CUDA(cudaStreamWaitEvent(d->cs, d->ev_geom, 0));
    if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[4], d->cs));
    float *geom = d->base + d->n_drgb;
    GSX_NCCL(g_nccl.AllReduce(geom, geom, d->n_geom, ncclFloat, ncclSum, d->comm, d->cs));
    if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[5], d->cs));
    GSX_CUDA(cudaEventRecord(d->ev_reduced, d->cs));
    GSX_CUDA(cudaStreamWaitEvent(st, d->ev_gathered, 0));
  }
  if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[6], st));
  rc = gsx_sh_grads_from_view_rgb(p->xyz, N, d->l_max, d->campos_slots.data(), d->n_views, d->base, (size_t)N * 3, d->sh_grad,
                                  d->sh_stride, d->band0_grad, GSX_GRAD_OVERWRITE, st);
  if (rc != GSX_OK) return rc;
  if (W > 1) GSX_CUDA(cudaStreamWaitEvent(st, d->ev_reduced, 0));
  if (adam) {   // the optimiser step as the epilogue of the exchange: every rank applies the same update to its replica
    gsx_adam_group_t grp[6];
    float *par[6] = {adam->params.xyz, adam->params.scale, adam->params.quaternion, adam->params.opacity, adam->params.band0, adam->params.sh};
    const float *gr[6] = {g.xyz, g.scale, g.quaternion, g.opacity, g.band0, g.sh};
    const int w[6] = {3, 3, 4, 1, 3, d->sh_stride};
    int ng = 0;
    for (int a = 0; a < 6; ++a) {
      if (w[a] == 0 || (a == 5 && d->l_max == 0)) continue;
      grp[ng].param = par[a]; grp[ng].grad = gr[a]; grp[ng].exp_avg = adam->exp_avg[a]; grp[ng].exp_avg_sq = adam->exp_avg_sq[a];
      grp[ng].width = w[a]; grp[ng].lr = adam->lr[a];
      ++ng;
    }
    rc = gsx_adam_step_fused(grp, ng, nullptr, N, adam->b1, adam->b2, adam->eps, adam->bias1, adam->bias2, nullptr, nullptr, nullptr, st);
    if (rc != GSX_OK) return rc;
  }
  if (d->timing) GSX_CUDA(cudaEventRecord(d->tm[7], st));
  if (host && sync) {   // never return with copies in flight
    cudaStreamSynchronize(d->cs_in); cudaStreamSynchronize(d->cs_out);
    GSX_CUDA(cudaStreamSynchronize(st));
    d->dl_pending[0] = d->dl_pending[1] = false;
  } else if (host) {    // asynchronous form: the caller's stream joins the downloads, so synchronising it covers them
    for (int i = 0; i < 2; ++i)
      if (d->dl_pending[i]) GSX_CUDA(cudaStreamWaitEvent(st, d->ev_dl[i], 0));
  }
  return GSX_OK;
}

int gsx_dp_step(gsx_dp_t *d, const gsx_view_t *views_all_host, const gsx_params_t *params, const float *const *grad_images_dev,
                const gsx_adam_epilogue_t *adam, void *stream) {
  if (!d || !views_all_host || !params || !grad_images_dev) { set_error("null argument"); return GSX_E_NULL; }
  return dp_step_impl(d, views_all_host, params, grad_images_dev, nullptr, nullptr, adam, (cudaStream_t)stream, false);
}

int gsx_dp_render_host(gsx_dp_t *d, const gsx_view_t *views_all_host, const gsx_params_t *params,
                       const float *const *grad_images_host, float *const *images_host, const gsx_adam_epilogue_t *adam,
                       void *stream) {
  if (!d || !views_all_host || !params || !grad_images_host) { set_error("null argument"); return GSX_E_NULL; }
  return dp_step_impl(d, views_all_host, params, nullptr, grad_images_host, images_host, adam, (cudaStream_t)stream, true);
}

// The same without the final synchronisation: everything (uploads, step, downloads) is enqueued and `stream` is made to
// wait for the downloads, so consecutive steps pipeline and one synchronisation of `stream` by the caller covers all of
// them.  The host buffers must stay untouched until then.
int gsx_dp_render_host_async(gsx_dp_t *d, const gsx_view_t *views_all_host, const gsx_params_t *params,
                             const float *const *grad_images_host, float *const *images_host, const gsx_adam_epilogue_t *adam,
                             void *stream) {
  if (!d || !views_all_host || !params || !grad_images_host) { set_error("null argument"); return GSX_E_NULL; }
  return dp_step_impl(d, views_all_host, params, nullptr, grad_images_host, images_host, adam, (cudaStream_t)stream, false);
}

}  // extern "C"
