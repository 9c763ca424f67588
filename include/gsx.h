/* gsx.h -- C ABI of the B200-native splatting hot path (libgsx.so).
 *
 * Drop-in boundary for the `gsplat_cuda` library of AndrewBoessen/3DGS: every entry point below replaces one
 * free function declared in the reference's include/gsplat_cuda/cuda_forward.cuh / cuda_backward.cuh (cited per
 * function, paths relative to /root/reference), or one of its two callers (cuda/raster.cu rasterize_image,
 * cuda/trainer.cu backward_pass).  Plain pointers and sizes only; all pointers are DEVICE pointers unless the
 * name says host; `stream` is a cudaStream_t passed as void*.  Every function returns 0 on success or a
 * GSX_E_* code (the reference prints to stderr and exit(1)s instead, cuda/checks.cuh:17-38; the C++ compat
 * shim include/gsplat_cuda/ restores that behaviour on top of these).
 *
 * Alignment: the quaternion arrays (parameters and gradients) and radius[N][4] are accessed as float4 and must be 16-byte
 * aligned, uv / uv gradients 8-byte aligned (GSX_E_ARG otherwise); everything else only needs float alignment.
 * Threads: a gsx_ctx is used by one host thread at a time; entry points that fall back to the library's internal cached
 * workspace (workspace == NULL) share one grow-only buffer per device under a mutex.
 *
 * Array layouts are the reference's: xyz[N][3], uv[N][2], quat[N][4] (w,x,y,z), scale[N][3] (log),
 * opacity[N] (logit), band0[N][3], sh[N][n-1][3], sigma[N][6], J[N][6], conic[N][3], radius[N][4]
 * (r_major,r_minor,sin,cos), view/proj row-major 4x4, image[H][W][3], weight[H][W], splats_per_pixel[H][W].
 */
#ifndef GSX_H_
#define GSX_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSX_OK 0
#define GSX_E_NULL 1      /* null pointer argument */
#define GSX_E_NOTDEV 2    /* pointer is not device memory */
#define GSX_E_ARG 3       /* bad scalar argument (l_max, sizes) */
#define GSX_E_CUDA 4      /* CUDA runtime error (see gsx_last_error) */
#define GSX_E_WORKSPACE 5 /* caller workspace too small */
#define GSX_E_CAPACITY 6  /* instance capacity exceeded */

#define GSX_TILE 16 /* TILE_SIZE_FWD / TILE_SIZE_BWD, cuda_forward.cuh:8, cuda_backward.cuh:8 */

/* flags */
#define GSX_SH_DSPH_CORRECT 1 /* F10: use sphericart's true [axis][k] derivative layout instead of the
                                 reference's [k][axis] reading.  Default 0 = reference behaviour. */
#define GSX_GRAD_OVERWRITE 2  /* gsx_backward: overwrite parameter gradients instead of accumulating */
#define GSX_GRAD_NO_SH 4      /* gsx_batch_backward_params: leave the band-0 / SH coefficient gradients out (they are
                                 rebuilt from the gathered per-view colour gradients, gsx_sh_grads_from_view_rgb) */
#define GSX_LISTS_CONTRIBUTING 8 /* gsx_view.flags, view-level path only: per-tile lists keep only the instances whose
                                 {alpha >= 1/255} footprint reaches the tile (the reference's 3-sigma OBB test keeps ~3x
                                 more).  Image and gradients are unchanged -- the dropped instances are exactly those
                                 compositing rejects for every pixel of the tile -- but sorted list, ranges and
                                 splats_per_pixel are then a filtered version of the reference's.  Default 0 =
                                 reference-identical lists. */
#define GSX_MAX_VIEWS_TOTAL 64 /* views of a whole (multi-GPU) batch accepted by gsx_sh_grads_from_view_rgb */

int gsx_version(void);
const char *gsx_last_error(void);

/* ---- F1..F8: cuda_forward.cuh ------------------------------------------------------------------------ */
/* cuda_forward.cuh:48  compute_camera_space_points */
int gsx_compute_camera_space_points(const float *xyz_w, const float *view, int N, float *xyz_c, void *stream);
/* cuda_forward.cuh:61  project_to_screen */
int gsx_project_to_screen(const float *xyz_c, const float *proj, int N, int width, int height, float *uv,
                          void *stream);
/* cuda_forward.cuh:76  cull_gaussians ; mask is bool[N], nonzero = keep */
int gsx_cull_gaussians(const float *uv, const float *xyz_c, int N, float near_thresh, int padding, int width,
                       int height, uint8_t *mask, void *stream);
/* cuda_forward.cuh:38  compute_sigma */
int gsx_compute_sigma(const float *quaternion, const float *scale, int N, float *sigma, void *stream);
/* cuda_forward.cuh:26  compute_conic ; radius is float4[N] */
int gsx_compute_conic(const float *xyz_c, const float *view, const float *sigma, float focal_x, float focal_y,
                      float tan_fovx, float tan_fovy, float mh_dist, int N, float *J, float *conic, float *radius,
                      void *stream);
/* cuda_forward.cuh:107 precompute_spherical_harmonics ; xyz is the WORLD position (Q11); campos by value */
int gsx_precompute_spherical_harmonics(const float *xyz_w, const float *sh, const float *band0, float campos_x,
                                       float campos_y, float campos_z, int l_max, int N, float *rgb,
                                       void *stream);
/* cuda_forward.cuh:92  get_sorted_gaussian_list, phase 1 (sorted_gaussians == nullptr): returns the number of
 * coarse candidate pairs C -- what the reference stores in `sorted_gaussian_bytes` -- and the exact instance
 * count K.  Synchronises `stream` (the reference does too, culling.cu:404). */
int gsx_sorted_list_size(const float *uv, const float *radius, int n_tiles_x, int n_tiles_y, int N,
                         long long *coarse_pairs_host, long long *num_instances_host, void *stream);
/* workspace bytes needed by gsx_get_sorted_gaussian_list for N Gaussians and up to `capacity` instances */
size_t gsx_sorted_list_workspace_bytes(int N, long long capacity);
/* cuda_forward.cuh:92  get_sorted_gaussian_list, phase 2.  sorted_gaussians has room for `capacity` ints
 * (the reference caller allocates C of them); ranges has n_tiles+1 ints.  Writes K into *num_instances_dev
 * (device int, may be null).  Order: (tile, depth, gaussian id) ascending -- the reference's order with its
 * nondeterministic ties canonicalised.  workspace may be null: the library then uses an internal cached one. */
int gsx_get_sorted_gaussian_list(const float *uv, const float *xyz_c, const float *radius, int n_tiles_x,
                                 int n_tiles_y, int N, long long capacity, int *sorted_gaussians, int *ranges,
                                 int *num_instances_dev, void *workspace, size_t workspace_bytes, void *stream);
/* cuda_forward.cuh:128 render_image */
int gsx_render_image(const float *uv, const float *opacity, const float *conic, const float *rgb,
                     float background, const int *sorted_splats, const int *splat_range_by_tile, int width,
                     int height, int *splats_per_pixel, float *weight_per_pixel, float *image, void *stream);

/* ---- F9..F15: cuda_backward.cuh ---------------------------------------------------------------------- */
/* cuda_backward.cuh:118 render_image_backward ; outputs accumulated (+=) */
int gsx_render_image_backward(const float *uv, const float *opacity, const float *conic, const float *rgb,
                              float background, const int *sorted_splats, const int *splat_range_by_tile,
                              const int *splats_per_pixel, const float *weight_per_pixel,
                              const float *grad_image, int width, int height, int N, float *grad_rgb,
                              float *grad_opacity, float *grad_uv, float *grad_conic, void *stream);
/* cuda_backward.cuh:92  precompute_spherical_harmonics_backward ; sh_grad/band0_grad (=), xyz_grad (+=) */
int gsx_precompute_spherical_harmonics_backward(const float *xyz_w, const float *band0, const float *sh,
                                                float campos_x, float campos_y, float campos_z,
                                                const float *rgb_grad, int l_max, int N, int flags,
                                                float *sh_grad, float *band0_grad, float *xyz_grad,
                                                void *stream);
/* cuda_backward.cuh:62  compute_conic_backward ; J_grad, sigma_grad (+=) */
int gsx_compute_conic_backward(const float *J, const float *sigma, const float *view, const float *conic,
                               const float *conic_grad, int N, float *J_grad, float *sigma_grad, void *stream);
/* cuda_backward.cuh:47  compute_projection_jacobian_backward ; xyz_c_grad (+=) */
int gsx_compute_projection_jacobian_backward(const float *xyz_c, float focal_x, float focal_y, float tan_fovx,
                                             float tan_fovy, const float *J_grad, int N, float *xyz_c_grad,
                                             void *stream);
/* cuda_backward.cuh:76  compute_sigma_backward ; quat_grad, scale_grad (=) */
int gsx_compute_sigma_backward(const float *quaternion, const float *scale, const float *sigma_grad, int N,
                               float *quat_grad, float *scale_grad, void *stream);
/* cuda_backward.cuh:21  project_to_screen_backward ; xyz_c_grad (+=) */
int gsx_project_to_screen_backward(const float *xyz_c, const float *proj, const float *uv_grad, int N, int width,
                                   int height, float *xyz_c_grad, void *stream);
/* cuda_backward.cuh:34  compute_camera_space_points_backward ; xyz_w_grad (+=) ; xyz_w unused */
int gsx_compute_camera_space_points_backward(const float *xyz_w, const float *view, const float *xyz_c_grad,
                                             int N, float *xyz_w_grad, void *stream);

/* ---- standalone sort primitives (hand-written onesweep) ------------------------------------------------- */
size_t gsx_sort_workspace_bytes(long long n);
/* Sorts n (key,value) pairs ascending on key bits [0, key_bits); stable.  Result is left in keys_out/vals_out
 * (keys_in/vals_in are clobbered).  n_dev, when non-null, is a device int holding the live count (<= n). */
int gsx_sort_pairs_u64(uint64_t *keys_in, uint64_t *keys_out, uint32_t *vals_in, uint32_t *vals_out, long long n,
                       const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, void *stream);
/* Same for 32-bit keys: the two sorts of the view-level path (32 depth bits per Gaussian, then ceil(log2 T) tile bits per
 * instance) are this one. */
int gsx_sort_pairs_u32(uint32_t *keys_in, uint32_t *keys_out, uint32_t *vals_in, uint32_t *vals_out, long long n,
                       const int *n_dev, int key_bits, void *workspace, size_t workspace_bytes, void *stream);
/* ranges[t] = first i with (keys[i] >> 32) >= t, ranges[n_tiles] = n.  n_dev as above. */
int gsx_tile_ranges(const uint64_t *sorted_keys, long long n, const int *n_dev, int n_tiles, int *ranges,
                    void *stream);

/* ---- view-level fused path: replaces rasterize_image (cuda/raster.cu:12-136) and backward_pass
 *      (cuda/trainer.cu:926-1015, minus fused_loss) without compaction copies or host round trips.
 *      Everything is indexed by GLOBAL Gaussian id so a batch of views (and GPUs) can sum gradients. ------ */
typedef struct gsx_view {
  float view[16];        /* row-major world->camera, trainer.cu:1322-1332 */
  float proj[16];        /* row-major projection, trainer.cu:1310-1320 */
  float campos[3];       /* Image::CamPos() */
  float focal_x, focal_y;   /* camera.params[0], params[1] (Q16) */
  float tan_fovx, tan_fovy; /* W/(2 fx), H/(2 fy), raster.cu:92-93 */
  int width, height;
  float near_thresh;     /* config near_thresh */
  int cull_padding;      /* config cull_mask_padding */
  float mh_dist;         /* config mh_dist */
  float background;      /* scalar grey background (Q4) */
  int l_max;             /* 0..3 */
  int flags;             /* GSX_SH_DSPH_CORRECT | GSX_GRAD_OVERWRITE | GSX_LISTS_CONTRIBUTING */
} gsx_view_t;

typedef struct gsx_params {   /* GaussianParameters, cuda_data.cuh:11-24 */
  const float *xyz, *scale, *quaternion, *opacity, *band0, *sh;
  int sh_stride;              /* floats per Gaussian in sh: 3*((l_alloc+1)^2-1); the trainer always allocates 45 */
} gsx_params_t;

typedef struct gsx_grads {    /* GaussianGradients, cuda_data.cuh:26-46 (parameter gradients only) */
  float *xyz, *scale, *quaternion, *opacity, *band0, *sh;
  float *uv;                  /* optional [N][2]: screen-space gradient as F9 leaves it (densification statistic) */
} gsx_grads_t;

typedef struct gsx_ctx gsx_ctx_t;   /* persistent per-device workspace (replaces ForwardPassData) */

typedef struct gsx_stats {
  int N, P;                   /* Gaussians in, Gaussians passing the cull */
  long long K;                /* tile instances */
  long long capacity;         /* current instance capacity */
} gsx_stats_t;

int gsx_ctx_create(gsx_ctx_t **ctx, int max_gaussians, int max_width, int max_height, long long instance_capacity);
int gsx_ctx_destroy(gsx_ctx_t *ctx);
/* Forward for one view.  Fully asynchronous on `stream`; if K exceeds the context's instance capacity the
 * overflow is detected at gsx_ctx_stats()/gsx_forward_sync and the caller re-runs after gsx_ctx_reserve. */
int gsx_forward(gsx_ctx_t *ctx, const gsx_view_t *view_host, const gsx_params_t *params, int N, void *stream);
/* Backward for the view last rendered by gsx_forward on this ctx.  grad_image[H][W][3] device. */
int gsx_backward(gsx_ctx_t *ctx, const gsx_view_t *view_host, const gsx_params_t *params, int N,
                 const float *grad_image, const gsx_grads_t *grads, void *stream);
/* Synchronises and reports P, K of the last forward; returns GSX_E_CAPACITY (with K = the capacity to reserve) if ANY view
 * of the current batch overflowed.  A view whose K exceeds the capacity emits, sorts and composites nothing: its image
 * is the background and it contributes no gradients until the caller reserves and re-runs (no stale data is read). */
int gsx_ctx_stats(gsx_ctx_t *ctx, gsx_stats_t *out_host, void *stream);
int gsx_ctx_reserve(gsx_ctx_t *ctx, long long instance_capacity);   /* capacity < 2^30; clears the overflow flag */
/* Overflow poll WITHOUT synchronisation (a pinned flag the device raises): GSX_E_CAPACITY and the K to reserve if any
 * view forwarded on this context since the last gsx_ctx_reserve exceeded the instance capacity, else GSX_OK.  Definitive
 * after the stream that ran the views was synchronised; a training loop calls it once per step. */
int gsx_ctx_overflow(gsx_ctx_t *ctx, long long *required_capacity_host);
/* Device views of the last forward's saved state (ForwardPassData, cuda_data.cuh:70-86), global-index. */
typedef struct gsx_state {
  const float *image, *weight_per_pixel; const int *splats_per_pixel;   /* [H][W][3], [H][W], [H][W] */
  const float *rec0;    /* [N][4] u, v, ex, ey  (ex,ey: alpha>=1/255 half-extents used for warp culling) */
  const float *rec1;    /* [N][4] conic a, b, c, sigmoid(opacity) */
  const float *rec2;    /* [N][4] r, g, b, opa*(1-opa) */
  const float *radius;  /* [N][4] r_major, r_minor, sin, cos */
  const float *depth;   /* [N] camera-space z */
  const uint8_t *mask;  /* [N] cull mask, 1 = keep */
  const int *tiles_touched;            /* [N] exact instance count per Gaussian */
  const int *sorted_gaussians;         /* [K] global Gaussian ids, (tile, depth, id) order */
  const uint64_t *sorted_keys;         /* always NULL: the two-level sort never materialises tile<<32 | depth keys;
                                          gsx_ctx_build_sorted_keys rebuilds them for inspection */
  const int *ranges;                   /* [T+1] */
  const int *num_instances;            /* device int: K */
  const float *packed_grads;           /* [N][12] compositing gradients (zero between views) */
} gsx_state_t;
int gsx_ctx_state(gsx_ctx_t *ctx, gsx_state_t *out_host);
/* [K] tile<<32 | order-preserving depth bits of the last forward, composed on demand (synchronises the device). */
int gsx_ctx_build_sorted_keys(gsx_ctx_t *ctx, const uint64_t **keys_dev_out);

/* ---- batch path: the n views of one training batch (n <= GSX_MAX_VIEWS) share ONE pass over the parameters in each
 *      direction (the reference reads them twice per view, raster.cu:26-104 / trainer.cu:941-1012).  Protocol:
 *        gsx_batch_reserve(ctx, n)                                   once
 *        gsx_batch_forward_params(ctx, views, n, params, N, s)       F1-F6 + tile counts of all views -> per-view slots
 *        for k in 0..n-1:
 *          gsx_batch_forward_view(ctx, k, s)                         F7 + F8 of view k; outputs via gsx_ctx_state
 *          gsx_batch_backward_view(ctx, k, grad_image, s)            F9 of view k (needs view k's sorted list: must
 *                                                                    directly follow its forward_view)
 *        gsx_batch_backward_params(ctx, params, N, grads, flags, s)  F10-F15 of all views, gradients summed over the
 *                                                                    batch ('=' with GSX_GRAD_OVERWRITE, else '+=')
 *      All asynchronous on `s`; no host round trip. ------------------------------------------------------------------ */
#define GSX_MAX_VIEWS 8
int gsx_batch_reserve(gsx_ctx_t *ctx, int n_slots);
int gsx_batch_forward_params(gsx_ctx_t *ctx, const gsx_view_t *views_host, int n_views, const gsx_params_t *params,
                             int N, void *stream);
int gsx_batch_forward_view(gsx_ctx_t *ctx, int slot, void *stream);
int gsx_batch_backward_view(gsx_ctx_t *ctx, int slot, const float *grad_image, void *stream);
int gsx_batch_backward_params(gsx_ctx_t *ctx, const gsx_params_t *params, int N, const gsx_grads_t *grads, int flags,
                              void *stream);
/* ---- CUDA-graph replay of a whole batch step.  gsx_batch_graph_capture records the batch protocol above
 *      -- per-Gaussian forward, then per view: bin + sort + composite, the view's loss gradient, compositing backward; then the
 *      per-Gaussian backward -- into a CUDA graph, side streams included; gsx_graph_launch replays it with one call per
 *      training step.  Loss gradients: grad_images_dev[k] (fixed device buffers whose CONTENTS the caller updates between
 *      launches), or, when gt_images_dev is given instead, computed inside the graph from each rendered image by
 *      gsx_fused_loss(image, gt_images_dev[k], ssim_weight) (mean loss of view k -> losses_dev[k] when non-NULL).
 *      images_out_dev[k] (optional) receives view k's rendered image.  `stream` must be a non-default stream; views, N and
 *      every pointer are baked in (capture again after gsx_ctx_reserve or when the views change). ------------------------ */
typedef struct gsx_graph gsx_graph_t;
int gsx_batch_graph_capture(gsx_ctx_t *ctx, const gsx_view_t *views_host, int n_views, const gsx_params_t *params, int N,
                            const float *const *grad_images_dev, const float *const *gt_images_dev, float ssim_weight,
                            float *const *images_out_dev, float *losses_dev, const gsx_grads_t *grads, int flags, void *stream,
                            gsx_graph_t **graph_out);
int gsx_graph_launch(gsx_graph_t *graph, void *stream);
int gsx_graph_destroy(gsx_graph_t *graph);

/* ---- view-parallel multi-GPU exchange.  A view's SH coefficient gradient is rank 1:
 *      d sh[k][c] = d rgb_c * Y_k(dir_view) (cuda/spherical_harmonics_backward.cu:100-104), so instead of allreducing the
 *      192 B/Gaussian band-0 + SH gradient, ranks all-gather the 12 B/Gaussian d rgb of each view and every rank rebuilds
 *      the batch sum locally; only the 44 B/Gaussian geometry gradients (xyz, scale, quaternion, opacity) are allreduced.
 *      gsx_batch_extract_view_rgb_grad: d rgb of the view in `slot` after its gsx_batch_backward_view, dense [N][3]
 *      (zeros where no gradient arrived).  gsx_sh_grads_from_view_rgb: drgb is [n_views][view_stride floats] holding
 *      [N][3] per view, campos_host is [n_views][3]; '=' with GSX_GRAD_OVERWRITE, else '+='. ------------------------- */
int gsx_batch_extract_view_rgb_grad(gsx_ctx_t *ctx, int slot, float *drgb, void *stream);
/* Fused extraction + all-gather over NVLink peer memory: the view's d rgb is stored straight into every rank's gather
 * buffer -- through the NVSwitch multicast address `multicast_drgb` (multimem.st) when non-NULL, else to each of the
 * n_peers (<= 8) peer pointers (this rank's own buffer included) with plain P2P stores.  The pointers address the view's
 * [N][3] block inside a symmetric allocation (cuMem/NVLS mapping owned by the caller, e.g. torch symmetric memory). */
int gsx_batch_extract_view_rgb_grad_peers(gsx_ctx_t *ctx, int slot, float *multicast_drgb, float *const *peer_drgb_host,
                                          int n_peers, void *stream);
/* Same transfer for a dense [N][3] block that was already extracted locally (gsx_batch_extract_view_rgb_grad): lets the
 * caller run the push on a side stream so the per-Gaussian backward is not queued behind NVLink back-pressure. */
int gsx_push_view_rgb_peers(const float *drgb_local, int N, float *multicast_drgb, float *const *peer_drgb_host,
                            int n_peers, void *stream);
/* In-switch (NVLS) sum-allreduce of n_floats (multiple of 4, 16-byte aligned) floats of a symmetric buffer through its
 * multicast address: rank r reduces slice r with multimem.ld_reduce and broadcasts it with multimem.st.  The caller
 * barriers across the ranks before and after. */
int gsx_allreduce_multimem_f32(float *multicast_ptr, size_t n_floats, int rank, int world, void *stream);
int gsx_sh_grads_from_view_rgb(const float *xyz_w, int N, int l_max, const float *campos_host, int n_views,
                               const float *drgb, size_t view_stride, float *sh_grad, int sh_stride, float *band0_grad,
                               int flags, void *stream);

/* Host-buffer batch step (bench.py's e2e leg): per view, H2D of grad_images_host[k] and D2H of the rendered image into
 * images_host[k] (pinned host memory, [H][W][3] floats), then the whole protocol above.  Synchronises. */
int gsx_batch_render_host(gsx_ctx_t *ctx, const gsx_view_t *views_host, int n_views, const gsx_params_t *params, int N,
                          const float *const *grad_images_host, float *const *images_host, const gsx_grads_t *grads,
                          int flags, void *stream);

/* Optional per-stage CUDA-event timing of gsx_forward/gsx_backward (used by bench.py's roofline leg, never inside
 * the timed throughput region).  ms_out[7] = preprocess_fwd, scan+emit, sort, ranges, composite_fwd, composite_bwd,
 * preprocess_bwd. */
int gsx_ctx_enable_timing(gsx_ctx_t *ctx, int on);
int gsx_ctx_timing(gsx_ctx_t *ctx, float *ms_out);
/* out_host[2] = {S, S+}: evaluated pixel-entries (sum of splats_per_pixel) and those with alpha >= 1/255 in the
 * last forward -- the units of the compositing roofline.  Synchronises. */
int gsx_ctx_count_work(gsx_ctx_t *ctx, unsigned long long *out_host, void *stream);
/* out_host[2] = {warp-hits, contributing warp-hits} of the compositing backward over the last forward's lists (the
 * (8x8 pixel block, entry) pairs that survive the extent-box cull / that have a contributing pixel).  Synchronises. */
int gsx_ctx_count_hits(gsx_ctx_t *ctx, const float *grad_image, unsigned long long *out_host, void *stream);
/* Measured FP32 FMA peak of the current device in TFLOP/s (roofline denominator of the compositing kernels). */
int gsx_measure_fp32_tflops(float *tflops_host);

/* Host-buffer entry point (what bench.py's e2e leg and a host-side caller use): copies the view + grad_image
 * from (pinned) host memory, runs forward+backward, copies the image back.  params/grads stay on the device as
 * in the reference trainer (CudaDataManager). */
int gsx_render_view_host(gsx_ctx_t *ctx, const gsx_view_t *view_host, const gsx_params_t *params, int N,
                         const float *grad_image_host, float *image_host, const gsx_grads_t *grads, void *stream);

/* ---- Adam on the path's global-index gradients (the step right after the backward) ----
 * gsx_adam_step replaces adam_step (include/gsplat_cuda/optimizer.cuh:27, cuda/optimizer.cu:30-44) on one array of
 * n_floats = N*S elements: m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr (m/bias1) / (sqrt(v/bias2) + eps);
 * a NaN gradient counts as 0 (optimizer.cu:14-16).
 * gsx_adam_step_fused replaces TrainerImpl::optimizer_step (cuda/trainer.cu:1027-1158) in ONE launch: every parameter
 * group [N][width] with its own learning rate, gradients indexed by the GLOBAL Gaussian index (as gsx_backward /
 * gsx_batch_backward_params write them; the reference compacts to [P] first).  mask [N] (the cull mask, uint8) or
 * null: a Gaussian with mask 0 keeps parameters and both moments untouched -- the effect of the reference's
 * compact -> adam_step -> scatter.  uv_grad [N][2] (8-byte aligned) or null: when given, uv_grad_accum[i] +=
 * |uv_grad[i]| and grad_accum_dur[i] += 1 for the same Gaussians (trainer.cu:1138-1157). */
#define GSX_ADAM_MAX_GROUPS 8
typedef struct {
  float *param;          /* [N][width] */
  const float *grad;     /* [N][width] */
  float *exp_avg;        /* [N][width] */
  float *exp_avg_sq;     /* [N][width] */
  int width;             /* floats per Gaussian: 3 xyz, 3 scale, 4 quaternion, 1 opacity, 3 band0, 3(n-1) sh */
  float lr;
} gsx_adam_group_t;
int gsx_adam_step(float *params, const float *param_grads, float *exp_avg, float *exp_avg_sq, float lr, float b1,
                  float b2, float eps, float bias1, float bias2, long long n_floats, void *stream);
int gsx_adam_step_fused(const gsx_adam_group_t *groups_host, int n_groups, const uint8_t *mask, int N, float b1,
                        float b2, float eps, float bias1, float bias2, const float *uv_grad, float *uv_grad_accum,
                        int *grad_accum_dur, void *stream);

/* ---- L1 + SSIM loss and its image gradient (the step right before gsx_backward) ----
 * Replaces fused_loss (include/gsplat_cuda/cuda_forward.cuh:144, cuda/loss.cu:440-471).  predicted / gt / image_grad are
 * [rows][cols][3] interleaved RGB; loss = mean over rows*cols*3 of (1-w)|p-g| + w(1 - ssim), 11x11 Gaussian window;
 * image_grad = d loss / d predicted (overwritten).  workspace: gsx_fused_loss_workspace_bytes(rows, cols) device bytes
 * (the reference allocates per call).  loss_dev (device float, may be null) receives the mean loss without any
 * synchronisation; loss_host (host float, may be null) additionally copies it back and synchronises the stream, which
 * is what the reference's return value does. */
size_t gsx_fused_loss_workspace_bytes(int rows, int cols);
int gsx_fused_loss(const float *predicted, const float *gt, int rows, int cols, float ssim_weight, float *image_grad,
                   void *workspace, size_t workspace_bytes, float *loss_dev, float *loss_host, void *stream);
/* gsx_fused_loss_static: NO counterpart in the reference -- added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same kernels,
 * arguments and results as gsx_fused_loss, but the workspace (loss sum + nine derivative maps) is a file-scope static
 * __device__ array inside libgsx.so instead of a caller buffer: the call performs no cudaMalloc, cudaFree or other
 * allocating call.  Limit: rows * cols <= GSX_FUSED_LOSS_STATIC_MAX_PIXELS (1 Mi pixels, e.g. 1280x720 or 1024x1024; the
 * array is 36 MiB + 256 B, held by every GPU that loads the library until the process exits); larger images are rejected
 * with GSX_E_WORKSPACE and a message -- use gsx_fused_loss.  All calls on a device share the one array, which keeps the
 * previous call's contents (the entry point clears the loss accumulator itself): calls that overlap on the GPU -- different
 * streams without an ordering between them -- race, so serialise them.  loss_host != NULL synchronises the stream (leave it
 * NULL under capture).  The first call on each device resolves the array's device address (cudaGetSymbolAddress); make it
 * outside stream capture. */
#define GSX_FUSED_LOSS_STATIC_MAX_PIXELS (1 << 20)
int gsx_fused_loss_static(const float *predicted, const float *gt, int rows, int cols, float ssim_weight, float *image_grad,
                          float *loss_dev, float *loss_host, void *stream);

/* ---- SURVEY 8f rank 4 (first half): Morton re-layout of the Gaussian arrays ----
 * gsx_compute_morton_codes replaces compute_morton_codes (include/gsplat_cuda/cuda_forward.cuh:186, cuda/culling.cu:42-68,
 * 345-359): 63-bit codes, 21 bits per axis quantised as (v - min) * (2097151 / (max - min)), z|y|x interleaved.
 * flags = 0 reproduces the reference's spread_bits (culling.cu:18-40) bit for bit -- its mask sequence keeps only
 * coordinate bits {0,2,3,9,10,12}, so the codes are a coarse key; GSX_MORTON_TRUE_INTERLEAVE gives the real 3-way
 * interleave (better locality; results then differ from the reference's by construction).
 * gsx_morton_order replaces the sequence + sort_by_key of TrainerImpl::sort_gaussians (cuda/trainer.cu:886-893): order[i]
 * = index of the Gaussian that comes i-th in Morton order (stable: equal codes keep index order); sorted_codes (may be
 * null) receives the codes in that order.  gsx_gather_rows replaces gather_with_stride (trainer.cu:897-923), out of
 * place: dst[i][0..width) = src[order[i]][0..width). */
#define GSX_MORTON_TRUE_INTERLEAVE 1
int gsx_compute_morton_codes(int N, const float *xyz, float x_max, float y_max, float z_max, float x_min, float y_min,
                             float z_min, int flags, uint64_t *codes, void *stream);
size_t gsx_morton_order_workspace_bytes(int N);
int gsx_morton_order(int N, const float *xyz, float x_max, float y_max, float z_max, float x_min, float y_min,
                     float z_min, int flags, int *order, uint64_t *sorted_codes, void *workspace,
                     size_t workspace_bytes, void *stream);
int gsx_gather_rows(const float *src, const int *order, int N, int width, float *dst, void *stream);
/* The same gather for up to 24 arrays (parameters and optimiser moments) in one launch; the three tables are host arrays. */
int gsx_gather_rows_multi(const float *const *src_host, float *const *dst_host, const int *width_host, int n_arrays,
                          const int *order, int N, void *stream);

/* ---- the view-parallel multi-GPU step, NCCL called from C++ -------------------------------------------
 * One process (or host thread) per GPU, parameters replicated, the n_views_total views of a batch sharded in blocks over
 * the ranks (rank r owns views [r * n/world, (r+1) * n/world)), gradients summed over all views.  The caller this serves is
 * the reference's training loop (cuda/trainer.cu:1350-1360: rasterize_image, backward_pass, optimizer_step), widened to a
 * batch; the reference itself is single-GPU.
 *   gsx_dp_unique_id(id)      rank 0: a 128-byte ncclUniqueId the caller hands to every rank (MPI, a file, a socket, ...)
 *   gsx_dp_init(...)          NCCL communicator over `world` ranks (one GPU each, current device), exchange buffers
 *   gsx_dp_step(...)          forward + backward of this rank's views through ctx's batch path, then the exchange: each
 *                             local view's dense d rgb is all-gathered as soon as its compositing backward is done (under
 *                             the next views), the 44 B/Gaussian geometry gradients are allreduced under the local rebuild
 *                             of the band-0 + SH gradients from all views' d rgb; with `adam` the optimiser step runs as
 *                             the epilogue on every rank.  Asynchronous on `stream`.  views_all_host: ALL views of the
 *                             batch in global order; grad_images_dev: this rank's views only.
 *   gsx_dp_render_host(...)   the same from pinned host buffers: loss-gradient images up, rendered images down, both
 *                             pipelined on copy streams under the compute.  Synchronises.
 *   gsx_dp_grads(dp, &g)      where the batch-summed gradients are (device memory owned by dp; valid after a step)
 *   gsx_dp_allreduce(...)     plain sum-allreduce of a caller buffer (the un-optimised baseline: 236 B/Gaussian at once)
 * NCCL is bound at run time (dlopen of libnccl.so.2, GSX_NCCL_LIB overrides): libgsx.so has no link-time dependency on it. */
#define GSX_DP_ID_BYTES 128
typedef struct gsx_dp gsx_dp_t;
typedef struct gsx_adam_epilogue {
  gsx_grads_t params;                   /* the parameter arrays to update in place (uv unused) */
  float *exp_avg[6], *exp_avg_sq[6];    /* Adam moments: xyz, scale, quaternion, opacity, band0, sh */
  float lr[6];
  float b1, b2, eps, bias1, bias2;      /* as gsx_adam_step_fused */
} gsx_adam_epilogue_t;
int gsx_dp_unique_id(void *id128_out);
int gsx_dp_init(gsx_dp_t **dp, gsx_ctx_t *ctx, const void *id128, int rank, int world, int N, int n_views_total, int l_max,
                int sh_stride);
int gsx_dp_destroy(gsx_dp_t *dp);
int gsx_dp_grads(gsx_dp_t *dp, gsx_grads_t *out_host);
int gsx_dp_step(gsx_dp_t *dp, const gsx_view_t *views_all_host, const gsx_params_t *params, const float *const *grad_images_dev,
                const gsx_adam_epilogue_t *adam, void *stream);
int gsx_dp_render_host(gsx_dp_t *dp, const gsx_view_t *views_all_host, const gsx_params_t *params,
                       const float *const *grad_images_host, float *const *images_host, const gsx_adam_epilogue_t *adam,
                       void *stream);
/* gsx_dp_render_host without the final synchronisation: `stream` is made to wait for the image downloads, so consecutive
 * steps pipeline and one synchronisation of `stream` covers them; the host buffers must stay untouched until then. */
int gsx_dp_render_host_async(gsx_dp_t *dp, const gsx_view_t *views_all_host, const gsx_params_t *params,
                             const float *const *grad_images_host, float *const *images_host, const gsx_adam_epilogue_t *adam,
                             void *stream);
int gsx_dp_allreduce(gsx_dp_t *dp, float *buf, size_t n_floats, void *stream);
int gsx_dp_enable_timing(gsx_dp_t *dp, int on);
/* ms_out[6] = views phase, gather tail after the last view, per-Gaussian backward, allreduce, SH rebuild (+ waits,
 * + Adam), whole step of the last timed step (synchronises; -1 where a phase does not exist, e.g. world == 1) */
int gsx_dp_timing(gsx_dp_t *dp, float *ms_out);

/* ---- SURVEY 8f rank 4 (second half): adaptive density control -- clone / split / prune ----
 * gsx_clone_gaussians / gsx_split_gaussians replace clone_gaussians / split_gaussians
 * (include/gsplat_cuda/adaptive_density.cuh:28,54, cuda/adaptive_density.cu:13-206) with the same argument meaning: mask
 * [N] selects the parents, write_ids [N] is the exclusive scan of the mask, the outputs are separate buffers holding one
 * (clone) or two (split) rows per selected parent at row write_ids[i] (* 2 + j).  A clone is an exact copy; a split child
 * keeps colour / opacity / rotation, gets scale log(exp(s) / scale_factor) and centre xyz + R(q) (n * exp(s)), n ~ N(0, I).
 * The normals come from Philox4x32-10 keyed by `seed` with counter (parent index, child): reproducible, independent of the
 * launch shape, no state array (the reference re-seeds a curand state per thread with time(NULL) at every call). */
int gsx_clone_gaussians(int N, int num_sh_coef, const uint8_t *mask, const int *write_ids, const float *xyz_in,
                        const float *rgb_in, const float *op_in, const float *scale_in, const float *quat_in,
                        const float *sh_in, float *xyz_out, float *rgb_out, float *op_out, float *scale_out,
                        float *quat_out, float *sh_out, void *stream);
int gsx_split_gaussians(int N, float scale_factor, int num_sh_coef, const uint8_t *mask, const int *write_ids,
                        const float *xyz_in, const float *rgb_in, const float *op_in, const float *scale_in,
                        const float *quat_in, const float *sh_in, float *xyz_out, float *rgb_out, float *op_out,
                        float *scale_out, float *quat_out, float *sh_out, unsigned long long seed, void *stream);
/* The whole step, TrainerImpl::adaptive_density_step (cuda/trainer.cu:416-775), in five launches and one 12-byte host
 * read.  gsx_densify_plan decides per Gaussian (same predicates: prune if the opacity logit is below
 * logit(delete_opacity_threshold) or the largest scale exceeds 0.1 * scene_extent unless it is about to be cloned / split;
 * clone if the mean |uv gradient| exceeds uv_grad_threshold and the largest scale is <= 0.01 * scene_extent, split if it
 * is larger), scans the decisions into destination rows inside `workspace`, and returns the counts (synchronises).
 * gsx_densify_apply moves every array of the set -- parameters and, when given, both Adam moments of each -- from `in`
 * to `out` (distinct buffers, sized for counts.n_out rows): survivors first in their old order (clone parents included),
 * then the clones, then the split children; appended Gaussians get zero moments.  GSX_E_CAPACITY (nothing written) when
 * n_out exceeds max_gaussians -- the reference skips the step with a warning (trainer.cu:579-585). */
typedef struct gsx_density_config {
  float uv_grad_threshold;          /* config uv_grad_threshold */
  float delete_opacity_threshold;   /* config delete_opacity_threshold, a probability in (0, 1) */
  float scene_extent;               /* computeMaxDiagonal of the camera positions */
  float split_scale_factor;         /* config split_scale_factor (1.6) */
  int max_gaussians;                /* capacity of the output arrays (0: unchecked) */
} gsx_density_config_t;
typedef struct gsx_density_counts {
  int n_in, n_prune, n_clone, n_split, n_out;   /* n_out = n_in - n_prune - n_split + n_clone + 2 n_split */
} gsx_density_counts_t;
typedef struct gsx_gaussian_set {   /* CudaDataManager's parameter + optimiser arrays (cuda_data.cuh:11-24, 55-68) */
  float *xyz, *scale, *quaternion, *opacity, *band0, *sh;
  int sh_stride;                    /* floats per row of sh (and of its moments) */
  int sh_width;                     /* floats of a row in use: 3 * ((l_max + 1)^2 - 1); 0 = no sh array */
  float *exp_avg[6], *exp_avg_sq[6];   /* Adam moments in the order xyz, scale, quaternion, opacity, band0, sh; NULL = absent */
} gsx_gaussian_set_t;
size_t gsx_densify_workspace_bytes(int N);
int gsx_densify_plan(const gsx_density_config_t *cfg, int N, const float *opacity, const float *scale,
                     const float *uv_grad_accum, const int *grad_accum_dur, void *workspace, size_t workspace_bytes,
                     gsx_density_counts_t *counts_host, void *stream);
int gsx_densify_apply(const gsx_density_config_t *cfg, const gsx_density_counts_t *counts, const gsx_gaussian_set_t *in,
                      const gsx_gaussian_set_t *out, const void *workspace, size_t workspace_bytes, unsigned long long seed,
                      void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GSX_H_ */
