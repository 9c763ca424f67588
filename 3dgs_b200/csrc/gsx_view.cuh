// gsx_view.cuh -- the per-view constants every per-Gaussian kernel needs, passed by value in kernel parameter space
// (no device copy of the camera, no host<->device traffic per view).
#pragma once
#include "gsx_common.cuh"

namespace gsx {

struct ViewConst {
  float view[12];   // rows 0..2 of the row-major world->camera matrix
  float proj[16];
  float3 campos;
  float fx, fy, tan_fovx, tan_fovy;
  int width, height, ntx, nty;
  float near_thresh;
  int padding;
  float mh_dist;
  int tight;        // GSX_LISTS_CONTRIBUTING: clip each Gaussian's tile rectangle to its alpha >= 1/255 extent
};

static inline ViewConst make_view_const(const gsx_view_t &v) {
  ViewConst c;
  for (int k = 0; k < 12; ++k) c.view[k] = v.view[k];
  for (int k = 0; k < 16; ++k) c.proj[k] = v.proj[k];
  c.campos = make_float3(v.campos[0], v.campos[1], v.campos[2]);
  c.fx = v.focal_x; c.fy = v.focal_y; c.tan_fovx = v.tan_fovx; c.tan_fovy = v.tan_fovy;
  c.width = v.width; c.height = v.height;
  c.ntx = (v.width + 15) / 16; c.nty = (v.height + 15) / 16;
  c.near_thresh = v.near_thresh; c.padding = v.cull_padding; c.mh_dist = v.mh_dist;
  c.tight = (v.flags & GSX_LISTS_CONTRIBUTING) ? 1 : 0;
  return c;
}

constexpr int GSX_MAX_BATCH = 8;   // views processed together by the batch kernels (batch.cu)

// per-view device buffers of one batch slot (api.cu owns them)
struct SlotPtrs {
  float4 *r0, *r1, *r2, *radius;
  float *depth;
  uint8_t *mask;
  int *touched;
  float4 *gacc;   // [N][3] float4
  uint32_t *ckey, *cid;   // [N] depth bits / ids of the Gaussians touching a tile, compacted (unordered)
  int *ccount;            // device int: how many
  uint2 *trec;            // [N] what the exact tile count found, for the emission (batch.cu / binning.cu k_emit_lists):
                          //   small rectangle: x = origin tile | rows << 16, y = hit mask over its <= 16 cells (tx-major)
                          //   large rectangle: x = start of its tile ids in BatchConst::tlist (0xffffffff: none), y = 1 << 31 | count
};

struct BatchConst {
  ViewConst v[GSX_MAX_BATCH];
  SlotPtrs s[GSX_MAX_BATCH];
  int nv;
  // Tile ids (16-bit) of the LARGE rectangles (> COOP_CELLS cells) of every view of the batch, append-only; small ones
  // travel as hit masks in SlotPtrs::trec.  tl_cap == 0: no buffer (large rectangles are re-enumerated by the emission).
  uint16_t *tlist;
  unsigned *tcursor;      // device counter: entries handed out so far
  unsigned tl_cap;        // entries in tlist
};

}  // namespace gsx
