// gsx_math.cuh -- per-Gaussian device math shared by the per-function (F1..F15) kernels and the fused
// preprocess kernels.  Arithmetic follows the reference expression by expression (citations per function,
// paths relative to /root/reference) so float results agree to rounding; discrete decisions (cull, window,
// SAT) use non-contracted __fmul_rn/__fadd_rn and agree bit for bit with oracle/golden.cpp.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsx {

struct Mat34 { float m[12]; };   // rows 0..2 of a row-major 4x4
struct Mat44 { float m[16]; };

// F1  cuda/projection.cu:41-44
__device__ __forceinline__ float3 cam_point(const float *v, float wx, float wy, float wz) {
  float3 c;
  c.x = v[0] * wx + v[1] * wy + v[2] * wz + v[3];
  c.y = v[4] * wx + v[5] * wy + v[6] * wz + v[7];
  c.z = v[8] * wx + v[9] * wy + v[10] * wz + v[11];
  return c;
}

// F2  cuda/projection.cu:87-97
__device__ __forceinline__ float2 screen_point(const float *p, float3 c, int width, int height) {
  const float xc = p[0] * c.x + p[1] * c.y + p[2] * c.z + p[3];
  const float yc = p[4] * c.x + p[5] * c.y + p[6] * c.z + p[7];
  const float wc = p[12] * c.x + p[13] * c.y + p[14] * c.z + p[15];
  const float xn = xc / (wc + 1e-6f);
  const float yn = yc / (wc + 1e-6f);
  float2 uv;
  uv.x = (xn * 0.5f + 0.5f) * width;
  uv.y = (yn * 0.5f + 0.5f) * height;
  return uv;
}

// F3  cuda/culling.cu:70-75 ; pure comparisons
__device__ __forceinline__ bool keep_gaussian(float u, float v, float z, float near_thresh, int padding, int width,
                                              int height) {
  return (z >= near_thresh) && (u >= (float)(-1 * padding)) && (u <= (float)(width + padding)) &&
         (v >= (float)(-1 * padding)) && (v <= (float)(height + padding));
}

// F4  cuda/gaussian.cu:14-74 ; also returns R, S for the backward
struct SigmaOut { float s[6]; };
__device__ __forceinline__ void sigma_from_qs(float qw, float qx, float qy, float qz, float sx_log, float sy_log,
                                              float sz_log, float *sig) {
  const float norm = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
  const float inv = 1.0f / (norm + 1e-6f);
  const float w = qw * inv, x = qx * inv, y = qy * inv, z = qz * inv;
  const float x2 = x * x, y2 = y * y, z2 = z * z, xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y,
              wz = w * z;
  const float r00 = 1.0f - 2.0f * (y2 + z2), r01 = 2.0f * (xy - wz), r02 = 2.0f * (xz + wy);
  const float r10 = 2.0f * (xy + wz), r11 = 1.0f - 2.0f * (x2 + z2), r12 = 2.0f * (yz - wx);
  const float r20 = 2.0f * (xz - wy), r21 = 2.0f * (yz + wx), r22 = 1.0f - 2.0f * (x2 + y2);
  const float sx = expf(sx_log), sy = expf(sy_log), sz = expf(sz_log);
  const float a00 = r00 * sx, a10 = r10 * sx, a20 = r20 * sx;
  const float a01 = r01 * sy, a11 = r11 * sy, a21 = r21 * sy;
  const float a02 = r02 * sz, a12 = r12 * sz, a22 = r22 * sz;
  sig[0] = a00 * a00 + a01 * a01 + a02 * a02;
  sig[1] = a00 * a10 + a01 * a11 + a02 * a12;
  sig[2] = a00 * a20 + a01 * a21 + a02 * a22;
  sig[3] = a10 * a10 + a11 * a11 + a12 * a12;
  sig[4] = a10 * a20 + a11 * a21 + a12 * a22;
  sig[5] = a20 * a20 + a21 * a21 + a22 * a22;
}

// F5a  cuda/gaussian.cu:187-217
__device__ __forceinline__ void projection_jacobian(float x, float y, float z, float fx, float fy, float tan_fovx,
                                                    float tan_fovy, float *J) {
  if (fabsf(z) < 1e-6f) {
#pragma unroll
    for (int k = 0; k < 6; ++k) J[k] = 0.0f;
    return;
  }
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float txtz = x / z, tytz = y / z;
  x = fminf(limx, fmaxf(-limx, txtz)) * z;
  y = fminf(limy, fmaxf(-limy, tytz)) * z;
  J[0] = fx / z;
  J[1] = 0.0f;
  J[2] = -(fx * x) / (z * z);
  J[3] = 0.0f;
  J[4] = fy / z;
  J[5] = -(fy * y) / (z * z);
}

// M = J W, V = Sigma M^T  (cuda/gaussian.cu:127-141, gaussian_backward.cu:147-161)
struct MV { float m00, m01, m02, m10, m11, m12, v00, v01, v10, v11, v20, v21; };
__device__ __forceinline__ MV compute_mv(const float *J, const float *S, const float *view) {
  const float w00 = view[0], w01 = view[1], w02 = view[2], w10 = view[4], w11 = view[5], w12 = view[6],
              w20 = view[8], w21 = view[9], w22 = view[10];
  MV o;
  o.m00 = J[0] * w00 + J[1] * w10 + J[2] * w20;
  o.m01 = J[0] * w01 + J[1] * w11 + J[2] * w21;
  o.m02 = J[0] * w02 + J[1] * w12 + J[2] * w22;
  o.m10 = J[3] * w00 + J[4] * w10 + J[5] * w20;
  o.m11 = J[3] * w01 + J[4] * w11 + J[5] * w21;
  o.m12 = J[3] * w02 + J[4] * w12 + J[5] * w22;
  const float s00 = S[0], s01 = S[1], s02 = S[2], s11 = S[3], s12 = S[4], s22 = S[5];
  o.v00 = s00 * o.m00 + s01 * o.m01 + s02 * o.m02;
  o.v01 = s00 * o.m10 + s01 * o.m11 + s02 * o.m12;
  o.v10 = s01 * o.m00 + s11 * o.m01 + s12 * o.m02;
  o.v11 = s01 * o.m10 + s11 * o.m11 + s12 * o.m12;
  o.v20 = s02 * o.m00 + s12 * o.m01 + s22 * o.m02;
  o.v21 = s02 * o.m10 + s12 * o.m11 + s22 * o.m12;
  return o;
}

// F5b  cuda/gaussian.cu:143-174
__device__ __forceinline__ void conic_and_radius(const float *J, const float *S, const float *view, float mh_dist,
                                                 float *conic, float4 *radius) {
  const MV t = compute_mv(J, S, view);
  const float cov00 = t.m00 * t.v00 + t.m01 * t.v10 + t.m02 * t.v20 + 0.3f;
  const float cov01 = t.m00 * t.v01 + t.m01 * t.v11 + t.m02 * t.v21;
  const float cov11 = t.m10 * t.v01 + t.m11 * t.v11 + t.m12 * t.v21 + 0.3f;
  const float det = cov00 * cov11 - cov01 * cov01;
  const float inv_det = 1.0f / det;
  conic[0] = cov11 * inv_det;
  conic[1] = -cov01 * inv_det;
  conic[2] = cov00 * inv_det;
  const float mid = 0.5f * (cov00 + cov11);
  const float lt = sqrtf(fmaxf(0.1f, mid * mid - det));
  const float l1 = mid + lt, l2 = mid - lt;
  float s, c;
  sincosf(0.5f * atan2f(2.0f * cov01, cov00 - cov11), &s, &c);
  *radius = make_float4(ceilf(mh_dist * sqrtf(l1)), ceilf(mh_dist * sqrtf(l2)), s, c);
}

// Real spherical harmonics for l <= 3 in sphericart's convention (k = l*l+l+m, l=1 order y,z,x;
// submodules/sphericart/sphericart/include/macros.hpp:51,76-79,133-141,161-172), evaluated at the normalised
// point as sphericart::cuda::SphericalHarmonics does (src/sphericart_cuda.cpp:51).  Closed forms replace the
// NVRTC-JIT sphericart kernel.  With GRAD, also the gradient w.r.t. the un-normalised input,
// (g - r (r.g)) * ir (src/sphericart_impl.cu:178-186), as separate x/y/z arrays.
template <int L, bool GRAD>
__device__ __forceinline__ void sph_eval(float px, float py, float pz, float *Y, float *gx, float *gy, float *gz) {
  const float ir2 = 1.0f / (px * px + py * py + pz * pz);
  const float ir = sqrtf(ir2);
  const float x = px * ir, y = py * ir, z = pz * ir;
  const float x2 = x * x, y2 = y * y, z2 = z * z;
  Y[0] = 0.282094791773878f;
  if (GRAD) { gx[0] = 0.f; gy[0] = 0.f; gz[0] = 0.f; }
  if (L >= 1) {
    const float c1 = 0.48860251190292f;
    Y[1] = c1 * y; Y[2] = c1 * z; Y[3] = c1 * x;
    if (GRAD) {
      gx[1] = 0.f; gy[1] = c1; gz[1] = 0.f;
      gx[2] = 0.f; gy[2] = 0.f; gz[2] = c1;
      gx[3] = c1; gy[3] = 0.f; gz[3] = 0.f;
    }
  }
  if (L >= 2) {
    const float a = 1.0925484305920792f, b = 0.31539156525252005f, c = 0.5462742152960396f;
    Y[4] = a * x * y; Y[5] = a * y * z; Y[6] = b * (2.f * z2 - x2 - y2); Y[7] = a * x * z; Y[8] = c * (x2 - y2);
    if (GRAD) {
      gx[4] = a * y;        gy[4] = a * x;        gz[4] = 0.f;
      gx[5] = 0.f;          gy[5] = a * z;        gz[5] = a * y;
      gx[6] = -2.f * b * x; gy[6] = -2.f * b * y; gz[6] = 4.f * b * z;
      gx[7] = a * z;        gy[7] = 0.f;          gz[7] = a * x;
      gx[8] = 2.f * c * x;  gy[8] = -2.f * c * y; gz[8] = 0.f;
    }
  }
  if (L >= 3) {
    const float a = 0.5900435899266435f, b = 2.890611442640554f, c = 0.4570457994644658f, d = 0.3731763325901154f,
                e = 1.445305721320277f;
    Y[9] = a * y * (3.f * x2 - y2);
    Y[10] = b * x * y * z;
    Y[11] = c * y * (4.f * z2 - x2 - y2);
    Y[12] = d * z * (2.f * z2 - 3.f * x2 - 3.f * y2);
    Y[13] = c * x * (4.f * z2 - x2 - y2);
    Y[14] = e * z * (x2 - y2);
    Y[15] = a * x * (x2 - 3.f * y2);
    if (GRAD) {
      gx[9] = 6.f * a * x * y;            gy[9] = 3.f * a * (x2 - y2);             gz[9] = 0.f;
      gx[10] = b * y * z;                 gy[10] = b * x * z;                      gz[10] = b * x * y;
      gx[11] = -2.f * c * x * y;          gy[11] = c * (4.f * z2 - x2 - 3.f * y2); gz[11] = 8.f * c * y * z;
      gx[12] = -6.f * d * x * z;          gy[12] = -6.f * d * y * z;               gz[12] = d * (6.f * z2 - 3.f * x2 - 3.f * y2);
      gx[13] = c * (4.f * z2 - 3.f * x2 - y2); gy[13] = -2.f * c * x * y;          gz[13] = 8.f * c * x * z;
      gx[14] = 2.f * e * x * z;           gy[14] = -2.f * e * y * z;               gz[14] = e * (x2 - y2);
      gx[15] = 3.f * a * (x2 - y2);       gy[15] = -6.f * a * x * y;               gz[15] = 0.f;
    }
  }
  if (GRAD) {
    constexpr int n = (L + 1) * (L + 1);
#pragma unroll
    for (int k = 0; k < n; ++k) {
      const float t = gx[k] * x + gy[k] * y + gz[k] * z;
      gx[k] = (gx[k] - x * t) * ir;
      gy[k] = (gy[k] - y * t) * ir;
      gz[k] = (gz[k] - z * t) * ir;
    }
  }
}

// F6  cuda/spherical_harmonics.cu:16-25 (dir) + :42-59 (rgb).  sh points at this Gaussian's (n-1)*3 floats.
template <int L>
__device__ __forceinline__ float3 sh_to_rgb(float px, float py, float pz, float3 campos, const float *band0,
                                            const float *__restrict__ sh) {
  const float dx = px - campos.x, dy = py - campos.y, dz = pz - campos.z;
  const float len = sqrtf(dx * dx + dy * dy + dz * dz) + 1e-9f;
  constexpr int n = (L + 1) * (L + 1);
  float Y[n];
  sph_eval<L, false>(dx / len, dy / len, dz / len, Y, nullptr, nullptr, nullptr);
  float r = band0[0] * Y[0] + 0.5f, g = band0[1] * Y[0] + 0.5f, b = band0[2] * Y[0] + 0.5f;
#pragma unroll
  for (int k = 0; k < n - 1; ++k) {
    r += sh[3 * k + 0] * Y[k + 1];
    g += sh[3 * k + 1] * Y[k + 1];
    b += sh[3 * k + 2] * Y[k + 1];
  }
  return make_float3(r, g, b);
}

// ---- F7 geometry: non-contracted arithmetic, bit-exact with oracle/golden.cpp -----------------------------
__device__ __forceinline__ float mul_(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_(float a, float b) { return __fsub_rn(a, b); }

struct Obb { float p[8]; };
// cuda/culling.cu:148-165
__device__ __forceinline__ Obb compute_obb(float u, float v, float r_major, float r_minor, float s, float c) {
  const float v1x = mul_(r_major, c), v1y = mul_(r_major, s), v2x = mul_(-r_minor, s), v2y = mul_(r_minor, c);
  Obb o;
  o.p[0] = sub_(sub_(u, v1x), v2x); o.p[1] = sub_(sub_(v, v1y), v2y);
  o.p[2] = sub_(add_(u, v1x), v2x); o.p[3] = sub_(add_(v, v1y), v2y);
  o.p[4] = add_(sub_(u, v1x), v2x); o.p[5] = add_(sub_(v, v1y), v2y);
  o.p[6] = add_(add_(u, v1x), v2x); o.p[7] = add_(add_(v, v1y), v2y);
  return o;
}
// Per-Gaussian part of the SAT (everything that does not depend on the tile), cuda/culling.cu:97-146
struct ObbSat {
  float min_x, max_x, min_y, max_y;   // axis 0/1 extents
  float a2x, a2y, lo2, hi2;           // major axis + OBB projection interval
  float a3x, a3y, lo3, hi3;           // minor axis + OBB projection interval
};
__device__ __forceinline__ ObbSat obb_sat_setup(const Obb &o) {
  ObbSat s;
  s.min_x = fminf(fminf(o.p[0], o.p[2]), fminf(o.p[4], o.p[6]));
  s.max_x = fmaxf(fmaxf(o.p[0], o.p[2]), fmaxf(o.p[4], o.p[6]));
  s.min_y = fminf(fminf(o.p[1], o.p[3]), fminf(o.p[5], o.p[7]));
  s.max_y = fmaxf(fmaxf(o.p[1], o.p[3]), fmaxf(o.p[5], o.p[7]));
  s.a2x = sub_(o.p[2], o.p[0]); s.a2y = sub_(o.p[3], o.p[1]);
  const float pr = add_(mul_(s.a2x, o.p[2]), mul_(s.a2y, o.p[3]));
  const float pl = add_(mul_(s.a2x, o.p[0]), mul_(s.a2y, o.p[1]));
  s.lo2 = fminf(pr, pl); s.hi2 = fmaxf(pr, pl);
  s.a3x = sub_(o.p[2], o.p[6]); s.a3y = sub_(o.p[3], o.p[7]);
  const float pt = add_(mul_(s.a3x, o.p[2]), mul_(s.a3y, o.p[3]));
  const float pb = add_(mul_(s.a3x, o.p[6]), mul_(s.a3y, o.p[7]));
  s.lo3 = fminf(pt, pb); s.hi3 = fmaxf(pt, pb);
  return s;
}
__device__ __forceinline__ bool sat_axis(float ax, float ay, float lo, float hi, float l, float r, float t, float b) {
  const float tl = add_(mul_(ax, l), mul_(ay, t)), tr = add_(mul_(ax, r), mul_(ay, t));
  const float bl = add_(mul_(ax, l), mul_(ay, b)), br = add_(mul_(ax, r), mul_(ay, b));
  const float mn = fminf(fminf(tl, tr), fminf(bl, br)), mx = fmaxf(fmaxf(tl, tr), fmaxf(bl, br));
  return !(mn > hi || mx < lo);
}
__device__ __forceinline__ bool sat_tile(const ObbSat &s, int tx, int ty) {
  const float l = (float)tx * 16.0f, r = (float)(tx + 1) * 16.0f, t = (float)ty * 16.0f, b = (float)(ty + 1) * 16.0f;
  if (s.min_x > r || s.max_x < l) return false;
  if (s.min_y > b || s.max_y < t) return false;
  if (!sat_axis(s.a2x, s.a2y, s.lo2, s.hi2, l, r, t, b)) return false;
  return sat_axis(s.a3x, s.a3y, s.lo3, s.hi3, l, r, t, b);
}
// candidate window, cuda/culling.cu:209-224
struct TileWindow { int x0, x1, y0, y1; };
__device__ __forceinline__ TileWindow candidate_window(float u, float v, float r_major, int ntx, int nty) {
  const int rt = (int)(add_(ceilf(mul_(r_major, 0.0625f)), 1.0f));
  const int ptx = (int)floorf(__fdiv_rn(u, 16.0f)), pty = (int)floorf(__fdiv_rn(v, 16.0f));
  TileWindow w;
  w.x0 = max(0, ptx - rt); w.x1 = min(ntx, ptx + rt + 1);
  w.y0 = max(0, pty - rt); w.y1 = min(nty, pty + rt + 1);
  return w;
}

// Exact sub-range of the candidate window that survives SAT axes 0 and 1 (the OBB's own AABB against the tile
// rows/columns).  The reference rejects tile column tx iff  min_x > 16(tx+1) || max_x < 16 tx  (culling.cu:103-110,
// inclusive bounds); with exact power-of-two scaling that is  ceil(min_x/16)-1 <= tx <= floor(max_x/16).  NaN
// bounds constrain nothing (fmaxf/fminf drop them), exactly as the reference's false comparisons do.  Small
// Gaussians thus test 1-4 tiles against axes 2 and 3 instead of >= 25.
struct TileRange { int x0, x1, y0, y1; };   // inclusive; empty when x1 < x0 or y1 < y0
__device__ __forceinline__ TileRange sat_aabb_range(const ObbSat &s, const TileWindow &w) {
  TileRange r;
  const float fx0 = (float)w.x0, fx1 = (float)(w.x1 - 1), fy0 = (float)w.y0, fy1 = (float)(w.y1 - 1);
  r.x0 = (int)fminf(fmaxf(ceilf(s.min_x * 0.0625f) - 1.0f, fx0), fx1 + 1.0f);
  r.x1 = (int)fmaxf(fminf(floorf(s.max_x * 0.0625f), fx1), fx0 - 1.0f);
  r.y0 = (int)fminf(fmaxf(ceilf(s.min_y * 0.0625f) - 1.0f, fy0), fy1 + 1.0f);
  r.y1 = (int)fmaxf(fminf(floorf(s.max_y * 0.0625f), fy1), fy0 - 1.0f);
  return r;
}
// GSX_LISTS_CONTRIBUTING: keep only the tile columns / rows the half-extents (ex, ey) of {alpha >= 1/255} around (u, v)
// can reach.  Compositing rejects an entry for a sub-tile [X0, X0+7] when u + ex < X0 or u - ex > X0 + 7
// (composite.cu); over the sub-tiles of tile column tx that leaves 16 tx <= u + ex and u - ex <= 16 tx + 15, and
// floor((u - ex)/16) <= tx is a superset of the latter -- so no (tile, entry) with a contributing pixel is dropped.
// NaN constrains nothing, +inf extents (conic not positive definite) keep everything, -inf extents (opacity below
// the threshold) leave an empty range.  The count (batch.cu) and the emission (binning.cu) both call this on the same
// stored floats, so the scanned offsets match the emitted instances.
__device__ __forceinline__ void clip_range_to_extent(TileRange &r, float u, float v, float ex, float ey) {
  const float big = 1.0e6f;
  r.x0 = max(r.x0, (int)fminf(fmaxf(floorf((u - ex) * 0.0625f), -big), big));
  r.x1 = min(r.x1, (int)fmaxf(fminf(floorf((u + ex) * 0.0625f), big), -big));
  r.y0 = max(r.y0, (int)fminf(fmaxf(floorf((v - ey) * 0.0625f), -big), big));
  r.y1 = min(r.y1, (int)fmaxf(fminf(floorf((v + ey) * 0.0625f), big), -big));
}
// Axes 2 and 3 only (axes 0/1 already decided by sat_aabb_range), separable form.  The reference takes the min / max of
// the four corner projections fl(X + Y), X in {ax l, ax r}, Y in {ay t, ay b} (culling.cu:112-146).  Rounding is
// monotone, so those extremes are exactly fl(min X + min Y) and fl(max X + max Y): the X part is per tile column, the Y
// part per tile row, and a tile costs 2 multiplies + 2 min/max + 2 adds per axis instead of 4 + 6 + 4 -- with the same
// bits as sat_axis (tests: per-tile lists bit-exact against the oracle and against the reference's own kernels).
struct SatCol { float x2mn, x2mx, x3mn, x3mx; };
__device__ __forceinline__ SatCol sat_col(const ObbSat &s, int tx) {
  const float l = (float)tx * 16.0f, r = (float)(tx + 1) * 16.0f;
  const float a = mul_(s.a2x, l), b = mul_(s.a2x, r), c = mul_(s.a3x, l), d = mul_(s.a3x, r);
  SatCol o;
  o.x2mn = fminf(a, b); o.x2mx = fmaxf(a, b); o.x3mn = fminf(c, d); o.x3mx = fmaxf(c, d);
  return o;
}
__device__ __forceinline__ bool sat_cell(const ObbSat &s, const SatCol &c, int ty) {
  const float t = (float)ty * 16.0f, b = (float)(ty + 1) * 16.0f;
  const float y2t = mul_(s.a2y, t), y2b = mul_(s.a2y, b);
  if (add_(c.x2mn, fminf(y2t, y2b)) > s.hi2 || add_(c.x2mx, fmaxf(y2t, y2b)) < s.lo2) return false;
  const float y3t = mul_(s.a3y, t), y3b = mul_(s.a3y, b);
  return !(add_(c.x3mn, fminf(y3t, y3b)) > s.hi3 || add_(c.x3mx, fmaxf(y3t, y3b)) < s.lo3);
}
__device__ __forceinline__ bool sat_tile_rot(const ObbSat &s, int tx, int ty) { return sat_cell(s, sat_col(s, tx), ty); }

// Warp-cooperative evaluation of ONE lane's tile rectangle (Gaussians covering many tiles would otherwise keep a
// single lane busy for hundreds of SAT tests while 31 wait): the leader's SAT constants and rectangle are broadcast,
// the 32 lanes test 32 cells per step (same tx-major order as the serial loop) and fn(c0, hit, ballot) sees each step.
// Must be called by all 32 lanes with the same `leader`.
#ifndef GSX_COOP_CELLS
#define GSX_COOP_CELLS 16
#endif
constexpr int COOP_CELLS = GSX_COOP_CELLS;   // rectangles with more cells than this are handled cooperatively
template <typename F>
__device__ __forceinline__ void coop_tile_rect(int leader, const ObbSat &so, const TileRange &tr, F &&fn) {
  const unsigned full = 0xffffffffu;
  ObbSat b;
  b.min_x = b.max_x = b.min_y = b.max_y = 0.0f;   // axes 0/1 were decided by sat_aabb_range
  b.a2x = __shfl_sync(full, so.a2x, leader); b.a2y = __shfl_sync(full, so.a2y, leader);
  b.lo2 = __shfl_sync(full, so.lo2, leader); b.hi2 = __shfl_sync(full, so.hi2, leader);
  b.a3x = __shfl_sync(full, so.a3x, leader); b.a3y = __shfl_sync(full, so.a3y, leader);
  b.lo3 = __shfl_sync(full, so.lo3, leader); b.hi3 = __shfl_sync(full, so.hi3, leader);
  const int x0 = __shfl_sync(full, tr.x0, leader), y0 = __shfl_sync(full, tr.y0, leader);
  const int ny = __shfl_sync(full, tr.y1 - tr.y0 + 1, leader);
  const int cells = __shfl_sync(full, (tr.x1 - tr.x0 + 1) * (tr.y1 - tr.y0 + 1), leader);
  const int lane = threadIdx.x & 31;
  for (int c0 = 0; c0 < cells; c0 += 32) {
    const int c = c0 + lane;
    bool hit = false;
    int tx = 0, ty = 0;
    if (c < cells) {
      tx = x0 + c / ny; ty = y0 + c - (c / ny) * ny;
      hit = sat_tile_rot(b, tx, ty);
    }
    fn(hit, __ballot_sync(full, hit), tx, ty);
  }
}
__device__ __forceinline__ int rect_cells(const TileRange &tr) {
  return (tr.x1 >= tr.x0 && tr.y1 >= tr.y0) ? (tr.x1 - tr.x0 + 1) * (tr.y1 - tr.y0 + 1) : 0;
}

// order-preserving map float -> uint32 (handles negative depths that only the per-function F7 can see)
__device__ __forceinline__ uint32_t float_sort_bits(float z) {
  const uint32_t b = __float_as_uint(z);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// ---- backward per-Gaussian math ----------------------------------------------------------------------------
// F11  cuda/gaussian_backward.cu:163-247 ; returns dSigma[6], dJ[6] (to be += by the caller)
__device__ __forceinline__ void conic_backward(const float *J, const float *S, const float *view, const float *conic,
                                               const float *gc, float *dS, float *dJ) {
  const MV t = compute_mv(J, S, view);
  const float c00 = conic[0], c01 = conic[1], c11 = conic[2];
  const float t00 = c00 * gc[0] + c01 * gc[1], t01 = c00 * gc[1] + c01 * gc[2];
  const float t10 = c01 * gc[0] + c11 * gc[1], t11 = c01 * gc[1] + c11 * gc[2];
  const float d00 = -(t00 * c00 + t01 * c01);
  const float d01 = -(t00 * c01 + t01 * c11);
  const float d11 = -(t10 * c01 + t11 * c11);
  const float dv00 = d00 * t.m00 + d01 * t.m10, dv01 = d01 * t.m00 + d11 * t.m10;
  const float dv10 = d00 * t.m01 + d01 * t.m11, dv11 = d01 * t.m01 + d11 * t.m11;
  const float dv20 = d00 * t.m02 + d01 * t.m12, dv21 = d01 * t.m02 + d11 * t.m12;
  dS[0] = dv00 * t.m00 + dv01 * t.m10;
  dS[1] = dv00 * t.m01 + dv01 * t.m11 + dv10 * t.m00 + dv11 * t.m10;
  dS[2] = dv00 * t.m02 + dv01 * t.m12 + dv20 * t.m00 + dv21 * t.m10;
  dS[3] = dv10 * t.m01 + dv11 * t.m11;
  dS[4] = dv10 * t.m02 + dv11 * t.m12 + dv20 * t.m01 + dv21 * t.m11;
  dS[5] = dv20 * t.m02 + dv21 * t.m12;
  const float s00 = S[0], s01 = S[1], s02 = S[2], s11 = S[3], s12 = S[4], s22 = S[5];
  const float dm00 = (d00 * t.v00 + d01 * t.v01) + (dv00 * s00 + dv10 * s01 + dv20 * s02);
  const float dm01 = (d00 * t.v10 + d01 * t.v11) + (dv00 * s01 + dv10 * s11 + dv20 * s12);
  const float dm02 = (d00 * t.v20 + d01 * t.v21) + (dv00 * s02 + dv10 * s12 + dv20 * s22);
  const float dm10 = (d01 * t.v00 + d11 * t.v01) + (dv01 * s00 + dv11 * s01 + dv21 * s02);
  const float dm11 = (d01 * t.v10 + d11 * t.v11) + (dv01 * s01 + dv11 * s11 + dv21 * s12);
  const float dm12 = (d01 * t.v20 + d11 * t.v21) + (dv01 * s02 + dv11 * s12 + dv21 * s22);
  const float w00 = view[0], w01 = view[1], w02 = view[2], w10 = view[4], w11 = view[5], w12 = view[6],
              w20 = view[8], w21 = view[9], w22 = view[10];
  dJ[0] = dm00 * w00 + dm01 * w01 + dm02 * w02;
  dJ[1] = dm00 * w10 + dm01 * w11 + dm02 * w12;
  dJ[2] = dm00 * w20 + dm01 * w21 + dm02 * w22;
  dJ[3] = dm10 * w00 + dm11 * w01 + dm12 * w02;
  dJ[4] = dm10 * w10 + dm11 * w11 + dm12 * w12;
  dJ[5] = dm10 * w20 + dm11 * w21 + dm12 * w22;
}

// F12  cuda/gaussian_backward.cu:20-77 ; returns false when |z|<1e-6 (reference leaves the output untouched)
__device__ __forceinline__ bool jacobian_backward(float x, float y, float z, float fx, float fy, float tan_fovx,
                                                  float tan_fovy, const float *dJ, float3 *out) {
  if (fabsf(z) < 1e-6f) return false;
  const float zi = 1.0f / (z + 1e-6f), zi2 = zi * zi, zi3 = zi2 * zi;
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float txtz = x * zi, tytz = y * zi;
  float dx = 0.f, dy = 0.f, dz = 0.f;
  dz += dJ[0] * (-fx * zi2);
  if (fabsf(txtz) <= limx) {
    dx += dJ[2] * (-fx * zi2);
    dz += dJ[2] * (2.0f * fx * x * zi3);
  } else {
    const float cx = (txtz > 0.0f ? limx : -limx);
    dz += dJ[2] * (fx * cx * zi2);
  }
  dz += dJ[4] * (-fy * zi2);
  if (fabsf(tytz) <= limy) {
    dy += dJ[5] * (-fy * zi2);
    dz += dJ[5] * (2.0f * fy * y * zi3);
  } else {
    const float cy = (tytz > 0.0f ? limy : -limy);
    dz += dJ[5] * (fy * cy * zi2);
  }
  *out = make_float3(dx, dy, dz);
  return true;
}

// F13  cuda/gaussian_backward.cu:277-414
__device__ __forceinline__ void sigma_backward(float qw, float qx, float qy, float qz, float s0, float s1, float s2,
                                               const float *g, float *dQ, float *dSc) {
  const float norm = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
  const float inv = 1.0f / (norm + 1e-6f);
  const float w = qw * inv, x = qx * inv, y = qy * inv, z = qz * inv;
  const float Sx = expf(s0), Sy = expf(s1), Sz = expf(s2);
  float R[9];
  R[0] = 1.0f - 2.0f * (y * y + z * z); R[1] = 2.0f * (x * y - w * z); R[2] = 2.0f * (x * z + w * y);
  R[3] = 2.0f * (x * y + w * z); R[4] = 1.0f - 2.0f * (x * x + z * z); R[5] = 2.0f * (y * z - w * x);
  R[6] = 2.0f * (x * z - w * y); R[7] = 2.0f * (y * z + w * x); R[8] = 1.0f - 2.0f * (x * x + y * y);
  float M[9];
#pragma unroll
  for (int r = 0; r < 3; ++r) { M[3 * r] = R[3 * r] * Sx; M[3 * r + 1] = R[3 * r + 1] * Sy; M[3 * r + 2] = R[3 * r + 2] * Sz; }
  float D[9];
  D[0] = g[0]; D[1] = 0.5f * g[1]; D[2] = 0.5f * g[2];
  D[3] = 0.5f * g[1]; D[4] = g[3]; D[5] = 0.5f * g[4];
  D[6] = 0.5f * g[2]; D[7] = 0.5f * g[4]; D[8] = g[5];
  float dM[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c)
      dM[3 * r + c] = 2.0f * (D[3 * r] * M[c] + D[3 * r + 1] * M[3 + c] + D[3 * r + 2] * M[6 + c]);
  float dR[9];
#pragma unroll
  for (int r = 0; r < 3; ++r) { dR[3 * r] = dM[3 * r] * Sx; dR[3 * r + 1] = dM[3 * r + 1] * Sy; dR[3 * r + 2] = dM[3 * r + 2] * Sz; }
  dSc[0] = (R[0] * dM[0] + R[3] * dM[3] + R[6] * dM[6]) * Sx;
  dSc[1] = (R[1] * dM[1] + R[4] * dM[4] + R[7] * dM[7]) * Sy;
  dSc[2] = (R[2] * dM[2] + R[5] * dM[5] + R[8] * dM[8]) * Sz;
  float dw = 0.f, dx = 0.f, dy = 0.f, dz = 0.f;
  dw += dR[1] * (-2.0f * z) + dR[2] * (2.0f * y);
  dx += dR[1] * (2.0f * y) + dR[2] * (2.0f * z);
  dy += dR[0] * (-4.0f * y) + dR[1] * (2.0f * x) + dR[2] * (2.0f * w);
  dz += dR[0] * (-4.0f * z) + dR[1] * (-2.0f * w) + dR[2] * (2.0f * x);
  dw += dR[3] * (2.0f * z) + dR[5] * (-2.0f * x);
  dx += dR[3] * (2.0f * y) + dR[4] * (-4.0f * x) + dR[5] * (-2.0f * w);
  dy += dR[3] * (2.0f * x) + dR[5] * (2.0f * z);
  dz += dR[3] * (2.0f * w) + dR[4] * (-4.0f * z) + dR[5] * (2.0f * y);
  dw += dR[6] * (-2.0f * y) + dR[7] * (2.0f * x);
  dx += dR[6] * (2.0f * z) + dR[7] * (2.0f * w) + dR[8] * (-4.0f * x);
  dy += dR[6] * (-2.0f * w) + dR[7] * (2.0f * z) + dR[8] * (-4.0f * y);
  dz += dR[6] * (2.0f * x) + dR[7] * (2.0f * y);
  const float dot = w * dw + x * dx + y * dy + z * dz;
  dQ[0] = inv * (dw - dot * w);
  dQ[1] = inv * (dx - dot * x);
  dQ[2] = inv * (dy - dot * y);
  dQ[3] = inv * (dz - dot * z);
}

// F14  cuda/projection_backward.cu:41-74 ; returns false when |w|<1e-6
__device__ __forceinline__ bool screen_backward(const float *p, float x, float y, float z, float gu, float gv,
                                                int width, int height, float3 *out) {
  const float xc = p[0] * x + p[1] * y + p[2] * z + p[3];
  const float yc = p[4] * x + p[5] * y + p[6] * z + p[7];
  const float wc = p[12] * x + p[13] * y + p[14] * z + p[15];
  if (fabsf(wc) < 1e-6f) return false;
  const float wi = 1.0f / wc, wi2 = wi * wi;
  const float dxn = gu * width * 0.5f, dyn = gv * height * 0.5f;
  const float dxc = dxn * wi, dyc = dyn * wi;
  const float dwc = -dxn * xc * wi2 - dyn * yc * wi2;
  const float dzc = 0.0f;
  out->x = p[0] * dxc + p[4] * dyc + p[8] * dzc + p[12] * dwc;
  out->y = p[1] * dxc + p[5] * dyc + p[9] * dzc + p[13] * dwc;
  out->z = p[2] * dxc + p[6] * dyc + p[10] * dzc + p[14] * dwc;
  return true;
}

// F15  cuda/projection_backward.cu:133-136
__device__ __forceinline__ float3 cam_point_backward(const float *v, float gx, float gy, float gz) {
  return make_float3(v[0] * gx + v[4] * gy + v[8] * gz, v[1] * gx + v[5] * gy + v[9] * gz,
                     v[2] * gx + v[6] * gy + v[10] * gz);
}

// F10  cuda/spherical_harmonics_backward.cu:28-166.  Writes coefficient gradients (=) through sh_grad/b0_grad
// (may be strided), returns d xyz_w (to be +=).  COMPAT reproduces the reference's [k][axis] reading of the
// [axis][k] derivative table (Q12).
template <int L>
__device__ __forceinline__ float3 sh_backward(float px, float py, float pz, float3 campos, const float *band0,
                                              const float *sh /* may alias sh_grad: row k is read before it is written */, float gr, float gg, float gb, bool compat,
                                              float *sh_grad, float *b0_grad, bool accumulate) {
  constexpr int n = (L + 1) * (L + 1);
  const float dx = px - campos.x, dy = py - campos.y, dz = pz - campos.z;
  const float len = sqrtf(dx * dx + dy * dy + dz * dz) + 1e-9f;
  const float dirx = dx / len, diry = dy / len, dirz = dz / len;
  float Y[n], D[3 * n];   // D = [axis][k]
  sph_eval<L, true>(dirx, diry, dirz, Y, D, D + n, D + 2 * n);
  if (accumulate) { b0_grad[0] += gr * Y[0]; b0_grad[1] += gg * Y[0]; b0_grad[2] += gb * Y[0]; }
  else { b0_grad[0] = gr * Y[0]; b0_grad[1] = gg * Y[0]; b0_grad[2] = gb * Y[0]; }
  float dR[3] = {0.f, 0.f, 0.f}, dG[3] = {0.f, 0.f, 0.f}, dB[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float R, G, B;
    if (k == 0) { R = band0[0]; G = band0[1]; B = band0[2]; }
    else {
      R = sh[3 * (k - 1)]; G = sh[3 * (k - 1) + 1]; B = sh[3 * (k - 1) + 2];
      float *o = sh_grad + 3 * (k - 1);
      if (accumulate) { o[0] += gr * Y[k]; o[1] += gg * Y[k]; o[2] += gb * Y[k]; }
      else { o[0] = gr * Y[k]; o[1] = gg * Y[k]; o[2] = gb * Y[k]; }
    }
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
      const float d = compat ? D[k * 3 + ax] : D[ax * n + k];
      dR[ax] += d * R; dG[ax] += d * G; dB[ax] += d * B;
    }
  }
  const float tx = gr * dR[0] + gg * dG[0] + gb * dB[0];
  const float ty = gr * dR[1] + gg * dG[1] + gb * dB[1];
  const float tz = gr * dR[2] + gg * dG[2] + gb * dB[2];
  const float dot = tx * dirx + ty * diry + tz * dirz;
  return make_float3((tx - dot * dirx) / len, (ty - dot * diry) / len, (tz - dot * dirz) / len);
}

// The position term of F10 alone (cuda/spherical_harmonics_backward.cu:106-164): d xyz_w through the view direction,
// without the coefficient gradients (the batch kernel accumulates those per Gaussian, batch.cu).  Same arithmetic
// and order as sh_backward.
template <int L>
__device__ __forceinline__ float3 sh_dir_backward(float px, float py, float pz, float3 campos, const float *band0,
                                                  const float *sh, float gr, float gg, float gb, bool compat) {
  constexpr int n = (L + 1) * (L + 1);
  const float dx = px - campos.x, dy = py - campos.y, dz = pz - campos.z;
  const float len = sqrtf(dx * dx + dy * dy + dz * dz) + 1e-9f;
  const float dirx = dx / len, diry = dy / len, dirz = dz / len;
  float Y[n], D[3 * n];   // D = [axis][k]
  sph_eval<L, true>(dirx, diry, dirz, Y, D, D + n, D + 2 * n);
  float dR[3] = {0.f, 0.f, 0.f}, dG[3] = {0.f, 0.f, 0.f}, dB[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < n; ++k) {
    float R, G, B;
    if (k == 0) { R = band0[0]; G = band0[1]; B = band0[2]; }
    else { R = sh[3 * (k - 1)]; G = sh[3 * (k - 1) + 1]; B = sh[3 * (k - 1) + 2]; }
#pragma unroll
    for (int ax = 0; ax < 3; ++ax) {
      const float d = compat ? D[k * 3 + ax] : D[ax * n + k];
      dR[ax] += d * R; dG[ax] += d * G; dB[ax] += d * B;
    }
  }
  const float tx = gr * dR[0] + gg * dG[0] + gb * dB[0];
  const float ty = gr * dR[1] + gg * dG[1] + gb * dB[1];
  const float tz = gr * dR[2] + gg * dG[2] + gb * dB[2];
  const float dot = tx * dirx + ty * diry + tz * dirz;
  return make_float3((tx - dot * dirx) / len, (ty - dot * diry) / len, (tz - dot * dirz) / len);
}

// Y_k of the view direction of a world point (the basis values F6 and the coefficient part of F10 use)
template <int L>
__device__ __forceinline__ void sh_basis_of_point(float px, float py, float pz, float3 campos, float *Y) {
  const float dx = px - campos.x, dy = py - campos.y, dz = pz - campos.z;
  const float len = sqrtf(dx * dx + dy * dy + dz * dz) + 1e-9f;
  sph_eval<L, false>(dx / len, dy / len, dz / len, Y, nullptr, nullptr, nullptr);
}

}  // namespace gsx
