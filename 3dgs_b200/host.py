"""Host-side mirror of the reference's operator interface for the splatting hot path.

* per-function wrappers named after the reference's `gsplat_cuda` functions
  (include/gsplat_cuda/cuda_forward.cuh, cuda_backward.cuh), taking CUDA torch tensors in the reference's layouts;
* `Rasterizer`: the view-level path -- what `rasterize_image` (cuda/raster.cu:12) and `TrainerImpl::backward_pass`
  (cuda/trainer.cu:926) do in the reference -- on a persistent context, with global-index gradients.

Everything goes through the C ABI of libgsx.so; torch only owns memory and streams.
"""
import ctypes as C

import torch

from . import _ffi
from ._ffi import GsxGrads, GsxParams, GsxState, GsxStats, GsxView, check, f32, lib


def _ptr(t):
    if t is None:
        return C.c_void_p(0)
    assert t.is_cuda and t.is_contiguous(), "expected a contiguous CUDA tensor"
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _chk_f32(*ts):
    for t in ts:
        assert t is None or t.dtype == torch.float32


# ------------------------------------------------------------------------------------------------ forward F1..F8
def compute_camera_space_points(xyz_w, view, xyz_c=None):
    n = xyz_w.shape[0]
    xyz_c = torch.empty((n, 3), device=xyz_w.device, dtype=torch.float32) if xyz_c is None else xyz_c
    check(lib().gsx_compute_camera_space_points(_ptr(xyz_w), _ptr(view), n, _ptr(xyz_c), _stream()))
    return xyz_c


def project_to_screen(xyz_c, proj, width, height, uv=None):
    n = xyz_c.shape[0]
    uv = torch.empty((n, 2), device=xyz_c.device, dtype=torch.float32) if uv is None else uv
    check(lib().gsx_project_to_screen(_ptr(xyz_c), _ptr(proj), n, int(width), int(height), _ptr(uv), _stream()))
    return uv


def cull_gaussians(uv, xyz_c, near_thresh, padding, width, height, mask=None):
    n = uv.shape[0]
    mask = torch.empty(n, device=uv.device, dtype=torch.uint8) if mask is None else mask
    check(lib().gsx_cull_gaussians(_ptr(uv), _ptr(xyz_c), n, f32(near_thresh), int(padding), int(width), int(height),
                                   _ptr(mask), _stream()))
    return mask


def compute_sigma(quaternion, scale, sigma=None):
    n = quaternion.shape[0]
    sigma = torch.empty((n, 6), device=quaternion.device, dtype=torch.float32) if sigma is None else sigma
    check(lib().gsx_compute_sigma(_ptr(quaternion), _ptr(scale), n, _ptr(sigma), _stream()))
    return sigma


def compute_conic(xyz_c, view, sigma, focal_x, focal_y, tan_fovx, tan_fovy, mh_dist):
    n = xyz_c.shape[0]
    dev = xyz_c.device
    J = torch.empty((n, 6), device=dev, dtype=torch.float32)
    conic = torch.empty((n, 3), device=dev, dtype=torch.float32)
    radius = torch.empty((n, 4), device=dev, dtype=torch.float32)
    check(lib().gsx_compute_conic(_ptr(xyz_c), _ptr(view), _ptr(sigma), f32(focal_x), f32(focal_y), f32(tan_fovx),
                                  f32(tan_fovy), f32(mh_dist), n, _ptr(J), _ptr(conic), _ptr(radius), _stream()))
    return J, conic, radius


def precompute_spherical_harmonics(xyz_w, sh, band0, campos, l_max, rgb=None):
    n = xyz_w.shape[0]
    rgb = torch.empty((n, 3), device=xyz_w.device, dtype=torch.float32) if rgb is None else rgb
    cx, cy, cz = [float(c) for c in campos]
    check(lib().gsx_precompute_spherical_harmonics(_ptr(xyz_w), _ptr(sh if l_max > 0 else None), _ptr(band0), f32(cx),
                                                   f32(cy), f32(cz), int(l_max), n, _ptr(rgb), _stream()))
    return rgb


def get_sorted_gaussian_list(uv, xyz_c, radius, n_tiles_x, n_tiles_y):
    """Both phases of the reference protocol.  Returns dict(C, K, sorted[K], ranges[T+1])."""
    n = uv.shape[0]
    dev = uv.device
    cc, kk = C.c_longlong(0), C.c_longlong(0)
    check(lib().gsx_sorted_list_size(_ptr(uv), _ptr(radius), int(n_tiles_x), int(n_tiles_y), n, C.byref(cc),
                                     C.byref(kk), _stream()))
    cap = max(int(cc.value), 1)       # the reference caller allocates C ints (cuda/raster.cu:115)
    srt = torch.zeros(cap, device=dev, dtype=torch.int32)
    ranges = torch.zeros(n_tiles_x * n_tiles_y + 1, device=dev, dtype=torch.int32)
    kdev = torch.zeros(1, device=dev, dtype=torch.int32)
    check(lib().gsx_get_sorted_gaussian_list(_ptr(uv), _ptr(xyz_c), _ptr(radius), int(n_tiles_x), int(n_tiles_y), n,
                                             C.c_longlong(cap), _ptr(srt), _ptr(ranges), _ptr(kdev), C.c_void_p(0),
                                             C.c_size_t(0), _stream()))
    K = int(kdev.item())
    assert K == kk.value
    return dict(C=int(cc.value), K=K, sorted=srt[:K], ranges=ranges)


def render_image(uv, opacity, conic, rgb, background, sorted_splats, ranges, width, height):
    dev = uv.device
    spp = torch.zeros((height, width), device=dev, dtype=torch.int32)
    wpp = torch.zeros((height, width), device=dev, dtype=torch.float32)
    img = torch.zeros((height, width, 3), device=dev, dtype=torch.float32)
    if sorted_splats.numel() == 0:
        sorted_splats = torch.zeros(1, device=dev, dtype=torch.int32)
    check(lib().gsx_render_image(_ptr(uv), _ptr(opacity), _ptr(conic), _ptr(rgb), f32(background), _ptr(sorted_splats),
                                 _ptr(ranges), int(width), int(height), _ptr(spp), _ptr(wpp), _ptr(img), _stream()))
    return img, wpp, spp


# ------------------------------------------------------------------------------------------------ backward F9..F15
def render_image_backward(uv, opacity, conic, rgb, background, sorted_splats, ranges, spp, wpp, grad_image, width,
                          height, grads=None):
    n = uv.shape[0]
    dev = uv.device
    if grads is None:
        grads = (torch.zeros((n, 3), device=dev), torch.zeros(n, device=dev), torch.zeros((n, 2), device=dev),
                 torch.zeros((n, 3), device=dev))
    g_rgb, g_opa, g_uv, g_con = grads
    if sorted_splats.numel() == 0:
        sorted_splats = torch.zeros(1, device=dev, dtype=torch.int32)
    check(lib().gsx_render_image_backward(_ptr(uv), _ptr(opacity), _ptr(conic), _ptr(rgb), f32(background),
                                          _ptr(sorted_splats), _ptr(ranges), _ptr(spp), _ptr(wpp), _ptr(grad_image),
                                          int(width), int(height), n, _ptr(g_rgb), _ptr(g_opa), _ptr(g_uv),
                                          _ptr(g_con), _stream()))
    return g_rgb, g_opa, g_uv, g_con


def precompute_spherical_harmonics_backward(xyz_w, band0, sh, campos, rgb_grad, l_max, xyz_grad=None, flags=0):
    n = xyz_w.shape[0]
    dev = xyz_w.device
    nc = (l_max + 1) ** 2
    sh_grad = torch.zeros((n, max(nc - 1, 1), 3), device=dev)
    b0_grad = torch.zeros((n, 3), device=dev)
    xyz_grad = torch.zeros((n, 3), device=dev) if xyz_grad is None else xyz_grad
    cx, cy, cz = [float(c) for c in campos]
    check(lib().gsx_precompute_spherical_harmonics_backward(
        _ptr(xyz_w), _ptr(band0), _ptr(sh if l_max > 0 else None), f32(cx), f32(cy), f32(cz), _ptr(rgb_grad),
        int(l_max), n, int(flags), _ptr(sh_grad if l_max > 0 else None), _ptr(b0_grad), _ptr(xyz_grad), _stream()))
    return sh_grad[:, :nc - 1], b0_grad, xyz_grad


def compute_conic_backward(J, sigma, view, conic, conic_grad, J_grad=None, sigma_grad=None):
    n = J.shape[0]
    J_grad = torch.zeros((n, 6), device=J.device) if J_grad is None else J_grad
    sigma_grad = torch.zeros((n, 6), device=J.device) if sigma_grad is None else sigma_grad
    check(lib().gsx_compute_conic_backward(_ptr(J), _ptr(sigma), _ptr(view), _ptr(conic), _ptr(conic_grad), n,
                                           _ptr(J_grad), _ptr(sigma_grad), _stream()))
    return J_grad, sigma_grad


def compute_projection_jacobian_backward(xyz_c, focal_x, focal_y, tan_fovx, tan_fovy, J_grad, xyz_c_grad=None):
    n = xyz_c.shape[0]
    xyz_c_grad = torch.zeros((n, 3), device=xyz_c.device) if xyz_c_grad is None else xyz_c_grad
    check(lib().gsx_compute_projection_jacobian_backward(_ptr(xyz_c), f32(focal_x), f32(focal_y), f32(tan_fovx),
                                                         f32(tan_fovy), _ptr(J_grad), n, _ptr(xyz_c_grad), _stream()))
    return xyz_c_grad


def compute_sigma_backward(quaternion, scale, sigma_grad):
    n = quaternion.shape[0]
    dq = torch.zeros((n, 4), device=quaternion.device)
    ds = torch.zeros((n, 3), device=quaternion.device)
    check(lib().gsx_compute_sigma_backward(_ptr(quaternion), _ptr(scale), _ptr(sigma_grad), n, _ptr(dq), _ptr(ds),
                                           _stream()))
    return dq, ds


def project_to_screen_backward(xyz_c, proj, uv_grad, width, height, xyz_c_grad=None):
    n = xyz_c.shape[0]
    xyz_c_grad = torch.zeros((n, 3), device=xyz_c.device) if xyz_c_grad is None else xyz_c_grad
    check(lib().gsx_project_to_screen_backward(_ptr(xyz_c), _ptr(proj), _ptr(uv_grad), n, int(width), int(height),
                                               _ptr(xyz_c_grad), _stream()))
    return xyz_c_grad


def compute_camera_space_points_backward(xyz_w, view, xyz_c_grad, xyz_w_grad=None):
    n = xyz_c_grad.shape[0]
    xyz_w_grad = torch.zeros((n, 3), device=xyz_c_grad.device) if xyz_w_grad is None else xyz_w_grad
    check(lib().gsx_compute_camera_space_points_backward(_ptr(xyz_w), _ptr(view), _ptr(xyz_c_grad), n,
                                                         _ptr(xyz_w_grad), _stream()))
    return xyz_w_grad


# ------------------------------------------------------------------------------------------------ sort primitives
# ------------------------------------------------------------------------------------------------ optimiser (8f rank 2)
def adam_step(params, param_grads, exp_avg, exp_avg_sq, lr, b1, b2, eps, bias1, bias2):
    """adam_step of the reference (include/gsplat_cuda/optimizer.cuh:27, cuda/optimizer.cu:30-44): every float of
    `params` is updated in place (N*S elements), moments in place."""
    _chk_f32(params, param_grads, exp_avg, exp_avg_sq)
    n = params.numel()
    assert param_grads.numel() == n and exp_avg.numel() == n and exp_avg_sq.numel() == n
    check(lib().gsx_adam_step(_ptr(params), _ptr(param_grads), _ptr(exp_avg), _ptr(exp_avg_sq), f32(lr), f32(b1), f32(b2),
                              f32(eps), f32(bias1), f32(bias2), C.c_longlong(n), _stream()))
    return params


def adam_step_fused(groups, mask, b1, b2, eps, bias1, bias2, uv_grad=None, uv_grad_accum=None, grad_accum_dur=None):
    """TrainerImpl::optimizer_step of the reference (cuda/trainer.cu:1027-1158) in one launch.
    groups: list of (param [N][w], grad [N][w], exp_avg, exp_avg_sq, lr), gradients indexed by the global Gaussian
    index; mask: uint8 [N] or None; Gaussians with mask 0 keep parameters and moments.  With uv_grad [N][2] the
    densification accumulators are advanced for the same Gaussians (trainer.cu:1138-1157)."""
    assert 0 < len(groups) <= _ffi.GSX_ADAM_MAX_GROUPS
    n = groups[0][0].shape[0]
    arr = (_ffi.GsxAdamGroup * len(groups))()
    for k, (p, g, m, v, lr) in enumerate(groups):
        _chk_f32(p, g, m, v)
        assert p.shape[0] == n and g.numel() == p.numel() and m.numel() == p.numel() and v.numel() == p.numel()
        for t in (p, g, m, v):
            assert t.is_cuda and t.is_contiguous()
        arr[k].param, arr[k].grad = p.data_ptr(), g.data_ptr()
        arr[k].exp_avg, arr[k].exp_avg_sq = m.data_ptr(), v.data_ptr()
        arr[k].width = p.numel() // n if n else 1
        arr[k].lr = float(lr)
    if mask is not None:
        assert mask.dtype in (torch.uint8, torch.bool) and mask.numel() == n
    if uv_grad is not None:
        assert uv_grad.dtype == torch.float32 and uv_grad.numel() == 2 * n
        assert uv_grad_accum.dtype == torch.float32 and grad_accum_dur.dtype == torch.int32
    check(lib().gsx_adam_step_fused(arr, len(groups), _ptr(mask), n, f32(b1), f32(b2), f32(eps), f32(bias1), f32(bias2),
                                    _ptr(uv_grad), _ptr(uv_grad_accum), _ptr(grad_accum_dur), _stream()))


# ------------------------------------------------------------------------------------------------ loss (8f rank 3)
_loss_ws = {}


def fused_loss(predicted, gt, ssim_weight, image_grad=None, sync=True):
    """fused_loss of the reference (include/gsplat_cuda/cuda_forward.cuh:144, cuda/loss.cu:440-471) on [H][W][3] CUDA
    tensors.  Returns (loss, image_grad); loss is a Python float when sync (the reference's behaviour), else a 1-element
    device tensor filled in stream order."""
    _chk_f32(predicted, gt, image_grad)
    rows, cols = predicted.shape[0], predicted.shape[1]
    assert predicted.shape == (rows, cols, 3) and gt.shape == predicted.shape
    image_grad = torch.empty_like(predicted) if image_grad is None else image_grad
    need = lib().gsx_fused_loss_workspace_bytes(rows, cols)
    key = (predicted.device.index, torch.cuda.current_stream().cuda_stream)
    ws = _loss_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = _loss_ws[key] = torch.empty(need, device=predicted.device, dtype=torch.uint8)
    loss_dev = torch.empty(1, device=predicted.device, dtype=torch.float32)
    loss_host = C.c_float(0)
    check(lib().gsx_fused_loss(_ptr(predicted), _ptr(gt), rows, cols, f32(ssim_weight), _ptr(image_grad), _ptr(ws),
                               C.c_size_t(ws.numel()), _ptr(loss_dev), C.byref(loss_host) if sync else None, _stream()))
    return (float(loss_host.value) if sync else loss_dev), image_grad


# ------------------------------------------------------------------------------------------------ Morton re-layout (8f rank 4)
def _bbox(xyz, bbox):
    if bbox is None:                                   # the reference reduces it with thrust (cuda/trainer.cu:874-884)
        mn, mx = xyz.amin(0).tolist(), xyz.amax(0).tolist()
        bbox = (mx[0], mx[1], mx[2], mn[0], mn[1], mn[2])
    return [f32(b) for b in bbox]


def compute_morton_codes(xyz, bbox=None, flags=0):
    """compute_morton_codes of the reference (cuda/culling.cu:345-359); bbox = (x_max, y_max, z_max, x_min, y_min, z_min)
    or None for the tight box; flags = 0 reproduces the reference's bit spread, GSX_MORTON_TRUE_INTERLEAVE (1) is the real
    interleave.  Returns int64 codes (uint64 bit patterns, 63 bits used)."""
    _chk_f32(xyz)
    n = xyz.shape[0]
    codes = torch.empty(n, device=xyz.device, dtype=torch.int64)
    check(lib().gsx_compute_morton_codes(n, _ptr(xyz), *_bbox(xyz, bbox), int(flags), _ptr(codes), _stream()))
    return codes


def morton_order(xyz, bbox=None, want_codes=False, flags=0):
    """The permutation TrainerImpl::sort_gaussians applies (cuda/trainer.cu:886-893): order[i] = Gaussian that comes i-th in
    Morton order, ties in index order."""
    _chk_f32(xyz)
    n = xyz.shape[0]
    order = torch.empty(n, device=xyz.device, dtype=torch.int32)
    codes = torch.empty(n, device=xyz.device, dtype=torch.int64) if want_codes else None
    wsb = lib().gsx_morton_order_workspace_bytes(n)
    ws = torch.empty(wsb, device=xyz.device, dtype=torch.uint8)
    check(lib().gsx_morton_order(n, _ptr(xyz), *_bbox(xyz, bbox), int(flags), _ptr(order), _ptr(codes), _ptr(ws), C.c_size_t(wsb),
                                 _stream()))
    return (order, codes) if want_codes else order


def gather_rows(src, order):
    """gather_with_stride of the reference (cuda/trainer.cu:897-923), out of place: dst[i] = src[order[i]]"""
    _chk_f32(src)
    n = order.numel()
    assert order.dtype == torch.int32 and src.shape[0] == n
    dst = torch.empty_like(src)
    check(lib().gsx_gather_rows(_ptr(src), _ptr(order), n, src.numel() // max(n, 1), _ptr(dst), _stream()))
    return dst


# ------------------------------------------------------------------------------------------------ densification (8f rank 4)
def clone_gaussians(mask, write_ids, arrays, n_out):
    """clone_gaussians of the reference (include/gsplat_cuda/adaptive_density.cuh:28): arrays = dict with xyz, rgb (band 0),
    opacity, scale, quaternion, sh ([N][k][3] or None); returns the same dict with n_out rows (selected parents copied to
    row write_ids[i], other rows zero)."""
    n = mask.numel()
    sh = arrays.get("sh")
    ncoef = 0 if sh is None else sh.numel() // (3 * n)
    out = {k: torch.zeros((n_out,) + tuple(v.shape[1:]), device=v.device, dtype=torch.float32) for k, v in arrays.items() if v is not None}
    g = lambda d, k: _ptr(d.get(k))
    check(lib().gsx_clone_gaussians(n, ncoef, _ptr(mask), _ptr(write_ids), g(arrays, "xyz"), g(arrays, "rgb"), g(arrays, "opacity"),
                                    g(arrays, "scale"), g(arrays, "quaternion"), g(arrays, "sh"), g(out, "xyz"), g(out, "rgb"),
                                    g(out, "opacity"), g(out, "scale"), g(out, "quaternion"), g(out, "sh"), _stream()))
    return out


def split_gaussians(mask, write_ids, arrays, n_out, scale_factor, seed):
    """split_gaussians of the reference (adaptive_density.cuh:54) with an explicit seed; n_out rows = 2 per selected parent"""
    n = mask.numel()
    sh = arrays.get("sh")
    ncoef = 0 if sh is None else sh.numel() // (3 * n)
    out = {k: torch.zeros((n_out,) + tuple(v.shape[1:]), device=v.device, dtype=torch.float32) for k, v in arrays.items() if v is not None}
    g = lambda d, k: _ptr(d.get(k))
    check(lib().gsx_split_gaussians(n, f32(scale_factor), ncoef, _ptr(mask), _ptr(write_ids), g(arrays, "xyz"), g(arrays, "rgb"),
                                    g(arrays, "opacity"), g(arrays, "scale"), g(arrays, "quaternion"), g(arrays, "sh"),
                                    g(out, "xyz"), g(out, "rgb"), g(out, "opacity"), g(out, "scale"), g(out, "quaternion"),
                                    g(out, "sh"), C.c_ulonglong(int(seed)), _stream()))
    return out


DENSIFY_NAMES = ("xyz", "scale", "quaternion", "opacity", "band0", "sh")


def _gaussian_set(params, moments, sh_width):
    s = _ffi.GsxGaussianSet()
    for k, nm in enumerate(DENSIFY_NAMES):
        t = params.get(nm)
        setattr(s, nm, t.data_ptr() if t is not None else None)
        mv = (moments or {}).get(nm)
        s.exp_avg[k] = mv[0].data_ptr() if mv is not None else None
        s.exp_avg_sq[k] = mv[1].data_ptr() if mv is not None else None
    sh = params.get("sh")
    s.sh_stride = 0 if sh is None else sh.numel() // max(sh.shape[0], 1)
    s.sh_width = int(sh_width)
    return s


def densify(params, moments, uv_grad_accum, grad_accum_dur, uv_grad_threshold, delete_opacity_threshold, scene_extent,
            split_scale_factor=1.6, max_gaussians=0, seed=0, l_max=3):
    """TrainerImpl::adaptive_density_step of the reference (cuda/trainer.cu:416-775) on device tensors.
    params: dict xyz [N][3], scale, quaternion, opacity [N], band0, sh [N][stride] (or None); moments: dict name -> (m, v)
    or None.  Returns (new params, new moments, counts dict); None, None, counts when the step would exceed max_gaussians
    (the reference skips it)."""
    n = params["xyz"].shape[0]
    dev = params["xyz"].device
    cfg = _ffi.GsxDensityConfig(float(uv_grad_threshold), float(delete_opacity_threshold), float(scene_extent),
                                float(split_scale_factor), int(max_gaussians))
    wsb = lib().gsx_densify_workspace_bytes(n)
    ws = torch.empty(wsb, device=dev, dtype=torch.uint8)
    cnt = _ffi.GsxDensityCounts()
    check(lib().gsx_densify_plan(C.byref(cfg), n, _ptr(params["opacity"]), _ptr(params["scale"]), _ptr(uv_grad_accum),
                                 _ptr(grad_accum_dur), _ptr(ws), C.c_size_t(wsb), C.byref(cnt), _stream()))
    counts = dict(n_in=cnt.n_in, n_prune=cnt.n_prune, n_clone=cnt.n_clone, n_split=cnt.n_split, n_out=cnt.n_out)
    if max_gaussians and cnt.n_out > max_gaussians:
        return None, None, counts
    new = lambda t: torch.empty((cnt.n_out,) + tuple(t.shape[1:]), device=dev, dtype=torch.float32)
    out_p = {k: (new(v) if v is not None else None) for k, v in params.items()}
    out_m = None if moments is None else {k: (new(m), new(v)) for k, (m, v) in moments.items()}
    shw = 3 * ((l_max + 1) ** 2 - 1) if params.get("sh") is not None else 0
    sin, sout = _gaussian_set(params, moments, shw), _gaussian_set(out_p, out_m, shw)
    check(lib().gsx_densify_apply(C.byref(cfg), C.byref(cnt), C.byref(sin), C.byref(sout), _ptr(ws), C.c_size_t(wsb),
                                  C.c_ulonglong(int(seed)), _stream()))
    return out_p, out_m, counts


def gather_rows_multi(srcs, order):
    """every array of a Gaussian set re-ordered in one launch (the 17 gather_with_stride calls of trainer.cu:897-923)"""
    n = order.numel()
    dsts = [torch.empty_like(t) for t in srcs]
    k = len(srcs)
    sp = (C.c_void_p * k)(*[t.data_ptr() for t in srcs])
    dp_ = (C.c_void_p * k)(*[t.data_ptr() for t in dsts])
    w = (C.c_int * k)(*[t.numel() // max(n, 1) for t in srcs])
    for t in srcs:
        _chk_f32(t)
        assert t.is_cuda and t.is_contiguous() and t.shape[0] == n
    check(lib().gsx_gather_rows_multi(sp, dp_, w, k, _ptr(order), n, _stream()))
    return dsts


def sort_pairs_u64(keys, vals, key_bits=64):
    """keys: int64 CUDA tensor holding uint64 bit patterns, vals: int32.  Returns sorted copies."""
    n = keys.numel()
    dev = keys.device
    if n == 0:
        return keys.clone(), vals.clone()
    k_in, v_in = keys.clone(), vals.clone()
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    wsb = lib().gsx_sort_workspace_bytes(C.c_longlong(max(n, 1)))
    ws = torch.empty(wsb, device=dev, dtype=torch.uint8)
    check(lib().gsx_sort_pairs_u64(_ptr(k_in), _ptr(k_out), _ptr(v_in), _ptr(v_out), C.c_longlong(n), C.c_void_p(0),
                                   int(key_bits), _ptr(ws), C.c_size_t(wsb), _stream()))
    return k_out, v_out


def tile_ranges(sorted_keys, n_tiles):
    n = sorted_keys.numel()
    ranges = torch.empty(n_tiles + 1, device=sorted_keys.device, dtype=torch.int32)
    check(lib().gsx_tile_ranges(_ptr(sorted_keys) if n else C.c_void_p(0), C.c_longlong(n), C.c_void_p(0), int(n_tiles),
                                _ptr(ranges), _stream()))
    return ranges


# ------------------------------------------------------------------------------------------------ view-level path
def make_view(view, proj, campos, focal_x, focal_y, width, height, near_thresh=0.3, cull_padding=100, mh_dist=3.0,
              background=0.0, l_max=3, flags=0):
    """Host-side camera block.  tan_fov as cuda/raster.cu:92-93."""
    v = GsxView()
    v.view[:] = [float(x) for x in view.reshape(-1)]
    v.proj[:] = [float(x) for x in proj.reshape(-1)]
    v.campos[:] = [float(x) for x in campos]
    v.focal_x, v.focal_y = float(focal_x), float(focal_y)
    v.tan_fovx, v.tan_fovy = width / (2.0 * focal_x), height / (2.0 * focal_y)
    v.width, v.height = int(width), int(height)
    v.near_thresh, v.cull_padding, v.mh_dist = float(near_thresh), int(cull_padding), float(mh_dist)
    v.background, v.l_max, v.flags = float(background), int(l_max), int(flags)
    return v


class Gaussians:
    """Device-resident parameters (GaussianParameters, cuda_data.cuh:11-24) + global-index gradient buffers."""

    def __init__(self, xyz, scale, quaternion, opacity, band0, sh, device="cuda"):
        to = lambda a: torch.as_tensor(a, dtype=torch.float32).contiguous().to(device)
        self.xyz, self.scale, self.quaternion = to(xyz), to(scale), to(quaternion)
        self.opacity, self.band0 = to(opacity).reshape(-1), to(band0)
        self.N = self.xyz.shape[0]
        if sh is None:
            self.sh = torch.zeros(1, device=device)
            self.sh_stride = 0
        else:
            self.sh = to(sh).reshape(self.N, -1)
            self.sh_stride = self.sh.shape[1]
        self.grads = None

    def params(self):
        p = GsxParams()
        p.xyz, p.scale, p.quaternion = self.xyz.data_ptr(), self.scale.data_ptr(), self.quaternion.data_ptr()
        p.opacity, p.band0, p.sh = self.opacity.data_ptr(), self.band0.data_ptr(), self.sh.data_ptr()
        p.sh_stride = self.sh_stride
        return p

    def zero_grads(self, with_uv=False):
        dev = self.xyz.device
        z = lambda *s: torch.zeros(s, device=dev, dtype=torch.float32)
        self.grads = dict(xyz=z(self.N, 3), scale=z(self.N, 3), quaternion=z(self.N, 4), opacity=z(self.N),
                          band0=z(self.N, 3), sh=z(self.N, max(self.sh_stride, 1)))
        if with_uv:
            self.grads["uv"] = z(self.N, 2)
        return self.grads

    def grads_struct(self):
        g = GsxGrads()
        d = self.grads
        g.xyz, g.scale, g.quaternion = d["xyz"].data_ptr(), d["scale"].data_ptr(), d["quaternion"].data_ptr()
        g.opacity, g.band0, g.sh = d["opacity"].data_ptr(), d["band0"].data_ptr(), d["sh"].data_ptr()
        g.uv = d["uv"].data_ptr() if "uv" in d else None
        return g


class Rasterizer:
    """Persistent context: forward (rasterize_image) + backward (backward_pass minus fused_loss)."""

    def __init__(self, max_gaussians, max_width, max_height, instance_capacity=0):
        self.h = C.c_void_p()
        check(lib().gsx_ctx_create(C.byref(self.h), int(max_gaussians), int(max_width), int(max_height),
                                   C.c_longlong(int(instance_capacity))))
        self.max_gaussians, self.max_width, self.max_height = max_gaussians, max_width, max_height
        self._view = None

    def close(self):
        if self.h:
            lib().gsx_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def forward(self, view, g):
        self._view, self._n = view, g.N
        self._p = g.params()
        check(lib().gsx_forward(self.h, C.byref(view), C.byref(self._p), g.N, _stream()))

    def backward(self, view, g, grad_image):
        if g.grads is None:
            g.zero_grads()
        gs = g.grads_struct()
        p = g.params()
        check(lib().gsx_backward(self.h, C.byref(view), C.byref(p), g.N, _ptr(grad_image), C.byref(gs), _stream()))

    def stats(self, raise_on_overflow=True):
        s = GsxStats()
        rc = lib().gsx_ctx_stats(self.h, C.byref(s), _stream())
        if rc != _ffi.GSX_OK and (raise_on_overflow or rc != _ffi.GSX_E_CAPACITY):
            check(rc)
        return dict(N=s.N, P=s.P, K=s.K, capacity=s.capacity, overflow=(rc == _ffi.GSX_E_CAPACITY))

    def reserve(self, capacity):
        check(lib().gsx_ctx_reserve(self.h, C.c_longlong(int(capacity))))

    def overflow(self):
        """0, or the instance count K that exceeded the capacity since the last reserve().  No synchronisation (a pinned
        flag raised by the device): a training loop polls it once per step and reserves + re-runs when it fires."""
        need = C.c_longlong(0)
        rc = lib().gsx_ctx_overflow(self.h, C.byref(need))
        if rc not in (_ffi.GSX_OK, _ffi.GSX_E_CAPACITY):
            check(rc)
        return int(need.value)

    def forward_checked(self, view, g):
        """forward + one sync to grow the instance capacity if it overflowed (first view of a scene)."""
        self.forward(view, g)
        st = self.stats(raise_on_overflow=False)
        if st["overflow"]:
            self.reserve(int(st["K"] * 1.25) + 1024)
            self.forward(view, g)
            st = self.stats()
        return st

    def render_view_host(self, view, g, grad_image_host, image_host):
        """Host-buffer entry point: H2D grad_image, forward, backward, D2H image (synchronises)."""
        if g.grads is None:
            g.zero_grads()
        gs = g.grads_struct()
        p = g.params()
        gi = C.c_void_p(grad_image_host.data_ptr()) if grad_image_host is not None else C.c_void_p(0)
        io = C.c_void_p(image_host.data_ptr()) if image_host is not None else C.c_void_p(0)
        check(lib().gsx_render_view_host(self.h, C.byref(view), C.byref(p), g.N, gi, io, C.byref(gs), _stream()))

    # ---- batch path (gsx_batch_*): one pass over the parameters per direction for all views of a batch ----------
    def batch_reserve(self, n_slots):
        check(lib().gsx_batch_reserve(self.h, int(n_slots)))

    def batch_forward_params(self, views, g):
        arr = (GsxView * len(views))(*views)
        self._bviews, self._n = arr, g.N
        self._p = g.params()
        check(lib().gsx_batch_forward_params(self.h, arr, len(views), C.byref(self._p), g.N, _stream()))

    def batch_forward_view(self, slot):
        self._view = self._bviews[slot]
        check(lib().gsx_batch_forward_view(self.h, int(slot), _stream()))

    def batch_backward_view(self, slot, grad_image):
        check(lib().gsx_batch_backward_view(self.h, int(slot), _ptr(grad_image), _stream()))

    def batch_extract_view_rgb_grad(self, slot, drgb):
        """dense [N][3] d rgb of the slot's view (after batch_backward_view): what ranks all-gather (dp.py)"""
        check(lib().gsx_batch_extract_view_rgb_grad(self.h, int(slot), _ptr(drgb), _stream()))

    def batch_backward_params(self, g, flags=_ffi.GSX_GRAD_OVERWRITE):
        if g.grads is None:
            g.zero_grads()
        gs = g.grads_struct()
        p = g.params()
        check(lib().gsx_batch_backward_params(self.h, C.byref(p), g.N, C.byref(gs), int(flags), _stream()))

    def batch_step(self, views, g, grad_images, flags=_ffi.GSX_GRAD_OVERWRITE, on_image=None):
        """fwd+bwd of every view of the batch, gradients summed over the views.  grad_images: one device tensor per view
        (or a single one used for all); on_image(slot) is called after each forward (the rendered image is in state())."""
        need = self.overflow()                          # raised by an EARLIER step (no sync): that step lost views
        if need:
            raise _ffi.GsxError(_ffi.GSX_E_CAPACITY, f"instance capacity exceeded in a previous step: K={need}; "
                                                     "reserve() and re-run it")
        self.batch_forward_params(views, g)
        for k in range(len(views)):
            self.batch_forward_view(k)
            if on_image is not None:
                on_image(k)
            gi = grad_images[k] if isinstance(grad_images, (list, tuple)) else grad_images
            self.batch_backward_view(k, gi)
        self.batch_backward_params(g, flags)

    def batch_graph_capture(self, views, g, grad_images=None, gt_images=None, ssim_weight=0.2, images_out=None, losses=None,
                            flags=_ffi.GSX_GRAD_OVERWRITE, stream=None):
        """The whole batch step as ONE CUDA graph (gsx_batch_graph_capture).  grad_images: per-view device tensors the caller
        refreshes between launches, or gt_images: the loss gradient is computed in the graph (L1 + SSIM).  Returns a handle
        for graph_launch(); capture happens on `stream` (a non-default torch stream; one is created when None)."""
        if g.grads is None:
            g.zero_grads()
        self._gstream = stream if stream is not None else torch.cuda.Stream()
        arr = (GsxView * len(views))(*views)
        self._bviews, self._n = arr, g.N
        self._view = arr[len(views) - 1]
        gs_, p = g.grads_struct(), g.params()
        mk = lambda ts: (C.c_void_p * len(views))(*[t.data_ptr() for t in ts]) if ts is not None else None
        h = C.c_void_p()
        self._gkeep = (grad_images, gt_images, images_out, losses)
        check(lib().gsx_batch_graph_capture(self.h, arr, len(views), C.byref(p), g.N, mk(grad_images), mk(gt_images), f32(ssim_weight),
                                            mk(images_out), _ptr(losses), C.byref(gs_), int(flags),
                                            C.c_void_p(self._gstream.cuda_stream), C.byref(h)))
        return h

    def graph_launch(self, h, stream=None):
        st = stream if stream is not None else self._gstream
        check(lib().gsx_graph_launch(h, C.c_void_p(st.cuda_stream)))

    @staticmethod
    def graph_destroy(h):
        lib().gsx_graph_destroy(h)

    def batch_render_host(self, views, g, grad_images_host, images_host, flags=_ffi.GSX_GRAD_OVERWRITE):
        """Host-buffer batch step: pinned host grad images in, pinned host images out (synchronises)."""
        if g.grads is None:
            g.zero_grads()
        arr = (GsxView * len(views))(*views)
        self._bviews, self._n = arr, g.N
        gs, p = g.grads_struct(), g.params()
        gi = (C.c_void_p * len(views))(*[t.data_ptr() for t in grad_images_host])
        io = (C.c_void_p * len(views))(*[t.data_ptr() for t in images_host])
        check(lib().gsx_batch_render_host(self.h, arr, len(views), C.byref(p), g.N, gi, io, C.byref(gs), int(flags),
                                          _stream()))
        self._view = arr[len(views) - 1]

    def enable_timing(self, on=True):
        check(lib().gsx_ctx_enable_timing(self.h, int(bool(on))))

    def timing(self):
        """ms per stage of the last timed forward+backward (synchronises)."""
        ms = (C.c_float * 7)()
        check(lib().gsx_ctx_timing(self.h, ms))
        names = ("preprocess_fwd", "scan_emit", "sort", "ranges", "composite_fwd", "composite_bwd", "preprocess_bwd")
        return dict(zip(names, [float(x) for x in ms]))

    def count_work(self):
        """(S, S+) of the last forward: evaluated pixel-entries and those with alpha >= 1/255."""
        out = (C.c_ulonglong * 2)()
        check(lib().gsx_ctx_count_work(self.h, out, _stream()))
        return int(out[0]), int(out[1])

    def count_hits(self, grad_image):
        """(warp-hits, contributing warp-hits) of the compositing backward over the last forward's lists"""
        out = (C.c_ulonglong * 2)()
        check(lib().gsx_ctx_count_hits(self.h, _ptr(grad_image), out, _stream()))
        return int(out[0]), int(out[1])

    def state_image(self):
        """zero-copy [H][W][3] view of the last rendered image (no synchronisation)"""
        s = GsxState()
        check(lib().gsx_ctx_state(self.h, C.byref(s)))
        return _from_ptr(s.image, (self._view.height, self._view.width, 3), torch.float32)

    def state(self):
        """Zero-copy torch views of the saved forward state (for tests / inspection)."""
        s = GsxState()
        check(lib().gsx_ctx_state(self.h, C.byref(s)))
        v = self._view
        n, W, H = self._n, v.width, v.height
        T = ((W + 15) // 16) * ((H + 15) // 16)
        K = int(_from_ptr(s.num_instances, (1,), torch.int32).item())
        out = dict(
            image=_from_ptr(s.image, (H, W, 3), torch.float32), weight_per_pixel=_from_ptr(s.weight_per_pixel, (H, W), torch.float32),
            splats_per_pixel=_from_ptr(s.splats_per_pixel, (H, W), torch.int32),
            rec0=_from_ptr(s.rec0, (n, 4), torch.float32), rec1=_from_ptr(s.rec1, (n, 4), torch.float32),
            rec2=_from_ptr(s.rec2, (n, 4), torch.float32), radius=_from_ptr(s.radius, (n, 4), torch.float32),
            depth=_from_ptr(s.depth, (n,), torch.float32), mask=_from_ptr(s.mask, (n,), torch.uint8),
            tiles_touched=_from_ptr(s.tiles_touched, (n,), torch.int32), ranges=_from_ptr(s.ranges, (T + 1,), torch.int32),
            packed_grads=_from_ptr(s.packed_grads, (n, 12), torch.float32), K=K)
        out["sorted_gaussians"] = _from_ptr(s.sorted_gaussians, (max(K, 1),), torch.int32)[:K]
        kp = C.c_void_p()
        check(lib().gsx_ctx_build_sorted_keys(self.h, C.byref(kp)))
        out["sorted_keys"] = (_from_ptr(kp.value, (max(K, 1),), torch.int64)[:K] if kp.value
                              else torch.zeros(0, dtype=torch.int64, device="cuda"))
        return out


class _CudaArray:
    def __init__(self, ptr, shape, dtype):
        itemsize = torch.empty(0, dtype=dtype).element_size()
        typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.uint8: "|u1", torch.int64: "<i8"}[dtype]
        strides = None
        self.__cuda_array_interface__ = dict(shape=tuple(shape), typestr=typestr, data=(int(ptr), False), version=2,
                                             strides=strides)
        self._itemsize = itemsize


def measure_fp32_tflops():
    out = C.c_float(0)
    check(lib().gsx_measure_fp32_tflops(C.byref(out)))
    return float(out.value)


def _from_ptr(ptr, shape, dtype):
    return torch.as_tensor(_CudaArray(ptr, shape, dtype), device="cuda")
