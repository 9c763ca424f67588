"""Test plumbing (not part of the product package): numpy in / numpy out adapter over host.py with the same function names as oracle/golden.py, so the
known-answer tests in tests/kat.py run unchanged against the CUDA path (through the C ABI)."""
import numpy as np
import torch

import importlib

H = importlib.import_module("3dgs_b200.host")
GSX_SH_DSPH_CORRECT = importlib.import_module("3dgs_b200._ffi").GSX_SH_DSPH_CORRECT


def _d(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).cuda().contiguous()


def _h(t):
    return t.detach().cpu().numpy()


def camera_space_points(xyz_w, view):
    return _h(H.compute_camera_space_points(_d(xyz_w), _d(view).reshape(16)))


def project_to_screen(xyz_c, proj, width, height):
    return _h(H.project_to_screen(_d(xyz_c), _d(proj).reshape(16), width, height))


def cull(uv, xyz_c, near_thresh, padding, width, height):
    return _h(H.cull_gaussians(_d(uv), _d(xyz_c), near_thresh, padding, width, height))


def sigma(quat, scale):
    return _h(H.compute_sigma(_d(quat), _d(scale)))


def conic(xyz_c, view, sig, fx, fy, tan_fovx, tan_fovy, mh_dist):
    return tuple(_h(t) for t in H.compute_conic(_d(xyz_c), _d(view).reshape(16), _d(sig), fx, fy, tan_fovx, tan_fovy, mh_dist))


def sh_forward(xyz_w, sh, band0, campos, l_max):
    return _h(H.precompute_spherical_harmonics(_d(xyz_w), _d(sh) if sh is not None and l_max > 0 else None, _d(band0),
                                               np.asarray(campos, np.float32), l_max))


def sorted_list(uv, xyz_c, radius, ntx, nty):
    r = H.get_sorted_gaussian_list(_d(uv), _d(xyz_c), _d(radius), ntx, nty)
    return dict(C=r["C"], K=r["K"], sorted=_h(r["sorted"]), ranges=_h(r["ranges"]))


def render(uv, opacity, con, rgb, bg, srt, ranges, W, H_):
    img, wpp, spp = H.render_image(_d(uv), _d(opacity), _d(con), _d(rgb), bg, _d(srt, torch.int32), _d(ranges, torch.int32), W, H_)
    return _h(img), _h(wpp), _h(spp)


def render_backward(uv, opacity, con, rgb, bg, srt, ranges, spp, wpp, grad_image, W, H_):
    out = H.render_image_backward(_d(uv), _d(opacity), _d(con), _d(rgb), bg, _d(srt, torch.int32), _d(ranges, torch.int32),
                                  _d(spp, torch.int32), _d(wpp), _d(grad_image), W, H_)
    return tuple(_h(t) for t in out)


def sh_backward(xyz_w, band0, sh, campos, rgb_grad, l_max, ref_compat_dsph=1, xyz_grad=None):
    flags = 0 if ref_compat_dsph else GSX_SH_DSPH_CORRECT
    out = H.precompute_spherical_harmonics_backward(
        _d(xyz_w), _d(band0), _d(sh) if sh is not None and l_max > 0 else None, np.asarray(campos, np.float32),
        _d(rgb_grad), l_max, xyz_grad=_d(xyz_grad) if xyz_grad is not None else None, flags=flags)
    return tuple(_h(t) for t in out)


def conic_backward(J, sig, view, con, conic_grad, J_grad=None, sigma_grad=None):
    out = H.compute_conic_backward(_d(J), _d(sig), _d(view).reshape(16), _d(con), _d(conic_grad),
                                   _d(J_grad) if J_grad is not None else None,
                                   _d(sigma_grad) if sigma_grad is not None else None)
    return tuple(_h(t) for t in out)


def jacobian_backward(xyz_c, fx, fy, tan_fovx, tan_fovy, J_grad, xyz_grad=None):
    return _h(H.compute_projection_jacobian_backward(_d(xyz_c), fx, fy, tan_fovx, tan_fovy, _d(J_grad),
                                                     _d(xyz_grad) if xyz_grad is not None else None))


def sigma_backward(quat, scale, sigma_grad):
    return tuple(_h(t) for t in H.compute_sigma_backward(_d(quat), _d(scale), _d(sigma_grad)))


def project_to_screen_backward(xyz_c, proj, uv_grad, width, height, xyz_grad=None):
    return _h(H.project_to_screen_backward(_d(xyz_c), _d(proj).reshape(16), _d(uv_grad), width, height,
                                           _d(xyz_grad) if xyz_grad is not None else None))


def camera_space_points_backward(view, xyz_c_grad, xyz_w_grad=None):
    g = _d(xyz_c_grad)
    return _h(H.compute_camera_space_points_backward(g, _d(view).reshape(16), g,
                                                     _d(xyz_w_grad) if xyz_w_grad is not None else None))


# ------------------------------------------------------------------------------------------------ optimiser (8f rank 2)
def adam_step(param, grad, exp_avg, exp_avg_sq, lr, b1, b2, eps, bias1, bias2):
    p, m, v = _d(param), _d(exp_avg), _d(exp_avg_sq)
    H.adam_step(p, _d(grad), m, v, lr, b1, b2, eps, bias1, bias2)
    return _h(p), _h(m), _h(v)


def adam_step_fused(groups, mask, b1, b2, eps, bias1, bias2, uv_grad=None, uv_grad_accum=None, grad_accum_dur=None):
    """groups: list of (param, grad, exp_avg, exp_avg_sq, lr) numpy arrays [N][w]; returns the list of updated
    (param, exp_avg, exp_avg_sq) and, with uv_grad, the advanced (uv_grad_accum, grad_accum_dur)."""
    dev = [(_d(p), _d(g), _d(m), _d(v), lr) for p, g, m, v, lr in groups]
    mk = None if mask is None else _d(mask, torch.uint8)
    acc = dur = None
    if uv_grad is not None:
        acc, dur = _d(uv_grad_accum), _d(grad_accum_dur, torch.int32)
    H.adam_step_fused(dev, mk, b1, b2, eps, bias1, bias2, _d(uv_grad) if uv_grad is not None else None, acc, dur)
    out = [(_h(p), _h(m), _h(v)) for p, _, m, v, _ in dev]
    return (out, (_h(acc), _h(dur))) if uv_grad is not None else out


# ------------------------------------------------------------------------------------------------ loss (8f rank 3)
def fused_loss(pred, gt, ssim_weight):
    loss, grad = H.fused_loss(_d(pred), _d(gt), ssim_weight)
    return loss, _h(grad)


# ------------------------------------------------------------------------------------------------ Morton re-layout (8f rank 4)
def morton_codes(xyz, bbox, true_interleave=False):
    return _h(H.compute_morton_codes(_d(xyz), bbox, flags=int(true_interleave))).view(np.uint64)


def morton_order(xyz, bbox, true_interleave=False):
    return _h(H.morton_order(_d(xyz), bbox, flags=int(true_interleave)))
