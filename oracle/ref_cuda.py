"""ctypes loader of oracle/_ref/libref3dgs.so: the REFERENCE's own CUDA kernels (built by oracle/build_ref.sh from the
sources under /root/reference in the build container).  TEST INFRASTRUCTURE ONLY (second oracle, same-box baseline B1);
torch owns the device memory.  The reference launches on the legacy default stream; callers synchronise around it."""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref3dgs.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
    return _lib


def _p(t):
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _f(x):
    return C.c_float(float(x))


def _new(shape, dev, dtype=torch.float32):
    return torch.zeros(shape, device=dev, dtype=dtype)


def camera_space_points(xyz_w, view):
    out = _new(xyz_w.shape, xyz_w.device)
    lib().ref_compute_camera_space_points(_p(xyz_w), _p(view), xyz_w.shape[0], _p(out))
    return out


def project_to_screen(xyz_c, proj, w, h):
    out = _new((xyz_c.shape[0], 2), xyz_c.device)
    lib().ref_project_to_screen(_p(xyz_c), _p(proj), xyz_c.shape[0], int(w), int(h), _p(out))
    return out


def cull(uv, xyz_c, near, pad, w, h):
    out = _new(uv.shape[0], uv.device, torch.uint8)
    lib().ref_cull_gaussians(_p(uv), _p(xyz_c), uv.shape[0], _f(near), int(pad), int(w), int(h), _p(out))
    return out


def sigma(q, s):
    out = _new((q.shape[0], 6), q.device)
    lib().ref_compute_sigma(_p(q), _p(s), q.shape[0], _p(out))
    return out


def conic(xyz_c, view, sig, fx, fy, tx, ty, mh):
    n, dev = xyz_c.shape[0], xyz_c.device
    J, con, rad = _new((n, 6), dev), _new((n, 3), dev), _new((n, 4), dev)
    lib().ref_compute_conic(_p(xyz_c), _p(view), _p(sig), _f(fx), _f(fy), _f(tx), _f(ty), _f(mh), n, _p(J), _p(con), _p(rad))
    return J, con, rad


def sorted_list(uv, xyz_c, radius, ntx, nty):
    n, dev = uv.shape[0], uv.device
    cnt = C.c_ulonglong(0)
    ranges = _new(ntx * nty + 1, dev, torch.int32)
    lib().ref_get_sorted_gaussian_list(_p(uv), _p(xyz_c), _p(radius), int(ntx), int(nty), n, C.byref(cnt), C.c_void_p(0), _p(ranges))
    coarse = int(cnt.value)
    srt = _new(max(coarse, 1), dev, torch.int32)
    lib().ref_get_sorted_gaussian_list(_p(uv), _p(xyz_c), _p(radius), int(ntx), int(nty), n, C.byref(cnt), _p(srt), _p(ranges))
    torch.cuda.synchronize()
    K = int(ranges[-1].item())
    return dict(C=coarse, K=K, sorted=srt[:K], ranges=ranges)


def render(uv, opacity, con, rgb, bg, srt, ranges, w, h):
    dev = uv.device
    spp, wpp, img = _new((h, w), dev, torch.int32), _new((h, w), dev), _new((h, w, 3), dev)
    lib().ref_render_image(_p(uv), _p(opacity), _p(con), _p(rgb), _f(bg), _p(srt), _p(ranges), int(w), int(h), _p(spp), _p(wpp), _p(img))
    return img, wpp, spp


def render_backward(uv, opacity, con, rgb, bg, srt, ranges, spp, wpp, gi, w, h):
    n, dev = uv.shape[0], uv.device
    g_rgb, g_opa, g_uv, g_con = _new((n, 3), dev), _new(n, dev), _new((n, 2), dev), _new((n, 3), dev)
    lib().ref_render_image_backward(_p(uv), _p(opacity), _p(con), _p(rgb), _f(bg), _p(srt), _p(ranges), _p(spp), _p(wpp), _p(gi),
                                    int(w), int(h), _p(g_rgb), _p(g_opa), _p(g_uv), _p(g_con))
    return g_rgb, g_opa, g_uv, g_con


def conic_backward(J, sig, view, con, g_con):
    n, dev = J.shape[0], J.device
    gJ, gS = _new((n, 6), dev), _new((n, 6), dev)
    lib().ref_compute_conic_backward(_p(J), _p(sig), _p(view), _p(con), _p(g_con), n, _p(gJ), _p(gS))
    return gJ, gS


def jacobian_backward(xyz_c, fx, fy, tx, ty, gJ, out=None):
    out = _new(xyz_c.shape, xyz_c.device) if out is None else out
    lib().ref_compute_projection_jacobian_backward(_p(xyz_c), _f(fx), _f(fy), _f(tx), _f(ty), _p(gJ), xyz_c.shape[0], _p(out))
    return out


def sigma_backward(q, s, gS):
    gq, gs = _new(q.shape, q.device), _new(s.shape, s.device)
    lib().ref_compute_sigma_backward(_p(q), _p(s), _p(gS), q.shape[0], _p(gq), _p(gs))
    return gq, gs


def project_to_screen_backward(xyz_c, proj, g_uv, w, h, out=None):
    out = _new(xyz_c.shape, xyz_c.device) if out is None else out
    lib().ref_project_to_screen_backward(_p(xyz_c), _p(proj), _p(g_uv), xyz_c.shape[0], int(w), int(h), _p(out))
    return out


def camera_space_points_backward(xyz_w, view, g_xyz_c, out=None):
    out = _new(g_xyz_c.shape, g_xyz_c.device) if out is None else out
    lib().ref_compute_camera_space_points_backward(_p(xyz_w), _p(view), _p(g_xyz_c), g_xyz_c.shape[0], _p(out))
    return out
