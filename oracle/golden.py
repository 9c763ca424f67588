"""ctypes loader for oracle/libgolden.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product path (3dgs_b200) never does.  All arrays are host numpy arrays (float32 / int32 / uint8).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "-s"])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libgolden.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.gold_state_new.restype = C.c_void_p
        _LIB.gold_state_get.restype = C.c_longlong
        _LIB.gold_state_copy_f.restype = C.c_longlong
        _LIB.gold_state_copy_i.restype = C.c_longlong
    return _LIB


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class GoldView(C.Structure):
    _fields_ = [("view", C.c_float * 16), ("proj", C.c_float * 16), ("campos", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("width", C.c_int), ("height", C.c_int), ("near_thresh", C.c_float), ("cull_padding", C.c_int),
                ("mh_dist", C.c_float), ("background", C.c_float), ("l_max", C.c_int),
                ("ref_compat_dsph", C.c_int)]


def make_view(view, proj, campos, fx, fy, width, height, near_thresh=0.3, cull_padding=100, mh_dist=3.0,
              background=0.0, l_max=3, ref_compat_dsph=1, tan_fovx=None, tan_fovy=None):
    v = GoldView()
    v.view[:] = [float(x) for x in np.asarray(view, np.float32).reshape(16)]
    v.proj[:] = [float(x) for x in np.asarray(proj, np.float32).reshape(16)]
    v.campos[:] = [float(x) for x in np.asarray(campos, np.float32).reshape(3)]
    v.fx, v.fy = fx, fy
    v.tan_fovx = width / (2.0 * fx) if tan_fovx is None else tan_fovx   # cuda/raster.cu:92-93
    v.tan_fovy = height / (2.0 * fy) if tan_fovy is None else tan_fovy
    v.width, v.height = width, height
    v.near_thresh, v.cull_padding, v.mh_dist = near_thresh, cull_padding, mh_dist
    v.background, v.l_max, v.ref_compat_dsph = background, l_max, ref_compat_dsph
    return v


# ----------------------------------------------------------------------------- per-function (F1..F15)
def camera_space_points(xyz_w, view):
    xyz_w, view = _f(xyz_w), _f(view).reshape(16)
    n = xyz_w.shape[0]
    out = np.empty((n, 3), np.float32)
    lib().gold_camera_space_points(_p(xyz_w), _p(view), n, _p(out))
    return out


def project_to_screen(xyz_c, proj, width, height):
    xyz_c, proj = _f(xyz_c), _f(proj).reshape(16)
    n = xyz_c.shape[0]
    out = np.empty((n, 2), np.float32)
    lib().gold_project_to_screen(_p(xyz_c), _p(proj), n, width, height, _p(out))
    return out


def cull(uv, xyz_c, near_thresh, padding, width, height):
    uv, xyz_c = _f(uv), _f(xyz_c)
    n = uv.shape[0]
    out = np.empty(n, np.uint8)
    lib().gold_cull(_p(uv), _p(xyz_c), n, C.c_float(near_thresh), int(padding), width, height, _p(out))
    return out


def sigma(quat, scale):
    quat, scale = _f(quat), _f(scale)
    n = quat.shape[0]
    out = np.empty((n, 6), np.float32)
    lib().gold_sigma(_p(quat), _p(scale), n, _p(out))
    return out


def conic(xyz_c, view, sig, fx, fy, tan_fovx, tan_fovy, mh_dist):
    xyz_c, view, sig = _f(xyz_c), _f(view).reshape(16), _f(sig)
    n = xyz_c.shape[0]
    J = np.empty((n, 6), np.float32)
    con = np.empty((n, 3), np.float32)
    rad = np.empty((n, 4), np.float32)
    lib().gold_conic(_p(xyz_c), _p(view), _p(sig), C.c_float(fx), C.c_float(fy), C.c_float(tan_fovx),
                     C.c_float(tan_fovy), C.c_float(mh_dist), n, _p(J), _p(con), _p(rad))
    return J, con, rad


def sph(xyz, l_max, grad=False):
    xyz = _f(xyz)
    n = xyz.shape[0]
    nc = (l_max + 1) ** 2
    Y = np.empty((n, nc), np.float32)
    dY = np.empty((n, 3, nc), np.float32) if grad else None
    lib().gold_sph(_p(xyz), n, l_max, _p(Y), _p(dY))
    return (Y, dY) if grad else Y


def sh_forward(xyz_w, sh, band0, campos, l_max):
    xyz_w, band0, campos = _f(xyz_w), _f(band0), _f(campos)
    n = xyz_w.shape[0]
    sh = _f(sh) if sh is not None and l_max > 0 else np.zeros(1, np.float32)
    out = np.empty((n, 3), np.float32)
    lib().gold_sh_forward(_p(xyz_w), _p(sh), _p(band0), _p(campos), l_max, n, _p(out))
    return out


def sorted_list(uv, xyz_c, radius, ntx, nty):
    """Returns dict(C, K, sorted, tiles, ranges)."""
    uv, xyz_c, radius = _f(uv), _f(xyz_c), _f(radius)
    n = uv.shape[0]
    cc, kk = C.c_longlong(0), C.c_longlong(0)
    lib().gold_sorted_list(_p(uv), _p(xyz_c), _p(radius), ntx, nty, n, C.c_longlong(0), C.byref(cc), C.byref(kk),
                           None, None, None)
    K = kk.value
    srt = np.empty(max(K, 1), np.int32)
    tiles = np.empty(max(K, 1), np.int32)
    ranges = np.zeros(ntx * nty + 1, np.int32)
    rc = lib().gold_sorted_list(_p(uv), _p(xyz_c), _p(radius), ntx, nty, n, C.c_longlong(K), C.byref(cc),
                                C.byref(kk), _p(srt), _p(tiles), _p(ranges))
    assert rc == 0
    return dict(C=cc.value, K=K, sorted=srt[:K], tiles=tiles[:K], ranges=ranges)


def render(uv, opacity, con, rgb, bg, srt, ranges, W, H):
    uv, opacity, con, rgb, srt, ranges = _f(uv), _f(opacity), _f(con), _f(rgb), _i(srt), _i(ranges)
    spp = np.zeros((H, W), np.int32)
    wpp = np.zeros((H, W), np.float32)
    img = np.zeros((H, W, 3), np.float32)
    if srt.size == 0:
        srt = np.zeros(1, np.int32)
    lib().gold_render(_p(uv), _p(opacity), _p(con), _p(rgb), C.c_float(bg), _p(srt), _p(ranges), W, H, _p(spp),
                      _p(wpp), _p(img))
    return img, wpp, spp


def render_backward(uv, opacity, con, rgb, bg, srt, ranges, spp, wpp, grad_image, W, H):
    uv, opacity, con, rgb, srt, ranges = _f(uv), _f(opacity), _f(con), _f(rgb), _i(srt), _i(ranges)
    spp, wpp, grad_image = _i(spp), _f(wpp), _f(grad_image)
    n = uv.shape[0]
    g_rgb = np.zeros((n, 3), np.float32)
    g_opa = np.zeros(n, np.float32)
    g_uv = np.zeros((n, 2), np.float32)
    g_con = np.zeros((n, 3), np.float32)
    if srt.size == 0:
        srt = np.zeros(1, np.int32)
    lib().gold_render_backward(_p(uv), _p(opacity), _p(con), _p(rgb), C.c_float(bg), _p(srt), _p(ranges), _p(spp),
                               _p(wpp), _p(grad_image), W, H, _p(g_rgb), _p(g_opa), _p(g_uv), _p(g_con))
    return g_rgb, g_opa, g_uv, g_con


def sh_backward(xyz_w, band0, sh, campos, rgb_grad, l_max, ref_compat_dsph=1, xyz_grad=None):
    xyz_w, band0, campos, rgb_grad = _f(xyz_w), _f(band0), _f(campos), _f(rgb_grad)
    n = xyz_w.shape[0]
    nc = (l_max + 1) ** 2
    sh = _f(sh) if sh is not None and l_max > 0 else np.zeros(1, np.float32)
    sh_grad = np.zeros((n, max(nc - 1, 1), 3), np.float32)
    b0_grad = np.zeros((n, 3), np.float32)
    xyz_grad = np.zeros((n, 3), np.float32) if xyz_grad is None else _f(xyz_grad).copy()
    lib().gold_sh_backward(_p(xyz_w), _p(band0), _p(sh), _p(campos), _p(rgb_grad), l_max, n, int(ref_compat_dsph),
                           _p(sh_grad), _p(b0_grad), _p(xyz_grad))
    return sh_grad[:, :nc - 1], b0_grad, xyz_grad


def conic_backward(J, sig, view, con, conic_grad, J_grad=None, sigma_grad=None):
    J, sig, view, con, conic_grad = _f(J), _f(sig), _f(view).reshape(16), _f(con), _f(conic_grad)
    n = J.shape[0]
    J_grad = np.zeros((n, 6), np.float32) if J_grad is None else _f(J_grad).copy()
    sigma_grad = np.zeros((n, 6), np.float32) if sigma_grad is None else _f(sigma_grad).copy()
    lib().gold_conic_backward(_p(J), _p(sig), _p(view), _p(con), _p(conic_grad), n, _p(J_grad), _p(sigma_grad))
    return J_grad, sigma_grad


def jacobian_backward(xyz_c, fx, fy, tan_fovx, tan_fovy, J_grad, xyz_grad=None):
    xyz_c, J_grad = _f(xyz_c), _f(J_grad)
    n = xyz_c.shape[0]
    xyz_grad = np.zeros((n, 3), np.float32) if xyz_grad is None else _f(xyz_grad).copy()
    lib().gold_jacobian_backward(_p(xyz_c), C.c_float(fx), C.c_float(fy), C.c_float(tan_fovx), C.c_float(tan_fovy),
                                 _p(J_grad), n, _p(xyz_grad))
    return xyz_grad


def sigma_backward(quat, scale, sigma_grad):
    quat, scale, sigma_grad = _f(quat), _f(scale), _f(sigma_grad)
    n = quat.shape[0]
    dq = np.zeros((n, 4), np.float32)
    ds = np.zeros((n, 3), np.float32)
    lib().gold_sigma_backward(_p(quat), _p(scale), _p(sigma_grad), n, _p(dq), _p(ds))
    return dq, ds


def project_to_screen_backward(xyz_c, proj, uv_grad, width, height, xyz_grad=None):
    xyz_c, proj, uv_grad = _f(xyz_c), _f(proj).reshape(16), _f(uv_grad)
    n = xyz_c.shape[0]
    xyz_grad = np.zeros((n, 3), np.float32) if xyz_grad is None else _f(xyz_grad).copy()
    lib().gold_project_to_screen_backward(_p(xyz_c), _p(proj), _p(uv_grad), n, width, height, _p(xyz_grad))
    return xyz_grad


def camera_space_points_backward(view, xyz_c_grad, xyz_w_grad=None):
    view, xyz_c_grad = _f(view).reshape(16), _f(xyz_c_grad)
    n = xyz_c_grad.shape[0]
    xyz_w_grad = np.zeros((n, 3), np.float32) if xyz_w_grad is None else _f(xyz_w_grad).copy()
    lib().gold_camera_space_points_backward(_p(view), _p(xyz_c_grad), n, _p(xyz_w_grad))
    return xyz_w_grad


# ----------------------------------------------------------------------------- view-level pipeline
class State:
    """Forward state (ForwardPassData of the reference, cuda_data.cuh:70-86) held on the C++ side."""
    _F = dict(xyz_c=0, uv=1, sigma=2, J=3, conic=4, radius=5, rgb=6, wpp=7, image=8)
    _I = dict(vis=0, sorted=1, ranges=2, spp=3, mask=4)

    def __init__(self):
        self.h = C.c_void_p(lib().gold_state_new())

    def __del__(self):
        try:
            lib().gold_state_free(self.h)
        except Exception:
            pass

    @property
    def N(self): return lib().gold_state_get(self.h, 0)
    @property
    def P(self): return lib().gold_state_get(self.h, 1)
    @property
    def K(self): return lib().gold_state_get(self.h, 2)
    @property
    def C(self): return lib().gold_state_get(self.h, 3)

    def f(self, name):
        n = lib().gold_state_copy_f(self.h, self._F[name], None)
        out = np.empty(n, np.float32)
        lib().gold_state_copy_f(self.h, self._F[name], _p(out))
        return out

    def i(self, name):
        n = lib().gold_state_copy_i(self.h, self._I[name], None)
        out = np.empty(n, np.int32)
        lib().gold_state_copy_i(self.h, self._I[name], _p(out))
        return out


def forward(vw, xyz, scale, quat, opacity, band0, sh):
    xyz, scale, quat, opacity, band0 = _f(xyz), _f(scale), _f(quat), _f(opacity), _f(band0)
    sh = _f(sh) if sh is not None and vw.l_max > 0 else np.zeros(1, np.float32)
    st = State()
    lib().gold_forward(C.byref(vw), xyz.shape[0], _p(xyz), _p(scale), _p(quat), _p(opacity), _p(band0), _p(sh), st.h)
    return st


def backward(vw, st, grad_image, grads=None, want_intermediates=False):
    """Accumulates into grads (dict of global-index arrays) and returns it."""
    N = st.N
    nc = (vw.l_max + 1) ** 2
    if grads is None:
        grads = dict(xyz=np.zeros((N, 3), np.float32), scale=np.zeros((N, 3), np.float32),
                     quat=np.zeros((N, 4), np.float32), opacity=np.zeros(N, np.float32),
                     band0=np.zeros((N, 3), np.float32), sh=np.zeros((N, max(nc - 1, 1), 3), np.float32))
    gi = _f(grad_image)
    P = st.P
    inter = None
    if want_intermediates:
        inter = dict(grad_rgb=np.zeros((max(P, 1), 3), np.float32), grad_opacity=np.zeros(max(P, 1), np.float32),
                     grad_uv=np.zeros((max(P, 1), 2), np.float32), grad_conic=np.zeros((max(P, 1), 3), np.float32))
    lib().gold_backward(C.byref(vw), st.h, _p(gi), _p(grads["xyz"]), _p(grads["scale"]), _p(grads["quat"]),
                        _p(grads["opacity"]), _p(grads["band0"]), _p(grads["sh"]),
                        _p(inter["grad_rgb"]) if inter else None, _p(inter["grad_opacity"]) if inter else None,
                        _p(inter["grad_uv"]) if inter else None, _p(inter["grad_conic"]) if inter else None)
    if want_intermediates:
        return grads, {k: v[:P] for k, v in inter.items()}
    return grads


# ------------------------------------------------------------------------------------------------ optimiser (8f rank 2)
def adam_step(param, grad, exp_avg, exp_avg_sq, lr, b1, b2, eps, bias1, bias2):
    """adam_kernel, cuda/optimizer.cu:6-28.  Returns (param, exp_avg, exp_avg_sq) as new arrays."""
    p, m, v = _f(param).copy(), _f(exp_avg).copy(), _f(exp_avg_sq).copy()
    g = _f(grad)
    lib().gold_adam_step(_p(p), _p(g), _p(m), _p(v), C.c_float(lr), C.c_float(b1), C.c_float(b2), C.c_float(eps),
                         C.c_float(bias1), C.c_float(bias2), C.c_longlong(p.size))
    return p, m, v


def adam_step_masked(param, grad, exp_avg, exp_avg_sq, mask, lr, b1, b2, eps, bias1, bias2):
    """One parameter group of TrainerImpl::optimizer_step (cuda/trainer.cu:1027-1085), global-index gradients."""
    p, m, v = _f(param).copy(), _f(exp_avg).copy(), _f(exp_avg_sq).copy()
    g = _f(grad)
    n = p.shape[0]
    width = p.size // n if n else 1
    mk = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
    lib().gold_adam_step_masked(_p(p), _p(g), _p(m), _p(v), width, _p(mk), n, C.c_float(lr), C.c_float(b1),
                                C.c_float(b2), C.c_float(eps), C.c_float(bias1), C.c_float(bias2))
    return p, m, v


def accumulate_uv_grad(uv_grad, mask, uv_grad_accum, grad_accum_dur):
    """cuda/trainer.cu:1138-1157.  Returns the advanced (uv_grad_accum, grad_accum_dur)."""
    acc, dur = _f(uv_grad_accum).copy(), _i(grad_accum_dur).copy()
    mk = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
    lib().gold_accumulate_uv_grad(_p(_f(uv_grad)), _p(mk), acc.shape[0], _p(acc), _p(dur))
    return acc, dur


# ------------------------------------------------------------------------------------------------ loss (8f rank 3)
def fused_loss(pred, gt, ssim_weight):
    """fused_loss, cuda/loss.cu:58-471, on [H][W][3] arrays.  Returns (mean loss, d loss / d pred)."""
    p, g = _f(pred), _f(gt)
    H, W = p.shape[0], p.shape[1]
    grad = np.empty_like(p)
    lib().gold_fused_loss.restype = C.c_float
    loss = lib().gold_fused_loss(_p(p), _p(g), H, W, C.c_float(ssim_weight), _p(grad))
    return float(loss), grad


# ------------------------------------------------------------------------------------------------ Morton re-layout (8f rank 4)
def morton_codes(xyz, bbox, true_interleave=False):
    """morton_codes_kernel, cuda/culling.cu:42-68; bbox = (x_max, y_max, z_max, x_min, y_min, z_min).  uint64 codes.
    Default: the reference's own (lossy) spread_bits; true_interleave: the real 3-way interleave."""
    p = _f(xyz)
    codes = np.empty(p.shape[0], np.uint64)
    lib().gold_morton_codes(p.shape[0], _p(p), *[C.c_float(b) for b in bbox], int(true_interleave), _p(codes))
    return codes


def morton_order(xyz, bbox, true_interleave=False):
    """sequence + sort_by_key of TrainerImpl::sort_gaussians (cuda/trainer.cu:886-893), ties in index order"""
    return np.argsort(morton_codes(xyz, bbox, true_interleave), kind="stable").astype(np.int32)


# ------------------------------------------------------------------------------------------------ densification (8f rank 4)
def philox4x32_10(ctr, key):
    """one Philox4x32-10 block (counter [4], key [2], uint32) -- pinned by the published known-answer vectors"""
    c, k, out = (C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), (C.c_uint32 * 4)()
    lib().gold_philox4x32_10_block(c, k, out)
    return [int(x) for x in out]


def density_decide(opacity, scale, uv_grad_accum, grad_accum_dur, uv_grad_threshold, delete_opacity_threshold, scene_extent):
    """decision functors of TrainerImpl::adaptive_density_step (cuda/trainer.cu:416-504): 0 keep, 1 prune, 2 clone, 3 split"""
    op, sc, acc, dur = _f(opacity), _f(scale), _f(uv_grad_accum), _i(grad_accum_dur)
    act = np.empty(op.shape[0], np.uint8)
    lib().gold_density_decide(op.shape[0], _p(op), _p(sc), _p(acc), _p(dur), C.c_float(uv_grad_threshold),
                              C.c_float(delete_opacity_threshold), C.c_float(scene_extent), _p(act))
    return act


def split_children(seed, i, xyz, scale, quat, scale_factor):
    """split_gaussians_kernel (cuda/adaptive_density.cu:69-153) for parent i with the Philox samples: (xyz [2][3], scale [3])"""
    x, s, q = _f(xyz), _f(scale), _f(quat)
    xo, so = np.empty((2, 3), np.float32), np.empty(3, np.float32)
    lib().gold_split_children(C.c_ulonglong(seed), C.c_uint32(i), _p(x), _p(s), _p(q), C.c_float(scale_factor), _p(xo), _p(so))
    return xo, so


def densify(params, moments, action, seed, scale_factor):
    """The compaction + append of adaptive_density_step (cuda/trainer.cu:655-775) given the per-Gaussian actions.
    params: dict name -> [N][w] array (xyz, scale, quaternion, opacity, band0, sh); moments: dict name -> (m, v) or None.
    Survivors (keep + clone parents) first in order, then the clones, then two children per split parent; appended
    Gaussians get zero moments (trainer.cu:694-746)."""
    action = np.asarray(action)
    keep = np.nonzero((action == 0) | (action == 2))[0]
    clone = np.nonzero(action == 2)[0]
    split = np.nonzero(action == 3)[0]
    out_p, out_m = {}, {}
    kids_xyz = np.zeros((split.size, 2, 3), np.float32)
    kids_scale = np.zeros((split.size, 3), np.float32)
    for r, i in enumerate(split):
        kids_xyz[r], kids_scale[r] = split_children(seed, int(i), params["xyz"][i], params["scale"][i], params["quaternion"][i], scale_factor)
    for nm, a in params.items():
        a2 = np.asarray(a, np.float32).reshape(action.size, -1)
        rep = np.repeat(a2[split], 2, axis=0)
        if nm == "xyz":
            rep = kids_xyz.reshape(-1, 3)
        elif nm == "scale":
            rep = np.repeat(kids_scale, 2, axis=0)
        out_p[nm] = np.concatenate([a2[keep], a2[clone], rep], axis=0)
        if moments is not None and nm in moments:
            out_m[nm] = tuple(np.concatenate([np.asarray(m, np.float32).reshape(action.size, -1)[keep],
                                              np.zeros((clone.size + 2 * split.size, a2.shape[1]), np.float32)], axis=0)
                              for m in moments[nm])
    return out_p, out_m
