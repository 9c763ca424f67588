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


def _ensure_unversioned_libcuda():
    """sphericart's CUDA backend dlopen()s "libcuda.so" (submodules/sphericart/sphericart/include/dynamic_cuda.hpp:228);
    a driver-only box may ship libcuda.so.1 alone.  oracle/_ref/shlib is on libref3dgs.so's RUNPATH: put a symlink to the
    driver library this process already mapped (torch loads it) there."""
    try:
        C.CDLL("libcuda.so")
        return
    except OSError:
        pass
    target = None
    try:
        with open("/proc/self/maps") as f:
            for ln in f:
                if "libcuda.so" in ln:
                    target = ln.split()[-1]
                    break
    except OSError:
        pass
    if target is None:
        for cand in ("/usr/lib/x86_64-linux-gnu/libcuda.so.1", "/usr/lib64/libcuda.so.1", "/usr/local/cuda/compat/libcuda.so.1"):
            if os.path.exists(cand):
                target = cand
                break
    if target:
        d = os.path.join(_HERE, "_ref", "shlib")
        os.makedirs(d, exist_ok=True)
        link = os.path.join(d, "libcuda.so")
        if not os.path.exists(link):
            try:
                os.symlink(target, link)
            except OSError:
                pass


def lib():
    global _lib
    if _lib is None:
        _ensure_unversioned_libcuda()
        _lib = C.CDLL(LIB_PATH)
    return _lib


_has_sh = None


def has_sh():
    """True when the reference's SH functions (sphericart CUDA backend: NVRTC JIT at first call) work on this box"""
    global _has_sh
    if _has_sh is None:
        _has_sh = False
        if available() and hasattr(lib(), "ref_precompute_spherical_harmonics") and torch.cuda.is_available():
            try:
                x = torch.tensor([[0.0, 0.0, 1.0]], device="cuda")
                out = sh_forward(x, None, torch.zeros((1, 3), device="cuda"), (0.0, 0.0, 0.0), 0)
                _has_sh = bool(abs(float(out[0, 0].item()) - 0.5) < 1e-6)
            except Exception:
                _has_sh = False
    return _has_sh


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


def sh_forward(xyz_w, sh, band0, campos, l_max):
    """precompute_spherical_harmonics (cuda/spherical_harmonics.cu:62-94)"""
    n, dev = xyz_w.shape[0], xyz_w.device
    out = _new((n, 3), dev)
    rc = lib().ref_precompute_spherical_harmonics(_p(xyz_w), _p(sh if l_max > 0 else None), _p(band0), _f(campos[0]), _f(campos[1]),
                                                  _f(campos[2]), int(l_max), n, _p(out))
    torch.cuda.synchronize()
    if rc != 0:
        raise RuntimeError("reference precompute_spherical_harmonics failed (sphericart / NVRTC)")
    return out


def sh_backward(xyz_w, band0, sh, campos, g_rgb, l_max, g_xyz=None):
    """precompute_spherical_harmonics_backward (cuda/spherical_harmonics_backward.cu:168-209): sh / band-0 gradients
    written, xyz gradient accumulated"""
    n, dev = xyz_w.shape[0], xyz_w.device
    nc = (l_max + 1) ** 2
    g_sh, g_b0 = _new((n, max(nc - 1, 1), 3), dev), _new((n, 3), dev)
    g_xyz = _new((n, 3), dev) if g_xyz is None else g_xyz
    rc = lib().ref_precompute_spherical_harmonics_backward(_p(xyz_w), _p(band0), _p(sh if l_max > 0 else None), _f(campos[0]),
                                                           _f(campos[1]), _f(campos[2]), _p(g_rgb), int(l_max), n,
                                                           _p(g_sh if l_max > 0 else None), _p(g_b0), _p(g_xyz))
    torch.cuda.synchronize()
    if rc != 0:
        raise RuntimeError("reference precompute_spherical_harmonics_backward failed (sphericart / NVRTC)")
    return g_sh[:, : nc - 1], g_b0, g_xyz


def replay_view(g, cam, W, H, l_max, near, pad, mh, grad_image, reps=2):
    """Same-box baseline B1: rasterize_image (cuda/raster.cu:26-135) + backward_pass (cuda/trainer.cu:941-1012) replayed
    call for call on one view through the reference's own kernels, CUDA-event timed per call.  g: dict of device tensors
    (xyz, scale, quaternion, opacity, band0, sh[N][n-1][3] or None).  In the reference's favour: its 16 thrust copy_if
    compactions are done by torch boolean indexing and reported separately.  SH degree l_max when has_sh(), else band 0
    computed with torch (the report says which).  Returns a dict (stage_ms, ms_per_view, P, C, K, sh_degree)."""
    xyz, scale, quat, opa, b0, sh = g["xyz"], g["scale"], g["quaternion"], g["opacity"].reshape(-1), g["band0"], g.get("sh")
    d = lambda a: torch.as_tensor(a, dtype=torch.float32).reshape(-1).cuda()
    view, proj = d(cam["view"]), d(cam["proj"])
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    tx, ty = W / (2 * cam["fx"]), H / (2 * cam["fy"])
    use_sh = has_sh()
    deg = l_max if use_sh else 0
    nsh = (deg + 1) ** 2 - 1

    def timed(acc, name, fn):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        acc[name] = acc.get(name, 0.0) + e0.elapsed_time(e1)
        return out

    def one_view(acc):
        xyz_c = timed(acc, "F1 camera_space_points", lambda: camera_space_points(xyz, view))
        uv = timed(acc, "F2 project_to_screen", lambda: project_to_screen(xyz_c, proj, W, H))
        mask = timed(acc, "F3 cull", lambda: cull(uv, xyz_c, near, pad, W, H))

        def compact():
            m = mask.bool()
            out = [t[m].contiguous() for t in (xyz_c, uv, scale, quat, opa, b0, xyz)]
            out.append(sh[m][:, :nsh].contiguous() if (sh is not None and nsh) else None)
            return out
        xyz_cc, uvc, sc, qc, oc, b0c, xyzw, shc = timed(acc, "compaction (torch indexing; reference: 8 thrust copy_if)", compact)
        if use_sh:
            rgb = timed(acc, "F6 spherical harmonics", lambda: sh_forward(xyzw, shc, b0c, cam["campos"], deg))
        else:
            rgb = (0.5 + 0.28209479 * b0c).contiguous()
        sig = timed(acc, "F4 sigma", lambda: sigma(qc, sc))
        J, con, rad = timed(acc, "F5 conic", lambda: conic(xyz_cc, view, sig, cam["fx"], cam["fy"], tx, ty, mh))
        lst = timed(acc, "F7 sorted list (both phases)", lambda: sorted_list(uvc, xyz_cc, rad, ntx, nty))
        img, wpp, spp = timed(acc, "F8 render", lambda: render(uvc, oc, con, rgb, 0.0, lst["sorted"], lst["ranges"], W, H))
        gr = timed(acc, "F9 render_backward", lambda: render_backward(uvc, oc, con, rgb, 0.0, lst["sorted"], lst["ranges"], spp, wpp, grad_image, W, H))
        gxw = None
        if use_sh:
            gxw = timed(acc, "F10 spherical harmonics backward", lambda: sh_backward(xyzw, b0c, shc, cam["campos"], gr[0], deg))[2]
        gJ, gS = timed(acc, "F11 conic_backward", lambda: conic_backward(J, sig, view, con, gr[3]))
        gx = timed(acc, "F12 jacobian_backward", lambda: jacobian_backward(xyz_cc, cam["fx"], cam["fy"], tx, ty, gJ))
        timed(acc, "F13 sigma_backward", lambda: sigma_backward(qc, sc, gS))
        gx = timed(acc, "F14 project_backward", lambda: project_to_screen_backward(xyz_cc, proj, gr[2], W, H, gx))
        timed(acc, "F15 camera_backward", lambda: camera_space_points_backward(xyzw, view, gx, gxw))
        return int(mask.sum().item()), lst["C"], lst["K"]

    out = {"kind": "reference CUDA kernels (oracle/_ref/libref3dgs.so, sm_100), rasterize_image + backward_pass call sequence",
           "sh_degree": deg, "sh": "sphericart CUDA backend (NVRTC)" if use_sh else "band 0 via torch (sphericart backend not loadable)"}
    try:
        one_view({})                                               # warm-up (NVRTC JIT, allocator)
        runs = []
        for _ in range(reps):
            acc = {}
            P, Cc, K = one_view(acc)
            runs.append(acc)
        # the reference allocates GBs of thrust temporaries inside F7 (96 M coarse pairs) and a sphericart object per SH
        # call: its times swing with the allocator's state, so the FASTEST replay is reported (in its favour) and the
        # spread beside it
        tot = [sum(r.values()) for r in runs]
        best = runs[int(min(range(reps), key=lambda i: tot[i]))]
        st = {k: round(v, 4) for k, v in best.items()}
        out.update(P=P, coarse_pairs_C=Cc, instances_K=K, stage_ms=st, ms_per_view=round(min(tot), 3), replays=reps,
                   ms_per_view_all_replays=[round(t, 3) for t in tot],
                   ms_per_view_kernels_only=round(sum(v for k, v in st.items() if not k.startswith("compaction")), 3))
    except Exception as e:                                         # e.g. out of memory in the coarse pair list
        out["error"] = repr(e)[:300]
    return out
