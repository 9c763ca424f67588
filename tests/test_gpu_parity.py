"""GPU parity tests proper: the CUDA path, called through the C ABI (3dgs_b200.host -> libgsx.so), against the
CPU oracle on the same seeded inputs.  Discrete outputs bit-exact; floats within the tolerances in tests/util.py."""
import os
import numpy as np
import pytest

import kat
import util
from util import G, S, gs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def B():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    import np_backend
    return np_backend


@pytest.mark.parametrize("check", kat.ALL, ids=[c.__name__ for c in kat.ALL])
def test_reference_kats_on_gpu(B, check):
    """every known-answer test of the reference (tests/cuda_forward_test.cpp, cuda_backward_test.cpp) through the C ABI"""
    check(B)


def _scene(N=20000, W=320, H=240, l_max=3, seed=11):
    cam = S.orbit_camera(W, H, azimuth_deg=30.0)
    g = S.synthetic_gaussians(N, l_max=l_max, seed=seed)
    return cam, g


# ------------------------------------------------------------------------------------------ per-function parity
def test_pointwise_forward_parity(B):
    cam, g = _scene()
    W, H = cam["width"], cam["height"]
    xyz_c = B.camera_space_points(g["xyz"], cam["view"])
    np.testing.assert_allclose(xyz_c, G.camera_space_points(g["xyz"], cam["view"]), rtol=1e-5, atol=1e-6)
    uv = B.project_to_screen(xyz_c, cam["proj"], W, H)
    np.testing.assert_allclose(uv, G.project_to_screen(xyz_c, cam["proj"], W, H), rtol=1e-5, atol=1e-3)
    m = B.cull(uv, xyz_c, S.NEAR_THRESH, S.CULL_PADDING, W, H)
    np.testing.assert_array_equal(m, G.cull(uv, xyz_c, S.NEAR_THRESH, S.CULL_PADDING, W, H))   # same inputs: exact
    sig = B.sigma(g["quaternion"], g["scale"])
    util.assert_grad_close(sig, G.sigma(g["quaternion"], g["scale"]), "sigma", rtol=2e-5)   # off-diagonals cancel
    tx, ty = W / (2 * cam["fx"]), H / (2 * cam["fy"])
    J, con, rad = B.conic(xyz_c, cam["view"], sig, cam["fx"], cam["fy"], tx, ty, S.MH_DIST)
    Jg, cong, radg = G.conic(xyz_c, cam["view"], sig, cam["fx"], cam["fy"], tx, ty, S.MH_DIST)
    np.testing.assert_allclose(J, Jg, rtol=1e-5, atol=1e-6)
    util.assert_grad_close(con, cong, "conic", rtol=1e-4)
    assert (rad[:, :2] != radg[:, :2]).mean() <= 1e-4      # ceil() radii: identical up to 1-ulp ties
    np.testing.assert_allclose(rad[:, 2:], radg[:, 2:], atol=2e-5)
    for l in (0, 1, 2, 3):
        nsh = (l + 1) ** 2 - 1
        sh = g["sh"][:, :nsh] if nsh else None
        rgb = B.sh_forward(g["xyz"], sh, g["band0"], cam["campos"], l)
        np.testing.assert_allclose(rgb, G.sh_forward(g["xyz"], sh, g["band0"], cam["campos"], l), rtol=1e-5, atol=2e-6)


def test_pointwise_backward_parity(B):
    cam, g = _scene(N=5000)
    W, H = cam["width"], cam["height"]
    rng = np.random.default_rng(3)
    N = g["xyz"].shape[0]
    xyz_c = G.camera_space_points(g["xyz"], cam["view"])
    sig = G.sigma(g["quaternion"], g["scale"])
    tx, ty = W / (2 * cam["fx"]), H / (2 * cam["fy"])
    J, con, _ = G.conic(xyz_c, cam["view"], sig, cam["fx"], cam["fy"], tx, ty, S.MH_DIST)
    gc = rng.standard_normal((N, 3)).astype(np.float32)
    a = B.conic_backward(J, sig, cam["view"], con, gc)
    b = G.conic_backward(J, sig, cam["view"], con, gc)
    util.assert_grad_close(a[0], b[0], "J_grad")
    util.assert_grad_close(a[1], b[1], "sigma_grad")
    gJ = rng.standard_normal((N, 6)).astype(np.float32)
    pre = rng.standard_normal((N, 3)).astype(np.float32)
    util.assert_grad_close(B.jacobian_backward(xyz_c, cam["fx"], cam["fy"], tx, ty, gJ, pre),
                           G.jacobian_backward(xyz_c, cam["fx"], cam["fy"], tx, ty, gJ, pre), "jacobian_backward(+=)")
    gS = rng.standard_normal((N, 6)).astype(np.float32)
    a = B.sigma_backward(g["quaternion"], g["scale"], gS)
    b = G.sigma_backward(g["quaternion"], g["scale"], gS)
    util.assert_grad_close(a[0], b[0], "quat_grad")
    util.assert_grad_close(a[1], b[1], "scale_grad")
    gu = rng.standard_normal((N, 2)).astype(np.float32)
    util.assert_grad_close(B.project_to_screen_backward(xyz_c, cam["proj"], gu, W, H, pre),
                           G.project_to_screen_backward(xyz_c, cam["proj"], gu, W, H, pre), "project_backward(+=)")
    util.assert_grad_close(B.camera_space_points_backward(cam["view"], pre, gJ[:, :3].copy()),
                           G.camera_space_points_backward(cam["view"], pre, gJ[:, :3].copy()), "cam_backward(+=)")
    grgb = rng.standard_normal((N, 3)).astype(np.float32)
    for l in (0, 1, 2, 3):
        nsh = (l + 1) ** 2 - 1
        sh = g["sh"][:, :nsh] if nsh else None
        for compat in (1, 0):
            a = B.sh_backward(g["xyz"], g["band0"], sh, cam["campos"], grgb, l, ref_compat_dsph=compat, xyz_grad=pre)
            b = G.sh_backward(g["xyz"], g["band0"], sh, cam["campos"], grgb, l, ref_compat_dsph=compat, xyz_grad=pre)
            if nsh:
                util.assert_grad_close(a[0], b[0], f"sh_grad l={l}")
            util.assert_grad_close(a[1], b[1], f"band0_grad l={l}")
            util.assert_grad_close(a[2], b[2], f"xyz_grad l={l} compat={compat}")


# ------------------------------------------------------------------------------------------ sort / ranges
@pytest.mark.parametrize("n,bits", [(0, 45), (1, 45), (4095, 45), (4096, 40), (4097, 64), (100_003, 45), (1_000_000, 47)])
def test_onesweep_sort_exact(n, bits):
    import torch
    H = gs.host
    rng = np.random.default_rng(n + bits)
    hi = rng.integers(0, 1 << (bits - 32), n, dtype=np.uint64)
    lo = rng.integers(0, 1 << 32, n, dtype=np.uint64)
    if n > 10:
        lo[: n // 3] = lo[0]          # many equal keys -> stability is observable
        hi[: n // 3] = hi[0]
    keys = (hi << np.uint64(32)) | lo
    vals = np.arange(n, dtype=np.int32)
    k = torch.as_tensor(keys.view(np.int64)).cuda()
    v = torch.as_tensor(vals).cuda()
    ks, vs = H.sort_pairs_u64(k, v, key_bits=bits)
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(ks.cpu().numpy().view(np.uint64), keys[order])
    np.testing.assert_array_equal(vs.cpu().numpy(), vals[order])      # stable: ties keep input order
    T = 1 << min(bits - 32, 14)
    rg = H.tile_ranges(ks, T).cpu().numpy() if n else H.tile_ranges(k, T).cpu().numpy()
    tiles = (keys[order] >> np.uint64(32)).astype(np.int64)
    np.testing.assert_array_equal(rg, np.searchsorted(np.minimum(tiles, T - 1), np.arange(T + 1), side="left"))


def test_sorted_list_bit_exact(B):
    """instance set, order and ranges vs the oracle on identical (uv, depth, radius): bit-exact, incl. ragged tiles,
    off-screen Gaussians, rotated boxes and equal depths (ties by Gaussian id)."""
    rng = np.random.default_rng(5)
    N, W, H = 30000, 333, 250
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    uv = np.stack([rng.uniform(-120, W + 120, N), rng.uniform(-120, H + 120, N)], 1).astype(np.float32)
    z = rng.uniform(0.3, 9.0, N).astype(np.float32)
    z[::7] = 2.5                                      # equal-depth ties
    xyz = np.stack([np.zeros(N), np.zeros(N), z], 1).astype(np.float32)
    rmaj = np.ceil(rng.exponential(6.0, N) + 1).astype(np.float32)
    rmin = np.ceil(rmaj * rng.uniform(0.1, 1.0, N)).astype(np.float32)
    th = rng.uniform(-np.pi / 2, np.pi / 2, N)
    rad = np.stack([rmaj, rmin, np.sin(th), np.cos(th)], 1).astype(np.float32)
    uv[:50] = np.round(uv[:50] / 16) * 16            # centres exactly on tile borders (inclusive SAT)
    rad[:50, 2:] = [0.0, 1.0]
    a = B.sorted_list(uv, xyz, rad, ntx, nty)
    b = G.sorted_list(uv, xyz, rad, ntx, nty)
    assert a["C"] == b["C"] and a["K"] == b["K"] and a["K"] > N
    np.testing.assert_array_equal(a["ranges"], b["ranges"])
    np.testing.assert_array_equal(a["sorted"], b["sorted"])
    # nothing on screen -> empty list, all-zero ranges (reference: 0-block launch over zero-initialised ranges, Q15)
    uv2 = uv.copy(); uv2[:] = -5000.0
    e = B.sorted_list(uv2, xyz, rad, ntx, nty)
    assert e["K"] == 0 and not e["ranges"].any()


@pytest.mark.parametrize("W,H,N,n_huge", [(333, 250, 30000, 40), (1920, 1080, 200_000, 200), (3840, 2160, 120_000, 300),
                                           (160, 16, 3000, 5)])
def test_direct_tile_binning_equals_sort_path_and_oracle(B, monkeypatch, W, H, N, n_huge):
    """The two second-level binning paths -- emission + radix sort (default) and the direct counting sort on the tile id
    (tilebin.cu, GSX_BINNING=direct) -- give the SAME list and ranges, bit for bit, and both equal the oracle's.  Cases: ragged tiles,
    screen-filling Gaussians (expanded by the whole warp in the middle of a chunk), equal depths, a 4K view (tile-row bands:
    more tiles than 16-bit counters per warp fit shared memory), a single tile row."""
    rng = np.random.default_rng(W * 7 + N)
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    uv = np.stack([rng.uniform(-120, W + 120, N), rng.uniform(-120, H + 120, N)], 1).astype(np.float32)
    z = rng.uniform(0.3, 9.0, N).astype(np.float32)
    z[::5] = 2.5                                      # equal-depth ties (ascending id inside a tile)
    xyz = np.stack([np.zeros(N), np.zeros(N), z], 1).astype(np.float32)
    rmaj = np.ceil(rng.exponential(8.0, N) + 1).astype(np.float32)
    huge = rng.choice(N, n_huge, replace=False)
    rmaj[huge] = np.ceil(rng.uniform(60, 0.6 * max(W, H), n_huge)).astype(np.float32)
    rmin = np.ceil(rmaj * rng.uniform(0.1, 1.0, N)).astype(np.float32)
    th = rng.uniform(-np.pi / 2, np.pi / 2, N)
    rad = np.stack([rmaj, rmin, np.sin(th), np.cos(th)], 1).astype(np.float32)
    monkeypatch.delenv("GSX_BINNING", raising=False)
    a = B.sorted_list(uv, xyz, rad, ntx, nty)
    monkeypatch.setenv("GSX_BINNING", "direct")
    d = B.sorted_list(uv, xyz, rad, ntx, nty)
    o = G.sorted_list(uv, xyz, rad, ntx, nty)
    assert a["K"] == d["K"] == o["K"] and d["K"] > 0
    np.testing.assert_array_equal(d["ranges"], o["ranges"])
    np.testing.assert_array_equal(d["sorted"], o["sorted"])
    np.testing.assert_array_equal(a["ranges"], d["ranges"])
    np.testing.assert_array_equal(a["sorted"], d["sorted"])


# ------------------------------------------------------------------------------------------ compositing
def _comp_inputs(N=4000, W=150, H=100, seed=2, big=False):
    rng = np.random.default_rng(seed)
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    uv = np.stack([rng.uniform(-10, W + 10, N), rng.uniform(-10, H + 10, N)], 1).astype(np.float32)
    z = rng.uniform(0.5, 9.0, N).astype(np.float32)
    xyz = np.stack([np.zeros(N), np.zeros(N), z], 1).astype(np.float32)
    s1 = rng.uniform(1.0, 12.0 if big else 5.0, N); s2 = s1 * rng.uniform(0.2, 1.0, N)
    th = rng.uniform(0, np.pi, N)
    c, s = np.cos(th), np.sin(th)
    cov00 = c * c * s1 ** 2 + s * s * s2 ** 2 + 0.3
    cov11 = s * s * s1 ** 2 + c * c * s2 ** 2 + 0.3
    cov01 = c * s * (s1 ** 2 - s2 ** 2)
    det = cov00 * cov11 - cov01 ** 2
    con = np.stack([cov11 / det, -cov01 / det, cov00 / det], 1).astype(np.float32)
    mid = 0.5 * (cov00 + cov11); lt = np.sqrt(np.maximum(0.1, mid * mid - det))
    ang = 0.5 * np.arctan2(2 * cov01, cov00 - cov11)
    rad = np.stack([np.ceil(3 * np.sqrt(mid + lt)), np.ceil(3 * np.sqrt(mid - lt)), np.sin(ang), np.cos(ang)], 1).astype(np.float32)
    opacity = (2.0 * rng.standard_normal(N)).astype(np.float32)
    opacity[:20] = 30.0       # sigmoid == 1.0f exactly -> Q7 (whole entry dropped in the backward)
    opacity[20:40] = -12.0    # never reaches 1/255
    rgb = rng.uniform(-0.2, 1.2, (N, 3)).astype(np.float32)
    lst = G.sorted_list(uv, xyz, rad, ntx, nty)
    return dict(W=W, H=H, uv=uv, opacity=opacity, conic=con, rgb=rgb, sorted=lst["sorted"], ranges=lst["ranges"])


@pytest.mark.parametrize("big,bg", [(False, 0.0), (True, 0.7)])
def test_render_forward_parity(B, big, bg):
    d = _comp_inputs(big=big, N=6000 if big else 4000)
    a = B.render(d["uv"], d["opacity"], d["conic"], d["rgb"], bg, d["sorted"], d["ranges"], d["W"], d["H"])
    b = G.render(d["uv"], d["opacity"], d["conic"], d["rgb"], bg, d["sorted"], d["ranges"], d["W"], d["H"])
    util.assert_image_close(a[0], b[0], "image")
    util.assert_image_close(a[1], b[1], "final transmittance")
    assert (a[2] != b[2]).mean() <= util.OUTLIER_FRAC          # splats_per_pixel: exact up to threshold flips
    if big:
        assert (b[2] < (d["ranges"][1:] - d["ranges"][:-1]).max()).any()   # early termination is exercised


@pytest.mark.parametrize("big,bg", [(False, 0.0), (True, 0.7)])
def test_render_backward_parity(B, big, bg):
    d = _comp_inputs(big=big, N=6000 if big else 4000, seed=9)
    img, wpp, spp = G.render(d["uv"], d["opacity"], d["conic"], d["rgb"], bg, d["sorted"], d["ranges"], d["W"], d["H"])
    gi = S.grad_image(d["W"], d["H"]) * 1e3
    a = B.render_backward(d["uv"], d["opacity"], d["conic"], d["rgb"], bg, d["sorted"], d["ranges"], spp, wpp, gi, d["W"], d["H"])
    b = G.render_backward(d["uv"], d["opacity"], d["conic"], d["rgb"], bg, d["sorted"], d["ranges"], spp, wpp, gi, d["W"], d["H"])
    for x, y, name in zip(a, b, ("grad_rgb", "grad_opacity", "grad_uv", "grad_conic")):
        util.assert_grad_close(x, y, name, outlier=2e-3)
    assert not a[0][:20].any() and not a[2][:20].any()     # Q7: opa == 1 -> every gradient of the entry dropped
    assert np.abs(b[0]).sum() > 0


# ------------------------------------------------------------------------------------------ view-level path
def _run_view(cam, g, l_max, background, flags=0, sh_alloc=None):
    import torch
    H = gs.host
    N = g["xyz"].shape[0]
    sh = g["sh"]
    if sh is not None and sh_alloc is not None:          # trainer-style: 45 floats allocated, l_max bands used
        full = np.zeros((N, (sh_alloc + 1) ** 2 - 1, 3), np.float32)
        full[:, : sh.shape[1]] = sh
        sh = full
    gd = H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], sh)
    r = H.Rasterizer(N, cam["width"], cam["height"], instance_capacity=1024)   # forces the grow-and-retry path
    v = util.gsx_view(cam, l_max, background, flags)
    stats = r.forward_checked(v, gd)
    torch.cuda.synchronize()
    return H, gd, r, v, stats


@pytest.mark.parametrize("l_max,sh_alloc,W,H", [(3, None, 320, 240), (1, 3, 203, 117), (0, None, 100, 80), (2, None, 64, 64)])
def test_view_forward_chain_consistent(B, l_max, sh_alloc, W, H):
    cam = S.orbit_camera(W, H, azimuth_deg=100.0)
    g = S.synthetic_gaussians(20000, l_max=l_max, seed=21)
    Hh, gd, r, v, stats = _run_view(cam, g, l_max, 0.3, sh_alloc=sh_alloc)
    st = {k: (t.cpu().numpy() if hasattr(t, "cpu") else t) for k, t in r.state().items()}
    gv = util.make_views(cam, l_max, 0.3)
    o = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    N = g["xyz"].shape[0]
    vis = o.i("vis")
    mask = st["mask"].astype(bool)
    # 1. cull mask and per-Gaussian floats vs the oracle's own chain
    assert (mask != o.i("mask").astype(bool)).sum() <= 2
    both = np.intersect1d(vis, np.nonzero(mask)[0])
    comp = np.searchsorted(vis, both)
    assert stats["P"] == mask.sum() and stats["K"] == st["K"]
    np.testing.assert_allclose(st["rec0"][both, :2], o.f("uv").reshape(-1, 2)[both], rtol=1e-5, atol=2e-3)
    np.testing.assert_allclose(st["depth"][both], o.f("xyz_c").reshape(-1, 3)[both, 2], rtol=1e-5, atol=1e-6)
    util.assert_grad_close(st["rec1"][both, :3], o.f("conic").reshape(-1, 3)[comp], "conic", rtol=2e-4)
    np.testing.assert_allclose(st["rec2"][both, :3], o.f("rgb").reshape(-1, 3)[comp], rtol=1e-5, atol=3e-6)
    rg = o.f("radius").reshape(-1, 4)[comp]
    assert (st["radius"][both, :2] != rg[:, :2]).mean() <= 2e-4
    opa = 1.0 / (1.0 + np.exp(-g["opacity"][both].astype(np.float64)))
    np.testing.assert_allclose(st["rec1"][both, 3], opa, rtol=1e-5)
    # 2. binning + sort + ranges: bit-exact vs the oracle fed with the GPU's own per-Gaussian values
    uv = np.where(mask[:, None], st["rec0"][:, :2], -1e6).astype(np.float32)
    rad = np.where(mask[:, None], st["radius"], np.array([0, 0, 0, 1], np.float32)).astype(np.float32)
    xyz_c = np.zeros((N, 3), np.float32); xyz_c[:, 2] = np.where(mask, st["depth"], 1.0)
    ntx, nty = (W + 15) // 16, (H + 15) // 16
    lst = G.sorted_list(uv, xyz_c, rad, ntx, nty)
    assert lst["K"] == st["K"] and st["K"] > 0
    np.testing.assert_array_equal(st["ranges"], lst["ranges"])
    np.testing.assert_array_equal(st["sorted_gaussians"], lst["sorted"])
    keys = st["sorted_keys"].view(np.uint64)
    np.testing.assert_array_equal((keys >> np.uint64(32)).astype(np.int32), lst["tiles"])
    assert (np.diff(keys.astype(np.uint64).view(np.int64)) >= 0).all()           # sortedness
    np.testing.assert_array_equal(st["tiles_touched"][mask].sum(), st["K"])
    # 3. compositing vs the oracle on the GPU's per-Gaussian values and list
    con = st["rec1"][:, :3].copy(); rgb = st["rec2"][:, :3].copy()
    img, wpp, spp = G.render(uv, g["opacity"], con, rgb, 0.3, lst["sorted"], lst["ranges"], W, H)
    util.assert_image_close(st["image"], img)
    util.assert_image_close(st["weight_per_pixel"], wpp, "final T")
    assert (st["splats_per_pixel"] != spp).mean() <= util.OUTLIER_FRAC
    # 4. and end to end vs the oracle's own chain (threshold flips bounded)
    assert util.close_frac(st["image"], o.f("image").reshape(H, W, 3), 1e-3, 1e-4) <= 2e-3
    r.close()


@pytest.mark.parametrize("l_max,sh_alloc,flags", [(3, None, 0), (3, None, 1), (1, 3, 0), (0, None, 0)])
def test_view_backward_parity(B, l_max, sh_alloc, flags):
    import torch
    W, H = 256, 192
    cam = S.orbit_camera(W, H, azimuth_deg=200.0)
    g = S.synthetic_gaussians(15000, l_max=l_max, seed=33)
    Hh, gd, r, v, stats = _run_view(cam, g, l_max, 0.5, flags=flags, sh_alloc=sh_alloc)
    st = {k: (t.cpu().numpy().copy() if hasattr(t, "cpu") else t) for k, t in r.state().items()}
    gi = S.grad_image(W, H) * 1e3
    gd.zero_grads(with_uv=True)
    gi_d = torch.as_tensor(gi).cuda()
    r.backward(v, gd, gi_d)
    r.backward(v, gd, gi_d)          # accumulate semantics: two identical views -> exactly twice the gradient
    torch.cuda.synchronize()
    got = {k: t.cpu().numpy() for k, t in gd.grads.items()}
    assert not r.state()["packed_grads"].any().item()       # the packed buffer is handed back zeroed
    # oracle chain F9 -> F10..F15 on the GPU's forward state (chain-consistent)
    N = g["xyz"].shape[0]
    mask = st["mask"].astype(bool)
    vis = np.nonzero(mask)[0]
    uv = st["rec0"][vis, :2].copy(); con = st["rec1"][vis, :3].copy(); rgb = st["rec2"][vis, :3].copy()
    inv = np.full(N, -1, np.int32); inv[vis] = np.arange(vis.size, dtype=np.int32)
    srt_c = inv[st["sorted_gaussians"]]
    assert (srt_c >= 0).all()
    g_rgb, g_opa, g_uv, g_con = G.render_backward(uv, g["opacity"][vis], con, rgb, 0.5, srt_c, st["ranges"],
                                                  st["splats_per_pixel"], st["weight_per_pixel"], gi, W, H)
    nsh = (l_max + 1) ** 2 - 1
    sh_c = g["sh"][vis, :nsh] if nsh else None
    xyz_c = G.camera_space_points(g["xyz"][vis], cam["view"])
    sig = G.sigma(g["quaternion"][vis], g["scale"][vis])
    tx, ty = W / (2 * cam["fx"]), H / (2 * cam["fy"])
    J, _, _ = G.conic(xyz_c, cam["view"], sig, cam["fx"], cam["fy"], tx, ty, S.MH_DIST)
    shg, b0g, xyzg = G.sh_backward(g["xyz"][vis], g["band0"][vis], sh_c, cam["campos"], g_rgb, l_max, ref_compat_dsph=0 if flags & 1 else 1)
    gJ, gS = G.conic_backward(J, sig, cam["view"], con, g_con)
    gxc = G.jacobian_backward(xyz_c, cam["fx"], cam["fy"], tx, ty, gJ)
    gq, gsc = G.sigma_backward(g["quaternion"][vis], g["scale"][vis], gS)
    gxc = G.project_to_screen_backward(xyz_c, cam["proj"], g_uv, W, H, gxc)
    xyzg = G.camera_space_points_backward(cam["view"], gxc, xyzg)
    exp = dict(xyz=util.scatter(xyzg, vis, N), scale=util.scatter(gsc, vis, N), quaternion=util.scatter(gq, vis, N),
               opacity=util.scatter(g_opa, vis, N), band0=util.scatter(b0g, vis, N), uv=util.scatter(g_uv, vis, N))
    for k, e in exp.items():
        util.assert_grad_close(got[k], 2.0 * e, f"d {k}", outlier=2e-3)
    if nsh:
        stride = gd.sh_stride
        e = np.zeros((N, stride), np.float32)
        e[vis, : nsh * 3] = shg.reshape(vis.size, -1)
        util.assert_grad_close(got["sh"], 2.0 * e, "d sh", outlier=2e-3)
    # overwrite mode: one call, '=' semantics, culled Gaussians get zeros
    v2 = util.gsx_view(cam, l_max, 0.5, flags | 2)
    for t in gd.grads.values():
        t.fill_(123.0)
    r.backward(v2, gd, gi_d)
    torch.cuda.synchronize()
    got1 = {k: t.cpu().numpy() for k, t in gd.grads.items()}
    util.assert_grad_close(got1["xyz"], exp["xyz"], "d xyz (overwrite)", outlier=2e-3)
    util.assert_grad_close(got1["opacity"], exp["opacity"], "d opacity (overwrite)", outlier=2e-3)
    if (~mask).any():
        assert not got1["scale"][~mask].any()
    r.close()


@pytest.mark.parametrize("l_max,sh_alloc,flags,overwrite", [(3, None, 0, True), (3, None, 1, False), (1, 3, 0, True), (0, None, 0, False)])
def test_batch_path_matches_per_view_and_oracle(B, l_max, sh_alloc, flags, overwrite):
    """gsx_batch_*: n views share one pass over the parameters per direction.  Per-view forward state must equal the
    single-view path's, and the batch gradients the sum over the views (per-view path and oracle)."""
    import torch
    W, H, N = 224, 160, 12000
    az = [10.0, 130.0, 250.0]
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in az]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=44)
    Hh = gs.host
    sh = g["sh"]
    if sh is not None and sh_alloc is not None:
        full = np.zeros((N, (sh_alloc + 1) ** 2 - 1, 3), np.float32)
        full[:, : sh.shape[1]] = sh
        sh = full
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], sh)
    gis = [S.grad_image(W, H, seed=70 + k) * 1e3 for k in range(len(az))]
    gis_d = [torch.as_tensor(x).cuda() for x in gis]
    bgs = [0.0, 0.4, 1.0]
    views = [util.gsx_view(c, l_max, bg, flags) for c, bg in zip(cams, bgs)]
    # ---- per-view path: reference for state + summed gradients
    r1 = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    gd.zero_grads(with_uv=True)
    states = []
    for v, gi in zip(views, gis_d):
        r1.forward(v, gd)
        st = {k: (t.cpu().numpy().copy() if hasattr(t, "cpu") else t) for k, t in r1.state().items()}
        states.append(st)
        r1.backward(v, gd, gi)
    torch.cuda.synchronize()
    want = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    r1.close()
    # ---- batch path
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(len(views))
    fill = 0.0 if not overwrite else 77.0
    for t in gd.grads.values():
        t.fill_(fill)
    got_states = []

    def grab(k):
        got_states.append({kk: (t.cpu().numpy().copy() if hasattr(t, "cpu") else t) for kk, t in r.state().items()})

    r.batch_step(views, gd, gis_d, flags=flags | (2 if overwrite else 0), on_image=grab)
    torch.cuda.synchronize()
    got = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    for k, (a, b) in enumerate(zip(got_states, states)):
        assert a["K"] == b["K"] and a["K"] > 0
        np.testing.assert_array_equal(a["mask"], b["mask"])
        np.testing.assert_array_equal(a["tiles_touched"], b["tiles_touched"])
        np.testing.assert_array_equal(a["ranges"], b["ranges"])
        np.testing.assert_array_equal(a["sorted_gaussians"], b["sorted_gaussians"])
        np.testing.assert_array_equal(a["sorted_keys"], b["sorted_keys"])
        m = b["mask"].astype(bool)
        for nm in ("rec0", "rec1", "rec2", "radius"):
            np.testing.assert_allclose(a[nm][m], b[nm][m], rtol=2e-6, atol=1e-6, err_msg=f"view {k} {nm}")
        util.assert_image_close(a["image"], b["image"], f"view {k} image")
        np.testing.assert_array_equal(a["splats_per_pixel"], b["splats_per_pixel"])
    nsh = ((l_max + 1) ** 2 - 1) * 3
    for k in want:
        a_, b_ = (got[k][:, :nsh], want[k][:, :nsh]) if k == "sh" else (got[k], want[k])   # '=' writes the used bands only
        if k == "sh" and nsh == 0:
            continue
        util.assert_grad_close(a_, b_, f"batch d {k} vs per-view sum", outlier=2e-3)
    # the slots' packed gradient buffers are handed back zeroed
    assert not r.state()["packed_grads"].any().item()
    # ---- and against the oracle's own chain, summed over the views (threshold flips bounded: norm-wise)
    tot = None
    for c, bg, gi in zip(cams, bgs, gis):
        gv = util.make_views(c, l_max, bg, flags)
        o = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
        tot = G.backward(gv, o, gi, grads=tot)
    for ko, kg in (("xyz", "xyz"), ("scale", "scale"), ("quat", "quaternion"), ("opacity", "opacity"), ("band0", "band0")):
        a, b = got[kg].reshape(-1), tot[ko].reshape(-1)
        assert np.linalg.norm(a - b) <= util.GRAD_RTOL * np.linalg.norm(b) + 1e-12, (kg, np.linalg.norm(a - b) / np.linalg.norm(b))
    if l_max > 0:
        nsh = ((l_max + 1) ** 2 - 1) * 3
        a = got["sh"][:, :nsh].reshape(-1); b = tot["sh"].reshape(N, -1)[:, :nsh].reshape(-1)
        assert np.linalg.norm(a - b) <= util.GRAD_RTOL * np.linalg.norm(b) + 1e-12, np.linalg.norm(a - b) / np.linalg.norm(b)
    r.close()


@pytest.mark.parametrize("l_max,sh_alloc", [(3, None), (1, 3), (0, None)])
def test_view_parallel_split_exchange_matches_fused(B, l_max, sh_alloc):
    """dp.py's exchange on one GPU: two emulated ranks each run their views with GSX_GRAD_NO_SH and extract the per-view
    d rgb; geometry gradients summed + SH gradients rebuilt from all views' d rgb must equal the fused batch step."""
    import torch
    W, H, N = 200, 144, int(os.environ.get("GSX_TEST_N", "9001"))   # odd: per-view d rgb blocks are not 16-byte aligned
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in (5.0, 95.0, 185.0, 275.0)]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=55)
    Hh, dp, ffi = gs.host, gs.dp, gs._ffi
    sh = g["sh"]
    if sh is not None and sh_alloc is not None:
        full = np.zeros((N, (sh_alloc + 1) ** 2 - 1, 3), np.float32)
        full[:, : sh.shape[1]] = sh
        sh = full
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], sh)
    gis = [torch.as_tensor(S.grad_image(W, H, seed=80 + k) * 1e3).cuda() for k in range(4)]
    views = [util.gsx_view(c, l_max, 0.2) for c in cams]
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(4)
    gd.zero_grads()
    r.batch_step(views, gd, gis, flags=ffi.GSX_GRAD_OVERWRITE)
    torch.cuda.synchronize()
    want = {k: t.clone() for k, t in gd.grads.items()}
    drgb = torch.zeros((4, N, 3), device="cuda")
    geom = {nm: None for nm, _ in dp.GEOM}
    for rank in range(2):
        mine = dp.shard_views(4, 2, rank)
        for t in gd.grads.values():
            t.fill_(55.0)
        r.batch_forward_params([views[k] for k in mine], gd)
        for j, k in enumerate(mine):
            r.batch_forward_view(j)
            r.batch_backward_view(j, gis[k])
            r.batch_extract_view_rgb_grad(j, drgb[k])
        r.batch_backward_params(gd, ffi.GSX_GRAD_OVERWRITE | ffi.GSX_GRAD_NO_SH)
        torch.cuda.synchronize()
        assert (gd.grads["band0"] == 55.0).all() and (gd.grads["sh"] == 55.0).all()   # left out, as documented
        for nm in geom:
            geom[nm] = gd.grads[nm].clone() if geom[nm] is None else geom[nm] + gd.grads[nm]
    dp.sh_grads_from_view_rgb(gd.xyz, [c["campos"] for c in cams], drgb, l_max, gd.grads["sh"], gd.grads["band0"],
                              gd.sh_stride, ffi.GSX_GRAD_OVERWRITE)
    torch.cuda.synchronize()
    nsh = ((l_max + 1) ** 2 - 1) * 3
    for nm in geom:
        util.assert_grad_close(geom[nm].cpu().numpy(), want[nm].cpu().numpy(), f"split d {nm}", outlier=2e-3)
    util.assert_grad_close(gd.grads["band0"].cpu().numpy(), want["band0"].cpu().numpy(), "split d band0", outlier=2e-3)
    if nsh:
        util.assert_grad_close(gd.grads["sh"][:, :nsh].cpu().numpy(), want["sh"][:, :nsh].cpu().numpy(), "split d sh", outlier=2e-3)
    # accumulate mode of the rebuild: a second call doubles
    dp.sh_grads_from_view_rgb(gd.xyz, [c["campos"] for c in cams], drgb, l_max, gd.grads["sh"], gd.grads["band0"],
                              gd.sh_stride, 0)
    torch.cuda.synchronize()
    util.assert_grad_close(gd.grads["band0"].cpu().numpy(), 2.0 * want["band0"].cpu().numpy(), "split d band0 x2", outlier=2e-3)
    r.close()


@pytest.mark.parametrize("binning", ["sort", "direct", "sort-nolists", "sort-shortlists", "sort-keys32", "sort-histpass"])
def test_view_sorted_list_depth_ties_canonical(B, monkeypatch, binning):
    """(both second-level binning paths: emission + radix sort, and the direct counting sort of tilebin.cu; the emission
    from the per-Gaussian forward's tile lists, without them (GSX_TILE_LISTS=0: SAT in the emission kernel), and with a list
    buffer that only holds 3000 entries, so that most Gaussians take the SAT fallback inside k_emit_lists)
    Every Gaussian three times (identical depth bits): the compacted depth keys arrive unordered, so the canonical
    tie order (ascending Gaussian id) exercises k_fix_ties.  Bit-exact against the oracle's list."""
    W, H = 160, 128
    cam = S.orbit_camera(W, H, azimuth_deg=70.0)
    g0 = S.synthetic_gaussians(2500, l_max=0, seed=77)
    rep = lambda a: None if a is None else np.concatenate([a, a, a], axis=0)
    g = {k: rep(v) for k, v in g0.items()}
    perm = np.random.default_rng(5).permutation(g["xyz"].shape[0])            # interleave the copies
    g = {k: (None if v is None else np.ascontiguousarray(v[perm])) for k, v in g.items()}
    monkeypatch.setenv("GSX_BINNING", binning.split("-")[0])
    if binning == "sort-nolists":
        monkeypatch.setenv("GSX_TILE_LISTS", "0")
    if binning == "sort-shortlists":
        monkeypatch.setenv("GSX_TILE_LISTS_CAP", "3000")
    if binning == "sort-keys32":                       # 32-bit tile keys through emission + tile sort (default: 16-bit)
        monkeypatch.setenv("GSX_KEYS16", "0")
    if binning == "sort-histpass":                     # separate histogram pass over the tile keys (default: the emission's)
        monkeypatch.setenv("GSX_EMIT_HIST", "0")
    Hh, gd, r, v, stats = _run_view(cam, g, 0, 0.0)
    st = {k: (t.cpu().numpy() if hasattr(t, "cpu") else t) for k, t in r.state().items()}
    N = g["xyz"].shape[0]
    mask = st["mask"].astype(bool)
    uv = np.where(mask[:, None], st["rec0"][:, :2], -1e6).astype(np.float32)
    rad = np.where(mask[:, None], st["radius"], np.array([0, 0, 0, 1], np.float32)).astype(np.float32)
    xyz_c = np.zeros((N, 3), np.float32); xyz_c[:, 2] = np.where(mask, st["depth"], 1.0)
    lst = G.sorted_list(uv, xyz_c, rad, (W + 15) // 16, (H + 15) // 16)
    assert lst["K"] == st["K"] and st["K"] > 0
    np.testing.assert_array_equal(st["ranges"], lst["ranges"])
    np.testing.assert_array_equal(st["sorted_gaussians"], lst["sorted"])
    keys = st["sorted_keys"].view(np.uint64)
    assert (np.diff(keys.view(np.int64)) >= 0).all()
    d = st["depth"][st["sorted_gaussians"]]
    same = (np.diff(keys.view(np.int64)) == 0)
    assert same.sum() > st["K"] // 4                                           # ties really are everywhere
    assert (np.diff(st["sorted_gaussians"])[same] > 0).all()                   # ... and broken by ascending id
    r.close()


def test_config1_fixture_and_companion(B):
    """C1: the reference's own COLMAP fixture (1 camera, 1 image, 1 point): the Gaussian survives the
    padded cull, touches no tile, the image is pure background and every gradient is zero; plus the non-degenerate
    companion (same camera, 1000 seeded Gaussians) against the oracle end to end."""
    import torch
    cam = S.test_data_camera()
    g1 = S.test_data_gaussian()
    Hh, gd, r, v, stats = _run_view(cam, g1, 0, 0.25)
    st = r.state()
    assert stats["P"] == 1 and stats["K"] == 0
    np.testing.assert_allclose(st["rec0"][0, :2].cpu().numpy(), [138.03, 79.78], atol=0.02)
    assert torch.all(st["image"] == 0.25) and not st["ranges"].any().item()
    gd.zero_grads()
    r.backward(v, gd, torch.ones((80, 100, 3), device="cuda"))
    torch.cuda.synchronize()
    assert all(not t.any().item() for t in gd.grads.values())
    r.close()
    g = S.frustum_gaussians(1000, cam, l_max=0, seed=4)
    Hh, gd, r, v, stats = _run_view(cam, g, 0, 0.25)
    gv = util.make_views(cam, 0, 0.25)
    o = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], None)
    st = r.state()
    assert stats["P"] == o.P and stats["K"] == o.K and o.K > 1000
    np.testing.assert_array_equal(st["ranges"].cpu().numpy(), o.i("ranges"))
    np.testing.assert_array_equal(st["sorted_gaussians"].cpu().numpy(), o.i("vis")[o.i("sorted")])
    util.assert_image_close(st["image"].cpu().numpy(), o.f("image").reshape(80, 100, 3))
    gi = S.grad_image(100, 80) * 1e3
    gd.zero_grads()
    r.backward(v, gd, torch.as_tensor(gi).cuda())
    torch.cuda.synchronize()
    og = G.backward(gv, o, gi)
    for k_o, k_g in (("xyz", "xyz"), ("scale", "scale"), ("quat", "quaternion"), ("opacity", "opacity"), ("band0", "band0")):
        util.assert_grad_close(gd.grads[k_g].cpu().numpy(), og[k_o], f"C1 companion d {k_g}", outlier=5e-3)
    r.close()


def test_empty_and_error_paths(B):
    import torch
    H = gs.host
    cam = S.orbit_camera(64, 48)
    g = S.synthetic_gaussians(256, l_max=0, seed=1)
    g["xyz"][:, 2] -= 100.0                                # everything behind the camera -> P = 0
    Hh, gd, r, v, stats = _run_view(cam, g, 0, 0.1)
    assert stats["P"] == 0 and stats["K"] == 0
    assert torch.all(r.state()["image"] == 0.1)             # the reference exits here (raster.cu:38-41); we render bg
    r.close()
    with pytest.raises(gs._ffi.GsxError):                   # non-device pointer -> status (reference: exit(1))
        gs._ffi.check(gs._ffi.lib().gsx_compute_sigma(None, None, 4, None, None))
    with pytest.raises(gs._ffi.GsxError):
        H.precompute_spherical_harmonics(torch.zeros((4, 3), device="cuda"), None, torch.zeros((4, 3), device="cuda"),
                                         (0, 0, 0), 7)


def test_full_size_properties_c2(B):
    """BASELINE config 2 at full size (1M Gaussians, 1297x840, deg 3): properties that need no oracle run --
    sortedness, ranges consistency, checksum of the id multiset, count conservation, idempotence, linearity of the
    backward in grad_image."""
    import torch
    N, W, H, l_max = S.CONFIGS["C2"]
    cam = S.orbit_camera(W, H, 0.0)
    g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + 2)
    Hh, gd, r, v, stats = _run_view(cam, g, l_max, 0.0)
    st = r.state()
    K = st["K"]
    assert K == int(st["tiles_touched"].sum().item()) and stats["P"] == int(st["mask"].sum().item())
    keys = st["sorted_keys"]
    assert bool((keys[1:] >= keys[:-1]).all())
    tiles = (keys >> 32).to(torch.int64)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    rg = st["ranges"].to(torch.int64)
    assert int(rg[0]) == 0 and int(rg[-1]) == K and bool((rg[1:] >= rg[:-1]).all())
    assert torch.equal(torch.searchsorted(tiles, torch.arange(T + 1, device="cuda")), rg)
    ids = st["sorted_gaussians"].to(torch.int64)
    assert torch.equal(torch.bincount(ids, minlength=N).to(torch.int32), st["tiles_touched"])   # multiset of ids
    img1 = st["image"].clone(); k1 = keys.clone()
    r.forward(v, gd); torch.cuda.synchronize()
    st2 = r.state()
    assert torch.equal(st2["image"], img1) and torch.equal(st2["sorted_keys"], k1)               # idempotent / deterministic
    assert 0.0 <= float(st2["weight_per_pixel"].min()) and float(st2["weight_per_pixel"].max()) <= 1.0
    gi = torch.as_tensor(S.grad_image(W, H) * 1e3).cuda()
    gd.zero_grads(); r.backward(v, gd, gi); torch.cuda.synchronize()
    g1 = {k: t.clone() for k, t in gd.grads.items()}
    gd.zero_grads(); r.backward(v, gd, 2.0 * gi); torch.cuda.synchronize()
    for k in g1:                                                                                  # linearity
        a, b = gd.grads[k], 2.0 * g1[k]
        assert float((a - b).norm()) <= 1e-3 * float(b.norm()) + 1e-12, k
    assert all(bool(torch.isfinite(t).all()) for t in gd.grads.values())
    r.close()


@pytest.mark.parametrize("cfg,backward", [("C2", True), ("C3", True), ("C5", False)])
def test_config_full_size_vs_oracle(B, cfg, backward):
    """BASELINE configs 2, 3 and 5 at FULL size against the CPU oracle (one view; C5 forward only, as the config says):
    P equal, K within a few 1-ulp radius ties, ranges + sorted ids BIT-EXACT against the oracle's binning fed with the
    GPU's own per-Gaussian values, image / final T within 1e-4 with every out-of-tolerance pixel an explained threshold
    flip, splats_per_pixel likewise, all six parameter-gradient arrays within 1e-3 norm-wise (oracle/parity.py).
    C4 gets the same report inside bench.py's cpu_baseline leg (`parity_c4`) and tests/test_ref_cuda.py."""
    import torch
    from oracle import parity
    G.lib().gold_set_threads(os.cpu_count() or 1)
    N, W, H, l_max = S.CONFIGS[cfg]
    cam = S.orbit_camera(W, H, 0.0)
    g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + int(cfg[1:]))
    Hh = gs.host
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    r = Hh.Rasterizer(N, W, H, instance_capacity=4 * N)
    v = util.gsx_view(cam, l_max, 0.0)
    r.forward_checked(v, gd)
    torch.cuda.synchronize()
    st = {k: (t.cpu().numpy() if hasattr(t, "cpu") else t) for k, t in r.state().items() if k != "sorted_keys"}
    gi = S.grad_image(W, H) * 1e3
    got = None
    if backward:
        gd.zero_grads()
        r.backward(v, gd, torch.as_tensor(gi).cuda())
        torch.cuda.synchronize()
        got = {k: t.cpu().numpy() for k, t in gd.grads.items()}
    gv = util.make_views(cam, l_max, 0.0)
    o = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    og = G.backward(gv, o, gi) if backward else None
    rep = parity.full_view_report(st, got, o, og, N, W, H, 0.0)
    print(cfg, "parity:", rep)
    parity.assert_report(rep)
    r.close()


@pytest.mark.parametrize("l_max", [3, 0])
def test_contributing_lists_same_image_and_gradients(B, l_max):
    """GSX_LISTS_CONTRIBUTING (view flag 8): per-tile lists clipped to each Gaussian's alpha >= 1/255 footprint.  The
    image must be BIT-identical to the reference-identical lists' (the same entries contribute, in the same order), the
    gradients equal up to the atomics' summation order, every per-tile list a subsequence of the exact one, and no
    pixel may have consumed an entry that the exact path did not."""
    import torch
    W, H, N = 224, 160, 12000
    az = [10.0, 130.0, 250.0]
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in az]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=45)
    Hh = gs.host
    gis_d = [torch.as_tensor(S.grad_image(W, H, seed=80 + k) * 1e3).cuda() for k in range(len(az))]
    bgs = [0.0, 0.4, 1.0]
    res = {}
    for tight in (0, 8):
        gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
        views = [util.gsx_view(c, l_max, bg, tight) for c, bg in zip(cams, bgs)]
        r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
        r.batch_reserve(len(views))
        states = []
        r.batch_step(views, gd, gis_d, flags=2,
                     on_image=lambda k: states.append({kk: (t.cpu().numpy().copy() if hasattr(t, "cpu") else t)
                                                       for kk, t in r.state().items()}))
        torch.cuda.synchronize()
        res[tight] = (states, {k: t.cpu().numpy().copy() for k, t in gd.grads.items()})
        r.close()
    (se, ge), (st, gt) = res[0], res[8]
    kept = 0
    for k, (a, b) in enumerate(zip(st, se)):
        assert 0 < a["K"] < b["K"], (a["K"], b["K"])
        kept += a["K"]
        np.testing.assert_array_equal(a["image"], b["image"], err_msg=f"view {k}: image not bit-identical")
        assert (a["splats_per_pixel"] <= b["splats_per_pixel"]).all()
        ra, rb = a["ranges"], b["ranges"]
        la, lb = a["sorted_gaussians"], b["sorted_gaussians"]
        for t in range(len(ra) - 1):
            xa, xb = la[ra[t]:ra[t + 1]], lb[rb[t]:rb[t + 1]]
            pos = {int(v): i for i, v in enumerate(xb)}
            idx = [pos[int(v)] for v in xa]          # KeyError = an instance the exact list does not have
            assert all(i < j for i, j in zip(idx, idx[1:])), f"view {k} tile {t}: order changed"
    nsh = ((l_max + 1) ** 2 - 1) * 3
    for k in ge:
        if k == "sh" and nsh == 0:
            continue
        a_, b_ = (gt[k][:, :nsh], ge[k][:, :nsh]) if k == "sh" else (gt[k], ge[k])
        util.assert_grad_close(a_, b_, f"d {k}: contributing-only vs exact lists", outlier=2e-3)
    print("instances kept:", kept, "of", sum(b["K"] for b in se))


@pytest.mark.parametrize("pinned", [True, False])
def test_host_buffer_entry_points_match_device_path(B, pinned):
    """gsx_batch_render_host (the call bench.py's e2e leg times: uploads / downloads pipelined on copy streams) and
    gsx_render_view_host must return the images and gradients of the device-buffer path, twice in a row (the staging
    buffers and events are re-used), with pinned and with pageable host memory."""
    import torch
    W, H, N, l_max = 224, 160, 12000, 3
    az = [10.0, 130.0, 250.0, 300.0]
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in az]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=46)
    Hh = gs.host
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    views = [util.gsx_view(c, l_max, bg) for c, bg in zip(cams, [0.0, 0.4, 1.0, 0.2])]
    mk = (lambda t: t.pin_memory()) if pinned else (lambda t: t)
    gis_h = [mk(torch.as_tensor(S.grad_image(W, H, seed=90 + k) * 1e3)) for k in range(len(az))]
    gis_d = [t.cuda() for t in gis_h]
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(len(views))
    want_img = []
    r.batch_step(views, gd, gis_d, flags=2, on_image=lambda k: want_img.append(r.state_image().cpu().numpy().copy()))
    torch.cuda.synchronize()
    want = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    for rep in range(2):
        imgs_h = [mk(torch.full((H, W, 3), -1.0, dtype=torch.float32)) for _ in views]
        for t in gd.grads.values():
            t.fill_(55.0)
        r.batch_render_host(views, gd, gis_h, imgs_h, flags=2)
        got = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
        for k, (a, b) in enumerate(zip(imgs_h, want_img)):
            np.testing.assert_array_equal(a.numpy(), b, err_msg=f"rep {rep} view {k}: host image")
        for k in want:
            util.assert_grad_close(got[k], want[k], f"rep {rep} d {k}: host-buffer batch vs device batch", outlier=2e-3)
    # single-view host entry point: same image, gradients accumulate onto the batch's
    img1 = mk(torch.empty((H, W, 3), dtype=torch.float32))
    for t in gd.grads.values():
        t.zero_()
    r.render_view_host(views[1], gd, gis_h[1], img1)
    np.testing.assert_array_equal(img1.numpy(), want_img[1])
    g1 = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    for t in gd.grads.values():
        t.zero_()
    r.forward(views[1], gd)
    r.backward(views[1], gd, gis_d[1])
    torch.cuda.synchronize()
    for k in g1:
        util.assert_grad_close(g1[k], gd.grads[k].cpu().numpy(), f"d {k}: render_view_host vs forward+backward", outlier=2e-3)
    r.close()


def test_batch_pipeline_call_orders(B):
    """The context bins views ahead on side streams (pre-binning of all views at gsx_batch_forward_params, emission + tile
    sort of view k+1 under the compositing of view k, two instance-buffer sets).  Whatever the call order -- in order,
    out of order, a view repeated, a batch abandoned half-way, a single-view forward in between -- every view's state
    must equal the in-order one's."""
    import torch
    W, H, N, l_max = 224, 160, 12000, 2
    az = [10.0, 100.0, 190.0, 280.0, 330.0]
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in az]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=47)
    Hh = gs.host
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    views = [util.gsx_view(c, l_max, 0.1 * k) for k, c in enumerate(cams)]
    gi = torch.as_tensor(S.grad_image(W, H, seed=99) * 1e3).cuda()
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(len(views))
    keys = ("K", "ranges", "sorted_gaussians", "sorted_keys", "splats_per_pixel", "image")

    def snap():
        st = r.state()
        return {k: (st[k].cpu().numpy().copy() if hasattr(st[k], "cpu") else st[k]) for k in keys}

    def same(a, b, what):
        assert a["K"] == b["K"], what
        for k in keys[1:]:
            np.testing.assert_array_equal(a[k][: a["K"]] if k.startswith("sorted") else a[k],
                                          b[k][: b["K"]] if k.startswith("sorted") else b[k], err_msg=f"{what}: {k}")

    # in order, forward + backward: the reference states and gradients
    want = []
    r.batch_step(views, gd, gi, flags=2, on_image=lambda k: want.append(snap()))
    torch.cuda.synchronize()
    want_g = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    # out of order, forward only
    r.batch_forward_params(views, gd)
    for k in (3, 0, 4, 1, 2):
        r.batch_forward_view(k)
        same(snap(), want[k], f"out of order, view {k}")
    # a view repeated, then the rest in order with backward
    r.batch_forward_params(views, gd)
    r.batch_forward_view(0)
    r.batch_forward_view(0)
    same(snap(), want[0], "repeated view 0")
    # abandoned half-way (views 1.. binned ahead, never consumed), a single-view forward in between, then a full batch
    r.batch_forward_params(views, gd)
    r.batch_forward_view(0)
    r.forward(views[2], gd)
    same(snap(), want[2], "single-view forward after an abandoned batch")
    got = []
    for t in gd.grads.values():
        t.fill_(3.0)
    r.batch_step(views, gd, gi, flags=2, on_image=lambda k: got.append(snap()))
    torch.cuda.synchronize()
    for k in range(len(views)):
        same(got[k], want[k], f"full batch after the detours, view {k}")
    for k in want_g:
        util.assert_grad_close(gd.grads[k].cpu().numpy(), want_g[k], f"d {k} after the detours", outlier=2e-3)
    r.close()


def test_capacity_overflow_is_safe_and_reported_for_every_view(B, monkeypatch):
    """a view whose K exceeds the instance capacity must (a) never touch the unwritten instance buffers --
    they are poisoned with 0xFF here -- (b) render background only and contribute no gradients, (c) be reported by
    gsx_ctx_stats for ANY view of the batch (not only the last one forwarded) and by the no-sync poll gsx_ctx_overflow;
    after reserve() the same batch must equal a never-overflowed run."""
    import torch
    monkeypatch.setenv("GSX_DEBUG_POISON", "1")
    W, H, N, l_max = 224, 160, 12000, 1
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in (20.0, 140.0, 260.0)]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=48)
    Hh = gs.host
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    views = [util.gsx_view(c, l_max, 0.25) for c in cams]
    gi = torch.as_tensor(S.grad_image(W, H, seed=98) * 1e3).cuda()
    big = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    big.batch_reserve(3)
    Ks = []
    big.batch_step(views, gd, gi, flags=2, on_image=lambda k: Ks.append(big.state()["K"]))
    torch.cuda.synchronize()
    want = {k: t.cpu().numpy().copy() for k, t in gd.grads.items()}
    want_img = big.state()["image"].cpu().numpy().copy()
    big.close()
    assert min(Ks) > 2000
    # capacity between the smallest and the largest K would be ideal; make only view 1 overflow by sizing for the others
    order = np.argsort(Ks)
    cap = (Ks[order[1]] + Ks[order[2]]) // 2 if Ks[order[1]] != Ks[order[2]] else Ks[order[2]] - 1
    r = Hh.Rasterizer(N, W, H, instance_capacity=cap)
    r.batch_reserve(3)
    assert r.overflow() == 0
    imgs = []
    for t in gd.grads.values():
        t.fill_(9.0)
    r.batch_step(views, gd, gi, flags=2, on_image=lambda k: imgs.append(r.state()["image"].cpu().numpy().copy()))
    torch.cuda.synchronize()                                # no illegal address although the buffers hold 0xFF
    over = int(order[2])
    assert (imgs[over] == np.float32(0.25)).all()           # the overflowed view composited nothing
    assert r.overflow() == Ks[over]                         # no-sync poll carries the K to reserve for
    st = r.stats(raise_on_overflow=False)
    assert st["overflow"] and st["K"] == Ks[over]           # reported although the LAST view fitted (when over != 2)
    assert all(bool(torch.isfinite(t).all()) for t in gd.grads.values())
    with pytest.raises(gs._ffi.GsxError):                   # the next step refuses to run on a lossy context
        r.batch_step(views, gd, gi, flags=2)
    r.reserve(st["K"] + 16)
    assert r.overflow() == 0
    r.batch_step(views, gd, gi, flags=2)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(r.state()["image"].cpu().numpy(), want_img)
    for k in want:
        util.assert_grad_close(gd.grads[k].cpu().numpy(), want[k], f"d {k} after reserve")
    r.close()
    # same protection on the per-function F7 entry point: capacity below K -> K reported, list empty, ranges zero
    cam, gg = _scene(N=3000, W=160, H=120, l_max=0)
    xyz_c = G.camera_space_points(gg["xyz"], cam["view"])
    uv = G.project_to_screen(xyz_c, cam["proj"], 160, 120)
    sig = G.sigma(gg["quaternion"], gg["scale"])
    _, _, rad = G.conic(xyz_c, cam["view"], sig, cam["fx"], cam["fy"], 160 / (2 * cam["fx"]), 120 / (2 * cam["fy"]), S.MH_DIST)
    full = B.sorted_list(uv, xyz_c, rad, 10, 8)
    assert full["K"] > 64
    d = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).cuda()
    srt = torch.full((64,), -1, dtype=torch.int32, device="cuda")
    ranges = torch.full((81,), -1, dtype=torch.int32, device="cuda")
    kdev = torch.zeros(1, dtype=torch.int32, device="cuda")
    import ctypes as C
    L = gs._ffi.lib()
    a_uv, a_xyz, a_rad = d(uv), d(xyz_c), d(rad)
    gs._ffi.check(L.gsx_get_sorted_gaussian_list(C.c_void_p(a_uv.data_ptr()), C.c_void_p(a_xyz.data_ptr()), C.c_void_p(a_rad.data_ptr()),
                                                 10, 8, 3000, C.c_longlong(64), C.c_void_p(srt.data_ptr()), C.c_void_p(ranges.data_ptr()),
                                                 C.c_void_p(kdev.data_ptr()), C.c_void_p(0), C.c_size_t(0), C.c_void_p(0)))
    torch.cuda.synchronize()
    assert int(kdev.item()) == full["K"] and not ranges.any().item() and (srt == -1).all().item()


def test_batch_graph_replay_matches_eager(B):
    """gsx_batch_graph_capture / gsx_graph_launch (CUDA-graph replay): the whole batch step -- side streams
    included -- captured once and replayed must give the eager path's images and gradients; with ground-truth images the loss
    gradient of every view is computed inside the graph (gsx_fused_loss) and must equal forward -> fused_loss -> backward."""
    import torch
    W, H, N, l_max = 224, 160, 12000, 2
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in (15.0, 135.0, 255.0)]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=49)
    Hh = gs.host
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    views = [util.gsx_view(c, l_max, 0.2) for c in cams]
    gis = [torch.as_tensor(S.grad_image(W, H, seed=60 + k) * 1e3).cuda() for k in range(3)]
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(3)
    imgs = []
    r.batch_step(views, gd, gis, flags=2, on_image=lambda k: imgs.append(r.state_image().clone()))
    torch.cuda.synchronize()
    want = {k: t.clone() for k, t in gd.grads.items()}
    # ---- given loss gradients
    out_imgs = [torch.zeros((H, W, 3), device="cuda") for _ in range(3)]
    h = r.batch_graph_capture(views, gd, grad_images=gis, images_out=out_imgs)
    for rep in range(2):
        for t in gd.grads.values():
            t.fill_(5.0)
        r.graph_launch(h)
        torch.cuda.synchronize()
        for k in range(3):
            assert torch.equal(out_imgs[k], imgs[k]), f"replay {rep}: image {k}"
        for k in want:
            util.assert_grad_close(gd.grads[k].cpu().numpy(), want[k].cpu().numpy(), f"replay {rep}: d {k}")
    # the caller refreshes the CONTENTS of the gradient images between launches
    for t in gis:
        t.mul_(2.0)
    r.graph_launch(h)
    torch.cuda.synchronize()
    util.assert_grad_close(gd.grads["xyz"].cpu().numpy(), 2.0 * want["xyz"].cpu().numpy(), "replay with refreshed loss gradients")
    Hh.Rasterizer.graph_destroy(h)
    assert r.overflow() == 0
    # ---- loss inside the graph
    gts = [(im + 0.05 * torch.randn_like(im)).clamp_(0, 1) for im in imgs]
    losses = torch.zeros(3, device="cuda")
    h2 = r.batch_graph_capture(views, gd, gt_images=gts, ssim_weight=0.2, losses=losses)
    r.graph_launch(h2)
    torch.cuda.synchronize()
    got = {k: t.clone() for k, t in gd.grads.items()}
    gis2, ls = [], []
    for im, gt in zip(imgs, gts):
        l_, gi_ = Hh.fused_loss(im, gt, 0.2)
        ls.append(l_); gis2.append(gi_)
    r.batch_step(views, gd, gis2, flags=2)
    torch.cuda.synchronize()
    np.testing.assert_allclose(losses.cpu().numpy(), ls, rtol=1e-6, atol=1e-7)
    for k in got:
        util.assert_grad_close(got[k].cpu().numpy(), gd.grads[k].cpu().numpy(), f"loss-in-graph d {k}")
    Hh.Rasterizer.graph_destroy(h2)
    r.close()


# ------------------------------------------------------------------------------------------ optimiser
_ADAM = dict(b1=0.9, b2=0.999, eps=1e-8)
# float tolerance of the Adam step: the device contracts b*m + (1-b)*g into a fused multiply-add, the oracle
# (-ffp-contract=off) does not: <= 1 ulp on each moment, a few ulp on the step, <= 1 ulp of the parameter
_ADAM_TOL_MOM = dict(rtol=4e-7, atol=1e-37)
_ADAM_TOL_PAR = dict(rtol=2.5e-7, atol=1e-7)   # steps are lr * O(0.1) = 1e-5 .. 1e-2 here


def _adam_arrays(rng, n, w):
    """moments consistent with the gradient history (m ~ g, v ~ g^2, as in a running optimiser), so that no sum in the
    update cancels and the fused-multiply-add difference stays at the ulp level the tolerances above state"""
    p = rng.normal(0, 1, (n, w)).astype(np.float32)
    g = (rng.normal(0, 1, (n, w)) * 10.0 ** rng.uniform(-6, 0, (n, 1))).astype(np.float32)
    m = (g * rng.uniform(0.5, 1.5, (n, w))).astype(np.float32)
    v = (g.astype(np.float64) ** 2 * rng.uniform(0.5, 1.5, (n, w))).astype(np.float32)
    return p, g, m, v


@pytest.mark.parametrize("n", [1, 3, 1024, 4099])
def test_adam_step_parity(B, n):
    """gsx_adam_step against the oracle's restatement of adam_kernel (cuda/optimizer.cu:6-28); NaN gradients count as 0"""
    rng = np.random.default_rng(100 + n)
    p, g, m, v = _adam_arrays(rng, n, 1)
    g[rng.random(g.shape) < 0.05] = np.nan
    it = 7
    bias1, bias2 = 1 - _ADAM["b1"] ** it, 1 - _ADAM["b2"] ** it
    got = B.adam_step(p, g, m, v, 1.6e-4, bias1=bias1, bias2=bias2, **_ADAM)
    exp = G.adam_step(p, g, m, v, 1.6e-4, bias1=bias1, bias2=bias2, **_ADAM)
    assert np.isfinite(got[0]).all()
    np.testing.assert_allclose(got[0], exp[0], **_ADAM_TOL_PAR)
    np.testing.assert_allclose(got[1], exp[1], **_ADAM_TOL_MOM)
    np.testing.assert_allclose(got[2], exp[2], **_ADAM_TOL_MOM)


def test_adam_step_unaligned_views():
    """arrays that start 4 bytes into an allocation take the scalar path and give the same bits as the float4 path"""
    import torch
    H = gs.host
    rng = np.random.default_rng(5)
    n = 10001
    p, g, m, v = (torch.from_numpy(a.reshape(-1)).cuda() for a in _adam_arrays(rng, n, 1))
    pa, ma, va = p.clone(), m.clone(), v.clone()
    H.adam_step(pa, g, ma, va, 1e-3, 0.9, 0.999, 1e-8, 0.1, 0.001)
    shifted = [torch.zeros(n + 1, device="cuda")[1:] for _ in range(4)]
    for dst, src in zip(shifted, (p, g, m, v)):
        dst.copy_(src)
    assert shifted[0].data_ptr() % 16 == 4
    H.adam_step(shifted[0], shifted[1], shifted[2], shifted[3], 1e-3, 0.9, 0.999, 1e-8, 0.1, 0.001)
    assert torch.equal(shifted[0], pa) and torch.equal(shifted[2], ma) and torch.equal(shifted[3], va)


@pytest.mark.parametrize("l_max,with_mask", [(0, True), (3, True), (3, False), (1, True)])
def test_adam_fused_parity(B, l_max, with_mask):
    """gsx_adam_step_fused (every parameter group + densification accumulators in one launch, global-index gradients,
    cull mask read in place) against the oracle's compact -> adam_step -> scatter (cuda/trainer.cu:1027-1158)"""
    rng = np.random.default_rng(40 + l_max)
    n = 10007
    widths = [3, 3, 1, 3, 4] + ([3 * ((l_max + 1) ** 2 - 1)] if l_max > 0 else [])
    lrs = [1.6e-4, 2.5e-3, 5e-2, 5e-3, 1e-3, 1.25e-4][:len(widths)]
    groups = [(*_adam_arrays(rng, n, w), lr) for w, lr in zip(widths, lrs)]
    mask = (rng.random(n) < 0.55).astype(np.uint8) if with_mask else None
    uv_grad = rng.normal(0, 1e-3, (n, 2)).astype(np.float32)
    acc0 = rng.random(n).astype(np.float32)
    dur0 = rng.integers(0, 100, n).astype(np.int32)
    it = 3
    bias1, bias2 = 1 - _ADAM["b1"] ** it, 1 - _ADAM["b2"] ** it
    got, (acc, dur) = B.adam_step_fused(groups, mask, bias1=bias1, bias2=bias2, uv_grad=uv_grad, uv_grad_accum=acc0,
                                        grad_accum_dur=dur0, **_ADAM)
    keep = np.ones(n, bool) if mask is None else mask.astype(bool)
    for (p, g, m, v, lr), (gp, gm, gv) in zip(groups, got):
        ep, em, ev = G.adam_step_masked(p, g, m, v, mask, lr, bias1=bias1, bias2=bias2, **_ADAM)
        np.testing.assert_allclose(gp, ep, **_ADAM_TOL_PAR)
        np.testing.assert_allclose(gm, em, **_ADAM_TOL_MOM)
        np.testing.assert_allclose(gv, ev, **_ADAM_TOL_MOM)
        # culled Gaussians: parameters and moments bit-identical to the input
        np.testing.assert_array_equal(gp[~keep], p[~keep])
        np.testing.assert_array_equal(gm[~keep], m[~keep])
        np.testing.assert_array_equal(gv[~keep], v[~keep])
    eacc, edur = G.accumulate_uv_grad(uv_grad, mask, acc0, dur0)
    np.testing.assert_allclose(acc, eacc, rtol=2e-7, atol=1e-9)
    np.testing.assert_array_equal(dur, edur)


def test_adam_fused_equals_per_array_at_size():
    """size-independent property at 1 M Gaussians, SH degree 3 (59 floats each): with every Gaussian visible the fused
    launch gives the same bits as one gsx_adam_step per array (the reference's call pattern, trainer.cu:1052-1124);
    with a mask, exactly the masked-out rows keep their bits"""
    import torch
    H = gs.host
    n = 1_000_000
    gen = torch.Generator(device="cuda").manual_seed(3)
    widths, lrs = [3, 3, 1, 3, 4, 45], [1.6e-4, 2.5e-3, 5e-2, 5e-3, 1e-3, 1.25e-4]
    mk = lambda w, s: torch.randn((n, w), device="cuda", generator=gen) * s
    groups = [(mk(w, 1.0), mk(w, 1e-3), mk(w, 1e-3), mk(w, 1e-3) ** 2, lr) for w, lr in zip(widths, lrs)]
    args = (0.9, 0.999, 1e-8, 1 - 0.9 ** 5, 1 - 0.999 ** 5)
    fused = [(p.clone(), g, m.clone(), v.clone(), lr) for p, g, m, v, lr in groups]
    H.adam_step_fused(fused, None, *args)
    for (p, g, m, v, lr), (fp, _, fm, fv, _) in zip(groups, fused):
        sp, sm, sv = p.clone(), m.clone(), v.clone()
        H.adam_step(sp, g, sm, sv, lr, *args)
        assert torch.equal(sp, fp) and torch.equal(sm, fm) and torch.equal(sv, fv)
    mask = (torch.rand(n, device="cuda", generator=gen) < 0.5).to(torch.uint8)
    masked = [(p.clone(), g, m.clone(), v.clone(), lr) for p, g, m, v, lr in groups]
    H.adam_step_fused(masked, mask, *args)
    keep = mask.bool()
    for (p, _, m, v, _), (fp, _, fm, fv, _), (mp, _, mm, mv, _) in zip(groups, fused, masked):
        assert torch.equal(mp[~keep], p[~keep]) and torch.equal(mm[~keep], m[~keep]) and torch.equal(mv[~keep], v[~keep])
        assert torch.equal(mp[keep], fp[keep]) and torch.equal(mm[keep], fm[keep]) and torch.equal(mv[keep], fv[keep])


def test_adam_rejects_host_pointers():
    """pointer checks of the reference (ASSERT_DEVICE_POINTER in adam_step, cuda/optimizer.cu:33-36) as a status code"""
    import ctypes as C
    from importlib import import_module
    ffi = import_module("3dgs_b200._ffi")
    host = np.zeros(16, np.float32)
    rc = ffi.lib().gsx_adam_step(host.ctypes.data_as(C.c_void_p), host.ctypes.data_as(C.c_void_p),
                                 host.ctypes.data_as(C.c_void_p), host.ctypes.data_as(C.c_void_p), C.c_float(1e-3),
                                 C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8), C.c_float(0.1), C.c_float(0.001),
                                 C.c_longlong(16), None)
    assert rc in (1, 2)


# ------------------------------------------------------------------------------------------ loss
def _loss_images(rng, H, W):
    yy, xx = np.mgrid[0:H, 0:W]
    base = 0.5 + 0.3 * np.sin(xx / 7.0)[..., None] * np.cos(yy / 5.0)[..., None] * np.array([1.0, 0.7, -0.5])
    pred = np.clip(base + rng.normal(0, 0.05, (H, W, 3)), 0, 1).astype(np.float32)
    gt = np.clip(base + rng.normal(0, 0.05, (H, W, 3)), 0, 1).astype(np.float32)
    return pred, gt


@pytest.mark.parametrize("H,W,w", [(16, 16, 0.2), (1, 1, 0.2), (7, 3, 0.5), (37, 45, 0.2), (80, 100, 0.2), (33, 129, 0.0),
                                   (131, 67, 1.0)])
def test_fused_loss_parity(B, H, W, w):
    """gsx_fused_loss against the oracle's restatement of fused_loss (cuda/loss.cu:58-471): ragged sizes, images smaller
    than the window, pure L1 and pure SSIM.  Tolerance: loss 2e-6 absolute (float sums in a different order); gradient
    1e-4 of its largest magnitude (fused multiply-adds on the device, none in the oracle)"""
    rng = np.random.default_rng(H * 1000 + W)
    pred, gt = _loss_images(rng, H, W)
    pred[0, 0] = gt[0, 0]                                       # the p == g subgradient (-1, cuda/loss.cu:421)
    loss, grad = B.fused_loss(pred, gt, w)
    eloss, egrad = G.fused_loss(pred, gt, w)
    assert abs(loss - eloss) < 2e-6, (loss, eloss)
    np.testing.assert_allclose(grad, egrad, rtol=0, atol=1e-4 * float(np.abs(egrad).max()))


def test_fused_loss_full_size_properties():
    """size-independent properties at 1920x1080: identical images give loss 0 (ssim = 1 everywhere) and the pure-L1
    subgradient; the device-resident loss equals the returned one; gradient of w-blend is linear in the SSIM part"""
    import torch
    Hh = gs.host
    H, W = 1080, 1920
    gen = torch.Generator(device="cuda").manual_seed(7)
    img = torch.rand((H, W, 3), device="cuda", generator=gen)
    loss, grad = Hh.fused_loss(img, img.clone(), 0.2)
    assert abs(loss) < 1e-6
    other = (img + 0.1 * torch.rand((H, W, 3), device="cuda", generator=gen)).clamp(0, 1)
    l_sync, g_sync = Hh.fused_loss(img, other, 0.2)
    l_dev, g_dev = Hh.fused_loss(img, other, 0.2, sync=False)
    torch.cuda.synchronize()
    assert abs(float(l_dev.item()) - l_sync) < 1e-6 and torch.allclose(g_dev, g_sync, rtol=0, atol=1e-12)
    l0, g0 = Hh.fused_loss(img, other, 0.0)
    l1, g1 = Hh.fused_loss(img, other, 1.0)
    n = H * W * 3
    torch.testing.assert_close(g0 * n, torch.where(img > other, 1.0, -1.0), rtol=0, atol=1e-6)   # pure L1: +-1 / n
    assert abs(l0 - float((img - other).abs().mean().item())) < 1e-6
    assert abs(l_sync - (0.8 * l0 + 0.2 * l1)) < 2e-6
    torch.testing.assert_close(g_sync, 0.8 * g0 + 0.2 * g1, rtol=0, atol=2e-6 * float(g_sync.abs().max().item()))


def test_fused_loss_rejects_small_workspace_and_host_pointers():
    import ctypes as C
    import torch
    from importlib import import_module
    ffi = import_module("3dgs_b200._ffi")
    a = torch.zeros((16, 16, 3), device="cuda")
    ws = torch.zeros(64, device="cuda", dtype=torch.uint8)
    p = lambda t: C.c_void_p(t.data_ptr())
    rc = ffi.lib().gsx_fused_loss(p(a), p(a), 16, 16, C.c_float(0.2), p(a), p(ws), C.c_size_t(64), None, None, None)
    assert rc == 5                                              # GSX_E_WORKSPACE
    host = np.zeros((16, 16, 3), np.float32)
    rc = ffi.lib().gsx_fused_loss(host.ctypes.data_as(C.c_void_p), p(a), 16, 16, C.c_float(0.2), p(a), p(ws),
                                  C.c_size_t(64), None, None, None)
    assert rc in (1, 2)


# ------------------------------------------------------------------------------------------ Morton re-layout
@pytest.mark.parametrize("true_interleave", [False, True])
def test_morton_codes_and_order_parity(B, true_interleave):
    """gsx_compute_morton_codes / gsx_morton_order against the oracle (cuda/culling.cu:18-68, trainer.cu:886-893) in both
    modes -- the reference's own lossy bit spread (default) and the real interleave: codes bit-exact, including points on
    and outside the box; order = stable argsort of the codes (equal codes keep index order)"""
    rng = np.random.default_rng(21)
    n = 200_003
    xyz = rng.normal(0, 2, (n, 3)).astype(np.float32)
    xyz[1000:1200] = xyz[0]                                  # duplicate positions: equal codes
    bbox = (float(xyz[:, 0].max()), float(xyz[:, 1].max()), float(xyz[:, 2].max()),
            float(xyz[:, 0].min()), float(xyz[:, 1].min()), float(xyz[:, 2].min()))
    t = true_interleave
    np.testing.assert_array_equal(B.morton_codes(xyz, bbox, t), G.morton_codes(xyz, bbox, t))
    np.testing.assert_array_equal(B.morton_order(xyz, bbox, t), G.morton_order(xyz, bbox, t))
    tight = (1.0, 1.0, 1.0, -1.0, -1.0, -1.0)              # many points outside: negative -> 0, beyond max wraps by the 21-bit mask
    np.testing.assert_array_equal(B.morton_codes(xyz, tight, t), G.morton_codes(xyz, tight, t))
    np.testing.assert_array_equal(B.morton_order(xyz, tight, t), G.morton_order(xyz, tight, t))


def test_gather_rows_multi_equals_per_array_gathers():
    """one launch for every array of a Gaussian set = the per-array gathers (and torch indexing)"""
    import torch
    Hh = gs.host
    n = 100_003
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    order = torch.randperm(n, device="cuda", generator=g).to(torch.int32)
    srcs = [torch.randn((n, w), device="cuda", generator=g) for w in (3, 3, 4, 1, 3, 45)]
    srcs.append(torch.randn(n, device="cuda", generator=g))
    outs = Hh.gather_rows_multi(srcs, order)
    for a, o in zip(srcs, outs):
        assert torch.equal(o, a[order.long()]) and torch.equal(o, Hh.gather_rows(a, order))


@pytest.mark.parametrize("flags", [0, 1])
def test_morton_relayout_properties_at_size(flags):
    """size-independent properties at 2 M Gaussians: the order is a permutation, the codes come out sorted, gathering the
    positions and recomputing their codes reproduces the sorted codes, and gather_rows equals torch indexing for every
    parameter width (1, 3, 4, 45)"""
    import torch
    Hh = gs.host
    n = 2_000_000
    gen = torch.Generator(device="cuda").manual_seed(4)
    xyz = torch.randn((n, 3), device="cuda", generator=gen)
    mn, mx = xyz.amin(0).tolist(), xyz.amax(0).tolist()
    bbox = (mx[0], mx[1], mx[2], mn[0], mn[1], mn[2])
    order, codes = Hh.morton_order(xyz, bbox, want_codes=True, flags=flags)
    assert torch.equal(torch.sort(order.long()).values, torch.arange(n, device="cuda"))
    assert bool((codes[1:] >= codes[:-1]).all().item())     # 63-bit codes: signed compare is the unsigned one
    moved = Hh.gather_rows(xyz, order)
    assert torch.equal(moved, xyz[order.long()])
    assert torch.equal(Hh.compute_morton_codes(moved, bbox, flags=flags), codes)
    for w in (1, 4, 45):
        a = torch.randn((n, w), device="cuda", generator=gen)
        assert torch.equal(Hh.gather_rows(a, order), a[order.long()])


# ------------------------------------------------------------------------------------------ densification
def _density_scene(rng, N, ext=10.0, l_max=1):
    """parameters + statistics whose decisions sit away from every threshold (largest scale in {0.3, 0.7, 1.3, 3} x the
    clone / max-scale thresholds), so expf rounding on the device cannot flip one"""
    logit = lambda p: np.log(p / (1 - p))
    fac = rng.choice([0.3, 0.7, 1.3, 3.0], N)
    big = rng.random(N) < 0.5
    smax = np.where(big, ext * 0.1 * fac * 0.9, ext * 0.01 * fac)
    scale = np.log(smax[:, None] * np.stack([np.ones(N), rng.uniform(0.2, 0.9, N), rng.uniform(0.05, 0.5, N)], 1)).astype(np.float32)
    scale = np.take_along_axis(scale, rng.permuted(np.tile(np.arange(3), (N, 1)), axis=1), axis=1)
    op = logit(np.where(rng.random(N) < 0.15, 0.001, rng.uniform(0.02, 0.98, N))).astype(np.float32)
    dur = rng.integers(0, 6, N).astype(np.int32)
    acc = (np.where(rng.random(N) < 0.5, 0.1, 2.0) * np.maximum(dur, 1)).astype(np.float32)
    q = rng.standard_normal((N, 4)).astype(np.float32)
    nsh = (l_max + 1) ** 2 - 1
    p = dict(xyz=rng.uniform(-3, 3, (N, 3)).astype(np.float32), scale=scale, quaternion=q, opacity=op,
             band0=rng.uniform(-1, 1, (N, 3)).astype(np.float32), sh=rng.normal(0, 0.2, (N, nsh * 3)).astype(np.float32))
    return p, acc, dur


def test_clone_and_split_entry_points_vs_oracle(B):
    """gsx_clone_gaussians / gsx_split_gaussians (the reference's two functions, adaptive_density.cuh:28,54): a clone is an
    exact copy at row write_ids[i]; split children copy colour / opacity / rotation / SH exactly, carry the scale
    log(exp(s) / f), and their centres equal the oracle's (same Philox samples; Box-Muller through libm vs device math:
    1e-5 of the parent's extent) -- and are reproducible from the seed."""
    import torch
    rng = np.random.default_rng(12)
    N = 5003
    p, _, _ = _density_scene(rng, N)
    mask = (rng.random(N) < 0.3)
    wid = (np.cumsum(mask) - mask).astype(np.int32)
    n_sel = int(mask.sum())
    d = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).cuda()
    Hh = gs.host
    arrays = dict(xyz=d(p["xyz"]), rgb=d(p["band0"]), opacity=d(p["opacity"]), scale=d(p["scale"]), quaternion=d(p["quaternion"]),
                  sh=d(p["sh"]).view(N, -1, 3))
    src = dict(xyz=p["xyz"], rgb=p["band0"], opacity=p["opacity"], scale=p["scale"], quaternion=p["quaternion"], sh=p["sh"])
    out = Hh.clone_gaussians(d(mask, torch.uint8), d(wid, torch.int32), arrays, n_sel)
    torch.cuda.synchronize()
    for k, v in out.items():
        np.testing.assert_array_equal(v.cpu().numpy().reshape(n_sel, -1), src[k].reshape(N, -1)[mask], err_msg=f"clone {k}")
    for seed in (99, 100):
        out = Hh.split_gaussians(d(mask, torch.uint8), d(wid, torch.int32), arrays, 2 * n_sel, 1.6, seed)
        torch.cuda.synchronize()
        o = {k: v.cpu().numpy().reshape(2 * n_sel, -1) for k, v in out.items()}
        for k in ("rgb", "opacity", "quaternion", "sh"):
            np.testing.assert_array_equal(o[k], np.repeat(src[k].reshape(N, -1)[mask], 2, axis=0), err_msg=f"split {k}")
        sel = np.nonzero(mask)[0]
        np.testing.assert_allclose(o["scale"], np.repeat(np.log(np.exp(p["scale"][sel].astype(np.float64)) / 1.6), 2, axis=0), atol=2e-6)
        for r in rng.choice(n_sel, 300, replace=False):
            i = int(sel[r])
            kids, sc = G.split_children(seed, i, p["xyz"][i], p["scale"][i], p["quaternion"][i], 1.6)
            ext = float(np.exp(p["scale"][i]).max())
            np.testing.assert_allclose(o["xyz"][2 * r:2 * r + 2], kids, atol=1e-5 * (ext * 6 + np.abs(p["xyz"][i]).max()))
        if seed == 99:
            first = o["xyz"].copy()
            again = Hh.split_gaussians(d(mask, torch.uint8), d(wid, torch.int32), arrays, 2 * n_sel, 1.6, 99)["xyz"].cpu().numpy()
            np.testing.assert_array_equal(again, first)                          # same seed, same samples
        else:
            assert not np.array_equal(o["xyz"], first)                           # another seed, another stream
    # distribution of the samples on the device: one anisotropic, rotated parent replicated 100k times
    M = 100_000
    one = {k: v[:1].expand(M, *v.shape[1:]).contiguous() for k, v in arrays.items()}
    allm, ids = torch.ones(M, dtype=torch.uint8, device="cuda"), torch.arange(M, dtype=torch.int32, device="cuda")
    pts = Hh.split_gaussians(allm, ids, one, 2 * M, 1.6, 5)["xyz"].cpu().numpy().astype(np.float64)
    q = p["quaternion"][0].astype(np.float64); q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    cov = R @ np.diag(np.exp(2 * p["scale"][0].astype(np.float64))) @ R.T
    sd = np.sqrt(np.abs(cov).max())
    np.testing.assert_allclose(pts.mean(0), p["xyz"][0], atol=5 * sd / np.sqrt(2 * M))
    np.testing.assert_allclose(np.cov(pts.T), cov, atol=0.02 * np.abs(cov).max())


@pytest.mark.parametrize("N,l_max,with_moments", [(50_003, 1, True), (4096, 3, False), (7, 0, True)])
def test_densify_step_vs_oracle(B, N, l_max, with_moments):
    """gsx_densify_plan + gsx_densify_apply = TrainerImpl::adaptive_density_step (cuda/trainer.cu:416-775): decisions and
    counts equal the oracle's, survivors / clones / split children land in the reference's order with every copied value
    bit-identical, children scales / centres as above, appended Gaussians get zero Adam moments; over max_gaussians the
    step is skipped."""
    import torch
    rng = np.random.default_rng(N)
    p, acc, dur = _density_scene(rng, N, l_max=l_max)
    if l_max == 0:
        p["sh"] = None
    act = G.density_decide(p["opacity"], p["scale"], acc, dur, 0.5, 0.005, 10.0)
    assert N < 100 or all((act == a).sum() > N // 50 for a in range(4))             # every branch is exercised
    d = lambda a, dt=torch.float32: torch.as_tensor(np.ascontiguousarray(a), dtype=dt).cuda()
    dp = {k: (d(v) if v is not None else None) for k, v in p.items()}
    mom_h = {k: (rng.normal(0, 1, v.shape).astype(np.float32), rng.random(v.shape).astype(np.float32)) for k, v in p.items() if v is not None} if with_moments else None
    mom_d = None if mom_h is None else {k: (d(m), d(v)) for k, (m, v) in mom_h.items()}
    Hh = gs.host
    out_p, out_m, cnt = Hh.densify(dp, mom_d, d(acc), d(dur, torch.int32), 0.5, 0.005, 10.0, 1.6, max_gaussians=0, seed=77, l_max=l_max)
    torch.cuda.synchronize()
    assert cnt == dict(n_in=N, n_prune=int((act == 1).sum()), n_clone=int((act == 2).sum()), n_split=int((act == 3).sum()),
                       n_out=int(((act == 0) | (act == 2)).sum() + (act == 2).sum() + 2 * (act == 3).sum()))
    ph = {k: (v.reshape(N, -1) if v is not None else None) for k, v in p.items()}
    ep, em = G.densify({k: v for k, v in ph.items() if v is not None}, None if mom_h is None else {k: (m.reshape(N, -1), v.reshape(N, -1)) for k, (m, v) in mom_h.items()}, act, 77, 1.6)
    n_front = int(((act == 0) | (act == 2)).sum()) + int((act == 2).sum())
    for k, e in ep.items():
        g = out_p[k].cpu().numpy().reshape(cnt["n_out"], -1)
        np.testing.assert_array_equal(g[:n_front], e[:n_front], err_msg=f"{k}: survivors + clones")
        if k == "xyz":
            ext = np.repeat(np.exp(ph["scale"][act == 3]).max(1), 2)[:, None]
            assert (np.abs(g[n_front:] - e[n_front:]) <= 1e-5 * (6 * ext + 3.0)).all(), "split centres"
        elif k == "scale":
            np.testing.assert_allclose(g[n_front:], e[n_front:], atol=2e-6, err_msg="split scales")
        else:
            np.testing.assert_array_equal(g[n_front:], e[n_front:], err_msg=f"{k}: split children")
    if with_moments:
        for k, (m, v) in em.items():
            np.testing.assert_array_equal(out_m[k][0].cpu().numpy().reshape(cnt["n_out"], -1), m, err_msg=f"exp_avg {k}")
            np.testing.assert_array_equal(out_m[k][1].cpu().numpy().reshape(cnt["n_out"], -1), v, err_msg=f"exp_avg_sq {k}")
    # capacity: the reference skips the step when it would exceed max_gaussians (trainer.cu:579-585)
    if cnt["n_out"] > 1:
        none_p, none_m, cnt2 = Hh.densify(dp, mom_d, d(acc), d(dur, torch.int32), 0.5, 0.005, 10.0, 1.6, max_gaussians=cnt["n_out"] - 1, seed=77, l_max=l_max)
        assert none_p is None and cnt2 == cnt
