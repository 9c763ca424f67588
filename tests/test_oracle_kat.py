"""Pins the CPU oracle against every known-answer test the reference holds for the hot path."""
import numpy as np
import pytest

import kat
from oracle import golden as G


@pytest.mark.parametrize("check", kat.ALL, ids=[c.__name__ for c in kat.ALL])
def test_oracle_kat(check):
    check(G)


def test_sh_backward_ref_compat_layout_reported():
    """Q12: with the reference's [k][axis] reading of sphericart's [axis][k] buffer, the position gradient differs
    from the finite-difference answer; record by how much (the reference's own test tolerance is 1e-2)."""
    xyz, g, b0, sh, fd = kat.sh_backward_case()
    _, _, xg_ok = G.sh_backward(xyz, b0, sh, np.zeros(3, np.float32), g, 2, ref_compat_dsph=0)
    _, _, xg_rc = G.sh_backward(xyz, b0, sh, np.zeros(3, np.float32), g, 2, ref_compat_dsph=1)
    err_ok = float(np.abs(xg_ok - fd).max())
    err_rc = float(np.abs(xg_rc - fd).max())
    print(f"SH d xyz max err vs FD: correct layout {err_ok:.3e}, reference-compatible layout {err_rc:.3e}")
    assert err_ok < 1e-3
    # l_max = 0: both layouts coincide (dY0 = 0)
    a = G.sh_backward(xyz, b0, None, np.zeros(3, np.float32), g, 0, ref_compat_dsph=0)[2]
    b = G.sh_backward(xyz, b0, None, np.zeros(3, np.float32), g, 0, ref_compat_dsph=1)[2]
    np.testing.assert_array_equal(a, b)


def test_sph_basis_pinned_to_sphericart_l3():
    """l = 3 is pinned by no reference test: pin the oracle's closed forms and normalised gradients
    against golden vectors produced by the vendored sphericart CPU code (oracle/gen_golden.py)."""
    import json, os
    with open(os.path.join(os.path.dirname(__file__), "golden", "sph_l3.json")) as f:
        d = json.load(f)
    xyz = np.array(d["xyz"], np.float64).reshape(-1, 3)
    sph = np.array(d["sph"]).reshape(-1, 16)
    dsph = np.array(d["dsph"]).reshape(-1, 3, 16)   # sphericart layout [sample][axis][k]
    Y, dY = G.sph(xyz.astype(np.float32), 3, grad=True)
    np.testing.assert_allclose(Y, sph, atol=2e-6)
    np.testing.assert_allclose(dY, dsph, atol=2e-5, rtol=1e-5)


def test_oracle_adam_masked_is_compact_step_scatter():
    """oracle self-check for the optimiser row: the masked form equals adam_step on the rows the
    mask keeps and leaves the others bit-identical (cuda/trainer.cu:1027-1085); accumulators per trainer.cu:1138-1157"""
    rng = np.random.default_rng(9)
    n, w = 257, 5
    p, g, m, v = (rng.normal(0, 1, (n, w)).astype(np.float32) for _ in range(4))
    v = v * v
    g[3, 2] = np.nan
    mask = (rng.random(n) < 0.5).astype(np.uint8)
    keep = mask.astype(bool)
    a = (1e-3, 0.9, 0.999, 1e-8, 0.19, 0.002)
    pm, mm, vm = G.adam_step_masked(p, g, m, v, mask, *a)
    pk, mk, vk = G.adam_step(p[keep], g[keep], m[keep], v[keep], *a)
    np.testing.assert_array_equal(pm[keep], pk)
    np.testing.assert_array_equal(mm[keep], mk)
    np.testing.assert_array_equal(vm[keep], vk)
    np.testing.assert_array_equal(pm[~keep], p[~keep])
    np.testing.assert_array_equal(mm[~keep], m[~keep])
    np.testing.assert_array_equal(vm[~keep], v[~keep])
    assert np.isfinite(pm).all()                      # the NaN gradient counted as 0 (cuda/optimizer.cu:14-16)
    pa, ma, va = G.adam_step_masked(p, g, m, v, None, *a)
    pf, mf, vf = G.adam_step(p, g, m, v, *a)
    np.testing.assert_array_equal(pa, pf)
    uv = rng.normal(0, 1, (n, 2)).astype(np.float32)
    acc, dur = G.accumulate_uv_grad(uv, mask, np.ones(n, np.float32), np.zeros(n, np.int32))
    np.testing.assert_array_equal(dur, mask.astype(np.int32))
    np.testing.assert_allclose(acc, 1 + keep * np.sqrt((uv.astype(np.float64) ** 2).sum(1)), rtol=3e-7)


def _ssim_l1_loss64(pred, gt, w):
    """float64 numpy statement of the loss (replicate-padded 11x11 Gaussian window), independent of the oracle's code"""
    from scipy.ndimage import correlate1d
    g = np.array([0.001028380123898387, 0.0075987582094967365, 0.036000773310661316, 0.10936068743467331,
                  0.21300552785396576, 0.26601171493530273, 0.21300552785396576, 0.10936068743467331,
                  0.036000773310661316, 0.0075987582094967365, 0.001028380123898387])
    blur = lambda a: correlate1d(correlate1d(a, g, axis=0, mode="nearest"), g, axis=1, mode="nearest")
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    mu1, mu2 = blur(pred), blur(gt)
    s1, s2, s12 = blur(pred * pred) - mu1 * mu1, blur(gt * gt) - mu2 * mu2, blur(pred * gt) - mu1 * mu2
    ssim = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))
    return float(((1 - w) * np.abs(pred - gt) + w * (1 - ssim)).mean())


def test_oracle_fused_loss_matches_float64_statement_and_finite_differences():
    """oracle pin for the loss row beyond the reference's uniform-image test: loss value against a
    float64 numpy/scipy statement, gradient against central differences of it at pixels >= 10 px from the border
    (nearer, the reference's zero-padded back-convolution is not the adjoint of its replicate-padded forward)"""
    rng = np.random.default_rng(3)
    H, W, w = 37, 45, 0.2
    yy, xx = np.mgrid[0:H, 0:W]
    base = 0.5 + 0.3 * np.sin(xx / 7.0)[..., None] * np.cos(yy / 5.0)[..., None] * np.array([1.0, 0.7, -0.5])
    pred = np.clip(base + rng.normal(0, 0.05, (H, W, 3)), 0, 1).astype(np.float32)
    gt = np.clip(base + rng.normal(0, 0.05, (H, W, 3)), 0, 1).astype(np.float32)
    loss, grad = G.fused_loss(pred, gt, w)
    p64, g64 = pred.astype(np.float64), gt.astype(np.float64)
    assert abs(loss - _ssim_l1_loss64(p64, g64, w)) < 2e-6
    for _ in range(12):
        y, x, c = int(rng.integers(10, H - 10)), int(rng.integers(10, W - 10)), int(rng.integers(0, 3))
        d = np.zeros_like(p64)
        d[y, x, c] = 1e-5
        if abs(p64[y, x, c] - g64[y, x, c]) < 2e-5:
            continue                                  # |p - g| kink inside the stencil
        fd = (_ssim_l1_loss64(p64 + d, g64, w) - _ssim_l1_loss64(p64 - d, g64, w)) / 2e-5
        assert abs(grad[y, x, c] - fd) < 2e-3 * abs(fd) + 1e-8, (y, x, c, grad[y, x, c], fd)
