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


# ------------------------------------------------------------------------------------------------ densification (8f rank 4)
def test_oracle_philox_known_answers():
    """The counter-based generator behind the split samples is Philox4x32-10 (Salmon, Moraes, Dror, Shaw, SC'11): pinned
    by the three known-answer vectors published with the Random123 library (kat_vectors: philox4x32 10)."""
    assert G.philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert G.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert G.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_oracle_split_children_reference_known_answers():
    """what the reference's own tests pin for split / clone (tests/adaptive_density_test.cpp:235-299): the children's
    log-scale is log(exp(s) / 1.6) on every axis for both children, everything else is copied; positions are random
    there (curand seeded with time(NULL)) and not asserted.  Here they must be distributed as N(xyz, R S^2 R^T)."""
    xyz, s = np.array([1.0, 2.0, 3.0], np.float32), np.log(np.array([2.0, 2.0, 2.0], np.float32))
    q = np.array([1.0, 0.0, 0.0, 0.0], np.float32)
    kids, sc = G.split_children(7, 0, xyz, s, q, 1.6)
    np.testing.assert_allclose(sc, np.log(2.0 / 1.6), atol=1e-6)                  # adaptive_density_test.cpp:281-290
    assert kids.shape == (2, 3) and np.isfinite(kids).all() and not np.array_equal(kids[0], kids[1])
    # distribution: anisotropic, rotated parent; 20000 parents' worth of children
    s2 = np.log(np.array([0.5, 0.1, 0.02], np.float32))
    q2 = np.array([0.9, 0.1, -0.3, 0.2], np.float32)                               # non-unit on purpose (normalised inside)
    pts = np.concatenate([G.split_children(123, i, xyz, s2, q2, 1.6)[0] for i in range(20000)])
    w, x, y, z = q2 / np.linalg.norm(q2)
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    cov = R @ np.diag(np.exp(2 * s2.astype(np.float64))) @ R.T
    np.testing.assert_allclose(pts.mean(0), xyz, atol=4 * 0.5 / np.sqrt(len(pts)))
    np.testing.assert_allclose(np.cov(pts.T), cov, atol=0.04 * np.abs(cov).max())


def test_oracle_density_decisions():
    """IdentifyPrune / IdentifyClone / IdentifySplit of cuda/trainer.cu:438-504 on hand-made rows (extent 10: max scale 1.0,
    clone threshold 0.1; uv gradient threshold 0.5; opacity threshold 0.005)."""
    ext, gthr, othr = 10.0, 0.5, 0.005
    logit = lambda p: np.log(p / (1 - p))
    rows = [  # opacity p, largest scale, mean grad (accum / dur), dur -> expected action
        (0.001, 0.05, 0.9, 4, 1),    # transparent -> prune, whatever the gradient
        (0.5, 0.05, 0.9, 4, 2),      # small + high gradient -> clone
        (0.5, 0.5, 0.9, 4, 3),       # large + high gradient -> split
        (0.5, 1.5, 0.9, 4, 3),       # above max scale but 1.5 / 1.6 <= 1.0 and about to be split -> split, not pruned
        (0.5, 1.7, 0.9, 4, 1),       # 1.7 / 1.6 > 1.0 -> pruned although the gradient is high
        (0.5, 1.2, 0.1, 4, 1),       # too large, low gradient -> prune
        (0.5, 0.5, 0.1, 4, 0),       # nothing to do
        (0.5, 0.05, 9.9, 0, 0),      # never visible (duration 0): mean gradient 0 -> keep
    ]
    op = np.array([logit(r[0]) for r in rows], np.float32)
    sc = np.log(np.array([[r[1], r[1] * 0.5, r[1] * 0.1] for r in rows], np.float32))
    acc = np.array([r[2] * max(r[3], 1) for r in rows], np.float32)
    dur = np.array([r[3] for r in rows], np.int32)
    act = G.density_decide(op, sc, acc, dur, gthr, othr, ext)
    np.testing.assert_array_equal(act, [r[4] for r in rows])
    # the compaction order of trainer.cu:655-775: survivors (clone parents included) in order, clones, split children
    params = dict(xyz=np.arange(24, dtype=np.float32).reshape(8, 3), scale=sc, quaternion=np.tile([1, 0, 0, 0], (8, 1)).astype(np.float32),
                  opacity=op.reshape(8, 1), band0=np.zeros((8, 3), np.float32))
    mom = {k: (np.ones_like(v), 2 * np.ones_like(v)) for k, v in params.items()}
    out_p, out_m = G.densify(params, mom, act, seed=5, scale_factor=1.6)
    assert out_p["xyz"].shape[0] == 3 + 1 + 2 * 2                                  # keep {1, 6, 7}, clone {1}, split {2, 3}
    np.testing.assert_array_equal(out_p["opacity"][:4, 0], op[[1, 6, 7, 1]])
    np.testing.assert_array_equal(out_p["xyz"][:4], params["xyz"][[1, 6, 7, 1]])
    np.testing.assert_allclose(out_p["scale"][4:6], np.tile(np.log(np.exp(sc[2]) / 1.6), (2, 1)), atol=1e-6)
    assert (out_m["xyz"][0][:3] == 1).all() and (out_m["xyz"][1][:3] == 2).all() and not out_m["xyz"][0][3:].any()
