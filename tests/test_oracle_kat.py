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
