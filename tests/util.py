"""Shared helpers for the parity tests: tolerances (stated in one place), scene plumbing, oracle chains."""
import importlib

import numpy as np

from oracle import golden as G

gs = importlib.import_module("3dgs_b200")
S = gs.scenes

# --- tolerances (BASELINE.json north_star / SURVEY 8c) -------------------------------------------------------
IMG_RTOL, IMG_ATOL = 1e-4, 2e-6      # rendered RGB / final transmittance: 1e-4 relative, small absolute floor
GRAD_RTOL = 1e-3                     # gradients: 1e-3 relative (atomic / warp summation order differs)
# alpha>1/255 and T<1e-4 are threshold decisions on ex2.approx vs expf; a flipped decision moves one pixel by up to
# 1/255.  Such pixels (and Gaussians dominated by one) are counted and bounded, never ignored silently.
OUTLIER_FRAC = 2e-4


def close_frac(a, b, rtol, atol):
    """fraction of elements violating |a-b| <= atol + rtol*|b|"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    bad = ~(np.abs(a - b) <= atol + rtol * np.abs(b))
    return float(bad.mean()) if bad.size else 0.0


def assert_image_close(a, b, what="image"):
    f = close_frac(a, b, IMG_RTOL, IMG_ATOL)
    assert f <= OUTLIER_FRAC, f"{what}: {f:.2e} of elements outside rtol {IMG_RTOL} (allowed {OUTLIER_FRAC})"
    return f


FLIP_GRAD_BOUND = 2.0 * 0.99 / 255.0   # see assert_grad_close


def assert_grad_close(a, b, what="grad", rtol=GRAD_RTOL, outlier=OUTLIER_FRAC):
    """The spec: norm-wise relative error <= rtol (1e-3).  Per element, |a-b| <= rtol*|b| + floor with floor =
    rtol*1e-2*max|b|; elements outside are not granted a blanket budget: they must be rare (<= `outlier`, counted) AND small
    -- a flipped alpha >= 1/255 / idx < count decision of ONE pixel (ex2.approx on the device vs expf in the oracle) removes
    or adds one pixel-entry whose alpha is 1/255, which perturbs that pixel's share of every gradient it feeds by at most
    ~alpha; so an outlier's absolute error is bounded by FLIP_GRAD_BOUND (2 x 0.99/255) of the array's largest magnitude.
    Anything larger is a defect, however rare."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if b.size == 0:
        return 0.0
    assert np.isfinite(a).all(), f"{what}: non-finite values"
    nb = np.linalg.norm(b)
    if nb > 0:
        rel = np.linalg.norm(a - b) / nb
        assert rel <= rtol, f"{what}: norm-wise relative error {rel:.3e} > {rtol}"
    else:
        assert np.abs(a).max() <= 1e-12, f"{what}: expected zeros"
    mx = np.abs(b).max()
    floor = rtol * 1e-2 * mx
    err = np.abs(a - b)
    bad = ~(err <= floor + rtol * np.abs(b))
    f = float(bad.mean())
    assert f <= max(outlier, 2.0 / b.size), f"{what}: {f:.2e} of elements outside rtol {rtol}"
    if bad.any():
        worst = float(err[bad].max())
        assert worst <= FLIP_GRAD_BOUND * mx, (f"{what}: {int(bad.sum())} outliers, worst |a-b| = {worst:.3e} = "
                                               f"{worst / mx:.3e} of max|b| > the one-pixel flip bound {FLIP_GRAD_BOUND:.3e}")
    return f


def make_views(cam, l_max, background=0.0, flags=0):
    """(gold view struct, gsx view struct) for one camera dict from scenes.py"""
    gv = G.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], cam["width"], cam["height"],
                     near_thresh=S.NEAR_THRESH, cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=background,
                     l_max=l_max, ref_compat_dsph=0 if (flags & 1) else 1)
    return gv


def gsx_view(cam, l_max, background=0.0, flags=0):
    H = gs.host
    return H.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], cam["width"], cam["height"],
                       near_thresh=S.NEAR_THRESH, cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST,
                       background=background, l_max=l_max, flags=flags)


def scatter(compact, vis, N, fill=0.0):
    out = np.full((N,) + compact.shape[1:], fill, compact.dtype)
    out[vis] = compact
    return out
