"""Full-view parity report: the CUDA path's saved forward state and gradients against the CPU oracle on the same
scene -- TEST INFRASTRUCTURE ONLY (used by tests/ and by bench.py's cpu_baseline leg, which already pays for the oracle
run of one C4 view; never by the product path).

Two comparisons, as north_star asks ("sort keys, tile ranges and per-tile lists bit-exact; RGB/alpha 1e-4; gradients 1e-3"):
  * chain-consistent F7: the oracle's binning + sort fed with the GPU's own per-Gaussian (uv, depth, radius) must give
    the GPU's ranges and sorted ids BIT-EXACTLY (discrete work on identical inputs);
  * end to end against the oracle's own chain: P, K, image / final T within 1e-4 relative, splats_per_pixel, and the six
    parameter-gradient arrays within 1e-3 norm-wise.

Out-of-tolerance pixels are not granted a blanket budget: each one is checked to be a threshold flip.  The compositing
loop takes two float-threshold decisions per (pixel, entry), alpha > 1/255 and T < 1e-4 (cuda/render.cu:74-76), on
ex2.approx (device) vs expf (oracle) values that differ in the last bits.  A flipped alpha decision moves the pixel by at
most (1/255) * T * |colour - what is behind| <= FLIP_ALPHA * (max |rgb| + |bg|); a flipped T decision makes one side stop
one entry earlier with T already < ~1e-4, so both final T are below FLIP_T and the colours differ by < FLIP_T * max |rgb|.
"""
import numpy as np

from . import golden as G

IMG_RTOL, IMG_ATOL = 1e-4, 2e-6
GRAD_RTOL = 1e-3
FLIP_ALPHA = 0.99 / 255.0 * 1.05      # alpha just above the threshold, 5 % slack for the float evaluation
FLIP_T = 1.05e-4


def _bad(a, b, rtol, atol):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return ~(np.abs(a - b) <= atol + rtol * np.abs(b))


def compare_pixels(img, wpp, spp, ref_img, ref_wpp, ref_spp, rgb_absmax, bg):
    """image / final T / splats_per_pixel of one view against a reference; every out-of-tolerance pixel must be explained
    by a threshold flip (see module docstring).  Returns the counted fractions and the worst flip magnitudes."""
    H, W = wpp.shape
    ref_img = ref_img.reshape(H, W, 3); ref_wpp = ref_wpp.reshape(H, W); ref_spp = ref_spp.reshape(H, W)
    bad_px = _bad(img, ref_img, IMG_RTOL, IMG_ATOL).any(axis=2)
    bad_t = _bad(wpp, ref_wpp, IMG_RTOL, IMG_ATOL)
    spp_ne = spp != ref_spp
    err = np.abs(img.astype(np.float64) - ref_img).max(axis=2)
    bound = FLIP_ALPHA * (rgb_absmax + abs(bg))
    out = dict(image_outlier_frac=float(bad_px.mean()), T_outlier_frac=float(bad_t.mean()), spp_mismatch_frac=float(spp_ne.mean()),
               image_outlier_max_abs=float(err[bad_px].max()) if bad_px.any() else 0.0, image_flip_bound=float(bound))
    # a pixel whose entry counts differ stopped one entry apart: both transmittances are already below the T threshold
    t_hi = np.maximum(wpp, ref_wpp)
    out["spp_mismatch_max_T"] = float(t_hi[spp_ne].max()) if spp_ne.any() else 0.0
    unexplained = bad_px & ~(err <= bound)
    out["image_unexplained"] = int(unexplained.sum())
    out["spp_unexplained"] = int((spp_ne & ~(t_hi <= FLIP_T)).sum())
    # an out-of-tolerance final T is the (1 - alpha) factor of one flipped entry: relative change <= FLIP_ALPHA
    relt = np.abs(wpp.astype(np.float64) - ref_wpp) / np.maximum(np.abs(ref_wpp), 1e-30)
    out["T_unexplained"] = int((bad_t & ~((relt <= FLIP_ALPHA) | (t_hi <= FLIP_T))).sum())
    return out


def compare_grads(got, want, names=(("xyz", "xyz"), ("scale", "scale"), ("quat", "quaternion"), ("opacity", "opacity"),
                                    ("band0", "band0"), ("sh", "sh"))):
    """norm-wise relative error per array (the spec's 1e-3) + per-element outlier count and worst outlier, relative to the
    array's largest magnitude"""
    out = {}
    for ko, kg in names:
        if ko not in want or kg not in got:
            continue
        b = np.asarray(want[ko], np.float64).reshape(-1)
        a = np.asarray(got[kg], np.float64).reshape(-1)[: b.size] if np.asarray(got[kg]).size != b.size else np.asarray(got[kg], np.float64).reshape(-1)
        nb = np.linalg.norm(b)
        rel = float(np.linalg.norm(a - b) / nb) if nb > 0 else float(np.abs(a).max())
        mx = float(np.abs(b).max()) if b.size else 0.0
        bad = ~(np.abs(a - b) <= GRAD_RTOL * np.abs(b) + GRAD_RTOL * 1e-2 * mx)
        out[kg] = dict(rel_norm=rel, outlier_frac=float(bad.mean()) if b.size else 0.0,
                       outlier_max_over_absmax=float(np.abs(a - b)[bad].max() / mx) if bad.any() and mx > 0 else 0.0)
    return out


def lists_bit_exact(state, N, W, H):
    """chain-consistent F7: oracle binning + sort on the GPU's own (uv, depth, radius) vs the GPU's ranges / sorted ids"""
    mask = state["mask"].astype(bool)
    uv = np.where(mask[:, None], state["rec0"][:, :2], -1e6).astype(np.float32)
    rad = np.where(mask[:, None], state["radius"], np.array([0, 0, 0, 1], np.float32)).astype(np.float32)
    xyz_c = np.zeros((N, 3), np.float32)
    xyz_c[:, 2] = np.where(mask, state["depth"], 1.0)
    lst = G.sorted_list(uv, xyz_c, rad, (W + 15) // 16, (H + 15) // 16)
    same_K = int(lst["K"]) == int(state["K"])
    return dict(K_oracle_on_gpu_values=int(lst["K"]), K_equal=same_K,
                ranges_bit_exact=bool(np.array_equal(lst["ranges"], state["ranges"])),
                sorted_ids_bit_exact=bool(same_K and np.array_equal(lst["sorted"], state["sorted_gaussians"])))


def full_view_report(state, got_grads, o, og, N, W, H, bg, lists=True):
    """state: numpy views of the GPU's forward state (host.Rasterizer.state()), got_grads: its gradients for this view
    alone (or None: forward only), o / og: oracle State and gradients of the same view."""
    rep = dict(P_gpu=int(state["mask"].sum()), P_oracle=int(o.P), K_gpu=int(state["K"]), K_oracle=int(o.K))
    if lists:
        rep.update(lists_bit_exact(state, N, W, H))
    rgb_absmax = float(np.abs(state["rec2"][state["mask"].astype(bool), :3]).max()) if rep["P_gpu"] else 0.0
    rep.update(compare_pixels(state["image"], state["weight_per_pixel"], state["splats_per_pixel"], o.f("image"), o.f("wpp"),
                              o.i("spp"), rgb_absmax, bg))
    if got_grads is not None and og is not None:
        rep["grads"] = compare_grads(got_grads, og)
    return rep


def assert_report(rep, max_outlier_frac=2e-4, grad_rtol=GRAD_RTOL, k_slack=4, max_spp_flip_frac=1e-3):
    """the bar: lists bit-exact on identical inputs; P equal and K within a few 1-ulp radius ties of the oracle's own chain;
    every out-of-tolerance pixel an explained flip and rare; gradients 1e-3 norm-wise"""
    if "ranges_bit_exact" in rep:
        assert rep["K_equal"] and rep["ranges_bit_exact"] and rep["sorted_ids_bit_exact"], rep
    assert abs(rep["P_gpu"] - rep["P_oracle"]) <= 2 and abs(rep["K_gpu"] - rep["K_oracle"]) <= max(k_slack, rep["K_oracle"] // 1_000_000), rep
    assert rep["image_unexplained"] == 0 and rep["spp_unexplained"] == 0 and rep["T_unexplained"] == 0, rep
    assert rep["image_outlier_frac"] <= max_outlier_frac and rep["T_outlier_frac"] <= max_outlier_frac, rep
    # entry counts differ where the T < 1e-4 decision flipped: every such pixel was checked above to sit AT the threshold
    # (both transmittances <= 1.05e-4), so only their number is bounded here -- measured 2.5e-4 at C4, where half of the
    # pixels saturate and each passes the threshold once
    assert rep["spp_mismatch_frac"] <= max_spp_flip_frac, rep
    for k, g in rep.get("grads", {}).items():
        assert g["rel_norm"] <= grad_rtol, (k, g)
