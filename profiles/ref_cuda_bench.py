"""Same-box baseline B1: the REFERENCE's own CUDA kernels (oracle/_ref/libref3dgs.so) replaying rasterize_image
(cuda/raster.cu:26-135) and backward_pass (cuda/trainer.cu:941-1012) call for call on one view of a bench scene,
CUDA-event timed.  Differences from the reference trainer, all in the reference's favour: the 16 thrust copy_if
compactions are done with torch boolean indexing outside the per-function timings (reported separately), and the two SH
functions (sphericart NVRTC backend, not buildable here) are replaced by a band-0 colour computed with torch.
usage: python profiles/ref_cuda_bench.py [C2|C3|C4] [reps]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import ref_cuda as R  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S = gs.scenes
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
N, W, H, l_max = S.CONFIGS[cfg]
g = S.synthetic_gaussians(N, l_max=0, seed=0x3D65 + int(cfg[1:]))
cam = S.orbit_camera(W, H, azimuth_deg=0.0)
d = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()
xyz, scale, quat, opa, b0 = d(g["xyz"]), d(g["scale"]), d(g["quaternion"]), d(g["opacity"]).reshape(-1), d(g["band0"])
view, proj = d(cam["view"].reshape(-1)), d(cam["proj"].reshape(-1))
gi = d(S.grad_image(W, H))
ntx, nty = (W + 15) // 16, (H + 15) // 16
tx, ty = W / (2 * cam["fx"]), H / (2 * cam["fy"])


def timed(acc, name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = fn()
    e1.record()
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + e0.elapsed_time(e1)
    return out


def one_view(acc):
    xyz_c = timed(acc, "F1 camera_space_points", lambda: R.camera_space_points(xyz, view))
    uv = timed(acc, "F2 project_to_screen", lambda: R.project_to_screen(xyz_c, proj, W, H))
    mask = timed(acc, "F3 cull", lambda: R.cull(uv, xyz_c, S.NEAR_THRESH, S.CULL_PADDING, W, H))

    def compact():
        m = mask.bool()
        return [t[m].contiguous() for t in (xyz_c, uv, scale, quat, opa, b0, xyz)]
    xyz_cc, uvc, sc, qc, oc, b0c, xyzw = timed(acc, "compaction (torch indexing; reference: 8 thrust copy_if)", compact)
    sig = timed(acc, "F4 sigma", lambda: R.sigma(qc, sc))
    J, con, rad = timed(acc, "F5 conic", lambda: R.conic(xyz_cc, view, sig, cam["fx"], cam["fy"], tx, ty, S.MH_DIST))
    rgb = (0.5 + 0.28209479 * b0c).contiguous()
    lst = timed(acc, "F7 sorted list (both phases)", lambda: R.sorted_list(uvc, xyz_cc, rad, ntx, nty))
    img, wpp, spp = timed(acc, "F8 render", lambda: R.render(uvc, oc, con, rgb, 0.0, lst["sorted"], lst["ranges"], W, H))
    gr = timed(acc, "F9 render_backward", lambda: R.render_backward(uvc, oc, con, rgb, 0.0, lst["sorted"], lst["ranges"], spp, wpp, gi, W, H))
    gJ, gS = timed(acc, "F11 conic_backward", lambda: R.conic_backward(J, sig, view, con, gr[3]))
    gx = timed(acc, "F12 jacobian_backward", lambda: R.jacobian_backward(xyz_cc, cam["fx"], cam["fy"], tx, ty, gJ))
    timed(acc, "F13 sigma_backward", lambda: R.sigma_backward(qc, sc, gS))
    gx = timed(acc, "F14 project_backward", lambda: R.project_to_screen_backward(xyz_cc, proj, gr[2], W, H, gx))
    timed(acc, "F15 camera_backward", lambda: R.camera_space_points_backward(xyzw, view, gx))
    return int(mask.sum().item()), lst["C"], lst["K"]


out = {"config": cfg, "N": N, "W": W, "H": H, "sh": "band 0 only (SH functions not buildable: sphericart NVRTC backend)"}
try:
    one_view({})                                                   # warm-up
    acc = {}
    for _ in range(reps):
        P, Cc, K = one_view(acc)
    st = {k: round(v / reps, 4) for k, v in acc.items()}
    out.update(P=P, coarse_pairs_C=Cc, instances_K=K, stage_ms=st, ms_per_view=round(sum(st.values()), 3),
               ms_per_view_kernels_only=round(sum(v for k, v in st.items() if not k.startswith("compaction")), 3))
except Exception as e:                                             # e.g. out of memory in the coarse pair list (H4)
    out["error"] = repr(e)[:300]
print(json.dumps(out))
