"""Stage timings (CUDA events inside libgsx) of the batch path at C4 for a batch of `nv` views.
usage: python profiles/stage_times.py [nv=8] [reps=3] [flags=2]   (flags 6 = GSX_GRAD_OVERWRITE | GSX_GRAD_NO_SH)"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, H = gs.scenes, gs.host
nv = int(sys.argv[1]) if len(sys.argv) > 1 else 8
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 2
N, W, Hh, l_max = S.CONFIGS["C4"]
g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + 4)
gd = H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
gd.zero_grads()
cams = [S.orbit_camera(W, Hh, azimuth_deg=45.0 * k) for k in range(nv)]
views = [H.make_view(c["view"], c["proj"], c["campos"], c["fx"], c["fy"], W, Hh, near_thresh=S.NEAR_THRESH,
                     cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max) for c in cams]
r = H.Rasterizer(N, W, Hh, instance_capacity=4 * N)
r.batch_reserve(nv)
gi = torch.as_tensor(S.grad_image(W, Hh)).cuda()
r.batch_step(views, gd, gi, flags=flags)
r.enable_timing(True)
acc = {}
for _ in range(reps):
    r.batch_forward_params(views, gd)
    for k in range(nv):
        r.batch_forward_view(k)
        r.batch_backward_view(k, gi)
        t = r.timing()
        for nm in ("scan_emit", "sort", "ranges", "composite_fwd", "composite_bwd"):
            acc[nm] = acc.get(nm, 0.0) + t[nm] / nv / reps
    r.batch_backward_params(gd, flags)
    t = r.timing()
    for nm in ("preprocess_fwd", "preprocess_bwd"):
        acc[nm] = acc.get(nm, 0.0) + t[nm] / reps
print(json.dumps({"nv": nv, "flags": flags, "ms": {k: round(v, 4) for k, v in acc.items()},
                  "ms_per_view": round(sum(v / (nv if k.startswith("preprocess") else 1) for k, v in acc.items()), 4)}))
