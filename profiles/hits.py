"""Work units of the compositing kernels on C4 view 0: S, S+ (pixel-entries the reference evaluates / that contribute) and
the warp-hits of this repo's kernels (8x8 pixel block x entry pairs surviving the extent cull / with a contributing pixel).
usage: python profiles/hits.py [C4]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, H = gs.scenes, gs.host
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
N, W, Hh, l_max = S.CONFIGS[cfg]
g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + int(cfg[1:]))
gd = H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
cam = S.orbit_camera(W, Hh, 0.0)
v = H.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, Hh, near_thresh=S.NEAR_THRESH,
                cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max)
r = H.Rasterizer(N, W, Hh, instance_capacity=4 * N)
st = r.forward_checked(v, gd)
gi = torch.as_tensor(S.grad_image(W, Hh)).cuda()
s, sp = r.count_work()
h, hc = r.count_hits(gi)
print(json.dumps(dict(config=cfg, P=st["P"], K=st["K"], S=s, S_plus=sp, warp_hits=h, contributing_warp_hits=hc,
                      pixel_evals=64 * h, pixel_evals_in_contributing_hits=64 * hc,
                      contributing_pixel_fraction_of_contributing_hits=round(sp / (64.0 * hc), 4) if hc else None)))
