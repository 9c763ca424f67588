"""Short C4 workload for ncu captures: `batches` batches of `views` views through the batch path (no timing legs).
Usage: python profiles/prof_driver.py [views=2] [batches=2]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, H = gs.scenes, gs.host
nv = int(sys.argv[1]) if len(sys.argv) > 1 else 2
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 2
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
for _ in range(nb):
    r.batch_step(views, gd, gi, flags=2)
# the step after the path: one fused Adam launch over the gradients just produced
names = ["xyz", "scale", "quaternion", "opacity", "band0", "sh"]
groups = [(getattr(gd, nm).view(N, -1), gd.grads[nm].view(N, -1), torch.zeros_like(getattr(gd, nm)).view(N, -1),
           torch.zeros_like(getattr(gd, nm)).view(N, -1), 1e-4) for nm in names]
vis = (torch.rand(N, device="cuda") < 0.52).to(torch.uint8)   # C4: 52 % of the Gaussians survive a view's cull
H.adam_step_fused(groups, vis, 0.9, 0.999, 1e-8, 0.1, 0.001, torch.zeros((N, 2), device="cuda"),
                  torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda", dtype=torch.int32))
# the step before the backward: loss + image gradient of the last rendered view
img = r.state_image().clone()
H.fused_loss(img, (img + 0.05 * torch.randn_like(img)).clamp_(0, 1), 0.2, sync=False)
torch.cuda.synchronize()
print("ok", r.stats())
