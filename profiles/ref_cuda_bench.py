"""Same-box baseline B1 on its own: the REFERENCE's CUDA kernels (oracle/_ref/libref3dgs.so) replaying rasterize_image +
backward_pass on one view of a bench scene (oracle/ref_cuda.py replay_view; bench.py runs the same leg inside its line).
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
g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + int(cfg[1:]))
cam = S.orbit_camera(W, H, azimuth_deg=0.0)
d = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()
gd = {k: (d(v) if v is not None else None) for k, v in g.items()}
out = R.replay_view(gd, cam, W, H, l_max, S.NEAR_THRESH, S.CULL_PADDING, S.MH_DIST, d(S.grad_image(W, H)), reps=reps)
out.update(config=cfg, N=N, W=W, H=H)
print(json.dumps(out))
