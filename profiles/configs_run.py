"""Per-config single-view timings through the view-level C ABI (gsx_forward / gsx_backward) for BASELINE.json's other
configs: C2 (1M, 1297x840), C3 (3M, 1237x822) fwd+bwd, C5 (10M, 3840x2160) forward only.  CUDA events, 3 warm-ups,
median of `reps`.  usage: python profiles/configs_run.py [reps]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, H = gs.scenes, gs.host
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 9
for cfg, do_bwd in (("C2", True), ("C3", True), ("C5", False)):
    N, W, Hh, l_max = S.CONFIGS[cfg]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + int(cfg[1:]))
    gd = H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    del g
    gd.zero_grads()
    cam = S.orbit_camera(W, Hh, azimuth_deg=45.0)
    v = H.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, Hh, near_thresh=S.NEAR_THRESH,
                    cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max, flags=2)
    r = H.Rasterizer(N, W, Hh, instance_capacity=4 * N)
    st = r.forward_checked(v, gd)
    gi = torch.as_tensor(S.grad_image(W, Hh)).cuda()
    r.enable_timing(True)
    fw, bw, stages = [], [], {}
    for it in range(reps + 3):
        e0, e1, e2 = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e0.record(); r.forward(v, gd); e1.record()
        if do_bwd:
            r.backward(v, gd, gi)
        e2.record(); torch.cuda.synchronize()
        if it >= 3:
            fw.append(e0.elapsed_time(e1)); bw.append(e1.elapsed_time(e2))
            for k, x in r.timing().items():
                stages.setdefault(k, []).append(x)
    s_, sp_ = r.count_work()
    out = dict(config=cfg, N=N, W=W, H=Hh, l_max=l_max, P=st["P"], K=st["K"], S=s_, S_plus=sp_,
               fwd_ms=round(float(np.median(fw)), 3), bwd_ms=round(float(np.median(bw)), 3) if do_bwd else None,
               stage_ms={k: round(float(np.median(x)), 4) for k, x in stages.items() if np.median(x) >= 0})
    print(json.dumps(out), flush=True)
    r.close(); del r, gd
    torch.cuda.empty_cache()
