import importlib, sys, os, torch, numpy as np
sys.path.insert(0, os.getcwd())
gs = importlib.import_module("3dgs_b200")
S, H = gs.scenes, gs.host
N, W, Hh, l_max = S.CONFIGS["C4"]
g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + 4)
gd = H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
c = S.orbit_camera(W, Hh, azimuth_deg=0.0)
v = H.make_view(c["view"], c["proj"], c["campos"], c["fx"], c["fy"], W, Hh, near_thresh=S.NEAR_THRESH, cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max)
r = H.Rasterizer(N, W, Hh, instance_capacity=4 * N)
r.forward_checked(v, gd)
st = r.state()
t = st["tiles_touched"].cpu().numpy(); d = st["depth"].cpu().numpy()
idx = np.nonzero(t > 0)[0]
order = idx[np.argsort(d[idx], kind="stable")]
cnt = t[order]
nb = (len(cnt) + 2047) // 2048
bs = np.add.reduceat(cnt, np.arange(0, len(cnt), 2048))
print("P'", len(cnt), "K", cnt.sum(), "blocks", nb, "block sum max", bs.max(), "mean", bs.mean(), "top5", np.sort(bs)[-5:], "big frac", (cnt > 16).mean(), "instances in big", cnt[cnt > 16].sum() / cnt.sum(), "max cnt", cnt.max())
ws = np.add.reduceat(cnt, np.arange(0, len(cnt), 32))
print("warp-chunk sums: max", ws.max(), "mean", ws.mean(), "p99", np.percentile(ws, 99))
