"""Bandwidth of the fused Adam launch at the C4 size: 6 M Gaussians x 59 floats, CUDA-event timed,
3 warm-up + 10 timed launches, arrays (4 x 1.4 GB) far larger than L2.  Prints one JSON line.
Usage: python profiles/adam_bw.py [tag]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
H = gs.host
N = 6_000_000
widths = [3, 3, 4, 1, 3, 45]
gen = torch.Generator(device="cuda").manual_seed(1)
mk = lambda w, s: torch.randn((N, w), device="cuda", generator=gen) * s
groups = [(mk(w, 1.0), mk(w, 1e-3), mk(w, 1e-3), mk(w, 1e-3) ** 2, 1e-4) for w in widths]
uvg, uva, dur = mk(2, 1e-4), torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda", dtype=torch.int32)
vis = (torch.rand(N, device="cuda", generator=gen) < 0.52).to(torch.uint8)
peak = 6650.0   # fallback of B200_PROFILING.md
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(mask):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(13):
        if it == 3:
            e0.record()
        H.adam_step_fused(groups, mask, 0.9, 0.999, 1e-8, 1 - 0.9 ** (it + 1), 1 - 0.999 ** (it + 1), uvg, uva, dur)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 10


full_bytes = 28 * 59 * N + 24 * N
ms_all, ms_vis = timed(None), timed(vis)
nv = int(vis.sum().item())
vis_bytes = 28 * 59 * nv + 24 * nv + N
print(json.dumps({"tag": sys.argv[1] if len(sys.argv) > 1 else "", "N": N, "floats_per_gaussian": 59,
                  "all_visible": {"ms": round(ms_all, 4), "bytes": full_bytes, "GBps": round(full_bytes / ms_all / 1e6, 1),
                                  "frac_of_hbm_peak": round(full_bytes / ms_all / 1e6 / peak, 4)},
                  "masked_52pct": {"ms": round(ms_vis, 4), "bytes": vis_bytes, "GBps": round(vis_bytes / ms_vis / 1e6, 1),
                                   "frac_of_hbm_peak": round(vis_bytes / ms_vis / 1e6 / peak, 4)},
                  "hbm_peak_GBps": peak}))
