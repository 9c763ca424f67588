"""Timing of the fused L1 + SSIM loss on one 1920x1080 view: CUDA events, 3 warm-up + 20 timed calls
(two launches each), loss left on the device.  Prints one JSON line.  Usage: python profiles/loss_bw.py [tag]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
H = gs.host
Hh, W = 1080, 1920
gen = torch.Generator(device="cuda").manual_seed(2)
pred = torch.rand((Hh, W, 3), device="cuda", generator=gen)
gt = (pred + 0.05 * torch.randn((Hh, W, 3), device="cuda", generator=gen)).clamp_(0, 1)
grad = torch.empty_like(pred)
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)      # larger than L2, written between calls
e = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(23)]
for a, b in e:
    flush.fill_(1)
    a.record()
    loss, _ = H.fused_loss(pred, gt, 0.2, image_grad=grad, sync=False)
    b.record()
torch.cuda.synchronize()
ms = sorted(a.elapsed_time(b) for a, b in e[3:])
med = ms[len(ms) // 2]
nbytes = 132 * W * Hh
print(json.dumps({"tag": sys.argv[1] if len(sys.argv) > 1 else "", "image": [W, Hh], "ms_median": round(med, 4), "ms_min": round(ms[0], 4),
                  "bytes": nbytes, "GBps": round(nbytes / med / 1e6, 1), "l2": "256 MB flush write between calls",
                  "loss": float(loss.item()), "grad_checksum": float(grad.double().abs().sum().item())}))
