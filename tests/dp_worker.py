"""Multi-GPU worker (run under torch.distributed.run, one rank per GPU): dp.ViewParallel over the available transports
(NCCL, and symmetric memory: multimem / p2p) against the fused single-GPU batch step on the same scene.
Prints one line 'DP_OK ...' on rank 0 if all gradients agree within the spec tolerance."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, H, dp, ffi = gs.scenes, gs.host, gs.dp, gs._ffi


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
    dist.init_process_group("nccl", device_id=dev)
    N, W, Hh, l_max, NV = 30011, 256, 176, 3, 8          # N not a multiple of 4 or 128: ragged tails everywhere
    g = S.synthetic_gaussians(N, l_max=l_max, seed=123)
    cams = [S.orbit_camera(W, Hh, azimuth_deg=45.0 * k) for k in range(NV)]
    views = [H.make_view(c["view"], c["proj"], c["campos"], c["fx"], c["fy"], W, Hh, near_thresh=S.NEAR_THRESH,
                         cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.1, l_max=l_max) for c in cams]
    gis = [torch.as_tensor(S.grad_image(W, Hh, seed=500 + k) * 1e3).to(dev) for k in range(NV)]
    mk = lambda: H.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"], device=dev)
    # reference: all 8 views fused on this GPU
    gd0 = mk()
    gd0.zero_grads()
    r0 = H.Rasterizer(N, W, Hh, instance_capacity=8 * N)
    r0.batch_reserve(NV)
    r0.batch_step(views, gd0, gis, flags=ffi.GSX_GRAD_OVERWRITE)
    torch.cuda.synchronize()
    want = {k: t.clone() for k, t in gd0.grads.items()}
    r0.close()
    report = []
    # the step behind the C ABI (gsx_dp_step / gsx_dp_render_host: NCCL called from C++), device and host-buffer entry points
    gd = mk()
    r = H.Rasterizer(N, W, Hh, instance_capacity=8 * N)
    nat = dp.DpNative(r, gd, views, l_max, world, rank)
    mine = dp.shard_views(NV, world, rank)
    for mode in ("step", "render_host"):
        if mode == "step":
            for _ in range(2):
                nat.step([gis[k] for k in mine])
        else:
            gih = [gis[k].cpu().pin_memory() for k in mine]
            imh = [torch.zeros((Hh, W, 3)).pin_memory() for _ in mine]
            nat.render_host(gih, imh)
            assert all(float(t.abs().sum()) > 0 for t in imh)
        torch.cuda.synchronize()
        dist.barrier()
        worst = 0.0
        for k, w in want.items():
            a, b = nat.grads[k].float().reshape(-1), w.reshape(-1)
            rel = float((a - b).norm() / (b.norm() + 1e-30))
            worst = max(worst, rel)
            assert rel <= 1e-3, f"gsx_dp_{mode}: d {k} rel err {rel}"
        report.append(f"capi-{mode}:{worst:.1e}")
    nat.close()
    r.close()
    # (symmetric memory?, d rgb gather, fused extraction?): NCCL only with both buffer layouts, then the hybrid default
    # (per-view NCCL gathers + in-switch allreduce) and the two peer-store variants
    for symmetric, gather, fused in ((False, "nccl", "1"), (False, "fused", "1"), (True, "nccl", "1"), (True, "fused", "1"),
                                     (True, "fused", "0")):
        os.environ["GSX_DP_FUSED"] = fused
        os.environ["GSX_DP_GATHER"] = gather
        gd = mk()
        gd.zero_grads()
        r = H.Rasterizer(N, W, Hh, instance_capacity=8 * N)
        vp = dp.ViewParallel(r, gd, views, [c["campos"] for c in cams], l_max, world, rank, symmetric=symmetric)
        mine = dp.shard_views(NV, world, rank)
        for _ in range(2):                                  # twice: buffers are reused across steps
            vp.step([gis[k] for k in mine])
        torch.cuda.synchronize()
        dist.barrier()
        worst = 0.0
        for k, w in want.items():
            a, b = gd.grads[k].float().reshape(-1), w.reshape(-1)
            rel = float((a - b).norm() / (b.norm() + 1e-30))
            worst = max(worst, rel)
            assert rel <= 1e-3, f"transport {vp.ex.transport}: d {k} rel err {rel}"
        how = ("-pergather" if vp.nccl_gather else "") if not symmetric else ("-hybrid" if vp.nccl_gather else ("-fused" if vp.fused_extract else "-push"))
        report.append(f"{vp.ex.transport}{how}:{worst:.1e}" + (f"(no symm: {vp.ex.why_not})" if symmetric and vp.ex.transport == "nccl" else ""))
        r.close()
    if rank == 0:
        print("DP_OK world=%d %s" % (world, " ".join(report)), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
