"""World-size-2 gloo test (CPU) of the view-parallel exchange in 3dgs_b200/dp.py: view sharding, all-gather of the per-view
d rgb, allreduce of the geometry gradients, local rebuild of the SH gradients (rank-1 identity).  The per-view compute is
the CPU oracle here (the checker); on GPUs the same GradExchange is driven by the CUDA batch path (dp.ViewParallel)."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N, W, H, L_MAX, N_VIEWS = 1500, 96, 64, 2, 4


def _scene():
    gs = importlib.import_module("3dgs_b200")
    S = gs.scenes
    g = S.synthetic_gaussians(N, l_max=L_MAX, seed=91)
    cams = [S.orbit_camera(W, H, azimuth_deg=90.0 * k + 15.0) for k in range(N_VIEWS)]
    gis = [S.grad_image(W, H, seed=300 + k) * 1e3 for k in range(N_VIEWS)]
    return gs, S, g, cams, gis


def _view_backward(S, g, cam, gi):
    """oracle fwd+bwd of one view -> (global-index grads dict, dense d rgb [N][3])"""
    from oracle import golden as G
    gv = G.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, H, near_thresh=S.NEAR_THRESH,
                     cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.1, l_max=L_MAX)
    o = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    grads, inter = G.backward(gv, o, gi, want_intermediates=True)
    drgb = np.zeros((N, 3), np.float32)
    drgb[o.i("vis")] = inter["grad_rgb"]
    return grads, drgb


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir, interleaved):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import golden as G
        gs, S, g, cams, gis = _scene()
        dp = gs.dp
        ex = dp.GradExchange(N, N_VIEWS, world, rank, "cpu", interleaved=interleaved)
        assert ex.slots_mine == ([rank, world + rank] if interleaved else ex.mine)
        assert ex.mine == dp.shard_views(N_VIEWS, world, rank) == list(range(rank * 2, rank * 2 + 2))
        geom = {nm: np.zeros((N, w) if w > 1 else (N,), np.float32) for nm, w in dp.GEOM}
        key = dict(xyz="xyz", scale="scale", quaternion="quat", opacity="opacity")
        for k, v in enumerate(ex.mine):
            grads, drgb = _view_backward(S, g, cams[v], gis[v])
            ex.local_drgb(k).copy_(torch.from_numpy(drgb))
            for nm in geom:
                geom[nm] += grads[key[nm]].reshape(geom[nm].shape)
        for nm in geom:
            ex.geom[nm].copy_(torch.from_numpy(geom[nm]))
        ex.all_gather_drgb()
        ex.all_reduce_geom()
        # local rebuild of the band-0 + SH gradients from every view's d rgb (what gsx_sh_grads_from_view_rgb does on GPU)
        nsh = (L_MAX + 1) ** 2 - 1
        sh_g, b0_g = np.zeros((N, nsh, 3), np.float32), np.zeros((N, 3), np.float32)
        for v in range(N_VIEWS):
            d = ex.drgb[ex.slot_of_view[v]].numpy()
            a, b, _ = G.sh_backward(g["xyz"], g["band0"], g["sh"][:, :nsh], cams[v]["campos"], d, L_MAX)
            sh_g += a.reshape(N, nsh, 3)
            b0_g += b
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), sh=sh_g, band0=b0_g, drgb=ex.drgb.numpy()[ex.slot_of_view],
                 **{nm: ex.geom[nm].numpy() for nm in geom})
    finally:
        dist.destroy_process_group()


def test_shard_views():
    gs = importlib.import_module("3dgs_b200")
    dp = gs.dp
    assert [dp.shard_views(8, 4, r) for r in range(4)] == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert dp.shard_views(8, 1, 0) == list(range(8)) and dp.shard_views(8, 8, 7) == [7]
    with pytest.raises(ValueError):
        dp.shard_views(8, 3, 0)
    with pytest.raises(ValueError):
        dp.shard_views(8, 2, 2)
    ex = dp.GradExchange(10, 4, 1, 0, "cpu")
    gathered, reduced = ex.link_bytes_per_step()
    assert gathered == 4 * 10 * 3 * 4 and reduced >= 10 * 11 * 4
    for nm, w in dp.GEOM:                                    # views of one flat buffer, 16-byte aligned segments
        assert ex.geom[nm].data_ptr() % 16 == 0 and ex.geom[nm].numel() == 10 * w


@pytest.mark.timeout(300)
@pytest.mark.parametrize("interleaved", [False, True])
def test_view_parallel_exchange_world2(tmp_path, interleaved):
    """blocked buffer layout (one all-gather after the views) and interleaved (one all-gather per local view)"""
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), interleaved), nprocs=world, join=True)
    gs, S, g, cams, gis = _scene()
    tot, drgb_all = None, []
    for v in range(N_VIEWS):
        grads, drgb = _view_backward(S, g, cams[v], gis[v])
        drgb_all.append(drgb)
        tot = grads if tot is None else {k: tot[k] + grads[k] for k in grads}
    nsh = (L_MAX + 1) ** 2 - 1
    outs = [np.load(os.path.join(str(tmp_path), f"rank{r}.npz")) for r in range(world)]
    for o in outs:
        # (the oracle's OpenMP reduction order differs between processes: compare to the parent's recomputation with a
        # tolerance, and the two ranks' gathered copies bit-exactly below)
        np.testing.assert_allclose(o["drgb"], np.stack(drgb_all), rtol=1e-3, atol=1e-5 * np.abs(np.stack(drgb_all)).max())
        for nm, ko in (("xyz", "xyz"), ("scale", "scale"), ("quaternion", "quat"), ("opacity", "opacity")):
            np.testing.assert_allclose(o[nm].reshape(-1), tot[ko].reshape(-1), rtol=1e-3, atol=1e-5 * np.abs(tot[ko]).max())
        np.testing.assert_allclose(o["band0"], tot["band0"], rtol=1e-3, atol=1e-5 * np.abs(tot["band0"]).max())
        np.testing.assert_allclose(o["sh"], tot["sh"].reshape(N, -1, 3)[:, :nsh], rtol=1e-3,
                                   atol=1e-5 * np.abs(tot["sh"]).max())
    for nm in ("drgb", "sh", "band0", "xyz", "scale", "quaternion", "opacity"):       # both ranks hold the same sums
        np.testing.assert_array_equal(outs[0][nm], outs[1][nm])


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_view_parallel_multi_gpu_transports():
    """>= 2 GPUs: dp.ViewParallel (NCCL and symmetric-memory transports) == the fused single-GPU batch (tests/dp_worker.py)."""
    import subprocess
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    n = 2
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                          "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
                          os.path.join(ROOT, "tests", "dp_worker.py")], capture_output=True, text=True, timeout=580)
    assert out.returncode == 0 and "DP_OK" in out.stdout, (out.stdout + out.stderr)[-3000:]
