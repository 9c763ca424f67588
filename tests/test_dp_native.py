"""The multi-GPU step behind the C ABI (gsx_dp_*, csrc/dp.cu: NCCL called from C++).
* tests/cpp/dp_smoke.cpp is a C++ program that uses only include/gsx.h + the CUDA runtime; it is run with world 1 on any GPU
  box and with world 2 (one process per GPU, NCCL id passed through a file) where two GPUs exist.
* DpNative (3dgs_b200/dp.py, a ctypes caller of the same entry points) at world 1: gradients equal the fused batch step, the
  Adam epilogue equals a separate gsx_adam_step_fused, the host-buffer step equals the device step."""
import importlib
import os
import subprocess
import tempfile

import numpy as np
import pytest

import util
from util import S, gs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "bin", "dp_smoke")


def test_dp_smoke_binary_uses_only_the_c_abi():
    if not os.path.exists(EXE):
        pytest.skip("tests/cpp/bin/dp_smoke not built (python __graft_entry__.py build)")
    out = subprocess.run(["ldd", "-r", EXE], capture_output=True, text=True)
    assert "libgsx.so" in out.stdout and "undefined symbol" not in out.stdout + out.stderr
    assert "nccl" not in out.stdout.lower() and "torch" not in out.stdout.lower()     # NCCL is bound at run time by libgsx.so
    src = open(os.path.join(ROOT, "tests", "cpp", "dp_smoke.cpp")).read()
    assert '#include "gsx.h"' in src and "torch" not in src and "nccl.h" not in src


def _run_ranks(world, n=20000, views=4):
    with tempfile.TemporaryDirectory() as td:
        idf = os.path.join(td, "nccl_id")
        procs = []
        for r in range(world):
            env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(r))
            procs.append(subprocess.Popen([EXE, str(r), str(world), idf, str(n), str(views)], stdout=subprocess.PIPE,
                                          stderr=subprocess.STDOUT, text=True, env=env))
        outs = []
        for p in procs:
            try:
                o, _ = p.communicate(timeout=240)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            outs.append((p.returncode, o))
    return outs


@pytest.mark.gpu
def test_dp_smoke_cpp_world1():
    if not os.path.exists(EXE):
        pytest.skip("tests/cpp/bin/dp_smoke not built")
    (rc, out), = _run_ranks(1)
    assert rc == 0 and "DP_SMOKE_OK rank 0/1" in out, out[-2000:]


@pytest.mark.gpu
def test_dp_smoke_cpp_world2():
    import torch
    if not os.path.exists(EXE):
        pytest.skip("tests/cpp/bin/dp_smoke not built")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    outs = _run_ranks(2)
    for r, (rc, out) in enumerate(outs):
        assert rc == 0 and f"DP_SMOKE_OK rank {r}/2" in out, out[-2000:]


@pytest.mark.gpu
@pytest.mark.parametrize("l_max", [3, 0])
def test_dp_native_world1_matches_batch_step_and_adam(l_max):
    import torch
    W, H, N = 224, 160, 12000
    cams = [S.orbit_camera(W, H, azimuth_deg=a) for a in (10.0, 130.0, 250.0)]
    g = S.synthetic_gaussians(N, l_max=l_max, seed=61)
    Hh, dp, ffi = gs.host, gs.dp, gs._ffi
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    gis = [torch.as_tensor(S.grad_image(W, H, seed=90 + k) * 1e3).cuda() for k in range(3)]
    views = [util.gsx_view(c, l_max, 0.3) for c in cams]
    r = Hh.Rasterizer(N, W, H, instance_capacity=8 * N)
    r.batch_reserve(3)
    gd.zero_grads()
    r.batch_step(views, gd, gis, flags=ffi.GSX_GRAD_OVERWRITE)
    torch.cuda.synchronize()
    want = {k: t.clone() for k, t in gd.grads.items()}
    nat = dp.DpNative(r, gd, views, l_max, 1, 0)
    nat.step(gis)
    torch.cuda.synchronize()
    names = ["xyz", "scale", "quaternion", "opacity", "band0"] + (["sh"] if l_max else [])
    for k in names:
        util.assert_grad_close(nat.grads[k].cpu().numpy().reshape(want[k].shape), want[k].cpu().numpy(), f"dp native d {k}")
    # host-buffer entry point: same gradients, images returned
    gih = [x.cpu().pin_memory() for x in gis]
    imh = [torch.zeros((H, W, 3)).pin_memory() for _ in gis]
    nat.render_host(gih, imh)
    for k in names:
        util.assert_grad_close(nat.grads[k].cpu().numpy().reshape(want[k].shape), want[k].cpu().numpy(), f"dp native host d {k}")
    r.forward(views[2], gd)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(imh[2].numpy(), r.state()["image"].cpu().numpy())
    # Adam as the epilogue of the exchange == a separate fused Adam launch on the same gradients
    par = {k: getattr(gd, k).clone() for k in names}
    mom = {k: (torch.zeros_like(par[k]), torch.zeros_like(par[k])) for k in names}
    lrs = dict(xyz=1.6e-4, scale=5e-3, quaternion=1e-3, opacity=5e-2, band0=2.5e-3, sh=1.25e-4)
    groups = [(par[k].clone().view(N, -1), want[k].view(N, -1), torch.zeros_like(par[k]).view(N, -1), torch.zeros_like(par[k]).view(N, -1), lrs[k]) for k in names]
    Hh.adam_step_fused(groups, None, 0.9, 0.999, 1e-8, 0.1, 0.001)
    saved = {k: getattr(gd, k) for k in names}
    ep = nat.adam_epilogue({k: par[k] for k in names}, mom, lrs, 0.9, 0.999, 1e-8, 0.1, 0.001)
    nat.step(gis, adam=ep)          # renders from gd's parameters, updates the copies in `par`
    torch.cuda.synchronize()
    for (p_ref, _, m_ref, v_ref, _), k in zip(groups, names):
        # the two gradient computations differ in their atomics' summation order (1e-6 relative): the first Adam step is
        # lr * g / (|g| + eps), so only gradients of magnitude ~eps can move visibly -- counted, not ignored
        assert util.close_frac(par[k].cpu().numpy().reshape(N, -1), p_ref.cpu().numpy(), 2e-6, 1e-7) <= 5e-4, f"adam epilogue {k}"
        # the first moment is (1 - b1) * g, linear in the gradient: the summation-order noise of g is bounded relative to the
        # array's scale, not to each element (an element that is a near-cancelling sum may differ by far more than 1e-3 of
        # itself: a per-element relative bound made this check depend on the run's atomic order)
        mg, mr = mom[k][0].cpu().numpy().reshape(N, -1).astype(np.float64), m_ref.cpu().numpy().astype(np.float64)
        assert np.linalg.norm(mg - mr) <= 1e-5 * np.linalg.norm(mr), f"adam epilogue m {k} (norm)"
        assert util.close_frac(mg, mr, 1e-3, 1e-5 * float(np.abs(mr).max())) == 0.0, f"adam epilogue m {k} (elements)"
    assert all(torch.equal(saved[k], getattr(gd, k)) for k in names)
    nat.close()
    r.close()
