"""gsx_fused_loss_static: the entry point that allocates nothing (its scratch is a file-scope static __device__ array inside
libgsx.so, include/gsx.h).  It has no counterpart in the reference; it is checked against the existing entry point,
gsx_fused_loss, which the oracle pins (tests/test_gpu_parity.py::test_fused_loss_parity).

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
nothing here depends on the state another test leaves, and the static array's previous contents must not matter."""
import ctypes as C
import importlib

import numpy as np
import pytest

ffi = importlib.import_module("3dgs_b200._ffi")

GSX_E_WORKSPACE = 5
MAX_PIXELS = 1 << 20   # GSX_FUSED_LOSS_STATIC_MAX_PIXELS


def _p(t):
    return C.c_void_p(t.data_ptr())


def _images(torch, H, W, seed):
    gen = torch.Generator(device="cuda").manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    base = 0.5 + 0.3 * (torch.sin(xx / 7.0) * torch.cos(yy / 5.0))[..., None] * torch.tensor([1.0, 0.7, -0.5], device="cuda")
    pred = (base + 0.05 * torch.randn((H, W, 3), device="cuda", generator=gen)).clamp(0, 1).contiguous()
    gt = (base + 0.05 * torch.randn((H, W, 3), device="cuda", generator=gen)).clamp(0, 1).contiguous()
    return pred.float(), gt.float()


@pytest.mark.gpu
def test_fused_loss_static_matches_fused_loss_on_two_calls_in_a_row():
    """Two calls in a row on different images and sizes (the second smaller, so stale derivative maps and a stale loss sum
    from the first sit in the array), each against gsx_fused_loss on the same inputs.  Tolerance: the loss tests' own --
    loss 2e-6 absolute (the CTA sums arrive by atomicAdd in any order), gradient 1e-4 of its largest magnitude (in fact
    the gradients are bit-identical: same kernels, no atomics on that path)."""
    import torch
    L = ffi.lib()
    torch.cuda.set_device(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    cases = [(720, 1280, 0.2, 11), (37, 45, 0.5, 12)]
    # every buffer of both calls exists before the first call: the calls themselves must not need any memory
    bufs = []
    for H, W, w, seed in cases:
        pred, gt = _images(torch, H, W, seed)
        need = L.gsx_fused_loss_workspace_bytes(H, W)
        bufs.append(dict(pred=pred, gt=gt, ws=torch.empty(need, device="cuda", dtype=torch.uint8),
                         g_ref=torch.empty_like(pred), g_new=torch.full_like(pred, float("nan")),
                         l_ref=torch.zeros(1, device="cuda"), l_new=torch.full((1,), float("nan"), device="cuda")))
    for (H, W, w, _), b in zip(cases, bufs):
        ffi.check(L.gsx_fused_loss(_p(b["pred"]), _p(b["gt"]), H, W, C.c_float(w), _p(b["g_ref"]), _p(b["ws"]),
                                   C.c_size_t(b["ws"].numel()), _p(b["l_ref"]), None, st))
    torch.cuda.synchronize()
    free = []
    for (H, W, w, _), b in zip(cases, bufs):                     # the two calls in a row
        ffi.check(L.gsx_fused_loss_static(_p(b["pred"]), _p(b["gt"]), H, W, C.c_float(w), _p(b["g_new"]), _p(b["l_new"]),
                                          None, st))
        free.append(torch.cuda.mem_get_info()[0])
    torch.cuda.synchronize()
    free.append(torch.cuda.mem_get_info()[0])
    assert free[1] == free[2] == free[0], f"free device memory moved across the no-alloc calls: {free}"
    for (H, W, w, _), b in zip(cases, bufs):
        l_ref, l_new = float(b["l_ref"].item()), float(b["l_new"].item())
        assert np.isfinite(l_new) and abs(l_new - l_ref) < 2e-6, (H, W, l_new, l_ref)
        g_ref, g_new = b["g_ref"].cpu().numpy(), b["g_new"].cpu().numpy()
        assert np.isfinite(g_new).all()
        np.testing.assert_allclose(g_new, g_ref, rtol=0, atol=1e-4 * float(np.abs(g_ref).max()))
    # the host-loss form (synchronises) without a loss_dev: the mean then lives in the static array itself
    H, W, w, _ = cases[1]
    b = bufs[1]
    host_loss = C.c_float(-1.0)
    ffi.check(L.gsx_fused_loss_static(_p(b["pred"]), _p(b["gt"]), H, W, C.c_float(w), _p(b["g_new"]), None,
                                      C.byref(host_loss), st))
    assert abs(host_loss.value - float(b["l_ref"].item())) < 2e-6


@pytest.mark.gpu
def test_fused_loss_static_rejects_sizes_beyond_the_static_scratch():
    import torch
    L = ffi.lib()
    torch.cuda.set_device(0)
    a = torch.zeros((16, 16, 3), device="cuda")
    rc = L.gsx_fused_loss_static(_p(a), _p(a), 1025, 1024, C.c_float(0.2), _p(a), None, None, None)   # 1 row too many
    assert rc == GSX_E_WORKSPACE and b"exceeds the static scratch" in L.gsx_last_error()
    assert 1024 * 1024 == MAX_PIXELS
    assert L.gsx_fused_loss_static(_p(a), _p(a), 0, 16, C.c_float(0.2), _p(a), None, None, None) == 3   # GSX_E_ARG
    host = np.zeros((16, 16, 3), np.float32)
    assert L.gsx_fused_loss_static(host.ctypes.data_as(C.c_void_p), _p(a), 16, 16, C.c_float(0.2), _p(a), None, None,
                                   None) in (1, 2)
