"""Head-to-head of libgsx.so's hand-written onesweep against the reference's sort kernel: cub::DeviceRadixSort::SortPairs of
the vendored CCCL 3.2.0 with its sm_100 tuning (Policy1000) -- oracle/_ref/libcubsort.so, test infrastructure.

Keys are the REAL ones of a bench scene (default C4, view 0): the (tile, Gaussian) instances in emission order (Gaussians
in depth order), i.e. exactly what the view-level path sorts.  Cases:
  tile_u32  : K instances, 32-bit tile-id keys (ceil(log2 T) significant bits) + 32-bit ids   -- the path's second sort
  depth_u32 : P' Gaussians, 32 depth bits + 32-bit ids                                       -- the path's first sort
  key_u64   : K instances, tile<<32|depth keys on 32+ceil(log2 T) bits + ids                 -- gsx_sort_pairs_u64
  ref_f64   : K instances, double keys over all 64 bits, what the reference itself runs (cuda/culling.cu:455-469), CUB only
Every case checks the two outputs are identical (keys and values: both sorts are stable), then times `reps` runs of each
with CUDA events (inputs restored before every run, outside the timed bracket).
usage: python profiles/sort_vs_cub.py [C4] [reps] > profiles/r2_sort_vs_cub.json"""
import ctypes as C
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

gs = importlib.import_module("3dgs_b200")
S, Hh = gs.scenes, gs.host
L = gs._ffi.lib()
cub = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libcubsort.so"))
cfg = sys.argv[1] if len(sys.argv) > 1 else "C4"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
N, W, H, l_max = S.CONFIGS[cfg]
dev = torch.device("cuda")
p = lambda t: C.c_void_p(t.data_ptr())
st = C.c_void_p(0)


def scene_keys():
    g = S.synthetic_gaussians(N, l_max=0, seed=0x3D65 + int(cfg[1:]))
    cam = S.orbit_camera(W, H, 0.0)
    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], None)
    r = Hh.Rasterizer(N, W, H, instance_capacity=4 * N)
    v = Hh.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, H, near_thresh=S.NEAR_THRESH,
                     cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=0)
    r.forward_checked(v, gd)
    s = r.state()
    ids = s["sorted_gaussians"].clone().to(torch.int64)
    rg = s["ranges"].to(torch.int64)
    tiles = torch.repeat_interleave(torch.arange(rg.numel() - 1, device=dev), rg[1:] - rg[:-1])
    depth = s["depth"].clone()
    dbits = depth.view(torch.int32).to(torch.int64) & 0xFFFFFFFF          # positive floats: bit pattern is order-preserving
    touched = s["tiles_touched"].clone()
    r.close()
    # emission order: Gaussians by (depth bits, id), each Gaussian's instances together
    order = torch.argsort(dbits[ids] * (1 << 23) + ids, stable=True)       # N < 2^23
    return tiles[order].contiguous(), ids[order].contiguous(), dbits, touched


def timed(fn, restore):
    for _ in range(2):
        restore(); fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        restore()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms)), float(min(ms))


def case(name, keys, vals, bits, kind):
    """kind: 'u32' | 'u64' (both libraries) | 'f64' (CUB only, the reference's call)"""
    n = keys.numel()
    kdt = {"u32": torch.int32, "u64": torch.int64, "f64": torch.float64}[kind]
    k0, v0 = keys.to(kdt).contiguous(), vals.to(torch.int32).contiguous()
    out = dict(case=name, n=n, key_bits=bits, key_bytes=k0.element_size())
    # ---- CUB
    ki, ko, vi, vo = k0.clone(), torch.empty_like(k0), v0.clone(), torch.empty_like(v0)
    need = C.c_size_t(0)
    if kind == "f64":
        rc = cub.cub_sort_pairs_f64(p(ki), p(ko), p(vi), p(vo), C.c_longlong(n), None, C.byref(need), st)
    else:
        f = cub.cub_sort_pairs_u32 if kind == "u32" else cub.cub_sort_pairs_u64
        rc = f(p(ki), p(ko), p(vi), p(vo), C.c_longlong(n), 0, bits, None, C.byref(need), st)
    assert rc == 0
    temp = torch.empty(max(need.value, 16), dtype=torch.uint8, device=dev)
    tb = C.c_size_t(temp.numel())

    def run_cub():
        if kind == "f64":
            rc_ = cub.cub_sort_pairs_f64(p(ki), p(ko), p(vi), p(vo), C.c_longlong(n), p(temp), C.byref(tb), st)
        else:
            rc_ = f(p(ki), p(ko), p(vi), p(vo), C.c_longlong(n), 0, bits, p(temp), C.byref(tb), st)
        assert rc_ == 0

    def restore_cub():
        ki.copy_(k0); vi.copy_(v0)
    med, best = timed(run_cub, restore_cub)
    passes_cub = None
    out["cub"] = dict(ms_median=round(med, 4), ms_best=round(best, 4), temp_bytes=int(need.value))
    ck, cv = ko.clone(), vo.clone()
    if kind != "f64":
        # ---- libgsx onesweep
        gi_, go_, gvi, gvo = k0.clone(), torch.empty_like(k0), v0.clone(), torch.empty_like(v0)
        wsb = L.gsx_sort_workspace_bytes(C.c_longlong(n))
        ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
        g = L.gsx_sort_pairs_u32 if kind == "u32" else L.gsx_sort_pairs_u64

        def run_gsx():
            gs._ffi.check(g(p(gi_), p(go_), p(gvi), p(gvo), C.c_longlong(n), None, bits, p(ws), C.c_size_t(wsb), st))

        def restore_gsx():
            gi_.copy_(k0); gvi.copy_(v0)
        med, best = timed(run_gsx, restore_gsx)
        passes = (bits + 7) // 8
        alg = n * (k0.element_size() + passes * 2 * (k0.element_size() + 4))   # histogram read + passes x (read + write)
        out["gsx"] = dict(ms_median=round(med, 4), ms_best=round(best, 4), passes=passes,
                          note="even pass counts include the copy to the documented output buffer" if passes % 2 == 0 else "")
        out["identical_output"] = bool(torch.equal(go_, ck) and torch.equal(gvo, cv))
        out["algorithmic_bytes"] = int(alg)
        for w in ("cub", "gsx"):
            out[w]["GBps_on_algorithmic_bytes"] = round(alg / out[w]["ms_median"] / 1e6, 1)
        out["gsx_over_cub_speed"] = round(out["cub"]["ms_median"] / out["gsx"]["ms_median"], 3)
    return out


tiles, ids, dbits, touched = scene_keys()
K = tiles.numel()
T = ((W + 15) // 16) * ((H + 15) // 16)
tb_ = max(1, int(np.ceil(np.log2(T))))
res = dict(config=cfg, N=N, K=K, T=T, tile_bits=tb_, reps=reps, cub_version=int(cub.cub_version()),
           how="CUDA events around one sort call, inputs restored outside the bracket, median of reps; keys = the real instances of view 0")
cases = [case("tile_u32", tiles, ids, tb_, "u32")]
live = torch.nonzero(touched > 0).flatten()
perm = live[torch.randperm(live.numel(), device=dev)]                     # the compacted depth list arrives unordered
cases.append(case("depth_u32", dbits[perm], perm, 32, "u32"))
k64 = (tiles << 32) | dbits[ids]
cases.append(case("key_u64", k64, ids, 32 + tb_, "u64"))
zmax = float(dbits.new_tensor(0).item())                                   # reference key: z + (max_z + 1) * tile as double
depth_f = dbits.to(torch.int32).view(torch.float32)
mz = float(depth_f[torch.unique(ids)].max().item()) + 1.0
cases.append(case("ref_f64 (what the reference runs: double keys, 64 bits)", depth_f[ids].double() + mz * tiles.double(), ids, 64, "f64"))
res["cases"] = cases
# the whole F7 sort work: reference = one 64-bit CUB sort of K doubles; this repo = depth sort of P' + tile sort of K
res["summary"] = dict(reference_f64_sort_ms=cases[3]["cub"]["ms_median"],
                      repo_two_level_ms=round(cases[0]["gsx"]["ms_median"] + cases[1]["gsx"]["ms_median"], 4),
                      cub_two_level_ms=round(cases[0]["cub"]["ms_median"] + cases[1]["cub"]["ms_median"], 4))
print(json.dumps(res))
