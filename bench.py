#!/usr/bin/env python
"""bench.py -- fwd+bwd views/s of the splatting hot path on BASELINE.json's config 4:
synthetic 6M Gaussians, SH deg 3, 1920x1080, a batch of 8 views sharded over the GPUs with an NCCL allreduce of
the per-Gaussian parameter gradients (one process per GPU; `python bench.py --gpus N --steps K --warmup W`, under
torchrun for N > 1).  One step = one batch of 8 views (fwd+bwd of every view + the gradient allreduce).

Prints ONE JSON line (rank 0).  `value` = views/s with inputs resident in HBM; `e2e` = the same metric through the
host-buffer entry point (gsx_render_view_host: H2D of the view's camera block and loss-gradient image from pinned
memory, D2H of the rendered image, every view); `roofline` = the dominant kernel against the measured peak;
`cpu_baseline` = the CPU oracle timed on this box's host cores on a bounded sample (rank 0, N = 1 only).
`--impl reference` times the reference's path on the host cores (the oracle port: the reference has no CPU path and
its CUDA build is timed separately, see BASELINE.md / DESIGN.md).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_VIEWS = 8
METRIC, UNIT = "fwd+bwd views/s (6M Gaussians, SH deg 3, 1920x1080, batch of 8 views)", "views/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C4", help="scene config from 3dgs_b200/scenes.py (C4 = the benchmarked one)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=150.0, help="budget for the CPU legs")
    return ap.parse_args()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p, self.lines, self.index = None, [], index

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                       "200", "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                      text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for ln in self.p.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_scene(cfg):
    gs = importlib.import_module("3dgs_b200")
    S = gs.scenes
    N, W, H, l_max = S.CONFIGS[cfg]
    idx = int(cfg[1:])
    g = S.synthetic_gaussians(N, l_max=l_max, seed=0x3D65 + idx)
    cams = [S.orbit_camera(W, H, azimuth_deg=45.0 * k) for k in range(N_VIEWS)]
    return gs, S, N, W, H, l_max, g, cams


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_threads():
    """every host thread for the CPU legs, whatever OMP_NUM_THREADS the launcher exported (torchrun sets it to 1)"""
    from oracle import golden as G
    n = os.cpu_count() or 1
    try:
        n = max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        pass
    G.lib().gold_set_threads(n)
    return G.lib().gold_max_threads()


def cpu_view_seconds(S, g, cam, W, H, l_max, gi, keep=False):
    """one full view fwd+bwd through the CPU oracle (all host threads)"""
    from oracle import golden as G
    gv = G.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, H, near_thresh=S.NEAR_THRESH,
                     cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max)
    t0 = time.perf_counter()
    st = G.forward(gv, g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"])
    og = G.backward(gv, st, gi)
    dt = time.perf_counter() - t0
    return (dt, st.P, st.K, st, og) if keep else (dt, st.P, st.K)


def run_reference_arm(args):
    """The reference's implementation of the path timed on the host cores.  The reference itself is CUDA-only (no CPU
    path exists), so this is the oracle port (kind 'port') with every host thread.  One step = ONE view of
    the same scene (a bounded sample of the 8-view batch); value = views/s."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import golden as G
    gs, S, N, W, H, l_max, g, cams = make_scene(args.config)
    cores = cpu_threads()                                          # not OMP_NUM_THREADS: torchrun exports 1
    gi = S.grad_image(W, H)
    for k in range(args.warmup):
        cpu_view_seconds(S, g, cams[k % N_VIEWS], W, H, l_max, gi)
    t0 = time.perf_counter()
    for k in range(args.steps):
        cpu_view_seconds(S, g, cams[k % N_VIEWS], W, H, l_max, gi)
    dt = time.perf_counter() - t0
    v = args.steps / dt
    sample = f"{args.steps} single views (fwd+bwd) of the {args.config} scene, one per step, {cores} OpenMP threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: synthetic {N} Gaussians, SH deg {l_max}, {W}x{H}; one view per step on the host CPU"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference has no CPU path (CUDA-only); this is the scalar C++ golden with OpenMP"}))


# ------------------------------------------------------------------------------------------------ GPU arm
def flat_grads(torch, N, sh_stride, device):
    """All parameter gradients as views of ONE flat buffer so the gradient exchange is a single NCCL allreduce
    (236 B/Gaussian at deg 3).  Segments are padded to 16 B."""
    seg = [("xyz", 3), ("scale", 3), ("quaternion", 4), ("opacity", 1), ("band0", 3), ("sh", max(sh_stride, 1))]
    offs, off = {}, 0
    for nm, w in seg:
        offs[nm] = off
        off += (N * w + 3) // 4 * 4
    flat = torch.zeros(off, device=device, dtype=torch.float32)
    views = {}
    for nm, w in seg:
        v = flat[offs[nm]: offs[nm] + N * w]
        views[nm] = v.view(N, w) if nm != "opacity" else v
    return flat, views


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference_arm(args)
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs CUDA devices: there is no CPU fallback for the product path"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL kernels on a high-priority stream: their CTAs are scheduled ahead of the (many, short) CTAs of the SH
        # rebuild kernel they overlap with (dp.py)
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
        dist.init_process_group("nccl", device_id=dev)
    assert N_VIEWS % world == 0, "the 8-view batch shards over 1, 2, 4 or 8 GPUs"
    gs, S, N, W, H, l_max, g, cams = make_scene(args.config)
    Hh = gs.host
    hbm_peak, peak_src = load_peaks()

    gd = Hh.Gaussians(g["xyz"], g["scale"], g["quaternion"], g["opacity"], g["band0"], g["sh"], device=dev)
    flat, gviews = flat_grads(torch, N, gd.sh_stride, dev)
    gd.grads = gviews
    my_views = gs.dp.shard_views(N_VIEWS, world, rank)             # blocked view sharding (dp.py)
    OVER = gs._ffi.GSX_GRAD_OVERWRITE
    mk = lambda cam, flags: Hh.make_view(cam["view"], cam["proj"], cam["campos"], cam["fx"], cam["fy"], W, H,
                                         near_thresh=S.NEAR_THRESH, cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST,
                                         background=0.0, l_max=l_max, flags=flags)
    # GSX_LISTS=contributing: per-tile lists clipped to the alpha >= 1/255 footprint (gsx.h GSX_LISTS_CONTRIBUTING);
    # default = the reference's lists
    VF = 8 if os.environ.get("GSX_LISTS", "reference") == "contributing" else 0
    vstructs = [mk(cams[k], VF) for k in my_views]
    nv = len(vstructs)
    r = Hh.Rasterizer(N, W, H, instance_capacity=4 * N)
    r.batch_reserve(nv)
    gi_host = [torch.as_tensor(S.grad_image(W, H)).pin_memory() for _ in range(nv)]
    img_host = [torch.empty((H, W, 3), dtype=torch.float32).pin_memory() for _ in range(nv)]
    gi_dev = gi_host[0].to(dev)

    # size the instance capacity once (the only host sync of the set-up)
    for v in vstructs:
        r.forward_checked(v, gd)
    for v in vstructs:                                             # re-check after any growth
        assert not r.forward_checked(v, gd)["overflow"]

    # One step = the rank's views of the batch through the batch path (parameters read once per direction, gradients
    # of the rank's views summed in one pass with '=' semantics: no 1.4 GB memset) + the gradient allreduce.
    all_views = [mk(c, VF) for c in cams]
    # GSX_DP_TRANSPORT=nccl forces the NCCL transport (default: symmetric memory -- multimem / p2p -- when available)
    # N > 1: the step lives behind the C ABI (gsx_dp_step / gsx_dp_render_host, csrc/dp.cu: NCCL called from C++);
    # GSX_DP_IMPL=python selects the Python-orchestrated transports of dp.ViewParallel (multimem / p2p kernels) for A/B
    native = world > 1 and os.environ.get("GSX_DP_IMPL", "native") != "python"
    nat = vp = None
    if native:
        nat = gs.dp.DpNative(r, gd, all_views, l_max, world, rank)
    else:
        vp = gs.dp.ViewParallel(r, gd, all_views, [c["campos"] for c in cams], l_max, world, rank,
                                symmetric=os.environ.get("GSX_DP_TRANSPORT", "auto") != "nccl")

    def step():
        if native:
            nat.step(gi_dev)
        else:
            vp.step(gi_dev)

    img_dev = torch.empty((H, W, 3), device=dev, dtype=torch.float32)

    def step_e2e():
        if world == 1:
            r.batch_render_host(vstructs, gd, gi_host, img_host, flags=OVER)
            return
        if native:
            # the library pipelines the copies; the asynchronous form lets consecutive steps overlap their launch overhead
            # (a rank with one 2 ms view per step is otherwise idle while the host enqueues the next step) -- the timed
            # region ends with a device synchronisation, which covers every download
            nat.render_host(gi_host, img_host, sync=False)
            return
        # host buffers in / out around the same exchange: per view H2D of the loss gradient, D2H of the image, on two copy
        # streams (uploads under the per-Gaussian forward, view k's download under its backward) like the library's
        # own gsx_batch_render_host does at N = 1
        cur = torch.cuda.current_stream()
        e2e_go.record(cur)
        with torch.cuda.stream(up_s):
            up_s.wait_event(e2e_go)
            for k_ in range(nv):
                gis_dev[k_].copy_(gi_host[k_], non_blocking=True)
                up_ev[k_].record(up_s)

        def grab(k_):
            if k_ >= 2:
                cur.wait_event(down_ev[k_ - 2])                    # the staging image is free again
            img_stage[k_ & 1].copy_(r.state_image())
            img_ev[k_].record(cur)
            with torch.cuda.stream(down_s):
                down_s.wait_event(img_ev[k_])
                img_host[k_].copy_(img_stage[k_ & 1], non_blocking=True)
                down_ev[k_].record(down_s)
            cur.wait_event(up_ev[k_])                              # this view's loss gradient has arrived
        vp.step(gis_dev, on_image=grab)
        torch.cuda.synchronize()

    if world > 1:
        up_s, down_s = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        gis_dev = [torch.empty((H, W, 3), device=dev, dtype=torch.float32) for _ in range(nv)]
        img_stage = [torch.empty((H, W, 3), device=dev, dtype=torch.float32) for _ in range(2)]
        mkev = lambda: [torch.cuda.Event() for _ in range(nv)]
        up_ev, img_ev, down_ev, e2e_go = mkev(), mkev(), mkev(), torch.cuda.Event()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    assert r.overflow() == 0, "a view exceeded the instance capacity inside the timed region (its work was skipped)"

    # e2e leg: host buffers in, host image out, every view
    for _ in range(2):
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    dt_e2e = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(dt_e2e, op=dist.ReduceOp.MAX)
    dt_e2e = float(dt_e2e.item())

    exch = None
    if world > 1 and native:                                       # per-phase times of one more step (outside the timed region)
        nat.enable_timing(True)
        step()
        barrier()
        exch = {k_: round(v_, 3) for k_, v_ in nat.phase_ms().items()}
        nat.enable_timing(False)
    elif world > 1:
        vp.timing = {}
        step()
        barrier()
        exch = {k_: round(v_, 3) for k_, v_ in vp.phase_ms().items()}
        vp.timing = None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline leg (rank 0): per-stage CUDA-event timing of this rank's views, outside the timed region --------
    r.enable_timing(True)
    acc, reps = {}, 3
    work = dict(P=0, Pt=0, K=0, S=0, Sp=0)
    per_view = ("scan_emit", "sort", "ranges", "composite_fwd", "composite_bwd")
    for rep in range(reps):
        r.batch_forward_params(vstructs, gd)
        for k_ in range(nv):
            r.batch_forward_view(k_)
            if rep == 0:
                s_, sp_ = r.count_work()
                stx = r.stats()
                work["P"] += stx["P"]; work["K"] += stx["K"]; work["S"] += s_; work["Sp"] += sp_
                work["Pt"] += int((r.state()["tiles_touched"] > 0).sum().item())
            r.batch_backward_view(k_, gi_dev)
            t_ = r.timing()
            for nm in per_view:
                acc[nm] = acc.get(nm, 0.0) + t_[nm] / nv
        r.batch_backward_params(gd, OVER)
        t_ = r.timing()
        for nm in ("preprocess_fwd", "preprocess_bwd"):            # batch kernels: one launch per batch
            acc[nm] = acc.get(nm, 0.0) + t_[nm]
    r.enable_timing(False)
    launch_ms = {k_: x / reps for k_, x in acc.items()}            # average duration of one launch (stage)
    P, Pt, K, Sx, Sp = (work[k_] / nv for k_ in ("P", "Pt", "K", "S", "Sp"))
    T = ((W + 15) // 16) * ((H + 15) // 16)
    D = 3 * ((l_max + 1) ** 2 - 1)
    passes = (max(1, int(np.ceil(np.log2(T)))) + 7) // 8            # tile-sort passes (8-bit digits over the tile id only)
    fp32_peak = Hh.measure_fp32_tflops()
    # algorithmic work per launch (SURVEY 8d / BASELINE.md section 5)
    PB_ = 44 + 4 * D                                                # parameter / gradient bytes per Gaussian (236 at deg 3)
    fused_hist = bool(os.environ.get("GSX_FUSED_HIST"))
    direct = os.environ.get("GSX_BINNING", "")[:1] == "d"
    records = os.environ.get("GSX_TILE_LISTS", "1")[:1] != "0" and T <= 65536 and not direct   # emission from the forward's tile records
    keys16 = (records and not fused_hist and os.environ.get("GSX_KEYS16", "1")[:1] != "0"
              and os.environ.get("GSX_RANGES", "")[:1] != "s")
    emit_hist = keys16 and os.environ.get("GSX_EMIT_HIST", "1")[:1] != "0"   # tile-sort histograms accumulated by the emission
    alg = {
        # batch kernels (one launch per batch of nv views): parameters once + per-view records / packed gradients
        "preprocess_fwd": ("hbm", PB_ * N + nv * (74 * P + 5 * (N - P) + (8 * Pt if records else 0))),
        "preprocess_bwd": ("hbm", PB_ * N + nv * (4 * N + 96 * Pt) + PB_ * N),
        # over the P' Gaussians that touch a tile (compacted by the per-Gaussian kernel): 4-pass depth sort (4 + 4*16), tie
        # fix-up (8), then -- tile records (default) -- one gather of the 8-byte record left in depth order (4 + 8 + 8), emission
        # from the records (4 + 8) + (2 + 4) B per instance; without them: scan through the permutation (20), emission with
        # the SAT (8 + 32) + 8 B per instance.  The 2-byte tile ids of large rectangles are not counted.
        "scan_emit": ("hbm", (108 * Pt + 6 * K) if records else (136 * Pt + 8 * K)),
        # tile sort on 16-bit keys (default): [histogram 2, unless the emission accumulated it] + per pass (2 + 4) in + (2 + 4) out, the last pass writes no keys;
        # on 32-bit keys: 4 + passes * (8 in + 8 out)
        "sort": ("hbm", (((0 if emit_hist else 2) + 12 * passes - 2) if keys16 else ((0 if fused_hist else 4) + 16 * passes)) * K),
        "ranges": ("hbm", 4 * K + 4 * (T + 1)),
        "composite_fwd": ("fp32", 18 * Sx),
        "composite_bwd": ("fp32", 10 * Sx + 46 * Sp),
    }
    stages = {}
    if launch_ms.get("ranges", 1.0) < 0.005:                       # tile ranges come out of the last sort pass (default): no stage
        alg.pop("ranges")
    for k_, (bound, w) in alg.items():
        t = launch_ms[k_] * 1e-3
        if bound == "hbm":
            a, pk, un = w / t / 1e9, hbm_peak, "GB/s"
        else:
            a, pk, un = w / t / 1e12, fp32_peak, "TFLOP/s"
        per_batch = k_ in ("preprocess_fwd", "preprocess_bwd")
        stages[k_] = {"ms": round(launch_ms[k_], 4), "launches_per_batch": 1 if per_batch else nv,
                      "ms_per_view": round(launch_ms[k_] / (nv if per_batch else 1), 4), "bound": bound,
                      "achieved": round(a, 2), "peak": round(pk, 2), "unit": un, "frac": round(a / pk, 4)}
    stage_ms = {k_: v_["ms_per_view"] for k_, v_ in stages.items()}
    top = max(stage_ms, key=stage_ms.get)
    roof = {k_: stages[top][k_] for k_ in ("bound", "achieved", "peak", "unit", "frac")}
    traffic, traffic_src = None, None                              # DRAM bytes per launch of that kernel from the newest committed ncu capture
    for fn in ("traffic_r2.json", "traffic_r1.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as f:
                tj = json.load(f)
            if tj.get(top) is not None:
                traffic, traffic_src = tj[top], f"profiles/{fn} ({tj.get('_capture', 'ncu --set full capture')}); not measured in this run"
                break
        except Exception:
            pass
    roof.update({"kernel": top, "traffic": traffic, "traffic_source": traffic_src, "ms": stages[top]["ms"],
                 "peak_source": peak_src if roof["bound"] == "hbm" else "FP32 FMA peak measured in this run (gsx_measure_fp32_tflops)",
                 "share_of_view": round(stage_ms[top] / sum(stage_ms.values()), 3)})

    # ---- next row, outside the timed region: the optimiser step that consumes these gradients --------
    # gsx_adam_step_fused: every parameter group + the densification accumulators in one launch over the global-index
    # gradient buffers (the reference: 16-19 compactions + 5-6 adam_step launches + 16-19 scatters, trainer.cu:1027-1158)
    next_rows = {}
    if world == 1:
        names = ["xyz", "scale", "quaternion", "opacity", "band0"] + (["sh"] if gd.sh_stride else [])
        lrs = dict(xyz=1.6e-4, scale=5e-3, quaternion=1e-3, opacity=5e-2, band0=2.5e-3, sh=1.25e-4)
        pcl = {nm: getattr(gd, nm).clone().view(N, -1) for nm in names}
        groups = [(pcl[nm], gviews[nm].view(N, -1), torch.zeros_like(pcl[nm]), torch.zeros_like(pcl[nm]), lrs[nm]) for nm in names]
        uvg = torch.randn((N, 2), device=dev) * 1e-4
        uva, dur = torch.zeros(N, device=dev), torch.zeros(N, device=dev, dtype=torch.int32)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps_a = 10
        for it in range(3 + reps_a):
            if it == 3:
                a0.record()
            Hh.adam_step_fused(groups, None, 0.9, 0.999, 1e-8, 1 - 0.9 ** (it + 1), 1 - 0.999 ** (it + 1), uvg, uva, dur)
        a1.record()
        torch.cuda.synchronize()
        adam_ms = a0.elapsed_time(a1) / reps_a
        adam_bytes = 28 * sum(p_.numel() for p_, *_ in groups) + 24 * N   # p,g,m,v read + p,m,v written; per Gaussian uv_grad 8 + accumulator 4+4 + counter 4+4
        assert int(dur[0].item()) == 3 + reps_a and bool(torch.isfinite(pcl["xyz"]).all().item())
        next_rows["optimizer_step"] = {"ms": round(adam_ms, 4), "launches": 1, "bound": "hbm", "bytes": adam_bytes,
                                       "achieved": round(adam_bytes / adam_ms / 1e6, 1), "peak": round(hbm_peak, 1), "unit": "GB/s",
                                       "frac": round(adam_bytes / adam_ms / 1e6 / hbm_peak, 4),
                                       "what": "gsx_adam_step_fused over all 6 parameter groups (59 floats/Gaussian) + densification accumulators, every Gaussian visible"}
        del pcl, groups, uvg, uva, dur

    # ---- next row, outside the timed region: L1 + SSIM loss of a rendered view -> grad_image ----------
    if world == 1:
        rendered = r.state_image().clone()
        gt_img = (rendered + 0.05 * torch.randn_like(rendered)).clamp_(0, 1)
        gl = torch.empty_like(rendered)
        l0_, l1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for it in range(13):
            if it == 3:
                l0_.record()
            loss_dev, _ = Hh.fused_loss(rendered, gt_img, 0.2, image_grad=gl, sync=False)
        l1_.record()
        torch.cuda.synchronize()
        loss_ms = l0_.elapsed_time(l1_) / 10
        loss_bytes = 132 * W * H          # fwd: 24 read + 36 written; bwd: 36 + 24 read + 12 written
        assert bool(torch.isfinite(gl).all().item()) and 0.0 <= float(loss_dev.item()) < 1.0
        next_rows["fused_loss"] = {"ms": round(loss_ms, 4), "launches": 2, "bound": "hbm", "bytes": loss_bytes,
                                   "achieved": round(loss_bytes / loss_ms / 1e6, 1), "peak": round(hbm_peak, 1), "unit": "GB/s",
                                   "frac": round(loss_bytes / loss_ms / 1e6 / hbm_peak, 4),
                                   "what": f"gsx_fused_loss (L1 + SSIM, 11x11 window) fwd+bwd on one {W}x{H} view, loss left on the device"}

    # ---- next row, outside the timed region: Morton re-layout and one adaptive-density step at N ----
    if world == 1:
        try:
            names = ["xyz", "scale", "quaternion", "opacity", "band0"] + (["sh"] if gd.sh_stride else [])
            arrs = [getattr(gd, nm).view(N, -1) for nm in names]
            ev = lambda: torch.cuda.Event(enable_timing=True)
            m0, m1, m2 = ev(), ev(), ev()
            bbox = None
            for it in range(3):
                if it == 1:
                    m0.record()
                order = Hh.morton_order(gd.xyz, bbox=(3.0, 3.0, 3.0, -3.0, -3.0, -3.0), flags=1)
                if it == 1:
                    m1.record()
                moved = Hh.gather_rows_multi(arrs + arrs + arrs, order)     # parameters + both Adam moments: 3 x 59 floats
                if it == 1:
                    m2.record()
                del moved
            torch.cuda.synchronize()
            ms_order, ms_gather = m0.elapsed_time(m1), m1.elapsed_time(m2)
            gbytes = N * (3 * 2 * 4 * sum(a.shape[1] for a in arrs) + 4 * len(arrs) * 3)
            obytes = N * (12 + 12 + 9 * 24)          # codes + ids written, 63 bits = 8 onesweep passes x 24 B + histogram read
            next_rows["morton_relayout"] = {
                "ms": round(ms_order + ms_gather, 4), "order_ms": round(ms_order, 4), "gather_ms": round(ms_gather, 4), "bound": "hbm",
                "gather_bytes": gbytes, "gather_GBps": round(gbytes / ms_gather / 1e6, 1), "gather_frac": round(gbytes / ms_gather / 1e6 / hbm_peak, 4),
                "order_bytes": obytes, "order_GBps": round(obytes / ms_order / 1e6, 1), "launches": "codes 1 + sort 11 + gather 1",
                "what": "gsx_morton_order (63-bit codes, 8-pass onesweep) + gsx_gather_rows_multi over 18 arrays (59 floats x 3 per Gaussian) in one launch; reference: thrust sort_by_key + 17 gathers (trainer.cu:871-924)"}
            del order
            # one adaptive-density step: ~5 % pruned, ~5 % cloned, ~5 % split
            rngd = torch.Generator(device=dev); rngd.manual_seed(7)
            u = torch.rand(N, device=dev, generator=rngd)
            dur = torch.full((N,), 4, device=dev, dtype=torch.int32)
            acc = torch.where(u < 0.10, 8.0, 0.4).to(torch.float32)        # mean |uv grad| 2.0 above / 0.1 below the 0.5 threshold
            opac = torch.where((u > 0.10) & (u < 0.15), -9.0, 1.0).to(torch.float32)
            extent = float(torch.exp(gd.scale).amax().item()) * 100.0 / 1.5   # clone threshold splits the high-gradient 10 % roughly in half
            med = float(torch.exp(gd.scale).amax(1).median().item())
            extent = med * 100.0
            pd_ = dict(xyz=gd.xyz, scale=gd.scale, quaternion=gd.quaternion, opacity=opac, band0=gd.band0, sh=(gd.sh if gd.sh_stride else None))
            momd = {k_: (torch.zeros_like(v_), torch.zeros_like(v_)) for k_, v_ in pd_.items() if v_ is not None}
            d0, d1 = ev(), ev()
            cnt = None
            for it in range(3):
                if it == 1:
                    d0.record()
                outp, outm, cnt = Hh.densify(pd_, momd, acc, dur, 0.5, 0.005, extent, 1.6, max_gaussians=0, seed=it, l_max=l_max)
                if it == 1:
                    d1.record()
                del outp, outm
            torch.cuda.synchronize()
            dms = d0.elapsed_time(d1)
            fl = sum(a.shape[1] for a in arrs)
            dbytes = 4 * fl * 3 * (N - cnt["n_prune"] + cnt["n_out"]) + N * (4 + 12 + 8 + 1 + 1 + 8 + 18 * 9)   # rows read + written (params + 2 moments), statistics, plan arrays
            next_rows["densify_step"] = {
                "ms": round(dms, 4), "bound": "hbm", "bytes": int(dbytes), "GBps": round(dbytes / dms / 1e6, 1), "frac": round(dbytes / dms / 1e6 / hbm_peak, 4),
                "counts": cnt, "launches": 5,
                "what": "gsx_densify_plan + gsx_densify_apply (decide, 3-kernel scan, one move of 18 arrays, split geometry; includes the 12-byte host read of the counts); reference: adaptive_density_step, trainer.cu:518-775"}
            del momd, pd_
        except Exception as e:
            next_rows["density_rows_error"] = repr(e)[:300]

    # ---- next row: the same batch step replayed as ONE CUDA graph (gsx_batch_graph_capture) ----------
    if world == 1:
        try:
            def timed_pair(rr, vs, gg, gi_list, reps_g=10):
                ev = lambda: torch.cuda.Event(enable_timing=True)
                gstream = torch.cuda.Stream()
                h_ = rr.batch_graph_capture(vs, gg, grad_images=gi_list, stream=gstream)
                with torch.cuda.stream(gstream):
                    for _ in range(2):
                        rr.graph_launch(h_, gstream)
                    a_, b_ = ev(), ev()
                    a_.record(gstream)
                    for _ in range(reps_g):
                        rr.graph_launch(h_, gstream)
                    b_.record(gstream)
                torch.cuda.synchronize()
                ms_graph = a_.elapsed_time(b_) / reps_g
                gsx_h.Rasterizer.graph_destroy(h_)
                for _ in range(2):
                    rr.batch_step(vs, gg, gi_list, flags=OVER)
                a_, b_ = ev(), ev()
                a_.record()
                for _ in range(reps_g):
                    rr.batch_step(vs, gg, gi_list, flags=OVER)
                b_.record()
                torch.cuda.synchronize()
                return ms_graph, a_.elapsed_time(b_) / reps_g
            gsx_h = Hh
            mg, me = timed_pair(r, vstructs, gd, [gi_dev] * nv)
            # a small scene, where the ~20 launches per view are not hidden behind long kernels
            Ns, Ws, Hs = 200_000, 480, 270
            gsm = S.synthetic_gaussians(Ns, l_max=l_max, seed=0x3D65 + 9)
            gds = Hh.Gaussians(gsm["xyz"], gsm["scale"], gsm["quaternion"], gsm["opacity"], gsm["band0"], gsm["sh"], device=dev)
            gds.zero_grads()
            rs_ = Hh.Rasterizer(Ns, Ws, Hs, instance_capacity=16 * Ns)
            rs_.batch_reserve(N_VIEWS)
            cams_s = [S.orbit_camera(Ws, Hs, azimuth_deg=45.0 * k) for k in range(N_VIEWS)]
            vs_s = [Hh.make_view(c_["view"], c_["proj"], c_["campos"], c_["fx"], c_["fy"], Ws, Hs, near_thresh=S.NEAR_THRESH,
                                 cull_padding=S.CULL_PADDING, mh_dist=S.MH_DIST, background=0.0, l_max=l_max) for c_ in cams_s]
            for v_ in vs_s:
                assert not rs_.forward_checked(v_, gds)["overflow"]
            gis_s = torch.as_tensor(S.grad_image(Ws, Hs)).to(dev)
            mgs, mes = timed_pair(rs_, vs_s, gds, [gis_s] * N_VIEWS)
            rs_.close()
            next_rows["graph_replay"] = {
                "c4_batch_ms": {"graph": round(mg, 3), "eager": round(me, 3)},
                "small_scene_batch_ms": {"graph": round(mgs, 3), "eager": round(mes, 3), "scene": f"{Ns} Gaussians, {Ws}x{Hs}, 8 views"},
                "what": "one cudaGraphLaunch per batch step (gsx_batch_graph_capture: per-Gaussian forward, 8 x (bin + sort + composite + backward), per-Gaussian backward; side streams captured) vs the same step enqueued call by call"}
        except Exception as e:
            next_rows["graph_replay_error"] = repr(e)[:300]

    views_per_s = N_VIEWS * args.steps / (ms_total * 1e-3)
    if world > 1:
        transport = "nccl from C++ (gsx_dp_step)" if native else vp.ex.transport
        ga_, ar_ = nat.link_bytes_per_step() if native else vp.ex.link_bytes_per_step()
        how = "NCCL all-gather" if native else {"nccl": "NCCL all-gather", "multimem": "in-kernel multicast stores (multimem.st)", "p2p": "in-kernel P2P stores"}[vp.ex.transport]
        how2 = "in-switch multimem allreduce kernel" if (not native and vp.ex.transport == "multimem") else "NCCL allreduce"
        xbytes = (f"{how} of {ga_ / 1e6:.0f} MB per-view d rgb + {how2} of {ar_ / 1e6:.0f} MB geometry "
                  f"gradients (SH gradients rebuilt locally; plain allreduce would be {flat.numel() * 4 / 1e6:.0f} MB)")
    else:
        xbytes = "none (single GPU)"
    # depth sort: histogram, digit scan, onesweep x4, tie fix-up; scan (record mode: gather + block sums; else x3); emission;
    # tile sort: histogram, digit scan, onesweep x passes; ranges: init + suffix-min fix-up; compositing x2
    launches_per_view = 2 + 4 + 1 + (2 if records else 3) + 1 + 1 + passes + 2 + 2 + (0 if (fused_hist or emit_hist) else 1)
    out = {
        "metric": METRIC, "value": views_per_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "ms_per_view_per_gpu": ms_total / args.steps / len(my_views),
        "config": {"workload": f"{args.config}: synthetic {N} Gaussians, SH deg {l_max}, {W}x{H}, batch of {N_VIEWS} views "
                               f"sharded over {world} GPU(s); per-tile lists: {'contributing-only (GSX_LISTS_CONTRIBUTING)' if VF else 'reference-identical'}; "
                               f"exchange per step: {xbytes}",
                   "views_per_step": N_VIEWS, "l2": "inputs larger than L2 (1.4 GB parameters + 1.4 GB gradients per view)",
                   "visible_P": int(P), "touching_P": int(Pt), "instances_K": int(K), "pixel_entries_S": int(Sx), "contributing_Splus": int(Sp)},
        "clocks": clocks,
        "e2e": {"value": N_VIEWS * args.steps / dt_e2e, "unit": UNIT,
                "h2d_bytes_per_step": N_VIEWS * (W * H * 3 * 4 + 200), "d2h_bytes_per_step": N_VIEWS * W * H * 3 * 4},
        "gpu_launches": (launches_per_view * nv + 2) * args.steps * world,
        "exchange_ms": exch, "exchange_transport": (transport if world > 1 else None), "roofline": roof, "stages": stages, "fp32_peak_tflops_measured": round(fp32_peak, 2),
        "next_rows": next_rows,
    }
    if world == 1 and not args.no_cpu_baseline:
        # ---- CPU legs (outside every timed region): the oracle on one view = cpu_baseline, and -- since that run IS the
        # oracle's answer for view 0 of the benchmarked scene -- the full-size parity report of the CUDA path against it
        from oracle import parity
        cores = cpu_threads()
        gi0 = gi_host[0].numpy()
        t, _, _, ost, ogr = cpu_view_seconds(S, g, cams[0], W, H, l_max, gi0, keep=True)
        out["cpu_baseline"] = {"value": 1.0 / t, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"1 view (fwd+bwd) of the same scene through oracle/libgolden.so, {cores} OpenMP threads, {t:.1f} s"}
        try:
            v0 = mk(cams[0], 0)                                    # reference-identical lists whatever GSX_LISTS says
            r.forward(v0, gd)
            torch.cuda.synchronize()
            st0 = {k_: (t_.cpu().numpy() if hasattr(t_, "cpu") else t_) for k_, t_ in r.state().items() if k_ != "sorted_keys"}
            gd.zero_grads()
            r.backward(v0, gd, gi_dev)
            torch.cuda.synchronize()
            got = {k_: t_.cpu().numpy() for k_, t_ in gd.grads.items()}
            rep = parity.full_view_report(st0, got, ost, ogr, N, W, H, 0.0)
            try:
                parity.assert_report(rep)
                rep["ok"] = True
            except AssertionError as e:
                rep["ok"], rep["why"] = False, str(e)[:300]
            rep["what"] = ("view 0 of the benchmarked scene, CUDA path vs CPU oracle: lists bit-exact on identical per-Gaussian inputs, "
                           "image / T 1e-4 with every outlier an explained threshold flip, gradients 1e-3 norm-wise (oracle/parity.py)")
            out["parity_" + args.config.lower()] = rep
            del st0, got
        except Exception as e:                                     # the bench line must survive a failing check
            out["parity_" + args.config.lower()] = {"ok": False, "why": repr(e)[:300]}
        del ost, ogr
        # ---- same-box baseline B1: the reference's own CUDA kernels on the same view (oracle/_ref, outside the timed region)
        try:
            from oracle import ref_cuda as RC
            if RC.available():
                td = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).to(dev)
                gref = dict(xyz=gd.xyz, scale=gd.scale, quaternion=gd.quaternion, opacity=gd.opacity, band0=gd.band0,
                            sh=(gd.sh.view(N, -1, 3) if gd.sh_stride else None))
                out["ref_cuda_baseline"] = RC.replay_view(gref, cams[0], W, H, l_max, S.NEAR_THRESH, S.CULL_PADDING, S.MH_DIST,
                                                          gi_dev, reps=3)
                b1 = out["ref_cuda_baseline"]
                if "ms_per_view" in b1:
                    b1["speedup_of_this_repo"] = round(b1["ms_per_view"] / (ms_total / args.steps / N_VIEWS), 2)
            else:
                out["ref_cuda_baseline"] = {"unavailable": "oracle/_ref/libref3dgs.so not built (needs /root/reference at build time)"}
        except Exception as e:
            out["ref_cuda_baseline"] = {"error": repr(e)[:300]}
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
