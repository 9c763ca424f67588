"""View-parallel data parallelism for the splatting hot path: one process per GPU, parameters replicated,
the views of a batch sharded over the ranks, per-Gaussian parameter gradients summed over all views of the batch.

The exchange is NOT a plain allreduce of the 236 B/Gaussian gradient block.  A view's SH coefficient gradient is rank 1,
d sh[k][c] = d rgb_c * Y_k(dir_view) (reference cuda/spherical_harmonics_backward.cu:100-104), so
  * every rank all-gathers the 12 B/Gaussian d rgb of each of its views (NCCL all-gather, issued as soon as the views'
    compositing backward is done, overlapping the per-Gaussian backward),
  * only the 44 B/Gaussian geometry gradients (xyz, scale, quaternion, opacity) are allreduced (NCCL, NVLink),
  * every rank rebuilds the batch-summed band-0 + SH gradients (192 B/Gaussian) from all views' d rgb locally
    (gsx_sh_grads_from_view_rgb), overlapping the allreduce.
Link bytes per Gaussian and step: 44 allreduced + 12 x views gathered, instead of 236 allreduced.

Two transports for the same exchange:
  * "nccl": NCCL all-gather + NCCL allreduce on a high-priority side stream (works everywhere);
  * "multimem" / "p2p" (GradExchange(..., symmetric=True), used when torch symmetric memory can map the buffers): the
    extraction kernel stores each view's d rgb straight into every rank's gather buffer through the NVSwitch multicast
    address (multimem.st) or per-peer NVLink stores -- the all-gather is fused into the producing kernel and overlaps
    the next views' compositing -- and the geometry gradients are summed in the switch by a hand-written NVLS kernel
    (multimem.ld_reduce + multimem.st, gsx_allreduce_multimem_f32).  Cross-rank ordering = symmetric-memory barriers.

`shard_views` and `GradExchange` are device-agnostic (the world-size-2 gloo tests drive them with CPU tensors);
`ViewParallel` is the GPU step on top of host.Rasterizer's batch path.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _ffi
from ._ffi import check, lib

GEOM = (("xyz", 3), ("scale", 3), ("quaternion", 4), ("opacity", 1))   # allreduced arrays, floats per Gaussian


def shard_views(n_views, world, rank):
    """Blocked assignment: rank r owns views [r*vpr, (r+1)*vpr).  The batch must divide evenly (8 views on 1/2/4/8 GPUs)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    if n_views % world:
        raise ValueError(f"{n_views} views do not shard evenly over {world} ranks")
    vpr = n_views // world
    return list(range(rank * vpr, (rank + 1) * vpr))


class GradExchange:
    """Buffers + collectives of one batch step.  drgb[v] is the dense d rgb of global view v ([N][3]); geom holds the
    allreduced arrays as views of one flat buffer (a single collective)."""

    def __init__(self, N, n_views, world, rank, device, group=None, symmetric=False, interleaved=False):
        """interleaved: buffer slot of (rank r, local view k) is k*world + r instead of the global view index r*vpr + k, so
        that the k-th views of all ranks are contiguous and can be all-gathered on their own (all_gather_view) as soon
        as they exist, under the compute of the following views.  slot_of_view maps global view -> slot."""
        self.N, self.n_views, self.world, self.rank, self.group = N, n_views, world, rank, group
        self.mine = shard_views(n_views, world, rank)
        self.interleaved = bool(interleaved) and world > 1
        vpr = n_views // world
        self.vpr = vpr
        slot = (lambda r, k: k * world + r) if self.interleaved else (lambda r, k: r * vpr + k)
        self.slots_mine = [slot(rank, k) for k in range(vpr)]
        self.slot_of_view = [slot(v // vpr, v % vpr) for v in range(n_views)]
        offs, off = {}, 0
        for nm, w in GEOM:
            offs[nm] = off
            off += (N * w + 3) // 4 * 4                     # 16-byte aligned segments
        n_drgb = (n_views * N * 3 + 3) // 4 * 4
        self.transport, self.hdl, self.why_not = "nccl", None, None
        base = None
        if symmetric and world > 1:
            try:                                            # one symmetric allocation: [drgb | geom_flat]
                import torch.distributed._symmetric_memory as symm_mem
                base = symm_mem.empty(n_drgb + off, dtype=torch.float32, device=device)
                self.hdl = symm_mem.rendezvous(base, group if group is not None else dist.group.WORLD)
                base.zero_()
                self.transport = "multimem" if (self.hdl.has_multicast_support and self.hdl.multicast_ptr) else "p2p"
            except Exception as e:                          # no symmetric memory on this system: NCCL transport
                self.why_not, self.hdl, base = repr(e)[:200], None, None
        if base is None:
            base = torch.zeros(n_drgb + off, device=device, dtype=torch.float32)
        self.base, self.n_drgb = base, n_drgb
        self.drgb = base[: n_views * N * 3].view(n_views, N, 3)
        self.geom_flat = base[n_drgb: n_drgb + off]
        self.geom = {}
        for nm, w in GEOM:
            v = self.geom_flat[offs[nm]: offs[nm] + N * w]
            self.geom[nm] = v.view(N, w) if w > 1 else v

    # ---- symmetric-memory transport ---------------------------------------------------------------------------
    def barrier(self, channel=0):
        """cross-rank barrier on the current stream (device side, symmetric-memory signal pads)"""
        self.hdl.barrier(channel=channel)

    def view_targets(self, k):
        """(multicast pointer or 0, peer pointers) of the [N][3] block of this rank's k-th view in every rank's buffer"""
        byte_off = self.slots_mine[k] * self.N * 3 * 4
        if self.transport == "multimem":
            return int(self.hdl.multicast_ptr) + byte_off, []
        return 0, [int(p) + byte_off for p in self.hdl.buffer_ptrs]

    def all_reduce_geom_multimem(self):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib().gsx_allreduce_multimem_f32(C.c_void_p(int(self.hdl.multicast_ptr) + self.n_drgb * 4),
                                               C.c_size_t(self.geom_flat.numel()), int(self.rank), int(self.world), st))

    def local_drgb(self, k):
        """buffer of this rank's k-th view"""
        return self.drgb[self.slots_mine[k]]

    def in_slot_order(self, per_view):
        """a per-global-view list (camera positions) re-ordered to the buffer's slot order"""
        out = [None] * self.n_views
        for v, x in enumerate(per_view):
            out[self.slot_of_view[v]] = x
        return out

    def all_gather_view(self, k):
        """all-gather of the ranks' k-th views alone (interleaved layout: one contiguous [world][N][3] block)"""
        assert self.interleaved
        blk = self.drgb[k * self.world:(k + 1) * self.world]
        try:
            dist.all_gather_into_tensor(blk.view(-1), blk[self.rank].view(-1), group=self.group)
        except (RuntimeError, NotImplementedError):         # backends without the flat variant
            parts = [blk[r].view(-1) for r in range(self.world)]
            dist.all_gather(parts, blk[self.rank].view(-1).clone(), group=self.group)

    def link_bytes_per_step(self):
        """payload bytes: (all-gathered, allreduced)"""
        return self.drgb.numel() * 4, self.geom_flat.numel() * 4

    def all_gather_drgb(self):
        if self.world == 1:
            return
        if self.interleaved:
            for k in range(self.vpr):
                self.all_gather_view(k)
            return
        flat = self.drgb.view(self.world, -1)               # blocked sharding: rank r's views are row r
        try:
            dist.all_gather_into_tensor(self.drgb.view(-1), flat[self.rank], group=self.group)
        except (RuntimeError, NotImplementedError):         # backends without the flat variant
            parts = [flat[r] for r in range(self.world)]
            dist.all_gather(parts, flat[self.rank].clone(), group=self.group)

    def all_reduce_geom(self):
        if self.world > 1:
            dist.all_reduce(self.geom_flat, group=self.group)


def sh_grads_from_view_rgb(xyz, campos_all, drgb, l_max, sh_grad, band0_grad, sh_stride, flags=_ffi.GSX_GRAD_OVERWRITE):
    """band-0 + SH coefficient gradients of the batch from every view's d rgb (C ABI: gsx_sh_grads_from_view_rgb)."""
    n_views, N = drgb.shape[0], drgb.shape[1]
    cp = (C.c_float * (3 * n_views))(*[float(x) for c in campos_all for x in c])
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib().gsx_sh_grads_from_view_rgb(
        C.c_void_p(xyz.data_ptr()), int(N), int(l_max), cp, int(n_views), C.c_void_p(drgb.data_ptr()),
        C.c_size_t(N * 3), C.c_void_p(sh_grad.data_ptr() if l_max > 0 else 0), int(sh_stride),
        C.c_void_p(band0_grad.data_ptr()), int(flags), st))


class DpNative:
    """The multi-GPU step behind the C ABI (gsx_dp_*, csrc/dp.cu): NCCL communicator, exchange buffers, streams and the
    whole step live in libgsx.so; this class only hands over pointers.  The 128-byte NCCL id travels from rank 0 to the
    others through torch.distributed (any backend) -- plumbing, as for an MPI or socket bootstrap in a C++ caller."""

    def __init__(self, rasterizer, gaussians, views_all, l_max, world, rank, group=None):
        import ctypes as C_
        self.r, self.g, self.world, self.rank = rasterizer, gaussians, world, rank
        self.views = (_ffi.GsxView * len(views_all))(*views_all)
        self.nv_local = len(views_all) // world
        dev = gaussians.xyz.device
        idb = torch.zeros(_ffi.GSX_DP_ID_BYTES, dtype=torch.uint8)
        if rank == 0:
            buf = (C_.c_ubyte * _ffi.GSX_DP_ID_BYTES)()
            check(lib().gsx_dp_unique_id(buf))
            idb = torch.tensor(list(buf), dtype=torch.uint8)
        if world > 1:
            t = idb.to(dev) if dist.get_backend(group) == "nccl" else idb
            dist.broadcast(t, src=0, group=group)
            idb = t.cpu()
        raw = (C_.c_ubyte * _ffi.GSX_DP_ID_BYTES)(*idb.tolist())
        self.h = C_.c_void_p()
        check(lib().gsx_dp_init(C_.byref(self.h), rasterizer.h, raw, int(rank), int(world), int(gaussians.N), len(views_all),
                                int(l_max), int(gaussians.sh_stride)))
        g = _ffi.GsxGrads()
        check(lib().gsx_dp_grads(self.h, C_.byref(g)))
        N = gaussians.N
        from .host import _from_ptr
        mk = lambda ptr, shape: _from_ptr(ptr, shape, torch.float32)
        self.grads = dict(xyz=mk(g.xyz, (N, 3)), scale=mk(g.scale, (N, 3)), quaternion=mk(g.quaternion, (N, 4)),
                          opacity=mk(g.opacity, (N,)), band0=mk(g.band0, (N, 3)),
                          sh=mk(g.sh, (N, max(gaussians.sh_stride, 1))) if gaussians.sh_stride else torch.zeros((N, 1), device=dev))
        gaussians.grads = self.grads                       # zero-copy views of the buffers the step writes

    def close(self):
        if self.h:
            lib().gsx_dp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, grad_images, adam=None):
        """grad_images: device tensors of this rank's views (one, or a list)"""
        gi = grad_images if isinstance(grad_images, (list, tuple)) else [grad_images] * self.nv_local
        arr = (C.c_void_p * self.nv_local)(*[t.data_ptr() for t in gi])
        p = self.g.params()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib().gsx_dp_step(self.h, self.views, C.byref(p), arr, C.byref(adam) if adam is not None else None, st))

    def render_host(self, grad_images_host, images_host, adam=None, sync=True):
        """pinned host loss-gradient images in, pinned host rendered images out.  sync=False: everything is enqueued and the
        current stream joins the downloads (gsx_dp_render_host_async): consecutive steps pipeline, the caller synchronises
        the stream once and leaves the host buffers alone until then."""
        gi = (C.c_void_p * self.nv_local)(*[t.data_ptr() for t in grad_images_host])
        io = (C.c_void_p * self.nv_local)(*[t.data_ptr() for t in images_host])
        p = self.g.params()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        fn = lib().gsx_dp_render_host if sync else lib().gsx_dp_render_host_async
        check(fn(self.h, self.views, C.byref(p), gi, io, C.byref(adam) if adam is not None else None, st))

    def allreduce(self, buf):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib().gsx_dp_allreduce(self.h, C.c_void_p(buf.data_ptr()), C.c_size_t(buf.numel()), st))

    def enable_timing(self, on=True):
        check(lib().gsx_dp_enable_timing(self.h, int(bool(on))))

    def phase_ms(self):
        ms = (C.c_float * 6)()
        check(lib().gsx_dp_timing(self.h, ms))
        return dict(zip(("views", "gather", "params_bwd", "reduce", "sh_rebuild", "total"), [float(x) for x in ms]))

    def link_bytes_per_step(self):
        N = self.g.N
        return len(self.views) * N * 3 * 4, 11 * N * 4

    def adam_epilogue(self, params, moments, lrs, b1, b2, eps, bias1, bias2):
        """gsx_adam_epilogue_t for step(): params / moments dicts keyed xyz, scale, quaternion, opacity, band0, sh"""
        a = _ffi.GsxAdamEpilogue()
        for k, nm in enumerate(("xyz", "scale", "quaternion", "opacity", "band0", "sh")):
            t = params.get(nm)
            setattr(a.params, nm, t.data_ptr() if t is not None else None)
            mv = moments.get(nm)
            a.exp_avg[k] = mv[0].data_ptr() if mv is not None else None
            a.exp_avg_sq[k] = mv[1].data_ptr() if mv is not None else None
            a.lr[k] = float(lrs.get(nm, 0.0))
        a.b1, a.b2, a.eps, a.bias1, a.bias2 = float(b1), float(b2), float(eps), float(bias1), float(bias2)
        return a


class ViewParallel:
    """One batch step on this rank's GPU: forward + backward of the rank's views, then the exchange above.
    `views_all` / `campos_all` describe ALL views of the batch (every rank needs every camera position)."""

    def __init__(self, rasterizer, gaussians, views_all, campos_all, l_max, world=1, rank=0, group=None, symmetric=True):
        self.r, self.g, self.l_max, self.world, self.rank = rasterizer, gaussians, l_max, world, rank
        self.views_all, self.campos_all = views_all, [tuple(float(x) for x in c) for c in campos_all]
        dev = gaussians.xyz.device
        self.mine = shard_views(len(views_all), world, rank)
        self.views = [views_all[k] for k in self.mine]
        rasterizer.batch_reserve(len(self.views))
        N = gaussians.N
        if gaussians.grads is None:
            gaussians.zero_grads()
        import os
        # symmetric transport: fuse the peer stores into the extraction kernel (1) or push from a side stream (0)
        self.fused_extract = os.environ.get("GSX_DP_FUSED", "1") == "1"
        # d rgb gather: "nccl" (default) = one NCCL all-gather per local view on the side stream, issued as soon as the
        # view's backward is done (interleaved buffer layout), under the compute of the following views; "fused" = peer
        # stores from the extraction kernel.  Measured on B200 (C4, v12 kernels): nccl 771 / 1323 / 1892 views/s at
        # 2 / 4 / 8 GPUs, fused 743 / 1248 / 1614 -- the stores of the fused kernel back-pressure the compositing.
        self.nccl_gather = os.environ.get("GSX_DP_GATHER", "nccl") == "nccl"
        if world > 1:
            self.ex = GradExchange(N, len(views_all), world, rank, dev, group, symmetric=symmetric,
                                   interleaved=self.nccl_gather)
            self.campos_buf = self.ex.in_slot_order(self.campos_all)
            for nm, _ in GEOM:                              # the geometry gradients live in the flat allreduce buffer
                gaussians.grads[nm] = self.ex.geom[nm]
            self.comm = torch.cuda.Stream(device=dev, priority=-1)
            self.ev_views, self.ev_geom = torch.cuda.Event(), torch.cuda.Event()
            self.ev_gathered, self.ev_reduced = torch.cuda.Event(), torch.cuda.Event()
        else:
            self.ex = None
        self.timing = None          # set to {} to collect per-phase CUDA-event times (ms) of the next step()

    def _mark(self, name, stream):
        if self.timing is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            self.timing.setdefault("_ev", {})[name] = e

    def phase_ms(self):
        """after a synchronize: durations of the phases of the last timed step"""
        ev = self.timing["_ev"]
        d = lambda a, b: ev[a].elapsed_time(ev[b])
        return dict(views=d("t0", "views"), gather=d("g0", "g1"), params_bwd=d("views", "bwd"), reduce=d("r0", "r1"),
                    sh_rebuild=d("s0", "s1"), total=d("t0", "end"))

    def step(self, grad_images, on_image=None):
        r, g = self.r, self.g
        OVER, NO_SH = _ffi.GSX_GRAD_OVERWRITE, _ffi.GSX_GRAD_NO_SH
        if self.world == 1:
            r.batch_step(self.views, g, grad_images, flags=OVER, on_image=on_image)
            return
        need = r.overflow()                                 # flag raised by an earlier step, read without a sync
        if need:
            raise _ffi.GsxError(_ffi.GSX_E_CAPACITY, f"instance capacity exceeded in a previous step: K={need}; reserve() and re-run it")
        if self.ex.transport != "nccl":
            return self._step_symmetric(grad_images, on_image)
        main = torch.cuda.current_stream()
        self._mark("t0", main)
        r.batch_forward_params(self.views, g)
        for k in range(len(self.views)):
            r.batch_forward_view(k)
            if on_image is not None:
                on_image(k)
            gi = grad_images[k] if isinstance(grad_images, (list, tuple)) else grad_images
            r.batch_backward_view(k, gi)
            r.batch_extract_view_rgb_grad(k, self.ex.local_drgb(k))
            if self.ex.interleaved:
                self._gather_view(k, main)
        self.ev_views.record(main)
        self._mark("views", main)
        with torch.cuda.stream(self.comm):                  # all-gather of d rgb overlaps the per-Gaussian backward
            self.comm.wait_event(self.ev_views)
            if not self.ex.interleaved:
                self._mark("g0", self.comm)
                self.ex.all_gather_drgb()
            self._mark("g1", self.comm)
            self.ev_gathered.record(self.comm)
        r.batch_backward_params(g, OVER | NO_SH)            # geometry gradients -> flat buffer
        self.ev_geom.record(main)
        self._mark("bwd", main)
        with torch.cuda.stream(self.comm):                  # allreduce overlaps the SH rebuild
            self.comm.wait_event(self.ev_geom)
            self._mark("r0", self.comm)
            self.ex.all_reduce_geom()
            self._mark("r1", self.comm)
            self.ev_reduced.record(self.comm)
        main.wait_event(self.ev_gathered)
        self._mark("s0", main)
        sh_grads_from_view_rgb(g.xyz, self.campos_buf, self.ex.drgb, self.l_max, g.grads["sh"], g.grads["band0"],
                               g.sh_stride, OVER)
        self._mark("s1", main)
        main.wait_event(self.ev_reduced)
        self._mark("end", main)

    def _gather_view(self, k, main):
        """side stream: all-gather of the ranks' k-th views, after this rank's extraction of view k"""
        self.ev_views.record(main)
        with torch.cuda.stream(self.comm):
            self.comm.wait_event(self.ev_views)
            if k == 0:
                self._mark("g0", self.comm)
            self.ex.all_gather_view(k)

    def _step_symmetric(self, grad_images, on_image):
        """Same exchange with the all-gather fused into the extraction kernel (stores into every rank's buffer over
        NVLink) and, with multicast, the geometry allreduce done in the switch by our own kernel."""
        r, g, ex = self.r, self.g, self.ex
        OVER, NO_SH = _ffi.GSX_GRAD_OVERWRITE, _ffi.GSX_GRAD_NO_SH
        main = torch.cuda.current_stream()
        st = C.c_void_p(main.cuda_stream)
        self._mark("t0", main)
        ex.barrier(0)                                       # every rank is done reading last step's buffers
        r.batch_forward_params(self.views, g)
        for k in range(len(self.views)):
            r.batch_forward_view(k)
            if on_image is not None:
                on_image(k)
            gi = grad_images[k] if isinstance(grad_images, (list, tuple)) else grad_images
            r.batch_backward_view(k, gi)
            mc, peers = ex.view_targets(k)
            arr = (C.c_void_p * max(len(peers), 1))(*peers) if peers else None
            if self.nccl_gather:
                r.batch_extract_view_rgb_grad(k, ex.local_drgb(k))
                self._gather_view(k, main)
            elif self.fused_extract:
                # extraction + all-gather in ONE kernel: rows go from the packed gradients straight to every rank
                check(lib().gsx_batch_extract_view_rgb_grad_peers(r.h, int(k), C.c_void_p(mc), arr, len(peers), st))
            else:
                # local extraction on the main stream, peer stores on the side stream: with many ranks pushing at once the
                # stores back-pressure (every rank receives (g-1) x 72 MB) and must not hold up the per-Gaussian backward
                r.batch_extract_view_rgb_grad(k, ex.local_drgb(k))
                self.ev_views.record(main)
                with torch.cuda.stream(self.comm):
                    self.comm.wait_event(self.ev_views)
                    if k == 0:
                        self._mark("g0", self.comm)
                    check(lib().gsx_push_view_rgb_peers(C.c_void_p(ex.local_drgb(k).data_ptr()), int(ex.N), C.c_void_p(mc), arr,
                                                        len(peers), C.c_void_p(self.comm.cuda_stream)))
        self._mark("views", main)
        if self.nccl_gather:
            with torch.cuda.stream(self.comm):
                self._mark("g1", self.comm)
                self.ev_gathered.record(self.comm)
        elif self.fused_extract:
            self._mark("g0", main); self._mark("g1", main)  # the gather has no phase of its own
        else:
            self._mark("g1", self.comm)
            self.ev_gathered.record(self.comm)
        r.batch_backward_params(g, OVER | NO_SH)            # geometry gradients -> this rank's part of the symmetric buffer
        self._mark("bwd", main)
        if self.nccl_gather or not self.fused_extract:
            main.wait_event(self.ev_gathered)
        ex.barrier(1)                                       # all d rgb stores and all geometry gradients have landed
        self.ev_geom.record(main)
        with torch.cuda.stream(self.comm):                  # allreduce overlaps the SH rebuild
            self.comm.wait_event(self.ev_geom)
            self._mark("r0", self.comm)
            if ex.transport == "multimem":
                ex.all_reduce_geom_multimem()
            else:
                ex.all_reduce_geom()
            self._mark("r1", self.comm)
            self.ev_reduced.record(self.comm)
        self._mark("s0", main)
        sh_grads_from_view_rgb(g.xyz, self.campos_buf, ex.drgb, self.l_max, g.grads["sh"], g.grads["band0"],
                               g.sh_stride, OVER)
        self._mark("s1", main)
        main.wait_event(self.ev_reduced)
        if ex.transport == "multimem":
            ex.barrier(2)                                   # every rank's slice of the sums has been broadcast
        self._mark("end", main)
