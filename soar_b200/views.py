"""View-sharded data parallelism for the render loop.

The reference is single-GPU: its view loop is a plain python loop over independent
`DiffGaussian.forward` calls (TS/renderer/gaussian_batch_renderer.py:273-323, gt views
:105-181). Each view only needs the (replicated) Gaussian parameters, so the loop shards
across ranks with no data-path exchange; the only collective is ONE sum-allreduce of the
Gaussian-parameter gradients per step, over a single flat fp32 buffer that the leaves'
`.grad` tensors alias (autograd accumulates into it in place, so there is no gather
copy before the collective). One process per GPU, torch.distributed (NCCL over
NVLink/NVSwitch on the B200 box; gloo in the CPU tests).
"""
from typing import Dict, List, Sequence

import torch
import torch.distributed as dist


def shard_views(n_views: int, rank: int, world_size: int, costs: Sequence[float] = None) -> List[int]:
    """Indices of the views rank `rank` renders. Without costs: round-robin. With per-view
    cost estimates (e.g. resolution x last frame's num_rendered): longest-processing-time
    greedy assignment, so the slowest rank finishes as early as possible."""
    if world_size <= 1:
        return list(range(n_views))
    if costs is None:
        return list(range(rank, n_views, world_size))
    order = sorted(range(n_views), key=lambda v: (-float(costs[v]), v))
    load = [0.0] * world_size
    mine = []
    for v in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        load[r] += float(costs[v])
        if r == rank:
            mine.append(v)
    return sorted(mine)


def iter_views(rasterizers, calls, streams=None):
    """Render several views of the same Gaussians back to back without idling the device.

    `rasterizers`: GaussianRasterizer modules (one per view); `calls`: one dict of forward()
    keyword arguments per view. The first phase of EVERY view (preprocess, depth sort, tile
    counts) is enqueued up front (GaussianRasterizer.begin), then the views are finished and
    yielded one at a time — `for out in iter_views(...): loss(out).backward()` keeps only one
    view's images and binning buffers alive. The reference's loop
    (TS/renderer/gaussian_batch_renderer.py:273-323) blocks the host once per view on
    num_rendered (rasterizer_impl.cu:250) with nothing queued behind it; here only the first
    view of a step waits. `streams`: see iter_view_pairs."""
    return iter_view_pairs(zip(rasterizers, calls), streams)


def view_streams(n, device=None):
    """n CUDA streams for iter_views(..., streams=...). Make them once and reuse them."""
    return [torch.cuda.Stream(device=device) for _ in range(n)]


def iter_view_pairs(pairs, streams=None):
    """iter_views over an iterable of (rasterizer, forward-kwargs) pairs. The iterable is consumed
    lazily: with a generator, view k's settings and tensors are built while the device already runs
    phase 1 of view k-1, so the first kernels of a step are enqueued as early as possible.

    `streams` (a list from view_streams): view k runs on streams[k % len(streams)], so the views of
    a step overlap on the device. The views are independent (the reference's loop is serial only
    because python is): a tile with a long list is one CTA's serial walk, and on a small image most
    of the 148 SMs idle behind the few longest tiles unless other views fill them; the sort's
    latency-bound passes of one view likewise hide behind another view's blend kernels. While the
    loop body of a view runs, torch's CURRENT stream is that view's stream: the loss and
    `backward()` issued in the body are enqueued there. When the generator is exhausted (or closed)
    the stream that was current at the start waits for all of them. Under FlatGrads.direct() the
    in-place additions of the views stay in view order (and the first one may store), see
    soar_rast_backward_ordered. Tensors made in the body belong to the view's stream; anything kept
    past the loop is safe to read afterwards."""
    if not streams:
        started = []
        for r, kw in pairs:
            r.begin(**kw)
            started.append((r, kw))
        for r, kw in started:
            yield r(**kw)
        return
    main = torch.cuda.current_stream()
    fork = torch.cuda.Event()
    fork.record(main)
    if _ACTIVE is not None:
        _ACTIVE.ordered = True
    started, used = [], []
    try:
        for k, (r, kw) in enumerate(pairs):
            s = streams[k % len(streams)]
            if k < len(streams):
                s.wait_event(fork)   # parameters written on the caller's stream are visible
                used.append(s)
            with torch.cuda.stream(s):
                r.begin(**kw)
            started.append((r, kw, s))
        for r, kw, s in started:
            with torch.cuda.stream(s):
                yield r(**kw)
    finally:
        for s in used:
            main.wait_stream(s)


# data_ptr of a registered leaf parameter -> (shape, its slice of a FlatGrads buffer). While a
# FlatGrads.direct() context is active, the rasterizer backward ADDS the gradients of registered
# leaves straight into these slices (SoarRastGrads.accumulate_mask) and hands autograd `None` for
# them: no per-view gradient tensors, no AccumulateGrad add kernels.
_DIRECT = {}
# the active FlatGrads.direct() context (None outside one)
_ACTIVE = None


def active_direct():
    """The FlatGrads.direct() context a forward pass runs under (kept in its autograd ctx)."""
    return _ACTIVE


def direct_sink(t):
    """The accumulation buffer registered for tensor `t`, or None."""
    if not _DIRECT or t is None or t.numel() == 0 or not t.requires_grad:
        return None
    hit = _DIRECT.get(t.data_ptr())
    if hit is None or hit[0] != tuple(t.shape) or not t.is_contiguous():
        return None
    return hit[1]


_NCCL_POOLS = []   # (backend, pool) of every registered flat buffer, dropped before the group goes away


def release_registered_buffers():
    """Deregisters the NCCL-registered flat buffers from their communicator. Must run before
    `dist.destroy_process_group()` (the pool's destructor would otherwise reach for an aborted
    communicator at interpreter exit); FlatGrads installs a wrapper around destroy_process_group
    that does it, so callers only need this when they tear NCCL down some other way."""
    while _NCCL_POOLS:
        backend, pool = _NCCL_POOLS.pop()
        try:
            backend.deregister_mem_pool(pool)
        except Exception:  # noqa: BLE001  (already gone with its communicator)
            pass


def _wrap_destroy_once():
    if getattr(dist.destroy_process_group, "_soar_wrapped", False):
        return
    inner = dist.destroy_process_group

    def destroy_process_group(*a, **kw):
        release_registered_buffers()
        return inner(*a, **kw)

    destroy_process_group._soar_wrapped = True
    dist.destroy_process_group = destroy_process_group


def _nccl_registered_zeros(n, dev):
    """The flat gradient buffer -> (tensor, pool or None, description). Under an NCCL process group of
    more than one rank it is carved from NCCL's own allocator (ncclMemAlloc) and registered with the
    communicator once as a symmetric window, so the per-step all-reduce runs on a user-registered
    buffer instead of staging through NCCL's internal ones: 118 MB over 8 B200s 0.378 -> 0.277 ms
    (profiles/r03c_*); with fewer than 4 ranks the plain allocation is faster and is kept.
    SOAR_FLAT_NCCL_POOL=0 forces a plain torch allocation, =1 registration without the symmetric
    window (0.341 ms at 8 ranks), =2 the window. Any failure of the (version-dependent) torch hooks falls back to
    the plain allocation with a warning: this is an allocation strategy, not a compute path."""
    import os
    import warnings
    mode = os.environ.get("SOAR_FLAT_NCCL_POOL", "auto")
    plain = lambda: torch.zeros(n, dtype=torch.float32, device=dev)
    if (mode == "0" or dev.type != "cuda" or not (dist.is_available() and dist.is_initialized())
            or dist.get_world_size() == 1 or dist.get_backend() != "nccl"):
        return plain(), None, "plain torch allocation"
    if mode == "auto":
        # measured on B200 / NVSwitch, 118 MB (profiles/r03c..e): plain / symmetric = 0.265 / 0.322 ms at
        # 2 ranks, 0.324 / 0.298 at 4, 0.378 / 0.277 at 8 — the window pays from 4 ranks on
        if dist.get_world_size() < 4:
            return plain(), None, "plain torch allocation (fewer than 4 ranks)"
        mode = "2"
    try:
        backend = dist.group.WORLD._get_backend(torch.device(dev))
        pool = torch.cuda.MemPool(backend.mem_allocator)
        with torch.cuda.use_mem_pool(pool, device=dev):
            flat = torch.zeros(n, dtype=torch.float32, device=dev)
        backend.register_mem_pool(pool, symm=(mode != "1"))
        _NCCL_POOLS.append((backend, pool))
        _wrap_destroy_once()
        return flat, pool, ("ncclMemAlloc pool registered with the communicator"
                            + (" as a symmetric window" if mode != "1" else ""))
    except Exception as e:  # noqa: BLE001
        warnings.warn(f"FlatGrads: NCCL-registered allocation unavailable ({type(e).__name__}: {e}); "
                      "using a plain allocation")
        return plain(), None, "plain torch allocation (registration failed)"


class FlatGrads:
    """One contiguous fp32 buffer holding the gradients of all Gaussian parameters.

    `attach(params)` points every leaf's `.grad` at its slice of the buffer; backward
    passes of any number of local views then accumulate straight into it.
    `allreduce()` sums the buffer over ranks with a single collective."""

    def __init__(self, params: Dict[str, torch.Tensor]):
        self.names = list(params.keys())
        self.params = params
        sizes = [params[k].numel() for k in self.names]
        dev = params[self.names[0]].device
        # every slice starts on a 16-byte boundary (the kernels move [P,4] and SH rows as float4:
        # a slice at an odd float offset would force their scalar paths); the pad floats stay zero
        self.slices = {}
        off = 0
        for k, n in zip(self.names, sizes):
            self.slices[k] = (off, off + n)
            off += (n + 3) // 4 * 4
        self.flat, self.nccl_pool, self.buffer_kind = _nccl_registered_zeros(off, dev)
        self._events = []   # pool for _Direct.order_events
        self.attach()

    def attach(self):
        for k in self.names:
            a, b = self.slices[k]
            self.params[k].grad = self.flat[a:b].view_as(self.params[k])

    def zero_(self):
        self.flat.zero_()

    def direct(self, overwrite=False):
        """Context manager: inside it, rasterizer backward passes accumulate the gradients of
        these parameters in place into the flat buffer (single stream). Gradients of anything else
        flow through autograd as usual. overwrite=False: the buffer must have been zeroed for the
        step. overwrite=True: the FIRST backward inside the context stores (every element of every
        registered parameter that takes part in it, zeros for culled Gaussians) and the following
        ones add — no zero-fill, and the first view's read-modify-write traffic is gone. Parameters
        that the first backward does not touch keep their old content: zero them yourself."""
        return _Direct(self, overwrite)

    def grad(self, name: str) -> torch.Tensor:
        a, b = self.slices[name]
        return self.flat[a:b].view_as(self.params[name])

    def nbytes(self) -> int:
        return self.flat.numel() * 4

    def allreduce(self, async_op: bool = False):
        """Sum over all ranks (no-op for a single process). Returns the work handle when async."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, async_op=async_op)


class _Direct:
    def __init__(self, fg, overwrite=False):
        self.fg = fg
        self.keys = []
        self.overwrite = overwrite
        self.stored = False
        # views on several streams (iter_view_pairs(streams=...)): the launches that add into the flat
        # buffer are chained through events, so the order of the additions is the order of the
        # backward calls whatever stream they run on
        self.ordered = False
        self.last_accumulated = None

    def order_events(self):
        """(event the next accumulation waits for or None, event it records), or None when the views
        share one stream."""
        if not self.ordered:
            return None
        pool = self.fg._events
        if self._n_events == len(pool):
            e = torch.cuda.Event()
            e.record()            # torch creates the CUDA event lazily: force it, the library re-records it
            pool.append(e)
        done = pool[self._n_events]
        self._n_events += 1
        after, self.last_accumulated = self.last_accumulated, done
        return after, done

    def first_store(self):
        """True exactly once per `direct(overwrite=True)` context: the first rasterizer backward of
        the views rendered under it STORES into the sinks (no zero-fill needed), later ones add."""
        if not self.overwrite or self.stored:
            return False
        self.stored = True
        return True

    def __enter__(self):
        global _ACTIVE
        if _ACTIVE is not None:
            raise RuntimeError("FlatGrads.direct() contexts do not nest")
        _ACTIVE = self
        self._n_events = 0
        for k in self.fg.names:
            p = self.fg.params[k]
            if p.is_cuda and p.is_contiguous():
                _DIRECT[p.data_ptr()] = (tuple(p.shape), self.fg.grad(k))
                self.keys.append(p.data_ptr())
        return self.fg

    def __exit__(self, *exc):
        global _ACTIVE
        _ACTIVE = None
        for k in self.keys:
            _DIRECT.pop(k, None)
        return False


# ----------------------------------------------------------------------------------------------
# All views of a step as ONE autograd node.
# ----------------------------------------------------------------------------------------------
def rasterize_views(settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                    rotations=None, cov3D_precomp=None, means2D=None, streams=None):
    """Render V views of the same Gaussians in one call.

    `settings`: V GaussianRasterizationSettings of one image size. The Gaussian tensors are the
    keyword arguments of GaussianRasterizer.forward (same meaning, same exactly-one-of rules);
    `means2D`, if given, is a [V, P, 3] tensor whose gradient receives each view's screen-space
    gradient (the reference's `viewspace_points`, one per view). Returns
    (color [V,3,H,W], normal [V,3,H,W], depth [V,1,H,W], opac [V,1,H,W], radii [V,P] int32): what
    the reference's loop builds with one rasterizer call per view and a torch.stack
    (TS/renderer/gaussian_batch_renderer.py:273-330), written in place.

    Why it exists: at 512x512 one view is ~0.3 ms of device work and the per-view python of an
    autograd node (argument checks, ~30 allocations, three ctypes calls, the engine) costs more than
    that, so a 32-view SDS step was bound by the host. Here the step is one node: one allocation per
    kind of buffer, the view structs built once and reused by the backward, the views spread over
    `streams` (view_streams) so they overlap on the device, the parameter gradients of all views
    added in view order into one buffer per parameter (or straight into a FlatGrads buffer under
    FlatGrads.direct()). Gradients w.r.t. the cameras (config[3] > 0) are not produced here: use the
    per-view GaussianRasterizer for that."""
    from diff_gaussian_rasterization import GaussianRasterizer
    shs, colors_precomp, scales, rotations, cov3D_precomp = GaussianRasterizer._normalise(
        shs, colors_precomp, scales, rotations, cov3D_precomp)
    if means2D is None:
        means2D = torch.Tensor([])
    return _RasterizeViews.apply(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                 cov3D_precomp, tuple(settings), streams)


class _RasterizeViews(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                settings, streams):
        import ctypes
        from diff_gaussian_rasterization import _C
        from soar_b200 import _capi
        lib = _capi.load()
        V = len(settings)
        if V == 0:
            raise RuntimeError("rasterize_views: no views")
        if means3D.dim() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        s0 = settings[0]
        H, W = int(s0.image_height), int(s0.image_width)
        for s in settings:
            if int(s.image_height) != H or int(s.image_width) != W:
                raise RuntimeError("rasterize_views: all views must have one image size")
        P, dev = means3D.size(0), means3D.device
        if means2D.numel() and tuple(means2D.shape) != (V, P, 3):
            raise RuntimeError("means2D must have dimensions (num_views, num_points, 3)")
        f32 = dict(dtype=torch.float32, device=dev)
        color, normal = torch.empty((V, 3, H, W), **f32), torch.empty((V, 3, H, W), **f32)
        depth, opac = torch.empty((V, 1, H, W), **f32), torch.empty((V, 1, H, W), **f32)
        radii = torch.empty((V, P), dtype=torch.int32, device=dev)
        ctx.V, ctx.P, ctx.H, ctx.W = V, P, H, W
        ctx.streams = streams
        ctx.has_means2D = bool(means2D.numel())
        ctx.mark_non_differentiable(radii)
        if P == 0:
            for t in (color, normal, depth, opac):
                t.zero_()
            ctx.state = None
            return color, normal, depth, opac, radii

        f = _C._f32
        means3D_, sh_, colors_ = f(means3D, "means3D"), f(sh, "sh"), f(colors_precomp, "colors_precomp")
        opac_, scales_ = f(opacities, "opacities"), f(scales, "scales")
        rot_, cov_ = f(rotations, "rotations"), f(cov3D_precomp, "cov3D_precomp")
        M = sh_.size(1) if sh_ is not None else 0
        keep = [means3D_, sh_, colors_, opac_, scales_, rot_, cov_]
        views = (_capi.SoarRastView * V)()
        for v, s in enumerate(settings):
            cam = [f(s.bg, "bg"), f(s.viewmatrix, "viewmatrix"), f(s.projmatrix, "projmatrix"),
                   f(s.prcppoint, "prcppoint"), f(s.patch_bbox, "patch_bbox"), f(s.campos, "campos"),
                   f(s.config, "config")]
            keep.append(cam)
            x = views[v]
            x.P, x.sh_degree, x.M, x.width, x.height = P, int(s.sh_degree), M, W, H
            x.tanfovx, x.tanfovy, x.scale_modifier = float(s.tanfovx), float(s.tanfovy), float(s.scale_modifier)
            x.prefiltered, x.render_front = int(bool(s.prefiltered)), int(bool(s.render_front))
            x.sort_descending, x.debug = int(bool(s.sort_descending)), 0
            x.bg, x.viewmatrix, x.projmatrix, x.prcppoint = (cam[0].data_ptr(), cam[1].data_ptr(),
                                                             cam[2].data_ptr(), cam[3].data_ptr())
            x.patch_bbox, x.campos, x.config = cam[4].data_ptr(), cam[5].data_ptr(), cam[6].data_ptr()
            x.means3D, x.shs, x.colors_precomp = means3D_.data_ptr(), _capi.ptr(sh_), _capi.ptr(colors_)
            x.opacities, x.scales = _capi.ptr(opac_), _capi.ptr(scales_)
            x.rotations, x.cov3D_precomp = _capi.ptr(rot_), _capi.ptr(cov_)
            x.config_host_flags = _C._config_flags(s.config)
            if x.config_host_flags & _capi.CFG_LRN_CAM:
                # this node has no camera inputs to hand the gradient to; dropping it silently would
                # train nothing. GaussianRasterizer (one view per call) returns camera gradients.
                raise RuntimeError("rasterize_views: config[3] > 0 (learnable camera) is not supported by "
                                   "the batched call; render such views through GaussianRasterizer")

        gb = (lib.soar_rast_geom_bytes(P) + 255) // 256 * 256
        ib = (lib.soar_rast_image_bytes(W, H) + 255) // 256 * 256
        byte = dict(dtype=torch.uint8, device=dev)
        geom, img = torch.empty((V, gb), **byte), torch.empty((V, ib), **byte)
        status = _PinnedCounts.get(V)
        geom_p, img_p, radii_p, status_p = geom.data_ptr(), img.data_ptr(), radii.data_ptr(), status.data_ptr()
        color_p, normal_p, depth_p, opac_p = color.data_ptr(), normal.data_ptr(), depth.data_ptr(), opac.data_ptr()
        main = torch.cuda.current_stream(dev)
        per_view = _view_streams_of(main, streams, V)
        handles = [ctypes.c_void_p(s.cuda_stream) for s in per_view]
        events = [torch.cuda.Event() for _ in range(V)]
        byref, check = ctypes.byref, _capi.check
        with torch.cuda.device(dev):
            for v in range(V):
                check(lib.soar_rast_forward_bin_async(byref(views[v]), geom_p + v * gb, gb, radii_p + v * P * 4,
                                                      status_p + v * 8, handles[v]),
                      "rasterize_views (preprocess)")
                events[v].record(per_view[v])
            binning, counts = [], []
            host = status.numpy()
            for v in range(V):
                events[v].synchronize()
                R, flag = int(host[v, 0]) & 0xffffffff, int(host[v, 1])
                if flag:
                    raise RuntimeError("rasterize_views (preprocess) failed: Point is filtered although "
                                       "prefiltered is set. This shouldn't happen!")
                b = torch.empty(lib.soar_rast_binning_bytes(R, W, H), **byte)
                binning.append(b)
                counts.append(R)
                check(lib.soar_rast_forward_render(byref(views[v]), geom_p + v * gb, b.data_ptr(), b.numel(),
                                                   img_p + v * ib, ib, R, color_p + v * 12 * H * W,
                                                   normal_p + v * 12 * H * W, depth_p + v * 4 * H * W,
                                                   opac_p + v * 4 * H * W, handles[v]),
                      "rasterize_views (render)")
            _join(main, per_view)
        _PinnedCounts.put(status)
        ctx.state = (views, keep, geom, img, binning, counts, radii, gb, ib, M, sh_)
        ctx.num_rendered = counts
        # leaves registered by FlatGrads.direct(): the backward adds straight into their slices
        ctx.direct = active_direct() if _DIRECT else None
        ctx.sinks = None
        if _DIRECT:
            sinks = {k: direct_sink(t) for k, t in (("means3D", means3D), ("sh", sh), ("colors", colors_precomp),
                                                    ("opacity", opacities), ("scales", scales),
                                                    ("rotations", rotations), ("cov3D", cov3D_precomp))}
            ctx.sinks = {k: t for k, t in sinks.items() if t is not None} or None
        ctx.needs = tuple(ctx.needs_input_grad[:8])
        return color, normal, depth, opac, radii

    @staticmethod
    def backward(ctx, g_color, g_normal, g_depth, g_opac, _):
        import ctypes
        from diff_gaussian_rasterization import _C
        from soar_b200 import _capi
        lib = _capi.load()
        V, P, H, W = ctx.V, ctx.P, ctx.H, ctx.W
        none = (None,) * 10
        if ctx.state is None:
            if P > 0 and getattr(ctx, "consumed", False):
                raise RuntimeError("rasterize_views: backward called a second time; the per-view scratch "
                                   "is released after the first backward (retain_graph is not supported)")
            return none
        views, keep, geom, img, binning, counts, radii, gb, ib, M, sh_ = ctx.state
        dev = geom.device
        f32 = dict(dtype=torch.float32, device=dev)
        f = _C._f32
        g_color, g_normal = f(g_color, "grad color"), f(g_normal, "grad normal")
        g_depth, g_opac = f(g_depth, "grad depth"), f(g_opac, "grad opac")
        sinks = ctx.sinks or {}
        # only a node that owns sinks may use up the step's one store: a sink-less node (detached or
        # derived inputs, e.g. an occupancy pass) running first would otherwise leave the owner adding
        # into a flat buffer nobody has written yet (the per-view path guards the same way)
        store_first = bool(sinks) and ctx.direct is not None and ctx.direct.first_store()

        def buf(name, shape):
            t = sinks.get(name)
            return (t, True) if t is not None else (torch.empty(shape, **f32), False)

        (d_means3D, k0), (d_colors, k1) = buf("means3D", (P, 3)), buf("colors", (P, 3))
        (d_opacity, k2), (d_cov3D, k3) = buf("opacity", (P, 1)), buf("cov3D", (P, 6))
        (d_sh, k4) = buf("sh", (P, M, 3)) if M else (None, False)
        (d_scales, k5), (d_rot, k6) = buf("scales", (P, 3)), buf("rotations", (P, 4))
        d_means2D = torch.empty((V, P, 3), **f32)
        d_cam = torch.zeros((V, 36), **f32)
        sb = (lib.soar_rast_backward_scratch_bytes(P) + 255) // 256 * 256
        scratch = torch.empty((V, sb), dtype=torch.uint8, device=dev)
        ALL = (_capi.ACC_COLORS | _capi.ACC_OPACITY | _capi.ACC_MEANS3D | _capi.ACC_COV3D | _capi.ACC_SH
               | _capi.ACC_SCALES | _capi.ACC_ROTATIONS)
        # a sink that already holds earlier views of the step is added to from the first view on
        sink_bits = ((_capi.ACC_MEANS3D if k0 else 0) | (_capi.ACC_COLORS if k1 else 0)
                     | (_capi.ACC_OPACITY if k2 else 0) | (_capi.ACC_COV3D if k3 else 0)
                     | (_capi.ACC_SH if k4 else 0) | (_capi.ACC_SCALES if k5 else 0)
                     | (_capi.ACC_ROTATIONS if k6 else 0))
        first_mask = 0 if store_first else sink_bits
        main = torch.cuda.current_stream(dev)
        per_view = _view_streams_of(main, ctx.streams, V)
        many = bool(ctx.streams)
        grads = _capi.SoarRastGrads()
        grads.dL_dcolors, grads.dL_dopacity = d_colors.data_ptr(), d_opacity.data_ptr()
        grads.dL_dmeans3D, grads.dL_dcov3D = d_means3D.data_ptr(), d_cov3D.data_ptr()
        grads.dL_dsh = d_sh.data_ptr() if d_sh is not None else None
        grads.dL_dscales, grads.dL_drotations = d_scales.data_ptr(), d_rot.data_ptr()
        geom_p, img_p, radii_p, scratch_p = geom.data_ptr(), img.data_ptr(), radii.data_ptr(), scratch.data_ptr()
        m2_p, cam_p = d_means2D.data_ptr(), d_cam.data_ptr()
        gc_p, gn_p, gd_p, go_p = g_color.data_ptr(), g_normal.data_ptr(), g_depth.data_ptr(), g_opac.data_ptr()
        # the additions of the views are chained through events: same order, same rounding on any
        # number of streams; under FlatGrads.direct() the chain continues the step's earlier views
        chain = None
        if ctx.direct is not None and (many or ctx.direct.ordered):
            ctx.direct.ordered = True
            chain = ctx.direct.order_events
        pool = _EVENTS.setdefault(dev.index, [])
        prev = None
        with torch.cuda.device(dev):
            for v in range(V):
                grads.dL_dmeans2D = m2_p + v * P * 12
                grads.dL_dviewmat, grads.dL_dprojmat = cam_p + v * 144, cam_p + v * 144 + 64
                grads.dL_dcampos = cam_p + v * 144 + 128
                grads.accumulate_mask = first_mask if v == 0 else ALL
                after = done = None
                if chain is not None:
                    after, done = chain()
                elif many:
                    if v == len(pool):
                        e = torch.cuda.Event()
                        e.record(main)
                        pool.append(e)
                    after, done = prev, pool[v]
                    prev = done
                _capi.check(lib.soar_rast_backward_ordered(
                    ctypes.byref(views[v]), radii_p + v * P * 4, geom_p + v * gb, binning[v].data_ptr(),
                    img_p + v * ib, counts[v], gc_p + v * 12 * H * W, gn_p + v * 12 * H * W,
                    gd_p + v * 4 * H * W, go_p + v * 4 * H * W, scratch_p + v * sb, sb, ctypes.byref(grads),
                    ctypes.c_void_p(after.cuda_event) if after is not None else None,
                    ctypes.c_void_p(done.cuda_event) if done is not None else None,
                    ctypes.c_void_p(per_view[v].cuda_stream)), "rasterize_views (backward)")
            _join(main, per_view)
        ctx.state = None
        ctx.consumed = True
        n = ctx.needs
        return (None if k0 or not n[0] else d_means3D,
                d_means2D if ctx.has_means2D and n[1] else None,
                None if k4 or not n[2] or d_sh is None else d_sh,
                None if k1 or not n[3] else d_colors,
                None if k2 or not n[4] else d_opacity,
                None if k5 or not n[5] else d_scales,
                None if k6 or not n[6] else d_rot,
                None if k3 or not n[7] else d_cov3D, None, None)


_EVENTS = {}   # device index -> events reused by the backward's ordering chain


def _view_streams_of(main, streams, V):
    """Stream of each view; the side streams first wait for everything enqueued on `main`."""
    if not streams:
        return [main] * V
    fork = torch.cuda.Event()
    fork.record(main)
    for s in streams[:V]:
        s.wait_event(fork)
    return [streams[v % len(streams)] for v in range(V)]


def _join(main, per_view):
    seen = set()
    for s in per_view:
        if s is not main and id(s) not in seen:
            seen.add(id(s))
            main.wait_stream(s)


class _PinnedCounts:
    """Page-locked [V, 2] int32 buffers for the asynchronous num_rendered read-backs."""
    _free = {}

    @classmethod
    def get(cls, V):
        lst = cls._free.get(V)
        return lst.pop() if lst else torch.zeros((V, 2), dtype=torch.int32).pin_memory()

    @classmethod
    def put(cls, t):
        cls._free.setdefault(t.size(0), []).append(t)
