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


def iter_views(rasterizers, calls):
    """Render several views of the same Gaussians back to back without idling the device.

    `rasterizers`: GaussianRasterizer modules (one per view); `calls`: one dict of forward()
    keyword arguments per view. The first phase of EVERY view (preprocess, depth sort, tile
    counts) is enqueued up front (GaussianRasterizer.begin), then the views are finished and
    yielded one at a time — `for out in iter_views(...): loss(out).backward()` keeps only one
    view's images and binning buffers alive. The reference's loop
    (TS/renderer/gaussian_batch_renderer.py:273-323) blocks the host once per view on
    num_rendered (rasterizer_impl.cu:250) with nothing queued behind it; here only the first
    view of a step waits."""
    return iter_view_pairs(zip(rasterizers, calls))


def iter_view_pairs(pairs):
    """iter_views over an iterable of (rasterizer, forward-kwargs) pairs. The iterable is consumed
    lazily: with a generator, view k's settings and tensors are built while the device already runs
    phase 1 of view k-1, so the first kernels of a step are enqueued as early as possible."""
    started = []
    for r, kw in pairs:
        r.begin(**kw)
        started.append((r, kw))
    for r, kw in started:
        yield r(**kw)


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
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
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
