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
        self.flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
        self.slices = {}
        off = 0
        for k, n in zip(self.names, sizes):
            self.slices[k] = (off, off + n)
            off += n
        self.attach()

    def attach(self):
        for k in self.names:
            a, b = self.slices[k]
            self.params[k].grad = self.flat[a:b].view_as(self.params[k])

    def zero_(self):
        self.flat.zero_()

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
