"""Optimizer step and densify / clone / split / prune of the Gaussian parameter groups on the sm_100a
kernels (soar_b200/csrc/densify.cu) — SURVEY.md §8 row f3. `GaussianStore` mirrors the part of the
reference's `GaussianBaseModel` that runs after every backward (TS/geometry/surfel_base.py), with the
same method names and argument meaning:

    training_setup / optimizer.step()      :596-681   -> training_setup(), step()
    add_densification_stats, max_radii2D    :1113-1135, :1214 -> add_densification_stats()
    prune_points / _prune_optimizer         :862-903   -> prune_points()
    cat_tensors_to_optimizer / densification_postfix :905-977 -> densification_postfix()
    densify_and_clone / densify_and_split   :979-1067  -> same names
    adaptive_prune / adaptive_densify       :1069-1111 -> same names

The reference does this with boolean-mask gathers, torch.cat and surgery on torch.optim.Adam's state
dict: about 8 launches per tensor per operation over 21 parameter/moment tensors + 6 statistics.
Here every operation is one launch over a table of tensors (Adam: all groups in one launch; prune /
clone / split: one row gather over parameters, moments and statistics together), and one host read
of the selected-row count (torch's boolean indexing synchronises the same way).
Out of scope: the tcnn attribute-field groups, `latent_pose`, learning-rate schedules.
"""
import ctypes
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import _capi

_lib = _capi.load()

# group name -> row width is taken from the tensors; the order is the reference's (:596-672)
GROUPS = ("xyz", "f_dc", "f_rest", "color", "opacity", "scaling", "rotation")
STATS = ("xyz_gradient_accum", "scale_gradient_accum", "rot_gradient_accum", "opac_gradient_accum",
         "denom", "max_radii2D")


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _rows(t):
    return t.shape[0], (t.numel() // t.shape[0] if t.shape[0] else int(torch.tensor(t.shape[1:]).prod()))


def allreduce_densify_stats(stats: Dict[str, torch.Tensor], group=None):
    """Makes the densification statistics of view-sharded replicas identical: the five
    accumulators and `denom` are SUMS over rendered views, so their all-reduce(SUM) over the ranks is
    what one process rendering every view would hold; `max_radii2D` is a running maximum, so it takes
    all-reduce(MAX). It has to be these statistics and not the raw screen-space gradient that is
    reduced: the per-view quantity is a norm (TS/geometry/surfel_base.py:1113-1117), which is not linear.
    Two collectives (one stacked SUM, one MAX) on whatever device the tensors live on; no-op without an
    initialised process group or with a single rank. Summation order over ranks is NCCL's / gloo's and
    the same on every rank, so all ranks end up with bit-identical statistics and hence take
    bit-identical clone / split / prune decisions."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    sums = [k for k in STATS if k != "max_radii2D"]
    packed = torch.stack([stats[k] for k in sums])
    dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    for i, k in enumerate(sums):
        stats[k].copy_(packed[i])
    dist.all_reduce(stats["max_radii2D"], op=dist.ReduceOp.MAX, group=group)


def replica_generator(device, seed: int, cycle: int) -> torch.Generator:
    """A generator every replica seeds identically for densify cycle number `cycle`: the split samples
    (torch.normal, TS/geometry/surfel_base.py:1000) then come out the same on every rank (same GPU
    model, same shapes, same seed), which keeps the replicas' parameters bit-identical after a split."""
    g = torch.Generator(device=device)
    g.manual_seed((int(seed) * 1_000_003 + int(cycle)) & 0x7FFFFFFFFFFFFFFF)
    return g


class GaussianStore:
    def __init__(self, xyz, features_dc, features_rest, colors, opacity, scaling, rotation,
                 surfel=True, percent_dense=0.01):
        """Parameters as the reference holds them: _xyz [N,3], _features_dc [N,1,3],
        _features_rest [N,R,3], _colors [N,3], _opacity [N,1] (logit), _scaling [N,3] (log),
        _rotation [N,4]. surfel = config[0] > 0."""
        if not xyz.is_cuda:
            raise RuntimeError("GaussianStore: CUDA tensors required (there is no CPU path)")
        f = lambda t: t.detach().to(torch.float32).contiguous().clone()
        self.p: Dict[str, torch.Tensor] = dict(xyz=f(xyz), f_dc=f(features_dc), f_rest=f(features_rest),
                                               color=f(colors), opacity=f(opacity), scaling=f(scaling),
                                               rotation=f(rotation))
        N = self.p["xyz"].shape[0]
        for k, t in self.p.items():
            if t.shape[0] != N:
                raise RuntimeError(f"GaussianStore: {k} has {t.shape[0]} rows, xyz has {N}")
        self.device = xyz.device
        self.surfel, self.percent_dense = bool(surfel), float(percent_dense)
        self.exp_avg: Dict[str, torch.Tensor] = {}
        self.exp_avg_sq: Dict[str, torch.Tensor] = {}
        self.step_count: Dict[str, int] = {}
        self.lr: Dict[str, float] = {k: 0.0 for k in GROUPS}
        self.betas, self.eps = (0.9, 0.999), 1e-15
        self.stats = {k: torch.zeros(N, dtype=torch.float32, device=self.device) for k in STATS}
        self.grad: Dict[str, Optional[torch.Tensor]] = {k: None for k in GROUPS}

    # ---- accessors with the reference's names --------------------------------------------------
    @property
    def num_points(self):
        return self.p["xyz"].shape[0]

    @property
    def get_xyz(self):
        return self.p["xyz"]

    @property
    def get_scaling(self):
        return torch.exp(self.p["scaling"])

    @property
    def get_opacity(self):
        return torch.sigmoid(self.p["opacity"])

    # ---- optimizer -------------------------------------------------------------------------------
    def training_setup(self, lrs: Dict[str, float], betas=(0.9, 0.999), eps=1e-15):
        """torch.optim.Adam(l, lr=0.0, eps=1e-15) over the per-Gaussian groups (:596-681): `lrs` maps
        group name -> learning rate (groups not named keep lr 0, like the reference's default)."""
        self.lr.update({k: float(v) for k, v in lrs.items()})
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        for k in GROUPS:
            self.exp_avg[k] = torch.zeros_like(self.p[k])
            self.exp_avg_sq[k] = torch.zeros_like(self.p[k])
            self.step_count[k] = 0

    def step(self, grads: Optional[Dict[str, torch.Tensor]] = None):
        """optimizer.step(): one fused launch over every group that has a gradient (a group without
        one is skipped and its step count does not advance, as in torch)."""
        grads = self.grad if grads is None else grads
        table = (_capi.SoarAdamGroup * _capi.ADAM_MAX_GROUPS)()
        keep, n = [], 0
        for k in GROUPS:
            g = grads.get(k)
            if g is None or self.p[k].numel() == 0:
                continue
            g = g.detach().to(torch.float32).contiguous()
            if g.numel() != self.p[k].numel():
                raise RuntimeError(f"GaussianStore.step: gradient of {k} has the wrong size")
            keep.append(g)
            self.step_count[k] += 1
            e = table[n]
            e.param, e.grad = self.p[k].data_ptr(), g.data_ptr()
            e.exp_avg, e.exp_avg_sq = self.exp_avg[k].data_ptr(), self.exp_avg_sq[k].data_ptr()
            e.numel, e.lr, e.step = self.p[k].numel(), self.lr[k], self.step_count[k]
            n += 1
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_adam_step(table, n, self.betas[0], self.betas[1], self.eps, _stream()),
                        "soar_adam_step")

    # ---- statistics --------------------------------------------------------------------------------
    def _stats_struct(self):
        s = _capi.SoarDensifyStats()
        for k in STATS:
            setattr(s, k, self.stats[k].data_ptr())
        return s

    def add_densification_stats(self, viewspace_grad, radii, scaling_grad=None):
        """One rendered view (update_states :1207-1216): max_radii2D = max(., radii) and, for the
        visible Gaussians (radii > 0), the five accumulators of add_densification_stats.
        viewspace_grad = viewspace_point_tensor.grad [N,3]; scaling_grad = _scaling.grad [N,3]."""
        N = self.num_points
        sg = self.grad["scaling"] if scaling_grad is None else scaling_grad
        f = lambda t: t.detach().to(torch.float32).contiguous()
        vg, sg = f(viewspace_grad), f(sg)
        rd = radii.detach().to(torch.int32).contiguous()
        s = self._stats_struct()
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_densify_add_stats(N, vg.data_ptr(), sg.data_ptr(),
                                                    self.p["rotation"].data_ptr(),
                                                    self.p["opacity"].data_ptr(), rd.data_ptr(),
                                                    ctypes.byref(s), _stream()), "soar_densify_add_stats")

    # ---- row machinery ---------------------------------------------------------------------------
    def _index_of(self, mask, invert=False):
        """Ascending rows with mask != 0 (or == 0): (index int32 [count], count). One host read."""
        N = mask.shape[0]
        m = mask.detach().reshape(-1).to(torch.uint8).contiguous()
        idx = torch.empty(max(N, 1), dtype=torch.int32, device=self.device)
        cnt = torch.zeros(1, dtype=torch.int32, device=self.device)
        ws = torch.empty(_lib.soar_mask_index_workspace_bytes(N), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_mask_to_index(m.data_ptr(), N, int(invert), idx.data_ptr(),
                                                cnt.data_ptr(), ws.data_ptr(), ws.numel(), _stream()),
                        "soar_mask_to_index")
        n = int(cnt.item())
        return idx[:n], n

    def _all_tensors(self):
        """(name, kind, tensor) of everything that has one row per Gaussian."""
        out = [(k, "p", self.p[k]) for k in GROUPS]
        if self.exp_avg:
            out += [(k, "m", self.exp_avg[k]) for k in GROUPS]
            out += [(k, "v", self.exp_avg_sq[k]) for k in GROUPS]
        out += [(k, "s", self.stats[k]) for k in STATS]
        return out

    def _store(self, name, kind, t):
        {"p": self.p, "m": self.exp_avg, "v": self.exp_avg_sq, "s": self.stats}[kind][name] = t

    def _gather(self, entries, index, n_index, n_out, dst_row0):
        """entries: (src tensor or None, dst tensor). One launch."""
        entries = [(s, d) for s, d in entries if d.numel() > 0]   # zero-width groups (f_rest at SH 0)
        n = len(entries)
        if n == 0:
            return
        src = (_capi.SoarRowTensor * _capi.ROWS_MAX_TENSORS)()
        dst = (_capi.SoarRowTensor * _capi.ROWS_MAX_TENSORS)()
        for i, (s, d) in enumerate(entries):
            w = d.numel() // d.shape[0] if d.shape[0] else 1
            src[i].data, src[i].width = (None if s is None else s.data_ptr()), w
            dst[i].data, dst[i].width = d.data_ptr(), w
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_rows_gather(src, dst, n, None if index is None else index.data_ptr(),
                                              n_index, n_out, dst_row0, _stream()), "soar_rows_gather")

    def prune_points(self, mask):
        """Remove the rows where mask is True from every parameter, both Adam moments and the six
        statistics (prune_points / _prune_optimizer, :862-903)."""
        idx, n_keep = self._index_of(mask, invert=True)
        if n_keep == self.num_points:
            return
        ents, news = [], []
        for name, kind, t in self._all_tensors():
            d = torch.empty((n_keep,) + tuple(t.shape[1:]), dtype=torch.float32, device=self.device)
            ents.append((t, d))
            news.append((name, kind, d))
        if n_keep:
            self._gather(ents, idx, n_keep, n_keep, 0)
        for name, kind, d in news:
            self._store(name, kind, d)
        self.grad = {k: None for k in GROUPS}

    def densification_postfix(self, new: Dict[str, torch.Tensor]):
        """Append rows (cat_tensors_to_optimizer, :905-977): parameters get `new[name]`, the Adam
        moments zeros, and all six statistics are RESET to zero for every row (:958-968)."""
        n_new, N = new["xyz"].shape[0], self.num_points
        ents = []
        for k in GROUPS:
            d = torch.empty((N + n_new,) + tuple(self.p[k].shape[1:]), dtype=torch.float32, device=self.device)
            d[:N] = self.p[k]
            d[N:] = new[k].reshape((n_new,) + tuple(self.p[k].shape[1:]))
            self.p[k] = d
            if self.exp_avg:
                for st in (self.exp_avg, self.exp_avg_sq):
                    z = torch.zeros((N + n_new,) + tuple(st[k].shape[1:]), dtype=torch.float32, device=self.device)
                    z[:N] = st[k]
                    st[k] = z
        self.stats = {k: torch.zeros(N + n_new, dtype=torch.float32, device=self.device) for k in STATS}
        self.grad = {k: None for k in GROUPS}

    def _append_selected(self, idx, n_sel, n_rep, override=None):
        """Rows idx (repeated n_rep times) of every parameter appended at the tail in ONE gather;
        moments of the new rows zero; statistics reset (densification_postfix). `override` replaces
        the gathered values of some groups (the split's new xyz / scaling)."""
        N, n_new = self.num_points, n_sel * n_rep
        ents, tail = [], []
        for k in GROUPS:
            d = torch.empty((N + n_new,) + tuple(self.p[k].shape[1:]), dtype=torch.float32, device=self.device)
            d[:N] = self.p[k]
            if override is not None and k in override:
                d[N:] = override[k]
            else:
                tail.append((self.p[k], d))
            ents.append((k, d))
        if n_new and tail:
            self._gather(tail, idx, n_sel, n_new, N)
        for k, d in ents:
            self.p[k] = d
            if self.exp_avg:
                for st in (self.exp_avg, self.exp_avg_sq):
                    z = torch.zeros((N + n_new,) + tuple(st[k].shape[1:]), dtype=torch.float32, device=self.device)
                    z[:N] = st[k]
                    st[k] = z
        self.stats = {k: torch.zeros(N + n_new, dtype=torch.float32, device=self.device) for k in STATS}
        self.grad = {k: None for k in GROUPS}

    # ---- the reference's densification entry points ----------------------------------------------
    def _select(self, mode, max_grad=0.0, extent=0.0, min_opacity=0.0):
        N = self.num_points
        a = torch.empty(N, dtype=torch.uint8, device=self.device)
        b = torch.empty(N, dtype=torch.uint8, device=self.device) if mode == 0 else None
        s = self._stats_struct()
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_densify_select(
                N, mode, ctypes.byref(s), self.p["scaling"].data_ptr(), self.p["opacity"].data_ptr(),
                float(max_grad), float(self.percent_dense * extent), float(min_opacity),
                float(0.5 * extent), float(1e-8 * extent ** 2), a.data_ptr(), _capi.ptr(b), _stream()),
                "soar_densify_select")
        return a, b

    def densify_and_clone(self, selected_pts_mask):
        """Append copies of the selected rows (:1036-1067). The mask comes from adaptive_densify."""
        idx, n = self._index_of(selected_pts_mask)
        self._append_selected(idx, n, 1)
        return n

    def densify_and_split(self, selected_pts_mask, N=2, samples=None, generator=None):
        """Replace every selected Gaussian by N samples of it (:979-1034): new centres
        R(q)·samples + xyz with samples ~ N(0, get_scaling), scales / (0.8 N), every other parameter
        repeated; the originals are pruned. `samples` [N·n_sel, 3] may be passed in (tests); by
        default they are drawn like the reference does, torch.normal(0, stds) on the device.
        `selected_pts_mask` covers the rows that existed BEFORE densify_and_clone appended its copies
        (the reference pads the gradient with zeros for the new rows, :984-986)."""
        n_before = selected_pts_mask.shape[0]
        idx, n_sel = self._index_of(selected_pts_mask)
        n_new = n_sel * N
        if samples is None:
            stds = self.get_scaling[idx.long()].repeat(N, 1)
            samples = torch.normal(mean=torch.zeros_like(stds), std=stds, generator=generator)
        samples = samples.detach().to(torch.float32).contiguous()
        new_xyz = torch.empty((n_new, 3), dtype=torch.float32, device=self.device)
        new_scaling = torch.empty((n_new, 3), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _capi.check(_lib.soar_densify_split_rows(
                self.p["xyz"].data_ptr(), self.p["scaling"].data_ptr(), self.p["rotation"].data_ptr(),
                _capi.ptr(idx), n_sel, N, _capi.ptr(samples), int(self.surfel), _capi.ptr(new_xyz),
                _capi.ptr(new_scaling), _stream()), "soar_densify_split_rows")
        total_before = self.num_points
        self._append_selected(idx, n_sel, N, override=dict(xyz=new_xyz, scaling=new_scaling))
        prune = torch.zeros(total_before + n_new, dtype=torch.uint8, device=self.device)
        prune[:n_before] = selected_pts_mask.reshape(-1).to(torch.uint8)
        self.prune_points(prune)
        return n_sel

    def adaptive_prune(self, min_opacity, extent):
        """:1069-1081 — prune low opacity, never-visible and degenerate-scale Gaussians."""
        mask, _ = self._select(1, extent=extent, min_opacity=min_opacity)
        self.prune_points(mask)
        return mask

    def adaptive_densify(self, max_grad, extent, N=2, samples=None, generator=None):
        """:1083-1111 — clone the small high-gradient Gaussians, split the large ones. Both masks come
        from the statistics as they are BEFORE the clone resets them (the reference computes
        `grad_pos` first and passes it to both calls)."""
        clone_mask, split_mask = self._select(0, max_grad=max_grad, extent=extent)
        n_clone = self.densify_and_clone(clone_mask)
        n_split = self.densify_and_split(split_mask, N=N, samples=samples, generator=generator)
        return n_clone, n_split

    # ---- view-sharded replicas -------------------------------------------------------
    def sync_stats(self, group=None):
        """All-reduce the statistics over the ranks (allreduce_densify_stats). Call it once, right
        before adaptive_densify / adaptive_prune, on every rank."""
        allreduce_densify_stats(self.stats, group)

    def densify_cycle(self, max_grad, min_opacity, extent, N=2, seed=0, group=None):
        """One replica-consistent prune + densify cycle: statistics all-reduced, split samples drawn
        from a generator seeded with (seed, cycle number), so every rank mutates its replica the same
        way and ends with the same P. Returns (n_clone, n_split, n_pruned). Rebuild anything sized by
        P afterwards (FlatGrads(store.p) for the gradient buffer)."""
        self.sync_stats(group)
        self.cycles = getattr(self, "cycles", 0) + 1
        gen = replica_generator(self.device, seed, self.cycles)
        # prune first: the densify postfix zeroes the statistics (max_radii2D included), and a prune
        # after it would see every Gaussian as never rendered
        mask = self.adaptive_prune(min_opacity, extent)
        n_clone, n_split = self.adaptive_densify(max_grad, extent, N=N, generator=gen)
        return n_clone, n_split, int(mask.sum().item())

    def replica_checksum(self):
        """[P, sum of the raw bits of every parameter, moment and statistic] as int64 — equal across
        ranks iff (up to hash collisions) the replicas are bit-identical."""
        tot = torch.zeros((), dtype=torch.int64, device=self.device)
        for i, (_, _, t) in enumerate(self._all_tensors()):
            if t.numel():
                bits = t.contiguous().view(torch.int32).to(torch.int64)
                w = torch.arange(bits.numel(), device=self.device, dtype=torch.int64).view_as(bits) % 8191 + 1 + i
                tot = tot + (bits * w).sum()
        return torch.stack([torch.tensor(self.num_points, dtype=torch.int64, device=self.device), tot])
