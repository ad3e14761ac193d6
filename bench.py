#!/usr/bin/env python
"""bench.py — splat fwd+bwd frames/sec (BASELINE.json metric) on 1..8 B200.

Workload (config.workload): BASELINE config 3 — single-view rasterizer forward + backward at
1024x1024, 500k Gaussians, SH degree 3, synthetic "random" distribution (soar_b200/synthetic.py).
A step = `views_per_rank` views (default 4, cycling over a ring of cameras) rendered fwd+bwd by
every rank with gradients accumulated into one flat buffer, then ONE NCCL sum-allreduce of that
buffer when N > 1 (weak scaling: per-GPU work fixed). value = frames of all ranks / max-over-ranks
device time.

  value   inputs (Gaussians, cameras, cotangents) resident in HBM; CUDA events around K steps.
  e2e     the same frames through the public API (GaussianRasterizer + autograd), with every
          step's cameras and target images copied host->device from pinned memory inside the
          timed region (double buffered one view ahead, across step boundaries too), an L2 loss
          on the four outputs (loss_and_cotangents: the same torch launches for both arms), and the
          loss scalar read back every step.
  roofline   dominant kernel by event time (soar_prof_* in the C-ABI library), algorithmic bytes
             from DESIGN.md / SURVEY.md §8d with the measured R, Pv.
  cpu_baseline  the CPU restatement of the rasterizer (oracle/, C + OpenMP) timed on the host cores
                on a bounded sample of the same workload (N=1, rank 0 only).

`--impl reference` runs the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libref_dgr.so, built
from /root/reference) through the same protocol on the same GPU — the reference has no CPU path
for the rasterizer (BASELINE.json north_star).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as dist

WORKLOADS = {
    # name: (P, W, H, sh_degree)
    "c3": (500_000, 1024, 1024, 3),
    "c2": (100_000, 512, 512, 0),
    "tiny": (20_000, 256, 256, 1),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="soar", choices=["soar", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS))
    ap.add_argument("--views-per-rank", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-lbs", action="store_true")
    ap.add_argument("--sequential", action="store_true",
                    help="one blocking forward+backward per view through autograd's own gradient "
                         "accumulation (no begin-all pipelining, no in-place accumulation)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1])); power.append(float(r[2]))
            except Exception:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm),
                "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------
CAMS_PER_RANK = 8
PARAM_KEYS = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations")


class Workload:
    def __init__(self, name, device, rank, world, views_per_rank):
        from soar_b200 import synthetic
        from soar_b200.views import FlatGrads
        self.name = name
        self.P, self.W, self.H, self.deg = WORKLOADS[name]
        self.device, self.rank, self.world, self.vpr = device, rank, world, views_per_rank
        self.pipelined = True
        g = synthetic.gaussians(self.P, sh_degree=self.deg, seed=0)
        self.params = {k: g[k].to(device).requires_grad_(True) for k in PARAM_KEYS if g.get(k) is not None}
        self.flat = FlatGrads(self.params)
        cams = synthetic.ring_cameras(CAMS_PER_RANK * world, self.W, self.H, seed=0)
        mine = cams[rank * CAMS_PER_RANK:(rank + 1) * CAMS_PER_RANK]
        self.cam_host = []   # pinned [35]: viewmatrix 16 | projmatrix 16 | campos 3
        self.cam_dev = []
        for c in mine:
            flat = torch.cat([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1), c["campos"]])
            self.cam_host.append(flat.pin_memory())
            self.cam_dev.append(flat.to(device))
        self.tanfovx, self.tanfovy = mine[0]["tanfovx"], mine[0]["tanfovy"]
        f32 = dict(dtype=torch.float32, device=device)
        self.bg = torch.zeros(3, **f32)
        self.patch_bbox = torch.tensor([0.0, 0.0, float(self.H), float(self.W)], **f32)
        self.prcppoint = torch.tensor([0.5, 0.5], **f32)
        self.config = torch.zeros(4, **f32)
        cot = synthetic.cotangents(self.W, self.H, seed=0)
        self.cot = {k: v.to(device) for k, v in cot.items()}
        # e2e: per-camera target images for the four outputs (8 channels), pinned on the host
        gen = torch.Generator().manual_seed(1234 + rank)
        self.tgt_host = [torch.rand((8, self.H, self.W), generator=gen).pin_memory()
                         for _ in range(min(CAMS_PER_RANK, views_per_rank * 2))]
        # e2e staging: double-buffered device copies filled by a side stream, so the host->device
        # copy of view k+1 overlaps the render of view k (both inside the timed region)
        self.tgt_dev = [torch.empty((8, self.H, self.W), **f32) for _ in range(2)]
        self.cam_stage = [torch.empty(35, **f32) for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=device)
        self.copied = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        # cameras of a step, double buffered over steps (the next step's are prefetched)
        self.cam_all = [[torch.empty(35, **f32) for _ in range(views_per_rank)] for _ in range(2)]
        self.cams_copied = [torch.cuda.Event() for _ in range(2)]
        self.cams_consumed = [torch.cuda.Event() for _ in range(2)]
        self.prefetched = None  # step whose first inputs are already on their way
        self.loss_host = torch.zeros(1).pin_memory()

    def scene_dict(self, cam_flat):
        """Flat scene dict in the layout tests/refdgr.py and tests/util.py use."""
        s = {k: (v.detach() if v is not None else None) for k, v in self.params.items()}
        s.setdefault("shs", None)
        s.setdefault("colors_precomp", None)
        s.update(viewmatrix=cam_flat[0:16].view(4, 4), projmatrix=cam_flat[16:32].view(4, 4),
                 campos=cam_flat[32:35], tanfovx=self.tanfovx, tanfovy=self.tanfovy,
                 image_height=self.H, image_width=self.W, bg=self.bg, scale_modifier=1.0,
                 patch_bbox=self.patch_bbox, prcppoint=self.prcppoint, sh_degree=self.deg,
                 prefiltered=False, render_front=False, sort_descending=False, debug=False,
                 config=self.config)
        return s

    def view_index(self, step, v):
        return (step * self.vpr + v) % CAMS_PER_RANK

    def h2d_bytes_per_step(self):
        return self.vpr * (35 * 4 + 8 * self.H * self.W * 4)


def loss_and_cotangents(color, normal, depth, opac, tgt):
    """The e2e step's loss — sum of the four mean squared errors against the 8-channel target — and
    its gradients w.r.t. the four images, written once and shared by both arms (same torch launches
    on either side of the comparison)."""
    loss, cots = None, []
    for a, b in ((color, tgt[0:3]), (normal, tgt[3:6]), (depth, tgt[6:7]), (opac, tgt[7:8])):
        d = a.detach() - b
        part = (d * d).mean()
        loss = part if loss is None else loss + part
        cots.append(d * (2.0 / d.numel()))
    return loss, cots


class SoarArm:
    """The product, through its public reference-compatible API."""
    name = "soar"

    def __init__(self, wl):
        from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
        self.wl = wl
        self.S, self.R = GaussianRasterizationSettings, GaussianRasterizer

    def _render(self, cam_flat):
        wl, p = self.wl, self.wl.params
        st = self.S(image_height=wl.H, image_width=wl.W, tanfovx=wl.tanfovx, tanfovy=wl.tanfovy,
                    bg=wl.bg, scale_modifier=1.0, viewmatrix=cam_flat[0:16].view(4, 4),
                    projmatrix=cam_flat[16:32].view(4, 4), patch_bbox=wl.patch_bbox,
                    prcppoint=wl.prcppoint, sh_degree=wl.deg, campos=cam_flat[32:35],
                    prefiltered=False, render_front=False, sort_descending=False, debug=False,
                    config=wl.config)
        means2D = torch.zeros_like(p["means3D"], requires_grad=True)
        return self.R(st)(means3D=p["means3D"], means2D=means2D, opacities=p["opacities"],
                          shs=p.get("shs"), colors_precomp=p.get("colors_precomp"),
                          scales=p["scales"], rotations=p["rotations"])

    def _rasterizer(self, cam_flat):
        wl = self.wl
        st = self.S(image_height=wl.H, image_width=wl.W, tanfovx=wl.tanfovx, tanfovy=wl.tanfovy,
                    bg=wl.bg, scale_modifier=1.0, viewmatrix=cam_flat[0:16].view(4, 4),
                    projmatrix=cam_flat[16:32].view(4, 4), patch_bbox=wl.patch_bbox,
                    prcppoint=wl.prcppoint, sh_degree=wl.deg, campos=cam_flat[32:35],
                    prefiltered=False, render_front=False, sort_descending=False, debug=False,
                    config=wl.config)
        p = wl.params
        kw = dict(means3D=p["means3D"], means2D=torch.zeros_like(p["means3D"], requires_grad=True),
                  opacities=p["opacities"], shs=p.get("shs"), colors_precomp=p.get("colors_precomp"),
                  scales=p["scales"], rotations=p["rotations"])
        return self.R(st), kw

    def views_resident(self, cams):
        """All views of a step: phase 1 of every view enqueued up front (views.iter_views),
        gradients accumulated in place in the flat buffer (FlatGrads.direct)."""
        from soar_b200.views import iter_views
        c = self.wl.cot
        rs, kws = zip(*(self._rasterizer(cam) for cam in cams))
        with self.wl.flat.direct():
            for color, normal, depth, opac, _ in iter_views(rs, kws):
                torch.autograd.backward([color, normal, depth, opac],
                                        [c["color"], c["normal"], c["depth"], c["opac"]])

    def view_resident(self, cam_flat):
        color, normal, depth, opac, _ = self._render(cam_flat)
        c = self.wl.cot
        torch.autograd.backward([color, normal, depth, opac], [c["color"], c["normal"], c["depth"], c["opac"]])

    def view_e2e(self, cam_flat, tgt):
        color, normal, depth, opac, _ = self._render(cam_flat)
        loss, cots = loss_and_cotangents(color, normal, depth, opac, tgt)
        torch.autograd.backward([color, normal, depth, opac], cots)
        return loss


class ReferenceArm:
    """The unmodified reference CUDA rasterizer (oracle/_ref/libref_dgr.so), driven the way its own
    torch glue drives it: zero-filled outputs and gradient tensors per call
    (DGR/rasterize_points.cu:61-66,133-147), legacy default stream."""
    name = "reference"

    def __init__(self, wl):
        import refdgr
        if not refdgr.available():
            raise RuntimeError("oracle/_ref/libref_dgr.so missing (built from /root/reference by oracle/build_ref.sh)")
        self.wl = wl
        self.ref = refdgr.RefRasterizer()
        self.map = {"means3D": "means3D", "shs": "sh", "colors_precomp": "colors",
                    "opacities": "opacity", "scales": "scales", "rotations": "rotations"}

    def _accumulate(self, g):
        for k in self.wl.params:
            self.wl.flat.grad(k).add_(g[self.map[k]].view_as(self.wl.params[k]))

    def view_resident(self, cam_flat):
        s = self.wl.scene_dict(cam_flat)
        self.ref.forward(s, sync=False)
        self._accumulate(self.ref.backward(self.wl.cot, sync=False))

    def view_e2e(self, cam_flat, tgt):
        s = self.wl.scene_dict(cam_flat)
        o = self.ref.forward(s, sync=False)
        loss, cots = loss_and_cotangents(o["color"], o["normal"], o["depth"], o["opac"], tgt)
        self._accumulate(self.ref.backward(dict(zip(("color", "normal", "depth", "opac"), cots)), sync=False))
        return loss


# ----------------------------------------------------------------------------------------------
# timed regions
# ----------------------------------------------------------------------------------------------
def barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms, world, device):
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def step_resident(arm, wl, step):
    wl.flat.zero_()
    if wl.pipelined and hasattr(arm, "views_resident"):
        arm.views_resident([wl.cam_dev[wl.view_index(step, v)] for v in range(wl.vpr)])
    else:
        for v in range(wl.vpr):
            arm.view_resident(wl.cam_dev[wl.view_index(step, v)])
    wl.flat.allreduce()


def _stage_view(wl, step, v, slot):
    """Enqueue the host->device copy of view v's camera and target images on the copy stream."""
    i = wl.view_index(step, v)
    with torch.cuda.stream(wl.copy_stream):
        wl.copy_stream.wait_event(wl.consumed[slot])  # the previous user of this slot is done
        wl.cam_stage[slot].copy_(wl.cam_host[i], non_blocking=True)
        wl.tgt_dev[slot].copy_(wl.tgt_host[i % len(wl.tgt_host)], non_blocking=True)
        wl.copied[slot].record(wl.copy_stream)


def _prefetch_pipelined(wl, step):
    """Enqueue, on the copy stream, the inputs step `step` needs first: its cameras (35 floats each,
    double buffered over steps) and the target images of its first view."""
    buf = step & 1
    with torch.cuda.stream(wl.copy_stream):
        wl.copy_stream.wait_event(wl.cams_consumed[buf])
        for v in range(wl.vpr):
            wl.cam_all[buf][v].copy_(wl.cam_host[wl.view_index(step, v)], non_blocking=True)
        wl.cams_copied[buf].record(wl.copy_stream)
        wl.copy_stream.wait_event(wl.consumed[0])
        wl.tgt_dev[0].copy_(wl.tgt_host[wl.view_index(step, 0) % len(wl.tgt_host)], non_blocking=True)
        wl.copied[0].record(wl.copy_stream)
    wl.prefetched = step


def step_e2e_pipelined(arm, wl, step):
    """The e2e step on the product's multi-view path: phase 1 of every view is enqueued up front;
    target images are double buffered on the copy stream, one view ahead — across the step boundary
    too (the next step's cameras and first target are requested before this step's loss is read
    back, like any input pipeline; every step still copies exactly one step's inputs inside the
    timed region); gradients accumulate in place in the flat buffer."""
    from soar_b200.views import iter_view_pairs
    wl.flat.zero_()
    main = torch.cuda.current_stream()
    if wl.prefetched != step:
        _prefetch_pipelined(wl, step)
    buf = step & 1

    def stage_target(v, slot):
        with torch.cuda.stream(wl.copy_stream):
            wl.copy_stream.wait_event(wl.consumed[slot])
            i = wl.view_index(step, v)
            wl.tgt_dev[slot].copy_(wl.tgt_host[i % len(wl.tgt_host)], non_blocking=True)
            wl.copied[slot].record(wl.copy_stream)

    main.wait_event(wl.cams_copied[buf])
    pairs = (arm._rasterizer(wl.cam_all[buf][v]) for v in range(wl.vpr))  # built lazily, view by view
    total = None
    with wl.flat.direct():
        for v, (color, normal, depth, opac, _) in enumerate(iter_view_pairs(pairs)):
            slot = v & 1
            if v + 1 < wl.vpr:
                stage_target(v + 1, slot ^ 1)
            main.wait_event(wl.copied[slot])
            loss, cots = loss_and_cotangents(color, normal, depth, opac, wl.tgt_dev[slot])
            torch.autograd.backward([color, normal, depth, opac], cots)
            wl.consumed[slot].record(main)
            total = loss if total is None else total + loss
    wl.cams_consumed[buf].record(main)
    wl.flat.allreduce()
    _prefetch_pipelined(wl, step + 1)
    wl.loss_host.copy_(total.reshape(1), non_blocking=True)
    torch.cuda.current_stream().synchronize()  # the device->host read of the step's result
    return float(wl.loss_host[0])


def step_e2e(arm, wl, step):
    if wl.pipelined and hasattr(arm, "views_resident"):
        return step_e2e_pipelined(arm, wl, step)
    wl.flat.zero_()
    total = None
    main = torch.cuda.current_stream()
    if wl.prefetched != step:
        _stage_view(wl, step, 0, 0)
    for v in range(wl.vpr):
        slot = v & 1
        if v + 1 < wl.vpr:
            _stage_view(wl, step, v + 1, slot ^ 1)
        main.wait_event(wl.copied[slot])
        loss = arm.view_e2e(wl.cam_stage[slot], wl.tgt_dev[slot])
        wl.consumed[slot].record(main)
        total = loss if total is None else total + loss
    wl.flat.allreduce()
    _stage_view(wl, step + 1, 0, 0)  # the next step's first view, same prefetch as the pipelined path
    wl.prefetched = step + 1
    wl.loss_host.copy_(total.reshape(1), non_blocking=True)
    torch.cuda.current_stream().synchronize()  # the device->host read of the step's result
    return float(wl.loss_host[0])


def timed(fn, arm, wl, steps, warmup, world, device):
    for s in range(warmup):
        fn(arm, wl, s)
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in range(steps):
        fn(arm, wl, warmup + s)
    e1.record()
    barrier(world)
    return max_over_ranks(e0.elapsed_time(e1), world, device)


def scene_stats(wl):
    """Measured P_visible and num_rendered per camera (outside any timed region)."""
    from diff_gaussian_rasterization import _C
    e = torch.Tensor([])
    p = wl.params
    Pv, R = [], []
    for cam in wl.cam_dev:
        out = _C.rasterize_gaussians(
            wl.bg, p["means3D"].detach(), p["colors_precomp"].detach() if "colors_precomp" in p else e,
            p["opacities"].detach(), p["scales"].detach(), p["rotations"].detach(), 1.0, e,
            cam[0:16].view(4, 4), cam[16:32].view(4, 4), wl.prcppoint, wl.patch_bbox, wl.tanfovx,
            wl.tanfovy, wl.H, wl.W, p["shs"].detach() if "shs" in p else e, wl.deg, cam[32:35],
            False, False, False, False, wl.config)
        R.append(int(out[0]))
        Pv.append(int((out[5] > 0).sum().item()))
    return sum(Pv) / len(Pv), sum(R) / len(R)


def algorithmic_bytes(wl, Pv, R):
    """Per-launch algorithmic bytes of every kernel."""
    P, Npix = wl.P, wl.W * wl.H
    T = ((wl.W + 15) // 16) * ((wl.H + 15) // 16)
    M = (wl.deg + 1) ** 2 if "shs" in wl.params else 0
    Cin = 44 + 12 * M if M else 56
    G = Cin + 12
    return {
        "preprocess_fwd": P * (Cin + 4) + Pv * 92,
        "scan_lookback": P * 8,
        "duplicate_keys": Pv * 20 + R * 12,
        "radix_hist": R * 8,
        "radix_hist_scan": 0,
        "radix_onesweep": R * 24,
        "tile_ranges": R * 8 + T * 8,
        "blend_fwd": R * (4 + 92) + Npix * 44,
        "blend_bwd": R * (4 + 92) + Npix * 44 + Pv * 52,
        "geom_bwd": Pv * 52 + P * Cin + P * G,
    }


def profile_pass(arm, wl, steps):
    from soar_b200 import _capi
    lib = _capi.load()
    lib.soar_prof_enable(1)
    for s in range(steps):
        step_resident(arm, wl, 1000 + s)
    n = 256  # >= the library's kernel-id count (soar_prof_summary fills at most this many)
    counts = (ctypes.c_int32 * n)()
    ms = (ctypes.c_float * n)()
    k = lib.soar_prof_summary(n, counts, ms)
    lib.soar_prof_enable(0)
    out = {}
    for i in range(k):
        if counts[i]:
            out[lib.soar_kernel_name(i).decode()] = (int(counts[i]), float(ms[i]))
    return out


def cpu_baseline(wl):
    """CPU restatement of the rasterizer (oracle/, test infrastructure) timed on the host cores on a
    bounded sample of the same workload."""
    try:
        from oracle import rast_oracle
        if not rast_oracle.available():
            raise RuntimeError("oracle/librast_oracle.so not built")
        return rast_oracle.bench_sample(wl.name, wl.P, wl.W, wl.H, wl.deg, frames=2)
    except Exception as ex:  # noqa: BLE001
        return {"value": None, "unit": "frames/s", "cores": os.cpu_count(), "kind": "port",
                "sample": f"unavailable: {ex}"}


def lbs_report(device, cpu=True, n=100_000):
    """BASELINE config 1: SMPL-X LBS skinning of 100k canonical Gaussians to one pose. GPU: the
    sm_100a kernels through the python API (SmplxPoser -> SkinWeightQuery -> skin_gaussians), CUDA
    events. CPU: the reference's PyTorch op chain restated in oracle/lbs_oracle.py (the reference's
    python cannot travel to this box), all host threads, synthetic 10475-vertex / 55-joint body."""
    from soar_b200 import synthetic
    from soar_b200.lbs import SkinWeightQuery, skin_gaussians
    from soar_b200.smplx_fwd import SmplxPoser
    m = synthetic.smplx_template()
    pose = synthetic.smplx_pose()
    gen = torch.Generator().manual_seed(7)
    V = m["v_template"].shape[0]
    pts = m["v_template"][torch.randint(0, V, (n,), generator=gen)] + 0.002 * torch.randn(n, 3, generator=gen)
    quats = torch.nn.functional.normalize(torch.randn(n, 4, generator=gen), dim=1)
    keys = ("v_template", "shapedirs", "posedirs", "J_regressor", "parents", "lbs_weights")
    poser = SmplxPoser(*(m[k].to(device) for k in keys))
    _, _, A0, _ = poser(pose["betas"].to(device), torch.zeros(165, device=device), None, None, want_verts=False)
    inv = torch.linalg.inv(A0)
    query = SkinWeightQuery(m["v_template"].to(device), m["lbs_weights"].to(device))
    d = {k: v.to(device) for k, v in pose.items()}
    pts_d, quats_d = pts.to(device), quats.to(device)

    def gpu_stage():
        _, _, _, bones = poser(d["betas"], d["full_pose"], d["transl"], inv, want_verts=False)
        w = query(pts_d)
        return skin_gaussians(pts_d, quats_d, w, bones, swizzle=True)

    for _ in range(3):
        out = gpu_stage()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 20
    e0.record()
    for _ in range(iters):
        out = gpu_stage()
    e1.record()
    torch.cuda.synchronize()
    rep = {"workload": f"c1: SMPL-X LBS of {n} canonical Gaussians to one pose (SMPL-X forward + 30-NN "
                       "skin-weight query + bone blend + skinning of points and quaternions)",
           "gpu_ms": e0.elapsed_time(e1) / iters}
    if cpu:
        try:
            from oracle import lbs_oracle
            inv_c = inv.cpu()
            lbs_oracle.lbs_stage_cpu(m, pose, inv_c, pts, quats)  # warm-up
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                p_c, q_c, split = lbs_oracle.lbs_stage_cpu(m, pose, inv_c, pts, quats)
                dt = (time.perf_counter() - t0) * 1e3
                if best is None or dt < best[0]:
                    best = (dt, split)
            rep.update(cpu_ms=best[0], cpu_split=best[1], cpu_cores=torch.get_num_threads(),
                       cpu_kind="port (oracle/lbs_oracle.py: torch restatement of the reference chain)",
                       max_abs_points_gpu_vs_cpu=float((out[0].cpu() - p_c).abs().max()),
                       speedup=best[0] / rep["gpu_ms"])
        except Exception as ex:  # noqa: BLE001
            rep["cpu_ms"] = None
            rep["cpu_error"] = str(ex)
    return rep


def postops_report(device, W=1024, H=1024, iters=50, cpu=True):
    """Row f2 stage report at the bench resolution: the renderer's post block (depth2normal +
    normal2curv + flips, fwd+bwd) — our two stencil launches against the reference's torch op chain
    (oracle/postops_oracle.py, pinned to the reference's functions) on the same B200 and on the host
    cores. CUDA events; inputs resident."""
    from soar_b200 import postops, synthetic
    from oracle import postops_oracle
    pb = synthetic.postops_problem(W, H, seed=0, touch_border=True)
    cam = pb["camera"]
    dev = {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in pb.items()}

    def run(fn, t):
        n = t["normal"].clone().requires_grad_(True)
        d = t["depth"].clone().requires_grad_(True)
        on, oc, od, _ = fn(n, d, t["opac"], cam)
        ((on * t["g_normal"]).sum() + (oc * t["g_curv"]).sum() + (od * t["g_depth_normal"]).sum()).backward()
        return n.grad, d.grad

    def time_gpu(fn):
        for _ in range(5):
            run(fn, dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(device)
        e0.record()
        for _ in range(iters):
            run(fn, dev)
        e1.record()
        torch.cuda.synchronize(device)
        return e0.elapsed_time(e1) / iters

    # kernel-only time of the two launches (soar_prof events)
    from soar_b200 import _capi
    lib = _capi.load()
    lib.soar_prof_enable(1)
    for _ in range(iters):
        run(postops.surfel_postprocess, dev)
    n = 256  # >= the library's kernel-id count (soar_prof_summary fills at most this many)
    counts, ms = (ctypes.c_int32 * n)(), (ctypes.c_float * n)()
    k = lib.soar_prof_summary(n, counts, ms)
    kern = {lib.soar_kernel_name(i).decode(): ms[i] / counts[i] for i in range(k) if counts[i]}
    lib.soar_prof_enable(0)
    ours = time_gpu(postops.surfel_postprocess)
    chain = time_gpu(postops_oracle.surfel_postprocess)
    npix = W * H
    out = {"workload": f"f2: surfel post block fwd+bwd at {W}x{H} (incl. the autograd loss glue both arms share)",
           "gpu_ms": ours, "torch_chain_same_gpu_ms": chain, "speedup_vs_torch_chain": chain / ours,
           "post_fwd_us": kern.get("post_fwd", 0) * 1e3, "post_bwd_us": kern.get("post_bwd", 0) * 1e3,
           "post_fwd_GBps": npix * 49 / (kern["post_fwd"] * 1e-3) / 1e9 if kern.get("post_fwd") else None,
           "post_bwd_GBps": npix * 61 / (kern["post_bwd"] * 1e-3) / 1e9 if kern.get("post_bwd") else None,
           "algorithmic_bytes_per_pixel": {"fwd": 49, "bwd": 61}}
    if cpu:
        run(postops_oracle.surfel_postprocess, pb)
        t0 = time.perf_counter()
        for _ in range(3):
            run(postops_oracle.surfel_postprocess, pb)
        out["cpu_ms"] = (time.perf_counter() - t0) / 3 * 1e3
        out["cpu_cores"] = torch.get_num_threads()
        out["cpu_kind"] = "port (oracle/postops_oracle.py, pinned to the reference's own functions)"
    return out


def densify_report(device, N=500_000, R=15, iters=20):
    """Row f3 stage report at the bench size (500k Gaussians, SH3-sized f_rest): the fused Adam
    step, the per-view statistics and one prune + clone + split cycle — our launches against the
    reference's way of doing it (torch.optim.Adam + oracle/densify_oracle.py's torch op chain, which
    is pinned to the reference's own methods) on the same B200. CUDA events; data resident."""
    from oracle import densify_oracle as O
    from soar_b200 import synthetic
    from soar_b200.densify import GROUPS, GaussianStore
    pb = synthetic.densify_problem(N, R=R, seed=0, n_adam=1, n_views=1)
    P = {k: v.to(device) for k, v in pb["params"].items()}
    G = {k: v.to(device) for k, v in pb["adam_grads"][0].items()}
    vg, radii, sg = (t.to(device) for t in pb["views"][0])
    nbytes = sum(v.numel() for v in P.values()) * 4

    def timed(fn, n=iters, setup=None):
        ts = []
        for i in range(n + 2):
            arg = setup() if setup else None
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(device)
            e0.record()
            fn(arg) if setup else fn()
            e1.record()
            torch.cuda.synchronize(device)
            if i >= 2:
                ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    st = GaussianStore(*(P[k] for k in GROUPS))
    st.training_setup(pb["lrs"])
    ours_adam = timed(lambda: st.step(G))
    tp = [torch.nn.Parameter(P[k].clone()) for k in GROUPS]
    for p_, k in zip(tp, GROUPS):
        p_.grad = G[k]
    opt = torch.optim.Adam([{"params": [p_], "lr": pb["lrs"][k]} for p_, k in zip(tp, GROUPS)], lr=0.0, eps=1e-15)
    ref_adam = timed(opt.step)
    ours_stats = timed(lambda: st.add_densification_stats(vg, radii, sg))
    ost = O.new_state(P)
    ref_stats = timed(lambda: O.add_densification_stats(ost, vg, radii, sg))

    def fresh_store():
        s_ = GaussianStore(*(P[k] for k in GROUPS))
        s_.training_setup(pb["lrs"])
        s_.step(G)
        s_.add_densification_stats(vg, radii, sg)
        return s_

    def fresh_state():
        o_ = O.new_state(P)
        O.adam_step(o_, G, pb["lrs"])
        O.add_densification_stats(o_, vg, radii, sg)
        return o_

    counts = {}

    def ours_cycle(s_):
        s_.adaptive_prune(0.1, 1.0)
        counts["clone_split"] = s_.adaptive_densify(0.8, 1.0)
        counts["n_after"] = s_.num_points

    def ref_cycle(o_):
        O.adaptive_prune(o_, 0.1, 1.0)
        _, split = O.densify_masks(o_, 0.8, 1.0)
        stds = torch.exp(o_["p"]["scaling"])[split].repeat(2, 1)
        O.adaptive_densify(o_, 0.8, 1.0, torch.normal(torch.zeros_like(stds), stds))

    ours_cyc = timed(ours_cycle, n=5, setup=fresh_store)
    ref_cyc = timed(ref_cycle, n=5, setup=fresh_state)
    return {"workload": f"f3: {N} Gaussians, f_rest [{N},{R},3]: Adam step over the 7 groups, per-view statistics, "
                        "prune + clone + split cycle",
            "adam_ms": ours_adam, "adam_torch_same_gpu_ms": ref_adam,
            "adam_GBps": nbytes * 7 / (ours_adam * 1e-3) / 1e9,   # p, g, m, v read + p, m, v written
            "stats_ms": ours_stats, "stats_torch_same_gpu_ms": ref_stats,
            "densify_cycle_ms": ours_cyc, "densify_cycle_torch_same_gpu_ms": ref_cyc,
            "densify_counts": counts}


def c5_report(device, N=2_000_000, W=1920, H=1080, iters=5):
    """BASELINE config 5 on one GPU (a parity-test configuration, reported for scale): one
    dance_0-shaped step — SMPL-X pose, 30-NN skin-weight query, skinning, surfel rasterizer pass +
    occupancy pass, post-ops, backward to the canonical Gaussians, and distCUDA2 of the cloud —
    through soar_b200.avatar.AvatarRenderer. CUDA events; per-kernel sums from soar_prof."""
    import math as _m
    from simple_knn._C import distCUDA2
    from soar_b200 import _capi, synthetic
    from soar_b200.avatar import AvatarCamera, AvatarRenderer
    from soar_b200.smplx_fwd import SmplxPoser
    m = synthetic.smplx_template(NB=10, seed=0)
    keys = ("v_template", "shapedirs", "posedirs", "J_regressor", "parents", "lbs_weights")
    poser = SmplxPoser(*(m[k].to(device) for k in keys))
    ps = {k: v.to(device) for k, v in synthetic.smplx_pose(NB=10, seed=0).items()}
    cano, _, A0, _ = poser(ps["betas"], torch.zeros_like(ps["full_pose"]), None)
    ar = AvatarRenderer(poser, cano, torch.linalg.inv(A0))
    g = {k: v.to(device) for k, v in synthetic.avatar_gaussians(N, cano.cpu(), seed=0).items()}
    c = synthetic.ring_cameras(1, W, H, seed=0, dist=(1.6, 2.0))[0]
    cam = AvatarCamera(2 * _m.atan(c["tanfovx"]), 2 * _m.atan(c["tanfovy"]), W, H,
                       c["viewmatrix"].to(device), c["projmatrix"].to(device), c["campos"].to(device),
                       torch.tensor([0.5, 0.5], device=device))
    pts = g["points"].requires_grad_(True)
    rot = g["rot"].requires_grad_(True)
    col = g["colors"].requires_grad_(True)
    bg = torch.zeros(3, device=device)
    gen = torch.Generator(device).manual_seed(0)
    cot = {k: torch.randn(ch, H, W, device=device, generator=gen) for k, ch in
           (("render", 3), ("normal", 3), ("depth", 1), ("mask", 1), ("curv", 1), ("pred_normal", 3))}
    stats = {}

    def step():
        d2 = distCUDA2(pts.detach())
        sc = (g["scale_jitter"] * d2.sqrt()[:, None]).clamp(1e-4, 0.02).requires_grad_(True)
        bones = ar.pose(ps["betas"], ps["full_pose"], ps["transl"])
        out = ar.render(pts, rot, ar.weights(pts), bones, col, sc, g["occ"], cam, bg)
        sum((out[k] * cot[k]).sum() for k in cot).backward()
        stats["visible"] = out["visibility_filter"]
        for t in (pts, rot, col):
            t.grad = None

    step()
    lib = _capi.load()
    lib.soar_prof_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    e0.record()
    for _ in range(iters):
        step()
    e1.record()
    torch.cuda.synchronize(device)
    n = 256  # >= the library's kernel-id count (soar_prof_summary fills at most this many)
    counts, ms = (ctypes.c_int32 * n)(), (ctypes.c_float * n)()
    k = lib.soar_prof_summary(n, counts, ms)
    kern = {lib.soar_kernel_name(i).decode(): round(ms[i] / iters, 4) for i in range(k) if counts[i]}
    lib.soar_prof_enable(0)
    return {"workload": f"c5: {N} skinned Gaussians at {W}x{H}: KNN + pose + weight query + skinning + "
                        "rasterizer pass + occupancy pass + post-ops + backward, 1 GPU",
            "ms_per_step": e0.elapsed_time(e1) / iters, "visible_fraction": float(stats["visible"].float().mean()),
            "kernel_ms_per_step": kern, "kernel_ms_sum": round(sum(kern.values()), 3)}


def main():
    args = parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the render path has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)

    wl = Workload(args.workload, device, rank, world, args.views_per_rank)
    wl.pipelined = not args.sequential
    arm = SoarArm(wl) if args.impl == "soar" else ReferenceArm(wl)
    Pv, R = scene_stats(wl)
    from soar_b200 import _capi
    lib = _capi.load()

    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    n0 = lib.soar_launch_count()
    ms = timed(step_resident, arm, wl, args.steps, args.warmup, world, device)
    launches = int(lib.soar_launch_count() - n0)
    clk = clocks.stop() if rank == 0 else None
    frames = args.steps * wl.vpr * world
    value = frames / (ms / 1e3)

    e2e = None
    if not args.no_e2e:
        ms_e = timed(step_e2e, arm, wl, args.steps, max(3, args.warmup // 2), world, device)
        e2e = {"value": frames / (ms_e / 1e3), "unit": "frames/s",
               "h2d_bytes_per_step": wl.h2d_bytes_per_step(), "d2h_bytes_per_step": 4,
               "ms_per_step": ms_e / args.steps}

    roofline, shares = None, None
    if args.impl == "soar":
        prof = profile_pass(arm, wl, min(args.steps, 10))
        tot = sum(t for _, t in prof.values())
        alg = algorithmic_bytes(wl, Pv, R)
        top = max(prof, key=lambda k: prof[k][1])
        cnt, t = prof[top]
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            with open(pk) as f:
                peak = float(json.load(f).get("hbm_gbs", peak))
            peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)"
        ach = alg.get(top, 0) / (t / cnt * 1e-3) / 1e9
        traffic, traffic_src = None, None
        if args.workload == "c3":  # measured DRAM bytes per launch from the committed ncu capture
            import glob
            tj = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")))
            if tj:
                with open(tj[-1]) as f:
                    td = json.load(f)
                traffic = td.get("bytes_per_launch", {}).get(top)
                traffic_src = os.path.basename(tj[-1])
        roofline = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "algorithmic_bytes": alg.get(top, 0), "avg_ms": t / cnt, "launches": cnt,
                    "peak_source": peak_src}
        shares = {k: {"launches": c, "ms": round(tt, 4), "share": round(tt / tot, 4),
                      "GBps": round(alg.get(k, 0) / (tt / c * 1e-3) / 1e9, 1) if tt > 0 else None}
                  for k, (c, tt) in sorted(prof.items(), key=lambda kv: -kv[1][1])}
        alg_frame = sum(alg[k] * (6 if k == "radix_onesweep" else 1) for k in alg)
        roofline["frame_algorithmic_bytes"] = alg_frame
        roofline["frame_frac_of_hbm_roofline"] = (alg_frame * value / world) / (peak * 1e9)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.impl == "soar":
        cpu = cpu_baseline(wl)

    lbs = None
    if rank == 0 and world == 1 and args.impl == "soar" and not args.no_lbs:
        lbs = lbs_report(device, cpu=not args.no_cpu_baseline)

    post = None
    if rank == 0 and world == 1 and args.impl == "soar" and not args.no_lbs:
        post = postops_report(device, wl.W, wl.H, cpu=not args.no_cpu_baseline)

    c5 = dens = None
    if rank == 0 and world == 1 and args.impl == "soar" and not args.no_lbs:
        c5 = c5_report(device)
        dens = densify_report(device)

    if rank == 0:
        line = {
            "metric": "splat fwd+bwd frames/sec", "value": value, "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "impl": args.impl,
            "config": {"workload": f"{args.workload}: single-view rasterizer fwd+bwd, {wl.W}x{wl.H}, "
                                   f"{wl.P} Gaussians, SH degree {wl.deg}",
                       "views_per_rank_per_step": wl.vpr, "global_views_per_step": wl.vpr * world,
                       "parallelism": f"view-sharded x{world}, 1 allreduce of {wl.flat.nbytes()} B/step" if world > 1 else "single GPU",
                       "view_loop": ("phase 1 of all views of a step enqueued up front "
                                     "(soar_b200.views.iter_views) + in-place gradient accumulation "
                                     "(FlatGrads.direct)") if (wl.pipelined and args.impl == "soar")
                       else "one blocking forward+backward per view, autograd gradient accumulation",
                       "P_visible": Pv, "num_rendered": R,
                       "l2": "inputs (>=118 MB params + per-view scratch) exceed the 126 MB L2; cameras cycle"},
            "clocks": clk, "e2e": e2e, "gpu_launches": launches if args.impl == "soar" else None,
            "roofline": roofline, "kernels": shares, "cpu_baseline": cpu, "lbs": lbs, "postops": post, "c5_step": c5, "densify": dens,
        }
        if args.impl == "reference":
            line["cpu_baseline"] = {"value": value, "unit": "frames/s", "cores": 0, "kind": "reference",
                                    "sample": "the reference has no CPU rasterizer: its unmodified CUDA code "
                                              "(oracle/_ref/libref_dgr.so) on the same B200, same workload"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
