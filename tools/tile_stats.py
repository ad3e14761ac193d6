#!/usr/bin/env python3
"""Tile list lengths of one view of a bench workload, and the single-view kernel times.
Usage: python tools/tile_stats.py <workload> [camera index]"""
import sys, os, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import bench, parity
name = sys.argv[1]; ci = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda", 0)
wl = bench.Workload(name, dev, 0, 1, 4)
s = wl.scene_dict(wl.cam_dev[ci])
d = parity.debug_tensors(s)
r = d["ranges"].long(); L = (r[:, 1] - r[:, 0]).clamp_min(0)
ne = L[L > 0].float()
q = lambda p: float(torch.quantile(ne, p))
print(f"{name}: R={d['R']} tiles={L.numel()} non-empty={ne.numel()} mean={ne.mean():.0f} p50={q(.5):.0f} p90={q(.9):.0f} p99={q(.99):.0f} max={ne.max():.0f}")
nc = d["n_contrib"].view(wl.H, wl.W)
TY, TX = (wl.H + 15) // 16, (wl.W + 15) // 16
pad = torch.zeros(TY * 16, TX * 16, dtype=nc.dtype, device=dev); pad[:wl.H, :wl.W] = nc
deep = pad.view(TY, 16, TX, 16).permute(0, 2, 1, 3).reshape(TY * TX, 256).max(1).values.float()
dn = deep[L > 0]
print(f"deepest n_contrib per non-empty tile: mean={dn.mean():.0f} p90={float(torch.quantile(dn,.9)):.0f} max={dn.max():.0f}; tiles walked to the end (deep > 0.9 len): {int((dn > 0.9 * ne).sum())}")
arm = bench.SoarArm(wl)
wl.streams = None
prof = bench.profile_pass(arm, wl, 3)
print({k: (c, round(t / c * 1e3, 1)) for k, (c, t) in sorted(prof.items(), key=lambda kv: -kv[1][1])})
