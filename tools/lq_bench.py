"""Times the skin-weight query (grid vs brute force) at N points on the synthetic template."""
import sys, torch
sys.path.insert(0, '.')
from soar_b200 import synthetic
from soar_b200.lbs import SkinWeightQuery
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
m = synthetic.smplx_template(NB=1, seed=0)
g = synthetic.avatar_gaussians(N, m["v_template"], seed=0)
pts = g["points"].cuda()
q = SkinWeightQuery(m["v_template"].cuda(), m["lbs_weights"].cuda())
import numpy as np
hdr = q.grid[:64].cpu().numpy()
print("grid h", hdr[12:16].view(np.float32), "dims", hdr[20:32].view(np.int32), "cells", hdr[36:40].view(np.int32))
MODES = sys.argv[2].split(",") if len(sys.argv) > 2 else ("cells", "per-point", "brute")
for mode in MODES:
    q.brute_force = mode == "brute"
    q.per_point = mode == "per-point"
    bf = mode
    q(pts); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): w = q(pts)
    e1.record(); torch.cuda.synchronize()
    print(mode, N, "points:", e0.elapsed_time(e1) / 3, "ms")
