#!/usr/bin/env python3
"""bench.AvatarWorkload: the overlapped step (views + distCUDA2 on separate streams, one backward) must
produce the gradients of the plain one-stream step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
dev = torch.device("cuda", 0)
wl = bench.AvatarWorkload(dev, 0, 1, N=300_000, W=960, H=540)
wl.step(allreduce=False, overlapped=False); torch.cuda.synchronize()
a = wl.flat.flat.clone(); d_a = wl.d2.clone()
for _ in range(3):
    wl.step(allreduce=False, overlapped=True)
torch.cuda.synchronize()
b = wl.flat.flat
err = float((a - b).abs().max() / a.abs().max())
print("c5 overlapped vs plain: max rel err %.3e, distCUDA2 identical %s" % (err, bool(torch.equal(d_a, wl.d2))))
assert err <= 1e-3 and torch.equal(d_a, wl.d2)
