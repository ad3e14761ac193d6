"""Times distCUDA2 at N avatar-distributed points (developer tool; `ncu -k regex:knn_search` target)."""
import sys, torch
sys.path.insert(0, '.')
from soar_b200 import synthetic
from simple_knn._C import distCUDA2
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
m = synthetic.smplx_template(NB=1, seed=0)
pts = synthetic.avatar_gaussians(N, m["v_template"], seed=0)["points"].cuda()
distCUDA2(pts); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): d = distCUDA2(pts)
e1.record(); torch.cuda.synchronize()
print("distCUDA2", N, "points:", e0.elapsed_time(e1) / 3, "ms; mean dist2", float(d.mean()))
