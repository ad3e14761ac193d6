"""Per-kernel times of one skin-weight query at N points (SOAR_LQ_HSCALE sweeps the grid's cell edge)."""
import ctypes, sys, torch
sys.path.insert(0, '.')
from soar_b200 import synthetic, _capi
from soar_b200.lbs import SkinWeightQuery
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
m = synthetic.smplx_template(NB=1, seed=0)
pts = synthetic.avatar_gaussians(N, m["v_template"], seed=0)["points"].cuda()
q = SkinWeightQuery(m["v_template"].cuda(), m["lbs_weights"].cuda())
q(pts); torch.cuda.synchronize()
lib = _capi.load()
lib.soar_prof_enable(1)
q(pts); torch.cuda.synchronize()
n = 256
counts, ms = (ctypes.c_int32 * n)(), (ctypes.c_float * n)()
k = lib.soar_prof_summary(n, counts, ms)
lib.soar_prof_enable(0)
print({lib.soar_kernel_name(i).decode(): (counts[i], round(ms[i], 3)) for i in range(k) if counts[i]})
