import os, sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from soar_b200 import synthetic
import test_rasterizer_gpu as T
s = T.to_device(synthetic.scene(100_000, 512, 512, sh_degree=0, seed=3, surfel=False, bg=(0., 0., 0.),
                                sort_descending=False, render_front=False))
cot = T.to_device(synthetic.cotangents(512, 512, seed=3))
res = {}
for mode in ("scalar", "fwd2"):
    if mode == "scalar": os.environ["SOAR_BLEND_FWD_SCALAR"] = "1"
    else: os.environ.pop("SOAR_BLEND_FWD_SCALAR", None)
    out, g = T.run_ours(s, cot)
    dbg = T._debug_tensors(s, out["radii"])
    res[mode] = (out["color"].clone(), dbg["n_contrib"].clone().reshape(512, 512), dbg["ranges"].clone())
ca, na, ra = res["scalar"]; cb, nb, rb = res["fwd2"]
d = (na != nb)
print("n_contrib differing pixels:", int(d.sum()), "of", d.numel(), "; color max abs", float((ca - cb).abs().max()))
ys, xs = torch.nonzero(d, as_tuple=True)
print("y%16 hist", torch.bincount(ys % 16, minlength=16).tolist())
print("x%16 hist", torch.bincount(xs % 16, minlength=16).tolist())
for k in range(min(8, len(ys))):
    y, x = int(ys[k]), int(xs[k])
    t = (y // 16) * 32 + x // 16
    print((y, x), "scalar", int(na[y, x]), "fwd2", int(nb[y, x]), "tile range len", int(ra.reshape(-1, 2)[t, 1] - ra.reshape(-1, 2)[t, 0]))
