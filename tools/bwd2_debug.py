import os, sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from soar_b200 import synthetic
import test_rasterizer_gpu as T
for name, P, W, H, deg, surfel, bg, desc, front in T.CASES[:6]:
    s = T.to_device(synthetic.scene(P, W, H, sh_degree=deg, seed=3, surfel=surfel, bg=bg, sort_descending=desc, render_front=front))
    cot = T.to_device(synthetic.cotangents(W, H, seed=3))
    res = {}
    for mode in ("scalar", "bwd2"):
        if mode == "scalar": os.environ["SOAR_BLEND_BWD_SCALAR"] = "1"
        else: os.environ.pop("SOAR_BLEND_BWD_SCALAR", None)
        out, g = T.run_ours(s, cot)
        res[mode] = {k: v.clone() for k, v in g.items() if torch.is_tensor(v)}
    worst = 0.0
    for k in res["scalar"]:
        a, b = res["scalar"][k], res["bwd2"][k]
        rel = float((a - b).abs().max() / (a.abs().max() + 1e-30))
        worst = max(worst, rel)
    print(name, "worst max-abs-diff / max|scalar| over gradient tensors:", f"{worst:.2e}")
