"""Runs every stage of the hot path once or twice (developer tool: the target of `ncu --set full
-k regex:...` captures and of `compute-sanitizer` runs; nothing is timed here).

  python tools/stages.py full   [stage ...]   BASELINE sizes (2M points, 1080p, 500k Gaussians)
  python tools/stages.py small  [stage ...]   sizes compute-sanitizer finishes in minutes

stages: rast knn query lbs smplx post densify occgrid (default: all)"""
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from soar_b200 import synthetic  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "small"
want = set(sys.argv[2:]) or {"rast", "knn", "query", "lbs", "smplx", "post", "densify", "occgrid"}
FULL = size == "full"
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
REPS = 2 if FULL else 1


def done(name):
    torch.cuda.synchronize()
    print("stage", name, "ok", flush=True)


if "rast" in want:
    from util import run_ours, to_device
    for (P, W, H, deg, surfel) in ([(500_000, 1024, 1024, 3, False), (200_000, 1024, 1024, None, True)] if FULL
                                   else [(20_000, 256, 256, 3, False), (10_000, 200, 136, None, True)]):
        cfg = (1.0, 1.0, 1.0, 0.0) if surfel else None
        s = to_device(synthetic.scene(P, W, H, sh_degree=deg, seed=3, surfel=surfel, config=cfg))
        cot = to_device(synthetic.cotangents(W, H, seed=3))
        for _ in range(REPS):
            run_ours(s, cot)
    done("rast")

m = synthetic.smplx_template(NB=10, seed=0)
N = 2_000_000 if FULL else 30_000
g = synthetic.avatar_gaussians(N, m["v_template"], seed=0)
pts = g["points"].to(dev)

if "knn" in want:
    from simple_knn._C import distCUDA2
    for _ in range(REPS):
        distCUDA2(pts)
    done("knn")

keys = ("v_template", "shapedirs", "posedirs", "J_regressor", "parents", "lbs_weights")
if want & {"query", "lbs", "smplx"}:
    from soar_b200.lbs import SkinWeightQuery, skin_gaussians
    from soar_b200.smplx_fwd import SmplxPoser
    poser = SmplxPoser(*(m[k].to(dev) for k in keys))
    ps = {k: v.to(dev) for k, v in synthetic.smplx_pose(NB=10, seed=0).items()}
    cano, _, A0, _ = poser(ps["betas"], torch.zeros_like(ps["full_pose"]), None)
    inv = torch.linalg.inv(A0)
    for _ in range(REPS):
        verts, _, _, bones = poser(ps["betas"], ps["full_pose"], ps["transl"], inv, want_verts=True)
    done("smplx")
    query = SkinWeightQuery(cano, m["lbs_weights"].to(dev))
    if "query" in want:
        for _ in range(REPS):
            w = query(pts)
        query.brute_force = True
        query(pts[: (100_000 if FULL else 2_000)])
        query.brute_force = False
        done("query")
    if "lbs" in want:
        w = query(pts)
        p = pts.clone().requires_grad_(True)
        q = g["rot"].to(dev).requires_grad_(True)
        for _ in range(REPS):
            out = skin_gaussians(p, q, w, bones, swizzle=True)
            (out[0].sum() + out[1].sum()).backward()
        done("lbs")

if "post" in want:
    from soar_b200 import postops
    W, H = (1024, 1024) if FULL else (200, 136)
    pb = synthetic.postops_problem(W, H, seed=0, touch_border=True)
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in pb.items()}
    for _ in range(REPS):
        n = t["normal"].clone().requires_grad_(True)
        d = t["depth"].clone().requires_grad_(True)
        on, oc, od, _ = postops.surfel_postprocess(n, d, t["opac"], pb["camera"])
        ((on * t["g_normal"]).sum() + (oc * t["g_curv"]).sum() + (od * t["g_depth_normal"]).sum()).backward()
    done("post")

if "densify" in want:
    from soar_b200.densify import GROUPS, GaussianStore
    Nd = 500_000 if FULL else 20_000
    pb = synthetic.densify_problem(Nd, R=15, seed=0, n_adam=1, n_views=1)
    P = {k: v.to(dev) for k, v in pb["params"].items()}
    G = {k: v.to(dev) for k, v in pb["adam_grads"][0].items()}
    vg, radii, sg = (x.to(dev) for x in pb["views"][0])
    st = GaussianStore(*(P[k] for k in GROUPS))
    st.training_setup(pb["lrs"])
    for _ in range(REPS):
        st.step(G)
        st.add_densification_stats(vg, radii, sg)
    st.adaptive_prune(0.1, 1.0)
    st.adaptive_densify(0.8, 1.0)
    done("densify")

if "occgrid" in want:
    import utils_cases as U
    from soar_b200 import cuda_utils
    OCC_KEYS = ("pos_min", "pos_max", "grid_len", "grid_dim", "pos", "rot", "scale", "opac", "cutoff")
    for name in (("large", "cut") if FULL else ("avatar", "cut")):
        c = U.occgrid_case(name)
        args = [torch.from_numpy(np.ascontiguousarray(c[k])).to(dev) for k in OCC_KEYS]
        cuda_utils.gaussian2occgrid(*args)
    c = U.contour_case("full_hd" if FULL else "mask")
    cuda_utils.contour_padding(torch.from_numpy(c["image"]).to(dev), torch.from_numpy(c["mask"]).to(dev), c["size"])
    done("occgrid")
print("stages: all done")
