"""The scripted sequence of tests/golden/make_golden_densify.py, replayed on the CPU oracle
(oracle/densify_oracle.py) or on the product (soar_b200.densify.GaussianStore), and the comparison
of a state against one recorded stage of the golden file."""
import glob
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "densify_*.npz")))
GROUPS = ("xyz", "f_dc", "f_rest", "color", "opacity", "scaling", "rotation")
STATS = ("xyz_gradient_accum", "scale_gradient_accum", "rot_gradient_accum", "opac_gradient_accum",
         "denom", "max_radii2D")
STAGES = ("adam", "stats", "prune", "densify", "post")


def load(name):
    from soar_b200 import synthetic
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    pb = synthetic.densify_problem(int(g["N"]), R=int(g["R"]), seed=int(g["seed"]))
    return g, pb


def post_inputs(g, device="cpu"):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
    return ({k: t(g[f"post_grad/{k}"]) for k in GROUPS}, (t(g["post_vg"]), t(g["post_radii"]), t(g["post_sg"])))


def compare(get, g, tag, rtol=2e-6, exact_rows=True):
    """get(kind, name) -> tensor of the state under test. Parameters / moments: |a-b| <= rtol·max|b|
    (+1e-12); statistics the same; row counts must match exactly."""
    worst = 0.0
    for kind, names in (("p", GROUPS), ("m", GROUPS), ("v", GROUPS), ("s", STATS)):
        for k in names:
            want = g[f"{tag}/{kind}/{k}"]
            got = get(kind, k).detach().cpu().numpy().reshape(want.shape) if want.size else None
            if want.size == 0:
                assert get(kind, k).shape[0] == want.shape[0], f"{tag}/{kind}/{k}: row count"
                continue
            assert got.shape == want.shape, f"{tag}/{kind}/{k}: shape {got.shape} vs {want.shape}"
            if not np.isfinite(want).all():
                fin = np.isfinite(want)
                assert np.array_equal(fin, np.isfinite(got)), f"{tag}/{kind}/{k}: non-finite pattern"
                want, got = np.where(fin, want, 0), np.where(fin, got, 0)
            err = np.abs(got.astype(np.float64) - want).max()
            tol = rtol * max(np.abs(want).max(), 1e-30) + 1e-12
            assert err <= tol, f"{tag}/{kind}/{k}: max-abs {err:.3e} > {tol:.3e}"
            worst = max(worst, err / max(np.abs(want).max(), 1e-30))
    return worst
