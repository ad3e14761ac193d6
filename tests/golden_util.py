"""Loading of the golden vectors under tests/golden/ (made by tests/golden/make_golden.py from the
unmodified reference CUDA code on a B200)."""
import glob
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RAST_CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "rast_*.npz")))
KNN_CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "knn_*.npz")))
IMAGES = ("color", "normal", "depth", "opac")


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def scene_of(d, as_torch=False):
    """Golden inputs -> scene dict in the layout of soar_b200.synthetic.scene."""
    s = {k[3:]: d[k] for k in d.files if k.startswith("in_")}
    for k in ("shs", "colors_precomp"):
        s.setdefault(k, None)
    for k in ("tanfovx", "tanfovy", "scale_modifier"):
        s[k] = float(s[k])
    for k in ("image_height", "image_width", "sh_degree"):
        s[k] = int(s[k])
    for k in ("prefiltered", "render_front", "sort_descending"):
        s[k] = bool(s[k])
    s["debug"] = False
    if as_torch:
        import torch
        s = {k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) and v.ndim else v)
             for k, v in s.items()}
    return s


def cot_of(d, as_torch=False):
    c = {k: d["cot_" + k] for k in IMAGES}
    if as_torch:
        import torch
        c = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in c.items()}
    return c


def grad_tol(d, name, rel=1e-3):
    """|a-b| <= rel * max|ref| + the reference's own run-to-run spread (its atomics are unordered)."""
    ref = d["grad_" + name]
    return rel * max(float(np.abs(ref).max()), 1e-30) + float(d["grad_spread_" + name])
