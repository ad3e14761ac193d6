#!/usr/bin/env python
"""Generates tests/golden/densify_*.npz by running a scripted sequence through the reference's OWN
methods — `_prune_optimizer`, `prune_points`, `cat_tensors_to_optimizer`, `densification_postfix`,
`densify_and_clone`, `densify_and_split`, `adaptive_prune`, `adaptive_densify`,
`add_densification_stats` (soar/threestudio-soar/geometry/surfel_base.py:847-1135) and
`build_rotation` (utils/general_utils.py:100-123) — with a real torch.optim.Adam configured like
`training_setup` (:596-681, the seven per-Gaussian groups, lr=0.0 default, eps=1e-15).

The reference module is not importable here (threestudio, pytorch3d, tinycudann, a CUDA device): the
method definitions are taken out of the class with `ast` at run time and exec'd UNCHANGED on a stub
object; `torch` is handed to them through a proxy that drops `device="cuda"` from torch.zeros and
records the torch.normal draw of densify_and_split so the sequence can be replayed. Nothing is copied
into the repo. Run in the build container:  python tests/golden/make_golden_densify.py
"""
import ast
import os
import sys

import numpy as np
import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
TS = "/root/reference/soar/threestudio-soar"
from soar_b200 import synthetic  # noqa: E402

METHODS = {"_prune_optimizer", "prune_points", "cat_tensors_to_optimizer", "densification_postfix",
           "densify_and_split", "densify_and_clone", "adaptive_prune", "adaptive_densify",
           "add_densification_stats", "get_scaling", "get_xyz", "get_opacity"}
GROUPS = ("xyz", "f_dc", "f_rest", "color", "opacity", "scaling", "rotation")
ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", color="_colors",
            opacity="_opacity", scaling="_scaling", rotation="_rotation")
STATS = ("xyz_gradient_accum", "scale_gradient_accum", "rot_gradient_accum", "opac_gradient_accum", "denom")


class TorchProxy:
    """`torch` for the extracted code: CPU instead of device="cuda", torch.normal recorded."""

    def __init__(self):
        self.samples = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def zeros(self, *a, **k):
        k.pop("device", None)
        return torch.zeros(*a, **k)

    def normal(self, *a, **k):
        out = torch.normal(*a, **k)
        self.samples.append(out.clone())
        return out


def reference_class(proxy):
    ns = {"torch": proxy, "nn": nn}
    tree = ast.parse(open(os.path.join(TS, "utils/general_utils.py")).read())
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == "build_rotation":
            exec(compile(ast.Module([node], []), "general_utils.py", "exec"), ns)
    tree = ast.parse(open(os.path.join(TS, "geometry/surfel_base.py")).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianSurfelModel")
    body = [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in METHODS]
    assert {n.name for n in body} == METHODS, METHODS - {n.name for n in body}
    ref = ast.ClassDef(name="Ref", bases=[], keywords=[], body=body, decorator_list=[])
    if hasattr(ref, "type_params"):
        ref.type_params = []
    mod = ast.fix_missing_locations(ast.Module([ref], []))
    exec(compile(mod, "surfel_base.py", "exec"), ns)
    return ns["Ref"]


def dump(obj, tag, out):
    for k in GROUPS:
        p = getattr(obj, ATTR[k])
        stt = obj.optimizer.state.get(p, {})
        out[f"{tag}/p/{k}"] = p.detach().numpy().copy()
        out[f"{tag}/m/{k}"] = stt["exp_avg"].numpy().copy() if stt else np.zeros_like(p.detach().numpy())
        out[f"{tag}/v/{k}"] = stt["exp_avg_sq"].numpy().copy() if stt else np.zeros_like(p.detach().numpy())
    for k in STATS:
        out[f"{tag}/s/{k}"] = getattr(obj, k).reshape(-1).numpy().copy()
    out[f"{tag}/s/max_radii2D"] = obj.max_radii2D.numpy().copy()


def main():
    torch.set_num_threads(1)
    for name, (N, R, seed, max_grad, extent) in {"densify_small": (600, 3, 0, 0.9, 1.0),
                                                 "densify_sh0": (257, 0, 1, 0.7, 2.0)}.items():
        proxy = TorchProxy()
        Ref = reference_class(proxy)
        pb = synthetic.densify_problem(N, R=R, seed=seed)
        obj = Ref()
        for k in GROUPS:
            setattr(obj, ATTR[k], nn.Parameter(pb["params"][k].clone()))
        obj.scaling_activation, obj.scaling_inverse_activation = torch.exp, torch.log
        obj.opacity_activation = torch.sigmoid
        obj.percent_dense = 0.01
        obj.config = torch.tensor([1.0, 1.0, 1.0, 0.0])
        for k in STATS:
            setattr(obj, k, torch.zeros((N, 1)))
        obj.max_radii2D = torch.zeros(N)
        obj.optimizer = torch.optim.Adam(
            [{"params": [getattr(obj, ATTR[k])], "lr": pb["lrs"][k], "name": k} for k in GROUPS],
            lr=0.0, eps=1e-15)
        out = dict(N=N, R=R, seed=seed, max_grad=max_grad, extent=extent)

        def adam(grads):
            for k in GROUPS:
                getattr(obj, ATTR[k]).grad = grads[k].clone()
            obj.optimizer.step()

        @torch.no_grad()                                   # update_states is decorated the same way (:1197)
        def views(vs):
            for vg, radii, sg in vs:                       # update_states :1207-1216
                obj._scaling.grad = sg.clone()
                vt = torch.zeros_like(vg)
                vt.grad = vg.clone()
                obj.max_radii2D = torch.max(obj.max_radii2D, radii.float())
                obj.add_densification_stats(vt, radii > 0)

        for g in pb["adam_grads"]:
            adam(g)
        dump(obj, "adam", out)
        views(pb["views"])
        dump(obj, "stats", out)
        with torch.no_grad():
            obj.adaptive_prune(0.1, extent)
        dump(obj, "prune", out)
        n1 = obj._xyz.shape[0]
        out["n_after_prune"] = n1
        with torch.no_grad():
            obj.adaptive_densify(max_grad, extent)
        dump(obj, "densify", out)
        out["samples"] = proxy.samples[0].numpy()
        n2 = obj._xyz.shape[0]
        # one optimizer step and one view on the densified set (new rows start from zero moments)
        rng_g, rng_v = pb["grads"](n2), pb["view"](n2)
        out.update({f"post_grad/{k}": v.numpy() for k, v in rng_g.items()})
        out.update(post_vg=rng_v[0].numpy(), post_radii=rng_v[1].numpy(), post_sg=rng_v[2].numpy())
        adam(rng_g)
        views([rng_v])
        dump(obj, "post", out)
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, os.path.getsize(path), "bytes; N", N, "->", n1, "->", n2,
              "split samples", tuple(out["samples"].shape))


if __name__ == "__main__":
    main()
