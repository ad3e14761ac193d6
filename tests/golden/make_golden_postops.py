#!/usr/bin/env python
"""Generates tests/golden/postops_*.npz by executing the reference's OWN `depth2normal`,
`normal2curv` and the renderer's post block (soar/threestudio-soar/renderer/
diff_gaussian_rasterizer.py:292-303, 355-448) on the synthetic images of
soar_b200.synthetic.postops_problem, forward and (through torch autograd) backward.

The reference module is not importable here (it imports threestudio, pytorch3d, ...): the three
function definitions are taken out of the file with `ast` at run time and exec'd unchanged; the
post block is the reference's statement sequence applied to those functions. Nothing is copied
into the repo. Run in the build container (the python reference cannot travel to the GPU box):
    python tests/golden/make_golden_postops.py
"""
import ast
import math
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_FILE = "/root/reference/soar/threestudio-soar/renderer/diff_gaussian_rasterizer.py"

from soar_b200 import synthetic  # noqa: E402

CASES = {
    # name: (W, H, seed, prcppoint, touch_border)
    "postops_centered": (96, 64, 0, (0.5, 0.5), False),
    "postops_border_offcenter": (75, 53, 1, (0.45, 0.58), True),
}


def reference_functions():
    src = open(REF_FILE).read()
    tree = ast.parse(src)
    want = {"fov2focal", "depth2normal", "normal2curv"}
    ns = {"torch": torch, "math": math}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in want:
            exec(compile(ast.Module([node], []), REF_FILE, "exec"), ns)
    assert want <= set(ns), "reference functions not found"
    return ns


def post_block(ns, rendered_normal, rendered_depth, rendered_opac, camera):
    """The reference's statements at :292-303, on clones (they write in place)."""
    rendered_normal = rendered_normal.clone()
    mask = rendered_opac > 1e-5
    normal_mask = mask.repeat(3, 1, 1)
    rendered_normal[~normal_mask] = rendered_normal[~normal_mask].detach()
    rendered_normal[1] *= -1
    rendered_normal[2] *= -1
    curv = ns["normal2curv"](rendered_normal, rendered_opac.detach() > 1e-5)
    rendered_normal = (rendered_normal + 1) / 2
    depth_normal = ns["depth2normal"](rendered_depth, rendered_opac.detach() > 1e-5, camera)
    depth_normal[1:] *= -1
    depth_normal = (depth_normal + 1) / 2
    return rendered_normal, curv, depth_normal, mask


def main():
    warnings.filterwarnings("ignore")
    torch.set_num_threads(1)
    ns = reference_functions()
    for name, (W, H, seed, pp, border) in CASES.items():
        pb = synthetic.postops_problem(W, H, seed=seed, prcppoint=pp, touch_border=border)
        cam = pb["camera"]
        mask = pb["opac"] > 1e-5
        out = dict(W=W, H=H, seed=seed, prcppoint=np.asarray(pp, np.float32), touch_border=border,
                   in_normal=pb["normal"].numpy(), in_depth=pb["depth"].numpy(),
                   in_opac=pb["opac"].numpy())
        # the two functions on their own
        d = pb["depth"].clone().requires_grad_(True)
        dn = ns["depth2normal"](d, mask, cam)
        (dn * pb["g_depth_normal"]).sum().backward()
        out["d2n"], out["d2n_grad_depth"] = dn.detach().numpy(), d.grad.numpy()
        n = pb["normal"].clone().requires_grad_(True)
        cv = ns["normal2curv"](n, mask)
        (cv * pb["g_curv"]).sum().backward()
        out["n2c"], out["n2c_grad_normal"] = cv.detach().numpy(), n.grad.numpy()
        # the renderer's post block
        n = pb["normal"].clone().requires_grad_(True)
        d = pb["depth"].clone().requires_grad_(True)
        on, oc, od, om = post_block(ns, n * 1.0, d, pb["opac"], cam)
        ((on * pb["g_normal"]).sum() + (oc * pb["g_curv"]).sum() + (od * pb["g_depth_normal"]).sum()).backward()
        out.update(post_normal=on.detach().numpy(), post_curv=oc.detach().numpy(),
                   post_depth_normal=od.detach().numpy(), post_mask=om.numpy(),
                   post_grad_normal=n.grad.numpy(), post_grad_depth=d.grad.numpy())
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(name, "->", path, os.path.getsize(path), "bytes; covered",
              float(mask.float().mean()))


if __name__ == "__main__":
    main()
