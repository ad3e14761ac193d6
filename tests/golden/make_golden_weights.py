#!/usr/bin/env python
"""Generates tests/golden/weights_*.npz and quat2mat.npz by executing the reference's OWN lines:

  * `SMPL_Guidance.query_weights_smpl` (soar/threestudio-soar/utils/smpl.py:618-637): the method is
    taken out of the class with `ast` at run time and exec'd unchanged on a stub `self`. The ONE
    substituted symbol is `knn_points` (pytorch3d.ops — an unvendored git dependency, absent offline):
    its result is fully specified (for every query the K nearest points, squared distances in
    ascending order, float32), so an exact stand-in computes it from the float32 coordinate
    differences — (dx*dx + dy*dy) + dz*dz, the left-to-right float32 sum that pytorch3d's kernels
    form — with a stable ascending sort (ties in distance -> lower vertex index). Everything after it
    (sqrt, the in-place clamp to [1e-4, 1], the inverse-distance weights, the normalisation, the
    gather and blend of the skin-weight rows) is the reference's own code.
  * `quaternion_to_matrix` as vendored at soar/threestudio-soar/data/uncond_multiview.py:2422-2451
    (pytorch3d's function, copied into the reference tree by its authors), exec'd unchanged.

`matrix_to_quaternion` (renderer/diff_gaussian_rasterizer.py:110-113, :146-148) has no copy in the
reference tree: it stays "parity unpinned" (DESIGN.md §6).

Run in the build container (the python reference cannot travel to the GPU box):
    python tests/golden/make_golden_weights.py
"""
import ast
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
SMPL_FILE = "/root/reference/soar/threestudio-soar/utils/smpl.py"
DATA_FILE = "/root/reference/soar/threestudio-soar/data/uncond_multiview.py"

from soar_b200 import synthetic  # noqa: E402


def knn_points(p1, p2, K=1, **kw):
    """Exact stand-in for pytorch3d.ops.knn_points(p1[1,N,3], p2[1,V,3], K) -> (dists[1,N,K] squared,
    ascending; idx[1,N,K]; None)."""
    assert p1.dim() == 3 and p1.shape[0] == 1 and p2.shape[0] == 1 and not kw
    x, y = p1[0].float(), p2[0].float()
    dists, idxs = [], []
    for s in range(0, x.shape[0], 2048):
        d = x[s:s + 2048, None, :] - y[None, :, :]
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        v, i = torch.sort(d2, dim=1, stable=True)
        dists.append(v[:, :K])
        idxs.append(i[:, :K])
    return torch.cat(dists)[None], torch.cat(idxs)[None], None


def reference_method(path, cls, name, ns):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls:
            for item in node.body:
                if isinstance(item, ast.FunctionDef) and item.name == name:
                    exec(compile(ast.Module([item], []), path, "exec"), ns)
                    return ns[name]
    raise RuntimeError(f"{cls}.{name} not found in {path}")


def reference_function(path, name, ns):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
            return ns[name]
    raise RuntimeError(f"{name} not found in {path}")


class Stub:
    """What query_weights_smpl reads from `self` (TS/utils/smpl.py:249-258)."""

    def __init__(self, cano_vertices, ori_lbs):
        self.cano_vertices = cano_vertices
        self.ori_lbs = ori_lbs


def template(V, seed, duplicates):
    m = synthetic.smplx_template(V=V, NB=1, seed=seed)
    verts, lw = m["v_template"].clone(), m["lbs_weights"].clone()
    if duplicates:
        verts[V // 2: V // 2 + 5] = verts[0]     # duplicated vertices: exact distance ties
    return verts, lw


def queries(verts, n, seed):
    """on-vertex, near-surface, off-body, far and duplicated-vertex queries."""
    gen = torch.Generator().manual_seed(seed)
    V = verts.shape[0]
    src = verts[torch.randint(0, V, (n,), generator=gen)]
    pts = src + 0.004 * torch.randn(n, 3, generator=gen)           # near the surface
    q = n // 5
    pts[:q] = src[:q]                                                # exactly on a vertex (clamp 1e-4)
    pts[q:2 * q] = src[q:2 * q] + 0.3 * torch.randn(q, 3, generator=gen)   # off the body
    pts[2 * q:2 * q + q // 2] = 4.0 * torch.randn(q // 2, 3, generator=gen)  # > 1 m away (clamp 1.0)
    pts[2 * q + q // 2] = verts[0]                                   # on the duplicated vertex
    return pts.contiguous()


CASES = {
    # name: (V, template seed, duplicated vertices, N queries, query seed)
    "weights_full": (synthetic.N_VERTS, 6, False, 600, 1),
    "weights_dups": (3000, 9, True, 500, 2),
    "weights_tiny": (40, 3, True, 64, 3),
}


def main():
    fn = reference_method(SMPL_FILE, "SMPL_Guidance", "query_weights_smpl",
                          {"torch": torch, "knn_points": knn_points})
    for name, (V, tseed, dups, N, qseed) in CASES.items():
        verts, lw = template(V, tseed, dups)
        pts = queries(verts, N, qseed)
        w = fn(Stub(verts, lw[None]), pts)               # ori_lbs is [1,V,J] in the reference (:258)
        assert w.shape == (N, lw.shape[1])
        digest = hashlib.sha256(verts.numpy().tobytes() + lw.numpy().tobytes()).hexdigest()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), points=pts.numpy(), weights=w.numpy(),
                            V=V, template_seed=tseed, duplicates=int(dups), template_sha256=digest)
        print(name, "N", N, "V", V, "row sums", float(w.sum(1).min()), float(w.sum(1).max()))
    q2m = reference_function(DATA_FILE, "quaternion_to_matrix", {"torch": torch})
    gen = torch.Generator().manual_seed(5)
    q = torch.randn(512, 4, generator=gen) * torch.rand(512, 1, generator=gen).mul(3).add(0.1)  # NOT unit
    q[0] = torch.tensor([1.0, 0.0, 0.0, 0.0])
    q[1] = torch.tensor([0.0, 0.0, 2.0, 0.0])
    R = q2m(q)
    np.savez_compressed(os.path.join(HERE, "quat2mat.npz"), quats=q.numpy(), matrices=R.numpy())
    print("quat2mat", tuple(R.shape))


if __name__ == "__main__":
    main()
