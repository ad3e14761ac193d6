#!/usr/bin/env python
"""Generates tests/golden/smplx_*.npz by running the reference's OWN vendored SMPL-X code
(`lbs()`, soar/threestudio-soar/utils/smplx/lbs.py:147-246, imported from /root/reference) on the
synthetic body model of soar_b200.synthetic.smplx_template. The reference ships no fixtures
(SURVEY.md §4); these pin oracle/smplx_oracle.py and the CUDA path (soar_smplx_forward).

Run in the build container (the python reference cannot travel to the GPU box):
    python tests/golden/make_golden_smplx.py
The big inputs (posedirs is 61 MB at full size) are NOT stored: they are regenerated from the
seed by the same deterministic generator (numpy PCG64); the small case stores its inputs too so a
generator change cannot go unnoticed."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF_UTILS = "/root/reference/soar/threestudio-soar/utils"
sys.path.insert(0, REF_UTILS)

from soar_b200 import synthetic  # noqa: E402
from smplx.lbs import lbs  # noqa: E402  (the reference's vendored smplx package)

CASES = {
    # name: (V, NB, template seed, pose seed, store inputs)
    "smplx_small": (96, 10, 1, 1, True),
    "smplx_full": (synthetic.N_VERTS, 20, 0, 0, False),
    "smplx_full_orient": (synthetic.N_VERTS, 20, 0, 5, False),
}


def main():
    torch.manual_seed(0)
    torch.set_num_threads(1)  # fixed reduction order inside the torch matmuls
    for name, (V, NB, tseed, pseed, store) in CASES.items():
        m = synthetic.smplx_template(V=V, NB=NB, seed=tseed)
        p = synthetic.smplx_pose(NB=NB, seed=pseed, global_orient_zero=(name != "smplx_full_orient"))
        verts, joints, A = lbs(p["betas"], p["full_pose"], m["v_template"], m["shapedirs"],
                               m["posedirs"], m["J_regressor"], m["parents"], m["lbs_weights"],
                               pose2rot=True, return_affine_mat=True)
        # SMPLX.forward tail (body_models.py:1380-1383): transl folded into verts, joints, A
        t = p["transl"]
        verts = verts + t.unsqueeze(1)
        joints = joints + t.unsqueeze(1)
        A = A.clone()
        A[:, :, :3, 3] += t.unsqueeze(1)
        # canonical pose for the bones test: zero pose, same betas (TS/utils/smpl.py:249)
        zero = torch.zeros_like(p["full_pose"])
        _, _, A0 = lbs(p["betas"], zero, m["v_template"], m["shapedirs"], m["posedirs"],
                       m["J_regressor"], m["parents"], m["lbs_weights"], pose2rot=True,
                       return_affine_mat=True)
        inv = torch.linalg.inv(A0)
        bones = torch.matmul(A, inv)  # TS/utils/smpl.py:609
        out = dict(V=V, NB=NB, template_seed=tseed, pose_seed=pseed,
                   verts=verts[0].numpy(), joints=joints[0].numpy(), A=A[0].numpy(),
                   inv_mats=inv[0].numpy(), bones=bones[0].numpy(),
                   betas=p["betas"].numpy(), full_pose=p["full_pose"].numpy(), transl=t.numpy())
        if store:
            out.update({"in_" + k: v.numpy() for k, v in m.items()})
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: wrote {path} ({os.path.getsize(path)} bytes)")


if __name__ == "__main__":
    main()
