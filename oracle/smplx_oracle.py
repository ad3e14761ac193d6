"""TEST INFRASTRUCTURE ONLY — numpy float32 restatement of the reference's vendored SMPL-X
forward for batch 1. Pinned by tests/golden/smplx_*.npz, which hold the outputs of the
reference's own `lbs()` (imported from /root/reference by tests/golden/make_golden_smplx.py).

Follows:
  * lbs()                    TS/utils/smplx/lbs.py:147-246
  * vertices2joints          :249-266      einsum("bik,ji->bjk")
  * blend_shapes             :269-290      einsum("bl,mkl->bmk")
  * batch_rodrigues          :293-328      angle = |v + 1e-8| (Q16)
  * transform_mat            :331-340
  * batch_rigid_transform    :343-396      A = G - pad(G @ [J;0])
  * SMPLX.forward tail       TS/utils/smplx/body_models.py:1380-1383  transl added to verts,
                             joints and A[:, :3, 3] (Q17)
  * bones                    TS/utils/smpl.py:609  cano2live = A @ inv_mats
"""
import numpy as np

f32 = np.float32


def batch_rodrigues(rot_vecs):
    rv = rot_vecs.astype(f32)
    angle = np.sqrt(((rv + f32(1e-8)) ** 2).sum(1, keepdims=True, dtype=f32)).astype(f32)
    d = (rv / angle).astype(f32)
    cos = np.cos(angle)[:, None].astype(f32)
    sin = np.sin(angle)[:, None].astype(f32)
    rx, ry, rz = d[:, 0], d[:, 1], d[:, 2]
    z = np.zeros_like(rx)
    K = np.stack([z, -rz, ry, rz, z, -rx, -ry, rx, z], 1).reshape(-1, 3, 3).astype(f32)
    return (np.eye(3, dtype=f32)[None] + sin * K + (f32(1) - cos) * (K @ K)).astype(f32)


def batch_rigid_transform(rot_mats, joints, parents):
    J = joints.shape[0]
    rel = joints.copy()
    rel[1:] -= joints[parents[1:]]
    T = np.zeros((J, 4, 4), dtype=f32)
    T[:, :3, :3] = rot_mats
    T[:, :3, 3] = rel
    T[:, 3, 3] = 1
    chain = [T[0]]
    for i in range(1, J):
        chain.append((chain[parents[i]] @ T[i]).astype(f32))
    G = np.stack(chain, 0)
    posed = G[:, :3, 3].copy()
    jh = np.concatenate([joints, np.zeros((J, 1), dtype=f32)], 1)[:, :, None]
    A = G.copy()
    A[:, :, 3] -= (G @ jh)[:, :, 0]
    return posed, A.astype(f32)


def lbs(betas, pose, v_template, shapedirs, posedirs, J_regressor, parents, lbs_weights):
    """betas [NB], pose [J*3] -> (verts [V,3], joints [J,3], A [J,4,4]); all float32."""
    betas, pose = betas.reshape(-1).astype(f32), pose.reshape(-1, 3).astype(f32)
    parents = np.asarray(parents, dtype=np.int64)
    v_shaped = (v_template + np.einsum("l,mkl->mk", betas, shapedirs)).astype(f32)
    Jl = np.einsum("ik,ji->jk", v_shaped, J_regressor).astype(f32)
    R = batch_rodrigues(pose)
    feat = (R[1:] - np.eye(3, dtype=f32)).reshape(-1)
    v_posed = ((feat @ posedirs).reshape(-1, 3) + v_shaped).astype(f32)
    posed, A = batch_rigid_transform(R, Jl, parents)
    T = (lbs_weights @ A.reshape(-1, 16)).reshape(-1, 4, 4).astype(f32)
    vh = np.concatenate([v_posed, np.ones((v_posed.shape[0], 1), dtype=f32)], 1)
    verts = np.einsum("vab,vb->va", T, vh)[:, :3].astype(f32)
    return verts, posed, A


def smplx_forward(model, betas, full_pose, transl=None, inv_mats=None):
    """-> dict(verts, joints, A[, bones]) with transl folded in like SMPLX.forward."""
    verts, joints, A = lbs(betas, full_pose, model["v_template"], model["shapedirs"],
                           model["posedirs"], model["J_regressor"], model["parents"],
                           model["lbs_weights"])
    if transl is not None:
        t = np.asarray(transl, dtype=f32).reshape(3)
        verts = verts + t
        joints = joints + t
        A = A.copy()
        A[:, :3, 3] += t
    out = dict(verts=verts, joints=joints, A=A)
    if inv_mats is not None:
        out["bones"] = (A @ inv_mats.astype(f32)).astype(f32)
    return out
