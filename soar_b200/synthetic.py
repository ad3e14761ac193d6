"""Deterministic synthetic inputs for tests and bench: SMPL-X assets and
videos are not available offline, so Gaussians, cameras, cotangents and an SMPL-X-shaped
skinning problem are generated from fixed seeds with numpy's PCG64 (bit-identical on
every machine), returned as CPU float32 torch tensors.

Cameras are built exactly like the reference does it
(TS/renderer/gaussian_batch_renderer.py:401-471: convert_pose, get_projection_matrix_gaussian,
get_cam_info_gaussian_cxcy), i.e. row-vector matrices that the kernels read column-major.
"""
import math

import numpy as np
import torch


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))


def look_at_c2w(eye, target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0)):
    """OpenGL-style camera-to-world (camera looks down -z, +y up), as threestudio produces."""
    eye, target, up = (np.asarray(v, dtype=np.float64) for v in (eye, target, up))
    f = target - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, up)
    r /= np.linalg.norm(r)
    u = np.cross(r, f)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = r, u, -f, eye
    return c2w


def camera(c2w, fovx, fovy, znear=0.1, zfar=100.0):
    """-> (world_view_transform [4,4], full_proj_transform [4,4], camera_center [3]) float32."""
    flip = np.diag([1.0, -1.0, -1.0, 1.0])
    w2c = np.linalg.inv(c2w @ flip)
    world_view = w2c.T.astype(np.float32)
    top = math.tan(fovy / 2) * znear
    right = math.tan(fovx / 2) * znear
    Pm = np.zeros((4, 4), dtype=np.float32)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = (zfar + znear) / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    full = (world_view @ Pm.T).astype(np.float32)
    center = np.linalg.inv(world_view.astype(np.float64))[3, :3].astype(np.float32)
    return _t(world_view), _t(full), _t(center)


def ring_cameras(n_views, W, H, seed=0, fovy_deg=40.0, dist=(2.5, 3.0), elev_deg=(-10.0, 20.0)):
    """n_views cameras on a ring of azimuths looking at the origin (SDS-style, §8d)."""
    rng = np.random.Generator(np.random.PCG64(seed + 7919))
    fovy = math.radians(fovy_deg)
    fovx = 2 * math.atan(math.tan(fovy / 2) * W / H)
    cams = []
    for v in range(n_views):
        az = 2 * math.pi * v / max(n_views, 1) + rng.uniform(-0.05, 0.05)
        el = math.radians(rng.uniform(*elev_deg))
        d = rng.uniform(*dist)
        eye = (d * math.cos(el) * math.sin(az), d * math.sin(el), d * math.cos(el) * math.cos(az))
        view, proj, center = camera(look_at_c2w(eye), fovx, fovy)
        cams.append(dict(viewmatrix=view, projmatrix=proj, campos=center,
                         tanfovx=math.tan(fovx / 2), tanfovy=math.tan(fovy / 2)))
    return cams


def gaussians(P, sh_degree=None, seed=0, surfel=False):
    """"random" distribution of §8d. sh_degree=None -> precomputed colours.
    surfel=True reproduces SOAR's usage: opacity 1, scale.z = -1e10 (Q7), isotropic in-plane."""
    rng = np.random.Generator(np.random.PCG64(seed))
    means = rng.uniform(-1.0, 1.0, size=(P, 3)) * 0.9
    log_s = rng.uniform(math.log(0.003), math.log(0.03), size=(P, 3))
    scales = np.exp(log_s)
    q = rng.normal(size=(P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-rng.normal(scale=1.5, size=(P, 1))))
    out = dict(means3D=_t(means), rotations=_t(q))
    if surfel:
        scales[:, 1] = scales[:, 0]
        scales[:, 2] = -1e10
        opac = np.ones((P, 1))
    out["scales"] = _t(scales)
    out["opacities"] = _t(opac)
    if sh_degree is None:
        out["colors_precomp"] = _t(rng.uniform(0.0, 1.0, size=(P, 3)))
        out["shs"] = None
    else:
        M = (sh_degree + 1) ** 2
        sh = rng.normal(scale=0.3, size=(P, M, 3))
        sh[:, 0, :] = rng.normal(scale=1.0, size=(P, 3))
        out["shs"] = _t(sh)
        out["colors_precomp"] = None
    return out


def cotangents(W, H, seed=0):
    """Fixed-seed N(0,1) cotangents for all four image outputs (§8d loss)."""
    rng = np.random.Generator(np.random.PCG64(seed + 104729))
    return dict(color=_t(rng.normal(size=(3, H, W))), normal=_t(rng.normal(size=(3, H, W))),
                depth=_t(rng.normal(size=(1, H, W))), opac=_t(rng.normal(size=(1, H, W))))


def scene(P, W, H, sh_degree=None, seed=0, surfel=False, config=None, bg=(0.0, 0.0, 0.0),
          sort_descending=False, render_front=False, view=0, n_views=1):
    """One view of a synthetic scene: everything GaussianRasterizationSettings and the
    rasterizer call need, as CPU tensors / python scalars."""
    g = gaussians(P, sh_degree=sh_degree, seed=seed, surfel=surfel)
    cam = ring_cameras(max(n_views, 1), W, H, seed=seed)[view]
    if config is None:
        config = (1.0, 1.0, 1.0, 0.0) if surfel else (0.0, 0.0, 0.0, 0.0)
    g.update(cam)
    g.update(dict(image_height=H, image_width=W, bg=_t(bg), scale_modifier=1.0,
                  patch_bbox=_t([0.0, 0.0, float(H), float(W)]), prcppoint=_t([0.5, 0.5]),
                  sh_degree=0 if sh_degree is None else sh_degree, prefiltered=False,
                  render_front=render_front, sort_descending=sort_descending, debug=False,
                  config=_t(config)))
    return g


# ---------------------------------------------------------------------------------------------
# SMPL-X-shaped skinning problem (55 joints, 10,475 canonical vertices)
# ---------------------------------------------------------------------------------------------
SMPLX_PARENTS = [-1, 0, 0, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 9, 9, 12, 13, 14, 16, 17, 18, 19, 15, 15,
                 15, 20, 25, 26, 20, 28, 29, 20, 31, 32, 20, 34, 35, 20, 37, 38, 21, 40, 41, 21, 43,
                 44, 21, 46, 47, 21, 49, 50, 21, 52, 53]
N_JOINTS = 55
N_VERTS = 10475


def _rodrigues(rv):
    ang = np.linalg.norm(rv + 1e-8, axis=-1, keepdims=True)  # Q16: eps added per component
    ax = rv / ang
    c, s = np.cos(ang)[..., None], np.sin(ang)[..., None]
    K = np.zeros(rv.shape[:-1] + (3, 3))
    K[..., 0, 1], K[..., 0, 2] = -ax[..., 2], ax[..., 1]
    K[..., 1, 0], K[..., 1, 2] = ax[..., 2], -ax[..., 0]
    K[..., 2, 0], K[..., 2, 1] = -ax[..., 1], ax[..., 0]
    return np.eye(3) + s * K + (1 - c) * (K @ K)


def skinning_problem(N, seed=0):
    """Synthetic humanoid: joints along a capsule skeleton, `N` canonical Gaussians near the
    bones, skin weights = top-4 softmax(-d^2/sigma^2) over joints (sigma 0.08 m), bone
    transforms cano->live from a random pose pushed through the SMPL-X kinematic tree
    (TS/utils/smplx/lbs.py:343-396) with transl (0, 0.3, 0) folded in (Q17)."""
    rng = np.random.Generator(np.random.PCG64(seed + 31337))
    J = N_JOINTS
    joints = np.zeros((J, 3))
    for j in range(1, J):
        p = SMPLX_PARENTS[j]
        step = rng.normal(size=3)
        step[1] = -abs(step[1]) if j in (1, 2, 4, 5, 7, 8, 10, 11) else abs(step[1])
        step /= np.linalg.norm(step)
        joints[j] = joints[p] + step * (0.12 if j < 22 else 0.03)
    seg = rng.integers(1, J, size=N)
    t = rng.uniform(size=(N, 1))
    pts = joints[np.asarray(SMPLX_PARENTS)[seg]] * (1 - t) + joints[seg] * t
    pts += rng.normal(scale=0.03, size=(N, 3))
    d2 = ((pts[:, None, :] - joints[None]) ** 2).sum(-1)
    logits = -d2 / (0.08 ** 2)
    kth = np.partition(logits, -4, axis=1)[:, -4][:, None]
    w = np.where(logits >= kth, np.exp(logits - logits.max(1, keepdims=True)), 0.0)
    w /= w.sum(1, keepdims=True)
    pose = rng.normal(scale=0.3, size=(J, 3))
    pose[0] = 0.0
    Rm = _rodrigues(pose)
    G = np.zeros((J, 4, 4))
    for j in range(J):
        local = np.eye(4)
        local[:3, :3] = Rm[j]
        local[:3, 3] = joints[j] - (joints[SMPLX_PARENTS[j]] if j else 0.0)
        G[j] = local if j == 0 else G[SMPLX_PARENTS[j]] @ local
    A = G.copy()
    A[:, :3, 3] -= np.einsum("jab,jb->ja", G[:, :3, :3], joints)
    A[:, :3, 3] += np.array([0.0, 0.3, 0.0])
    q = rng.normal(size=(N, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return dict(points=_t(pts), quats=_t(q), weights=_t(w), bones=_t(A))


def _skeleton(rng):
    """55 rest-pose joint positions of a capsule humanoid along the SMPL-X kinematic tree."""
    J = N_JOINTS
    joints = np.zeros((J, 3))
    for j in range(1, J):
        p = SMPLX_PARENTS[j]
        step = rng.normal(size=3)
        step[1] = -abs(step[1]) if j in (1, 2, 4, 5, 7, 8, 10, 11) else abs(step[1])
        step /= np.linalg.norm(step)
        joints[j] = joints[p] + step * (0.12 if j < 22 else 0.03)
    return joints


def smplx_template(V=N_VERTS, NB=20, seed=0):
    """SMPL-X-shaped body model with synthetic assets (§8d; the real SMPLX_*.npz are behind a
    login): v_template [V,3] on capsules around the skeleton, shapedirs [V,3,NB] ~ N(0,0.01^2),
    posedirs [(J-1)*9, V*3] ~ N(0,0.001^2), J_regressor [J,V] = row-normalised indicator of each
    joint's 64 nearest vertices, parents [J], lbs_weights [V,J] = top-4 softmax(-d^2/0.08^2)."""
    rng = np.random.Generator(np.random.PCG64(seed + 4241))
    J = N_JOINTS
    joints = _skeleton(rng)
    seg = rng.integers(1, J, size=V)
    t = rng.uniform(size=(V, 1))
    verts = joints[np.asarray(SMPLX_PARENTS)[seg]] * (1 - t) + joints[seg] * t
    verts += rng.normal(scale=0.03, size=(V, 3))
    d2 = ((verts[:, None, :] - joints[None]) ** 2).sum(-1)
    logits = -d2 / (0.08 ** 2)
    kth = np.partition(logits, -4, axis=1)[:, -4][:, None]
    w = np.where(logits >= kth, np.exp(logits - logits.max(1, keepdims=True)), 0.0)
    w /= w.sum(1, keepdims=True)
    k = min(64, V)
    near = np.argpartition(d2, k - 1, axis=0)[:k]  # [k, J]
    Jreg = np.zeros((J, V))
    for j in range(J):
        Jreg[j, near[:, j]] = 1.0 / k
    shapedirs = rng.normal(scale=0.01, size=(V, 3, NB))
    posedirs = rng.normal(scale=0.001, size=((J - 1) * 9, V * 3))
    return dict(v_template=_t(verts), shapedirs=_t(shapedirs), posedirs=_t(posedirs),
                J_regressor=_t(Jreg), parents=torch.tensor(SMPLX_PARENTS, dtype=torch.int64),
                lbs_weights=_t(w))


def smplx_pose(NB=20, seed=0, global_orient_zero=True):
    """betas [1,NB] ~ N(0,0.5^2), full_pose [1,165] axis-angles ~ N(0,0.3^2), transl (0,0.3,0)."""
    rng = np.random.Generator(np.random.PCG64(seed + 90001))
    betas = rng.normal(scale=0.5, size=(1, NB))
    pose = rng.normal(scale=0.3, size=(N_JOINTS, 3))
    if global_orient_zero:
        pose[0] = 0.0
    return dict(betas=_t(betas), full_pose=_t(pose.reshape(1, -1)), transl=_t([[0.0, 0.3, 0.0]]))


class ImageCamera:
    """The five camera attributes the image-space post-ops read (the reference passes its
    Camera NamedTuple, TS/renderer/gaussian_batch_renderer.py:73-83)."""

    def __init__(self, W, H, fovx, fovy, prcppoint=(0.5, 0.5)):
        self.image_width, self.image_height = int(W), int(H)
        self.FoVx, self.FoVy = float(fovx), float(fovy)
        self.prcppoint = _t(np.asarray(prcppoint))


def postops_problem(W, H, seed=0, prcppoint=(0.5, 0.5), fovy_deg=40.0, touch_border=False):
    """Rasterizer-shaped image outputs for the post-ops (row f2): a bumpy blob seen from ~2.7 m —
    `opac` [1,H,W] exactly 0 outside a wobbly silhouette (with a hole) and in (0.2, 1] inside,
    `depth` [1,H,W] smooth + noise inside, 0 outside, `normal` [3,H,W] roughly unit inside.
    touch_border=True lets the silhouette run over the image border (replicate-padding paths).
    Plus N(0,1) cotangents for normal / curv / depth_normal."""
    rng = np.random.Generator(np.random.PCG64(seed + 7919))
    ys, xs = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    cx, cy = (0.5 * W, 0.5 * H) if not touch_border else (0.3 * W, 0.65 * H)
    rad = (0.36 if not touch_border else 0.6) * min(W, H)
    ang = np.arctan2(ys - cy, xs - cx)
    rr = np.hypot(xs - cx, ys - cy) / (rad * (1 + 0.15 * np.sin(5 * ang + rng.uniform(0, 6.28))))
    inside = (rr < 1.0) & ~(np.hypot(xs - cx - 0.3 * rad, ys - cy + 0.2 * rad) < 0.12 * rad)
    inside &= rng.uniform(size=(H, W)) > 0.01          # pin holes
    z = np.sqrt(np.clip(1.0 - rr * rr, 0.0, None))
    depth = 2.7 - 0.5 * z + 0.02 * np.sin(xs * 0.21) * np.cos(ys * 0.17) + rng.normal(size=(H, W)) * 2e-3
    n = np.stack([(xs - cx) / rad, (ys - cy) / rad, -z - 0.2], 0) + rng.normal(size=(3, H, W)) * 0.05
    n /= np.linalg.norm(n, axis=0, keepdims=True)
    n *= 1.0 + rng.normal(size=(1, H, W)) * 0.02       # blended normals are not exactly unit
    opac = np.where(inside, np.clip(1.2 - rr, 0.2, 1.0), 0.0)
    fovy = math.radians(fovy_deg)
    fovx = 2 * math.atan(math.tan(fovy / 2) * W / H)
    return dict(
        normal=_t(n * inside), depth=_t((depth * inside)[None]), opac=_t(opac[None]),
        camera=ImageCamera(W, H, fovx, fovy, prcppoint),
        g_normal=_t(rng.normal(size=(3, H, W))), g_curv=_t(rng.normal(size=(1, H, W))),
        g_depth_normal=_t(rng.normal(size=(3, H, W))))


def avatar_gaussians(N, v_template, seed=0):
    """"avatar" distribution of §8d on the synthetic template: canonical Gaussian centres on the
    body surface (a random template vertex + N(0, 4 mm) jitter), random unit quaternions (the
    rotation PARAMETER, scaled so it is not unit: the renderer normalises it), colours U(0,1),
    occupancy sigmoid(N(0,1)). The in-plane scale is derived from distCUDA2 by the caller
    (U(0.5,1.5)·sqrt(dist2) clipped to [1e-4, 0.02]), `scale_jitter` holds the U(0.5,1.5) factor."""
    rng = np.random.Generator(np.random.PCG64(seed + 60013))
    vt = v_template.numpy() if torch.is_tensor(v_template) else np.asarray(v_template)
    idx = rng.integers(0, vt.shape[0], size=N)
    pts = vt[idx] + rng.normal(scale=0.004, size=(N, 3))
    q = rng.normal(size=(N, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    q *= rng.uniform(0.8, 1.25, size=(N, 1))
    return dict(points=_t(pts), rot=_t(q), colors=_t(rng.uniform(0.0, 1.0, size=(N, 3))),
                occ=_t(1.0 / (1.0 + np.exp(-rng.normal(size=(N, 1))))),
                scale_jitter=_t(rng.uniform(0.5, 1.5, size=(N, 1))))


def densify_problem(N, R=3, seed=0, n_adam=3, n_views=2):
    """Per-Gaussian parameter groups as the reference holds them (row f3) plus a scripted sequence of
    gradients: xyz [N,3], f_dc [N,1,3], f_rest [N,R,3], color [N,3], opacity [N,1] (logit), scaling
    [N,3] (log, z = -1e10 like the surfels), rotation [N,4]; `adam_grads`: n_adam dicts of gradients;
    `views`: n_views × (viewspace_grad [N,3], radii [N] int32 with ~30 % culled, scaling_grad [N,3])."""
    rng = np.random.Generator(np.random.PCG64(seed + 15485863))
    sc = np.log(np.exp(rng.uniform(math.log(0.002), math.log(0.05), size=(N, 3))))
    sc[:, 2] = -1e10
    q = rng.normal(size=(N, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    q *= rng.uniform(0.8, 1.25, size=(N, 1))
    params = dict(xyz=_t(rng.uniform(-1, 1, size=(N, 3))), f_dc=_t(rng.normal(size=(N, 1, 3))),
                  f_rest=_t(rng.normal(scale=0.3, size=(N, R, 3))), color=_t(rng.normal(size=(N, 3))),
                  opacity=_t(rng.normal(loc=-1.0, scale=2.0, size=(N, 1))), scaling=_t(sc), rotation=_t(q))
    shapes = {k: tuple(v.shape[1:]) for k, v in params.items()}

    def grads(n):
        g = {k: _t(rng.normal(scale=0.1, size=(n,) + s)) for k, s in shapes.items()}
        g["scaling"][:, 2] = 0.0
        return g

    def view(n):
        radii = rng.integers(1, 40, size=n).astype(np.int32)
        radii[rng.uniform(size=n) < 0.3] = 0
        sg = rng.normal(scale=1e-3, size=(n, 3))
        sg[:, 2] = 0.0
        return _t(rng.normal(size=(n, 3))), torch.from_numpy(radii), _t(sg)

    return dict(params=params, adam_grads=[grads(N) for _ in range(n_adam)],
                views=[view(N) for _ in range(n_views)], grads=grads, view=view,
                lrs=dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=1.25e-4, color=2.5e-3, opacity=0.05,
                         scaling=5e-3, rotation=1e-3))
