"""GPU parity of the composed hot path (BASELINE configs 4 and 5): SMPL-X pose -> skin-weight query
-> skinning -> surfel rasterizer (+ occupancy pass) -> image-space post-ops -> backward, through
soar_b200.avatar.AvatarRenderer (the mirror of the reference's DiffGaussian.forward).

Small sizes: the whole chain, forward and gradients, against the chain of checkers — the torch
restatement of the skinning (oracle/lbs_oracle.py), the UNMODIFIED reference CUDA rasterizer
(oracle/_ref) and the post-op oracle (pinned to the reference's functions).
Full sizes (2M Gaussians at 1920x1080; 32 SDS views at 512²): the rasterizer stage against the
reference CUDA code on the skinned Gaussians, plus size-independent properties."""
import math

import numpy as np
import pytest
import torch

import refdgr
from soar_b200 import synthetic
from util import rel_err

pytestmark = pytest.mark.gpu
IMG_TOL, GRAD_TOL = 1e-4, 1e-3


def _avatar(N, W, H, seed=0, V=synthetic.N_VERTS, n_views=1, sds=False):
    from simple_knn._C import distCUDA2
    from soar_b200.avatar import AvatarCamera, AvatarRenderer
    from soar_b200.smplx_fwd import SmplxPoser
    m = synthetic.smplx_template(V=V, NB=10, seed=seed)
    keys = ("v_template", "shapedirs", "posedirs", "J_regressor", "parents", "lbs_weights")
    poser = SmplxPoser(*(m[k].cuda() for k in keys))
    ps = synthetic.smplx_pose(NB=10, seed=seed)
    zero = torch.zeros_like(ps["full_pose"])
    cano_verts, _, A0, _ = poser(ps["betas"].cuda(), zero.cuda(), None)        # TS/utils/smpl.py:249
    inv = torch.linalg.inv(A0)
    ar = AvatarRenderer(poser, cano_verts, inv)
    g = synthetic.avatar_gaussians(N, cano_verts.cpu(), seed=seed)
    pts = g["points"].cuda()
    d2 = distCUDA2(pts)
    scale2d = (g["scale_jitter"].cuda() * d2.sqrt()[:, None]).clamp(1e-4, 0.02)
    cams = []
    for c in synthetic.ring_cameras(n_views, W, H, seed=seed, dist=(1.6, 2.0)):
        fovy = 2 * math.atan(c["tanfovy"]); fovx = 2 * math.atan(c["tanfovx"])
        cams.append(AvatarCamera(fovx, fovy, W, H, c["viewmatrix"].cuda(), c["projmatrix"].cuda(),
                                 c["campos"].cuda(), torch.tensor([0.5, 0.5], device="cuda")))
    bones = ar.pose(ps["betas"].cuda(), ps["full_pose"].cuda(), ps["transl"].cuda())
    return ar, g, pts, scale2d, bones, cams


def _cots(W, H, seed):
    gen = torch.Generator("cuda").manual_seed(seed)
    r = lambda c: torch.randn(c, H, W, device="cuda", generator=gen)
    return dict(render=r(3), normal=r(3), depth=r(1), mask=r(1), curv=r(1), pred_normal=r(3), occ=r(3))


def _loss(out, cot):
    return sum((out[k] * cot[k]).sum() for k in cot if out.get(k) is not None)


def _scene_of(ar, cam, means, rots, scales, colors, bg, front, descending=None):
    """front = the render_front SETTING; descending defaults to the occupancy / legacy pairing."""
    descending = (not front) if descending is None else descending
    return dict(means3D=means, rotations=rots, scales=scales, colors_precomp=colors, shs=None,
                opacities=torch.ones((means.shape[0], 1), device="cuda"), bg=bg,
                viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
                campos=cam.camera_center, prcppoint=cam.prcppoint, patch_bbox=cam.full_patch(),
                tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5),
                image_height=cam.image_height, image_width=cam.image_width, scale_modifier=1.0,
                sh_degree=0, prefiltered=False, render_front=front, sort_descending=descending,
                debug=False, config=ar.config)


def _ours(ar, g, pts, scale2d, bones, cam, cot, sds, weights=None, render_front=True):
    leaves = dict(points=pts.clone().requires_grad_(True), rot=g["rot"].cuda().requires_grad_(True),
                  colors=g["colors"].cuda().requires_grad_(True),
                  scale2d=scale2d.clone().requires_grad_(True))
    w = ar.weights(leaves["points"]) if weights is None else weights
    bg = torch.zeros(3, device="cuda")
    out = ar.render(leaves["points"], leaves["rot"], w, bones, leaves["colors"], leaves["scale2d"],
                    g["occ"].cuda(), cam, bg, sds_view=sds, render_front=render_front)
    _loss(out, cot).backward()
    return out, {k: v.grad for k, v in leaves.items()}, w


def _mute_degenerate(ar, g, pts, scale2d, bones, cam, cot, sds, render_front=True):
    """Zero the pred_normal cotangent on degenerate depth2normal stencils: there the reference's
    F.normalize backward makes 1e12-scale terms whose cancellation is order-dependent noise
    (oracle/postops_oracle.py header), which would swamp a 1e-3 comparison of the chain."""
    with torch.no_grad():
        o = ar.render(pts, g["rot"].cuda(), ar.weights(pts), bones, g["colors"].cuda(), scale2d,
                      g["occ"].cuda(), cam, torch.zeros(3, device="cuda"), sds_view=sds, render_occ=False,
                      render_front=render_front)
        deg = (o["mask"][0] > 1e-5) & ((o["pred_normal"] * 2 - 1).norm(dim=0) < 0.5)
        cot["pred_normal"][:, deg] = 0
    return deg


def _checker_chain(ar, g, pts, scale2d, bones, cam, cot, sds, render_front=True):
    """The same step from the checkers: torch skinning oracle (CPU, autograd) -> reference CUDA
    rasterizer fwd/bwd -> post-op oracle (torch on the device, autograd)."""
    from oracle import lbs_oracle, postops_oracle
    p = pts.cpu().clone().requires_grad_(True)
    r = g["rot"].clone().requires_grad_(True)
    w = lbs_oracle.query_weights(p.detach(), ar.query.verts.cpu(), ar.query.weights.cpu())
    sp, sq, _ = lbs_oracle.skin(p, torch.nn.functional.normalize(r, dim=-1), w, bones.cpu(), sds)
    scales = scale2d.repeat(1, 3)
    scales[:, 2] = -1e10
    bg = torch.zeros(3, device="cuda")
    # The rasterizer checker gets the product's skinned Gaussians (compared with the oracle's right
    # here): a 1e-6 difference in a centre may swap two depth-sorted neighbours or move a depth
    # sample, which the differencing post-ops amplify — that is conditioning, not a parity error.
    from soar_b200.lbs import skin_gaussians
    with torch.no_grad():
        op, oq, _ = skin_gaussians(pts, torch.nn.functional.normalize(g["rot"].cuda(), dim=-1),
                                   ar.weights(pts), bones, swizzle=sds)
    assert (op.cpu() - sp.detach()).abs().max().item() <= 5e-6, "skinned points differ from the oracle"
    assert (oq.cpu() - sq.detach()).abs().max().item() <= 5e-5, "skinned quaternions differ from the oracle"
    # the main pass: render_front SETTING False, sorted descending iff the render_front ARGUMENT is
    # False (TS/renderer/diff_gaussian_rasterizer.py:186-187)
    s = _scene_of(ar, cam, op.contiguous(), oq.contiguous(), scales, g["colors"].cuda(), bg, front=False,
                  descending=not render_front)
    ref = refdgr.RefRasterizer()
    ro = ref.forward(s)
    img = {k: ro[k].clone().requires_grad_(True) for k in ("color", "normal", "depth", "opac")}
    on, oc, od, _ = postops_oracle.surfel_postprocess(img["normal"], img["depth"], img["opac"], cam)
    out = dict(render=img["color"], normal=on, depth=img["depth"], mask=img["opac"], curv=oc,
               pred_normal=od)
    _loss(out, {k: v for k, v in cot.items() if k != "occ"}).backward()
    gr = ref.backward(dict(color=img["color"].grad, normal=img["normal"].grad,
                           depth=img["depth"].grad, opac=img["opac"].grad))
    torch.autograd.backward([sp, sq], [gr["means3D"].cpu(), gr["rotations"].cpu()])
    # occupancy pass (forward only; its inputs are detached in the reference, :281-291)
    so = _scene_of(ar, cam, s["means3D"], s["rotations"], scales, g["occ"].cuda().repeat(1, 3), bg, front=True)
    ref2 = refdgr.RefRasterizer()
    occ = ref2.forward(so)["color"]
    grads = dict(points=p.grad, rot=r.grad, colors=gr["colors"].cpu(),
                 scale2d=gr["scales"][:, :2].sum(1, keepdim=True).cpu())
    outs = {k: v.detach() for k, v in out.items()}
    outs.update(occ=occ, radii=ro["radii"], weights=w)
    ref.close(); ref2.close()
    return outs, grads


@pytest.mark.parametrize("render_front", [True, False])
@pytest.mark.parametrize("sds", [False, True])
def test_full_chain_matches_checker_chain(sds, render_front):
    if not refdgr.available():
        pytest.fail("oracle/_ref/libref_dgr.so missing: run oracle/build_ref.sh where /root/reference is mounted")
    W, H, N = 320, 256, 20_000
    ar, g, pts, scale2d, bones, cams = _avatar(N, W, H, seed=3, V=2000)
    cot = _cots(W, H, seed=11)
    _mute_degenerate(ar, g, pts, scale2d, bones, cams[0], cot, sds, render_front)
    out, grads, w = _ours(ar, g, pts, scale2d, bones, cams[0], cot, sds, render_front=render_front)
    ro, rg = _checker_chain(ar, g, pts, scale2d, bones, cams[0], cot, sds, render_front)
    assert (w.cpu() - ro["weights"]).abs().max().item() <= 2e-5
    assert torch.equal(out["radii"], ro["radii"]), "radii differ from the reference"
    assert (out["radii"] > 0).float().mean().item() > 0.2   # random surfel orientations: about half face away
    for k in ("render", "normal", "depth", "mask", "curv", "pred_normal", "occ"):
        err = (out[k].detach() - ro[k]).abs().max().item()
        tol = IMG_TOL * (4 if k == "curv" else 1)   # curvature sums four differences of 1e-4-accurate normals
        assert err <= tol, f"{k}: max-abs {err:.2e}"
    for k in ("points", "rot", "colors", "scale2d"):
        e = rel_err(grads[k].cpu(), rg[k])
        assert e <= GRAD_TOL, f"dL/d{k}: relative {e:.2e}"


def test_c5_dance_step_2m_gaussians_1080p():
    """BASELINE config 5: 2M skinned Gaussians at 1920x1080 — pose, weight query, skinning, both
    rasterizer passes, post-ops, backward, KNN (distCUDA2, inside _avatar). The rasterizer stage is
    checked against the reference CUDA code on the same skinned Gaussians; the rest by properties."""
    from simple_knn._C import distCUDA2
    W, H, N = 1920, 1080, 2_000_000
    ar, g, pts, scale2d, bones, cams = _avatar(N, W, H, seed=0)
    cot = _cots(W, H, seed=5)
    _mute_degenerate(ar, g, pts, scale2d, bones, cams[0], cot, False)
    out, grads, w = _ours(ar, g, pts, scale2d, bones, cams[0], cot, False)
    # skin weights: rows are convex combinations of convex rows
    assert (w >= 0).all() and (w.sum(1) - 1).abs().max().item() < 1e-4
    vis = out["visibility_filter"]
    assert vis.float().mean().item() > 0.2
    for k, v in grads.items():
        assert torch.isfinite(v).all(), f"dL/d{k} not finite"
    assert (grads["colors"][~vis] == 0).all() and (grads["scale2d"][~vis] == 0).all()
    assert out["mask"].min().item() >= 0 and out["mask"].max().item() <= 1
    # the rasterizer stage vs the reference on the skinned Gaussians: the whole contract (integer
    # state bit-exact, the four images, every gradient tensor with per-element statistics)
    if not refdgr.available():
        pytest.fail("oracle/_ref/libref_dgr.so missing: run oracle/build_ref.sh where /root/reference is mounted")
    from parity import compare_with_reference
    from soar_b200.lbs import skin_gaussians
    with torch.no_grad():
        sp, sq, _ = skin_gaussians(pts, torch.nn.functional.normalize(g["rot"].cuda(), dim=-1), w, bones)
    scales = scale2d.repeat(1, 3); scales[:, 2] = -1e10
    s = _scene_of(ar, cams[0], sp, sq, scales, g["colors"].cuda(), torch.zeros(3, device="cuda"), False,
                  descending=False)
    cot4 = {k: v.cuda() for k, v in synthetic.cotangents(W, H, seed=5).items()}
    o, ro = compare_with_reference("c5_raster_2M_1080p", s, cot4)
    assert torch.equal(out["radii"], ro["radii"])
    # ... and the chain's own images are that rasterizer pass
    for ours, theirs in (("render", "color"), ("depth", "depth"), ("mask", "opac")):
        err = (out[ours].detach() - ro[theirs]).abs().max().item()
        assert err <= IMG_TOL, f"{ours}: {err:.2e}"
    flipped = ro["normal"].clone(); flipped[1:] *= -1        # the post block: (n * (1,-1,-1) + 1) / 2
    err = (out["normal"].detach() - (flipped + 1) / 2).abs().max().item()
    assert err <= IMG_TOL, f"normal: {err:.2e}"
    del o, ro
    # KNN on the skinned cloud (the densify re-initialisation reads it): positive, finite
    d2 = distCUDA2(pts)
    assert torch.isfinite(d2).all() and (d2 >= 0).all() and d2.shape == (N,)


def test_c4_sds_step_32_views_matches_reference():
    """BASELINE config 4: 32 SDS views at 512² of one pose. The multi-view path (iter_views + in-place
    accumulation — what a rank does with its shard of the views) against the UNMODIFIED reference:
    every view's four images, and the gradient summed over the 32 views (reference passes summed in
    float64). shard_views must split the 32 views evenly over 2/4/8 ranks."""
    if not refdgr.available():
        pytest.fail("oracle/_ref/libref_dgr.so missing: run oracle/build_ref.sh where /root/reference is mounted")
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from soar_b200.lbs import skin_gaussians
    from soar_b200.views import FlatGrads, iter_views, shard_views
    W = H = 512
    N, V = 100_000, 32
    ar, g, pts, scale2d, bones, cams = _avatar(N, W, H, seed=1, n_views=V)
    with torch.no_grad():
        w = ar.weights(pts)
        sp, sq, _ = skin_gaussians(pts, torch.nn.functional.normalize(g["rot"].cuda(), dim=-1), w, bones, swizzle=True)
    scales = scale2d.repeat(1, 3); scales[:, 2] = -1e10
    params = dict(means3D=sp.clone().requires_grad_(True), colors_precomp=g["colors"].cuda().requires_grad_(True),
                  scales=scales.clone().requires_grad_(True), rotations=sq.clone().requires_grad_(True))
    ones = torch.ones((N, 1), device="cuda")
    bg = torch.zeros(3, device="cuda")
    cot = _cots(W, H, seed=2)
    patch = cams[0].full_patch()

    def rast(cam):
        return GaussianRasterizer(GaussianRasterizationSettings(
            image_height=H, image_width=W, tanfovx=math.tan(cam.FoVx / 2), tanfovy=math.tan(cam.FoVy / 2),
            bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
            patch_bbox=patch, prcppoint=cam.prcppoint, sh_degree=0, campos=cam.camera_center,
            prefiltered=False, render_front=False, sort_descending=False, debug=False, config=ar.config))

    def kw():
        return dict(means3D=params["means3D"], means2D=torch.zeros_like(sp, requires_grad=True),
                    opacities=ones, colors_precomp=params["colors_precomp"], scales=params["scales"],
                    rotations=params["rotations"])

    def loss(o):
        return (o[0] * cot["render"]).sum() + (o[1] * cot["normal"]).sum() + (o[2] * cot["depth"]).sum() \
            + (o[3] * cot["mask"]).sum()

    # the multi-view step; no zero-fill: the first backward of the step stores
    flat = FlatGrads(params)
    flat.flat.fill_(float("nan"))   # whatever the buffer held must not survive the first view
    images = []
    with flat.direct(overwrite=True):
        for o in iter_views([rast(c) for c in cams], [kw() for _ in cams]):
            images.append([t.detach().clone() for t in o[:4]])
            loss(o).backward()
    # the reference, view by view
    ref = refdgr.RefRasterizer()
    rcot = dict(color=cot["render"], normal=cot["normal"], depth=cot["depth"], opac=cot["mask"])
    want = {k: torch.zeros_like(v, dtype=torch.float64) for k, v in params.items()}
    names = dict(means3D="means3D", colors_precomp="colors", scales="scales", rotations="rotations")
    worst = 0.0
    for v, cam in enumerate(cams):
        s = _scene_of(ar, cam, sp, sq, scales, g["colors"].cuda(), bg, False, descending=False)
        ro = ref.forward(s)
        for img, k in zip(images[v], ("color", "normal", "depth", "opac")):
            err = (img - ro[k]).abs().max().item()
            worst = max(worst, err)
            assert err <= IMG_TOL, f"view {v}: {k} max-abs {err:.2e}"
        rg = ref.backward(rcot)
        for k in params:
            want[k] += rg[names[k]].double().view_as(want[k])
    ref.close()
    report = []
    for k in params:
        e = rel_err(flat.grad(k).double().view_as(want[k]), want[k])
        report.append(f"{k}: {e:.2e}")
        assert e <= GRAD_TOL, f"{k}: {e:.2e}"
    print(f"[c4_32_views] worst image max-abs {worst:.2e}; summed gradients vs reference: " + " | ".join(report))
    for world in (2, 4, 8):
        shards = [shard_views(V, r, world) for r in range(world)]
        assert sorted(sum(shards, [])) == list(range(V)) and {len(s) for s in shards} == {V // world}
