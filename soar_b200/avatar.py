"""Skin + render one view of the avatar — the hot-path content of the reference's
`DiffGaussian.forward` (TS/renderer/diff_gaussian_rasterizer.py:52-318) composed from the
sm_100a pieces of this package, with the reference's return dictionary:

    SMPL-X forward + cano->live bones       TS/utils/smpl.py:601-609        SmplxPoser
    skin-weight query (detached)            :611, :618-637                  SkinWeightQuery
    bone blend, skin points / quaternions   :613; renderer :103-114/:138-149  skin_gaussians
    rasterizer pass + occupancy pass        renderer :173-291               GaussianRasterizer
    mask fix-ups, normal2curv, depth2normal :292-303                        surfel_postprocess

What is NOT here (out of scope): the tcnn attribute field (:88-102) — colours,
in-plane scale and occupancy arrive as tensors — the trainer, losses and guidance.
A training step poses the body once (`AvatarRenderer.pose`) and queries the skin weights once
(`weights`), then renders its views; the reference re-runs both for every view.
"""
import math
from typing import NamedTuple, Optional

import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

from .lbs import SkinWeightQuery, skin_gaussians
from .postops import surfel_postprocess
from .smplx_fwd import SmplxPoser


class AvatarCamera(NamedTuple):
    """The fields of the reference's Camera (TS/renderer/gaussian_batch_renderer.py:73-83) that the
    render path reads."""
    FoVx: float
    FoVy: float
    image_width: int
    image_height: int
    world_view_transform: torch.Tensor
    full_proj_transform: torch.Tensor
    camera_center: torch.Tensor
    prcppoint: torch.Tensor

    def full_patch(self):
        """random_patch(-1, -1) of the reference camera: the whole image as (x0, y0, h, w)."""
        return torch.tensor([0.0, 0.0, float(self.image_height), float(self.image_width)],
                            dtype=torch.float32, device=self.world_view_transform.device)


class AvatarRenderer:
    def __init__(self, poser: SmplxPoser, cano_vertices, inv_mats, config=(1.0, 1.0, 1.0, 0.0)):
        """poser: the body model; cano_vertices [V,3]: the canonical (zero-pose, shaped) vertices the
        skin weights are queried against (TS/utils/smpl.py:249-258); inv_mats [J,4,4]: inverse
        canonical joint transforms (:609); config: the surfel rasterizer's 4 flags (pc.config)."""
        self.poser = poser
        self.query = SkinWeightQuery(cano_vertices, poser.lbs_weights)
        self.inv_mats = inv_mats.detach().to(device=poser.device, dtype=torch.float32).reshape(-1, 4, 4)
        self.config = torch.tensor(config, dtype=torch.float32, device=poser.device)
        self._full_patch = {}  # (H, W) -> device tensor, built once per image size

    # ---- once per pose / per step -------------------------------------------------------------
    def pose(self, betas, full_pose, transl=None):
        """-> cano->live bone transforms [J,4,4] (live_smpl.A @ inv_mats, transl folded in)."""
        return self.poser(betas, full_pose, transl, self.inv_mats, want_verts=False)[3]

    def weights(self, points):
        """-> detached skin weights [N,J] of the canonical Gaussians."""
        return self.query(points)

    # ---- once per view ------------------------------------------------------------------------
    def _patch(self, camera, patch_bbox):
        if patch_bbox is not None:
            return patch_bbox
        key = (int(camera.image_height), int(camera.image_width))
        if key not in self._full_patch:
            self._full_patch[key] = camera.full_patch()
        return self._full_patch[key]

    def render(self, points, rot, weights, bones, colors, scale2d, occ, camera: AvatarCamera,
               bg_color, sds_view=False, scaling_modifier=1.0, render_occ=True,
               screenspace_points: Optional[torch.Tensor] = None, render_front=True,
               patch_bbox: Optional[torch.Tensor] = None):
        """points [N,3], rot [N,4] (unnormalised parameter), weights [N,J], bones [J,4,4],
        colors [N,3], scale2d [N,1] (linear), occ [N,1] in (0,1). Returns the reference's dict.
        render_front: the reference's argument of the same name (default True, :60): the main pass
        always has the SETTING render_front=False and sorts `descending = not render_front` (:186-187) —
        front view = ascending depth order; the back-normal pass of gaussian_batch_renderer.py:167
        passes render_front=False. patch_bbox: (h0, w0, h1, w1) as camera.random_patch() returns it
        (:182); None = the whole image."""
        return self.render_finish(self.render_begin(
            points, rot, weights, bones, colors, scale2d, occ, camera, bg_color, sds_view, scaling_modifier,
            render_occ, screenspace_points, render_front, patch_bbox))

    def render_begin(self, points, rot, weights, bones, colors, scale2d, occ, camera: AvatarCamera,
                     bg_color, sds_view=False, scaling_modifier=1.0, render_occ=True,
                     screenspace_points: Optional[torch.Tensor] = None, render_front=True,
                     patch_bbox: Optional[torch.Tensor] = None):
        """First half of render(): skinning and the first phase (preprocess, depth order, tile counts)
        of BOTH rasterizer passes are enqueued on the current stream without waiting for the device
        (GaussianRasterizer.begin). A step that renders several views calls render_begin for all of
        them (each under its own stream if it wants them to overlap) before the first render_finish:
        the host then never sits in a num_rendered read-back with nothing queued behind it."""
        rot = torch.nn.functional.normalize(rot, dim=-1)                                  # :74
        means3D, rotations, _ = skin_gaussians(points, rot, weights, bones, swizzle=sds_view)
        if screenspace_points is None:                                                    # :155-164
            screenspace_points = torch.zeros_like(points, requires_grad=True) + 0
            try:
                screenspace_points.retain_grad()
            except Exception:
                pass
        common = dict(
            image_height=int(camera.image_height), image_width=int(camera.image_width),
            tanfovx=math.tan(camera.FoVx * 0.5), tanfovy=math.tan(camera.FoVy * 0.5), bg=bg_color,
            scale_modifier=scaling_modifier, viewmatrix=camera.world_view_transform,
            projmatrix=camera.full_proj_transform, patch_bbox=self._patch(camera, patch_bbox),
            prcppoint=camera.prcppoint, sh_degree=0, campos=camera.camera_center, prefiltered=False,
            debug=False, config=self.config)
        rasterizer = GaussianRasterizer(GaussianRasterizationSettings(
            render_front=False, sort_descending=not render_front, **common))              # :173-191
        scales = scale2d.repeat(1, 3)
        scales[..., -1] = -1e10                                                            # :239 (Q7)
        ones = torch.ones((points.shape[0], 1), dtype=torch.float32, device=points.device)
        main_kw = dict(means3D=means3D, means2D=screenspace_points, shs=None, colors_precomp=colors,
                       opacities=ones, scales=scales, rotations=rotations, cov3D_precomp=None)
        rasterizer.begin(**main_kw)
        occ_pass = None
        if render_occ:                                                                     # :281-291
            rasterizer_occ = GaussianRasterizer(GaussianRasterizationSettings(
                render_front=True, sort_descending=False, **common))
            occ_kw = dict(means3D=means3D.detach(), means2D=screenspace_points.detach(), shs=None,
                          colors_precomp=occ.repeat(1, 3), opacities=ones, scales=scales.detach(),
                          rotations=rotations.detach(), cov3D_precomp=None)
            rasterizer_occ.begin(**occ_kw)
            occ_pass = (rasterizer_occ, occ_kw)
        return (rasterizer, main_kw, occ_pass, camera, screenspace_points)

    def render_finish(self, pending):
        """Second half of render(): both passes' second phase, the post block, the return dict."""
        rasterizer, main_kw, occ_pass, camera, screenspace_points = pending
        image, normal, depth, opac, radii = rasterizer(**main_kw)                          # :254-263
        rendered_occ = None
        if occ_pass is not None:
            rendered_occ = occ_pass[0](**occ_pass[1])[0]
        normal_out, curv, depth_normal, _ = surfel_postprocess(normal, depth, opac, camera)  # :292-303
        return {"render": image, "normal": normal_out, "depth": depth, "pred_normal": depth_normal,
                "mask": opac, "occ": rendered_occ, "curv": curv,
                "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
                "radii": radii}
