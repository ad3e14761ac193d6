"""Drop-in replacement for the reference package `diff_gaussian_rasterization`
(DGR/diff_gaussian_rasterization/__init__.py): same names, same argument order,
same return values, so `from diff_gaussian_rasterization import
GaussianRasterizationSettings, GaussianRasterizer`
(TS/renderer/diff_gaussian_rasterizer.py:6-9) keeps working unchanged. The
kernels underneath are the sm_100a ones in soar_b200/csrc, reached through the
C-ABI in include/soar_b200.h.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    """Field order is part of the API (reference :267-284)."""

    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    patch_bbox: torch.Tensor
    prcppoint: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    render_front: bool
    sort_descending: bool
    debug: bool
    config: torch.Tensor


def _snapshot(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_snapshot(fn, args, debug, dump_name, phase):
    """debug=True keeps the reference's behaviour (:105-126, :210-233): inputs are
    copied to the host first and written to snapshot_{fw,bw}.dump if the call raises."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {phase}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    """autograd bridge (reference :58-264)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, viewmatrix, projmatrix, campos, raster_settings):
        s = raster_settings
        args = (s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier,
                cov3Ds_precomp, viewmatrix, projmatrix, s.prcppoint, s.patch_bbox, s.tanfovx,
                s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree, campos, s.prefiltered,
                s.render_front, s.sort_descending, s.debug, s.config)
        (num_rendered, color, normal, depth, opac, radii, geom, binning, img) = _call_with_snapshot(
            _C.rasterize_gaussians, args, s.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, normal, depth, opac, radii

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_normal, grad_out_depth, grad_out_opac, _):
        s = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning,
         img) = ctx.saved_tensors
        args = (s.bg, means3D, radii, colors_precomp, scales, rotations, s.scale_modifier,
                cov3Ds_precomp, s.viewmatrix, s.projmatrix, s.prcppoint, s.patch_bbox, s.tanfovx,
                s.tanfovy, grad_out_color, grad_out_normal, grad_out_depth, grad_out_opac, sh,
                s.sh_degree, s.campos, geom, ctx.num_rendered, binning, img, s.debug, s.config)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations, g_view,
         g_proj, g_campos) = _call_with_snapshot(
            _C.rasterize_gaussians_backward, args, s.debug, "snapshot_bw.dump", "backward")
        # one gradient per forward input, in forward order (reference :249-262)
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov3D,
                g_view, g_proj, g_campos, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, viewmatrix, projmatrix, campos, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, viewmatrix, projmatrix, campos,
                                     raster_settings)


class GaussianRasterizer(nn.Module):
    """nn.Module facade (reference :287-356)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            s = self.raster_settings
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        s = self.raster_settings
        if (shs is None) == (colors_precomp is None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                have_sr and cov3D_precomp is not None):
            raise Exception(
                "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales,
                                   rotations, cov3D_precomp, s.viewmatrix, s.projmatrix, s.campos, s)
