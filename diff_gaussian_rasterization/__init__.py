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
                cov3Ds_precomp, viewmatrix, projmatrix, campos, raster_settings, pending=None):
        s = raster_settings
        if pending is not None:
            # phase 1 was enqueued earlier by GaussianRasterizer.begin(): only wait + phase 2
            (num_rendered, color, normal, depth, opac, radii, geom, binning, img) = \
                _C.rasterize_gaussians_finish(pending)
        else:
            args = _forward_args(s, means3D, sh, colors_precomp, opacities, scales, rotations,
                                 cov3Ds_precomp, viewmatrix, projmatrix, campos)
            (num_rendered, color, normal, depth, opac, radii, geom, binning, img) = _call_with_snapshot(
                _C.rasterize_gaussians, args, s.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = s
        ctx.num_rendered = num_rendered
        # leaves registered by soar_b200.views.FlatGrads.direct(): accumulate in place in backward
        ctx.sinks = ctx.direct = None
        from soar_b200.views import _DIRECT, active_direct, direct_sink
        if _DIRECT:
            ctx.direct = active_direct()
            sinks = {k: direct_sink(t) for k, t in (("means3D", means3D), ("sh", sh),
                                                    ("colors", colors_precomp), ("opacity", opacities),
                                                    ("scales", scales), ("rotations", rotations),
                                                    ("cov3D", cov3Ds_precomp))}
            ctx.sinks = {k: v for k, v in sinks.items() if v is not None} or None
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
        if ctx.sinks:
            args = args + (ctx.sinks, ctx.direct is not None and ctx.direct.first_store(),
                           ctx.direct.order_events() if ctx.direct is not None else None)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations, g_view,
         g_proj, g_campos) = _call_with_snapshot(
            _C.rasterize_gaussians_backward, args, s.debug, "snapshot_bw.dump", "backward")
        if ctx.sinks:
            # already added to the caller's buffers by the kernel: nothing for autograd to add
            k = ctx.sinks
            g_means3D = None if "means3D" in k else g_means3D
            g_sh = None if "sh" in k else g_sh
            g_colors = None if "colors" in k else g_colors
            g_opacities = None if "opacity" in k else g_opacities
            g_scales = None if "scales" in k else g_scales
            g_rotations = None if "rotations" in k else g_rotations
            g_cov3D = None if "cov3D" in k else g_cov3D
        # one gradient per forward input, in forward order (reference :249-262)
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov3D,
                g_view, g_proj, g_campos, None, None)


def _forward_args(s, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                  viewmatrix, projmatrix, campos):
    """The 24 positional arguments of _C.rasterize_gaussians (reference :77-102)."""
    return (s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier,
            cov3Ds_precomp, viewmatrix, projmatrix, s.prcppoint, s.patch_bbox, s.tanfovx,
            s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree, campos, s.prefiltered,
            s.render_front, s.sort_descending, s.debug, s.config)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, viewmatrix, projmatrix, campos, raster_settings,
                        pending=None):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, viewmatrix, projmatrix, campos,
                                     raster_settings, pending)


class GaussianRasterizer(nn.Module):
    """nn.Module facade (reference :287-356)."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings
        self._pending = None

    def begin(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
              rotations=None, cov3D_precomp=None):
        """Extension (not in the reference): enqueue the first phase of forward() — preprocess,
        depth order, tile counts — without waiting for the device. The next forward() call with
        the SAME arguments picks it up. A multi-view step calls begin() on all its rasterizers
        before the first forward(), so only the first view ever waits for the device
        (soar_b200.views.render_views does exactly that)."""
        shs, colors_precomp, scales, rotations, cov3D_precomp = self._normalise(
            shs, colors_precomp, scales, rotations, cov3D_precomp)
        s = self.raster_settings
        with torch.no_grad():
            pend = _C.rasterize_gaussians_begin(*_forward_args(
                s, means3D.detach(), shs.detach(), colors_precomp.detach(), opacities.detach(),
                scales.detach(), rotations.detach(), cov3D_precomp.detach(), s.viewmatrix,
                s.projmatrix, s.campos))
        self._pending = (pend, self._signature(s, means3D, opacities, shs, colors_precomp, scales,
                                               rotations, cov3D_precomp))
        return self

    @staticmethod
    def _signature(settings, *tensors):
        """What begin() rendered from: the settings object and, per tensor argument, storage address,
        shape and version counter (an in-place update between begin() and forward() bumps it)."""
        return (id(settings),) + tuple((t.data_ptr(), tuple(t.shape), t._version) for t in tensors)

    @staticmethod
    def _normalise(shs, colors_precomp, scales, rotations, cov3D_precomp):
        if (shs is None) == (colors_precomp is None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                have_sr and cov3D_precomp is not None):
            raise Exception(
                "Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        empty = torch.Tensor([])
        return (empty if shs is None else shs, empty if colors_precomp is None else colors_precomp,
                empty if scales is None else scales, empty if rotations is None else rotations,
                empty if cov3D_precomp is None else cov3D_precomp)

    def markVisible(self, positions):
        with torch.no_grad():
            s = self.raster_settings
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        s = self.raster_settings
        shs, colors_precomp, scales, rotations, cov3D_precomp = self._normalise(
            shs, colors_precomp, scales, rotations, cov3D_precomp)
        pending = None
        if self._pending is not None:
            pending, sig = self._pending
            self._pending = None
            if sig != self._signature(s, means3D, opacities, shs, colors_precomp, scales, rotations,
                                      cov3D_precomp):
                raise RuntimeError("GaussianRasterizer.forward: arguments (or raster_settings) differ "
                                   "from the begin() call, or were modified in place since")
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales,
                                   rotations, cov3D_precomp, s.viewmatrix, s.projmatrix, s.campos, s,
                                   pending)
