"""Image-space post-ops of SOAR's surfel renderer on the sm_100a stencil kernels
(soar_b200/csrc/postops.cu) — SURVEY.md §8 row f2. Same names and argument meaning as the
reference's torch op chains:

    depth2normal(depth, mask, camera)    TS/renderer/diff_gaussian_rasterizer.py:359-430
    normal2curv(normal, mask)            :433-448
    surfel_postprocess(...)              the renderer's post block, :292-303

One launch forward, one launch backward (the reference: ~25 torch kernels, a CPU
`torch.inverse` and a blocking `.to(device)` per render, :381-382). Gradients flow to `depth`
and `normal`; masks are non-differentiable (the reference passes `rendered_opac.detach() > 1e-5`).
"""
import ctypes
import math

import numpy as np
import torch

from . import _capi

_lib = _capi.load()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def fov2focal(fov, pixels):
    """TS/renderer/diff_gaussian_rasterizer.py:355-356."""
    return pixels / (2 * math.tan(fov / 2))


# principal points already read back from the device: (data_ptr, version) -> (px, py). A camera's
# prcppoint is a CUDA tensor in the reference (TS/renderer/gaussian_batch_renderer.py:59-82) that
# never changes after the camera is built; reading it once per camera instead of once per render
# keeps the render path free of blocking device->host copies.
_PRCP_CACHE = {}


def _prcp_host(pp):
    if not torch.is_tensor(pp):
        return np.asarray(pp, dtype=np.float32)
    if not pp.is_cuda:
        return pp.detach().float().numpy()
    key = (pp.data_ptr(), pp._version, pp.device.index)
    hit = _PRCP_CACHE.get(key)
    # the weak reference guards against a different tensor re-using a freed address
    if hit is not None and hit[0]() is pp:
        return hit[1]
    val = pp.detach().float().cpu().numpy()
    if len(_PRCP_CACHE) > 4096:
        _PRCP_CACHE.clear()
    import weakref
    _PRCP_CACHE[key] = (weakref.ref(pp), val)
    return val


def camera_constants(camera):
    """(cx, cy, kinv_x, kinv_y) exactly as depth2normal derives them (:375-382): the principal
    point in pixels as a float32 product, and the inverse of K = diag(fov2focal(FoVy, H),
    fov2focal(FoVx, W)) in float32 — the reference's own pairing of K00 with x."""
    pp = np.asarray(_prcp_host(camera.prcppoint), dtype=np.float32)
    cx = np.float32(pp[0] * np.float32(camera.image_width))
    cy = np.float32(pp[1] * np.float32(camera.image_height))
    k00 = np.float32(fov2focal(camera.FoVy, camera.image_height))
    k11 = np.float32(fov2focal(camera.FoVx, camera.image_width))
    return float(cx), float(cy), float(np.float32(1.0) / k00), float(np.float32(1.0) / k11)


def _args(W, H, consts, surfel, mask, opac, normal, depth, threshold=1e-5):
    a = _capi.SoarPostArgs()
    a.width, a.height = W, H
    a.cx, a.cy, a.kinv_x, a.kinv_y = consts
    a.mask_threshold = threshold
    a.surfel = int(surfel)
    a.mask = _capi.ptr(mask)
    a.opac = _capi.ptr(opac)
    a.normal = _capi.ptr(normal)
    a.depth = _capi.ptr(depth)
    return a


def _f32(t):
    return t.detach().float().contiguous()


class _PostOps(torch.autograd.Function):
    """mode 0: depth2normal, 1: normal2curv, 2: the surfel post block (all outputs)."""

    @staticmethod
    def forward(ctx, normal, depth, mask_src, consts, mode):
        ref = normal if normal is not None else depth
        H, W = ref.shape[-2], ref.shape[-1]
        dev = ref.device
        surfel = mode == 2
        normal_c = _f32(normal).reshape(3, H, W) if normal is not None else None
        depth_c = _f32(depth).reshape(H, W) if depth is not None else None
        if surfel:
            mask_c, opac_c = None, _f32(mask_src).reshape(H, W)
        else:
            mask_c, opac_c = mask_src.detach().reshape(H, W).to(torch.uint8).contiguous(), None
        out_n = torch.empty((3, H, W), dtype=torch.float32, device=dev) if surfel else None
        out_c = torch.empty((1, H, W), dtype=torch.float32, device=dev) if normal is not None else None
        out_d = torch.empty((3, H, W), dtype=torch.float32, device=dev) if depth is not None else None
        out_m = torch.empty((1, H, W), dtype=torch.bool, device=dev) if surfel else None
        a = _args(W, H, consts, surfel, mask_c, opac_c, normal_c, depth_c)
        with torch.cuda.device(dev):
            _capi.check(_lib.soar_post_forward(ctypes.byref(a), _capi.ptr(out_n), _capi.ptr(out_c),
                                               _capi.ptr(out_d), _capi.ptr(out_m), _stream()),
                        "soar_post_forward")
        ctx.saved = (normal_c, depth_c, mask_c, opac_c)
        ctx.consts, ctx.mode, ctx.shape = consts, mode, (H, W)
        ctx.in_shapes = (None if normal is None else normal.shape, None if depth is None else depth.shape)
        if surfel:
            ctx.mark_non_differentiable(out_m)
            return out_n, out_c, out_d, out_m
        return out_d if mode == 0 else out_c

    @staticmethod
    def backward(ctx, *gs):
        normal_c, depth_c, mask_c, opac_c = ctx.saved
        H, W = ctx.shape
        surfel = ctx.mode == 2
        g_n = g_c = g_d = None
        if surfel:
            g_n, g_c, g_d = gs[0], gs[1], gs[2]
        elif ctx.mode == 0:
            g_d = gs[0]
        else:
            g_c = gs[0]
        g_n, g_c, g_d = (None if g is None else g.float().contiguous() for g in (g_n, g_c, g_d))
        dev = (normal_c if normal_c is not None else depth_c).device
        need_n = normal_c is not None and ctx.needs_input_grad[0] and (g_c is not None or g_n is not None)
        need_d = depth_c is not None and ctx.needs_input_grad[1] and g_d is not None
        d_normal = torch.empty((3, H, W), dtype=torch.float32, device=dev) if need_n else None
        d_depth = torch.empty((H, W), dtype=torch.float32, device=dev) if need_d else None
        if need_n or need_d:
            a = _args(W, H, ctx.consts, surfel, mask_c, opac_c, normal_c, depth_c)
            with torch.cuda.device(dev):
                _capi.check(_lib.soar_post_backward(ctypes.byref(a), _capi.ptr(g_n), _capi.ptr(g_c),
                                                    _capi.ptr(g_d), _capi.ptr(d_normal),
                                                    _capi.ptr(d_depth), _stream()),
                            "soar_post_backward")
        sn, sd = ctx.in_shapes
        return (d_normal.reshape(sn) if need_n else None, d_depth.reshape(sd) if need_d else None,
                None, None, None)


def _need_cuda(t, who):
    if not t.is_cuda:
        raise RuntimeError(f"{who}: CUDA tensors required (there is no CPU path)")


def depth2normal(depth, mask, camera):
    """depth [1,H,W], mask [1,H,W] bool, camera with prcppoint / image_width / image_height /
    FoVx / FoVy -> unit normals [3,H,W], zero outside the mask (reference :359-430)."""
    _need_cuda(depth, "depth2normal")
    return _PostOps.apply(None, depth, mask, camera_constants(camera), 0)


def normal2curv(normal, mask):
    """normal [3,H,W], mask [1,H,W] bool -> curvature [1,H,W] (reference :433-448)."""
    _need_cuda(normal, "normal2curv")
    return _PostOps.apply(normal, None, mask, (0.0, 0.0, 1.0, 1.0), 1)


def surfel_postprocess(rendered_normal, rendered_depth, rendered_opac, camera):
    """The renderer's post block (reference :292-303) on the rasterizer's outputs:
    returns (normal, curv, depth_normal, mask) = ((n'+1)/2, normal2curv(n', mask),
    (depth2normal(depth, mask)·(1,-1,-1)+1)/2, opac > 1e-5) with n' = normal·(1,-1,-1) and the
    normal's gradient cut outside the mask. Unlike the reference it does not modify
    `rendered_normal` in place."""
    _need_cuda(rendered_normal, "surfel_postprocess")
    return _PostOps.apply(rendered_normal, rendered_depth, rendered_opac, camera_constants(camera), 2)
