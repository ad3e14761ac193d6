"""Drop-in for the reference's pybind module `diff_gaussian_rasterization._C`
(DGR/ext.cpp:15-19): rasterize_gaussians / rasterize_gaussians_backward /
mark_visible with the reference's positional signatures
(DGR/rasterize_points.h:18-50), implemented on libsoar_b200.so through ctypes.

What the reference's torch glue does (DGR/rasterize_points.cu:35-205) and this
file mirrors: allocate the outputs and the three opaque scratch buffers as
torch tensors (so the caching allocator owns them and the autograd ctx can keep
them between forward and backward), then hand raw pointers to the C side.
Differences, on purpose: outputs and gradients are torch.empty — the kernels
write every element, the reference pre-fills 4 images and 15 gradient tensors
with zeros — and launches go to torch's current stream instead of the legacy
default stream.
"""
import ctypes
import threading

import torch

from soar_b200 import _capi

_lib = _capi.load()


def _f32(t, name):
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _config_flags(config):
    """SoarRastView.config_host_flags for a config tensor. The reference's config lives on the device
    (pc.config, TS/renderer/diff_gaussian_rasterizer.py:200); its four flags are read back ONCE per
    tensor object and version counter and remembered on the tensor itself, so the launchers can pick
    the specialised blend kernels on the host. Anything that cannot be tracked that way (None, a
    tensor that refuses attributes) reports 'unknown' and the kernels branch on the device copy."""
    if config is None or config.numel() < 4:
        return 0
    ver = config._version
    hit = getattr(config, "_soar_cfg", None)
    if hit is not None and hit[0] == ver and hit[1] == config.data_ptr():
        return hit[2]
    c = config.detach().reshape(-1)[:4].float().cpu()  # one small blocking read per config tensor
    flags = _capi.CFG_KNOWN
    for bit, v in zip((_capi.CFG_SURFACE, _capi.CFG_NORMALIZE_DEPTH, _capi.CFG_PERPIX_DEPTH,
                       _capi.CFG_LRN_CAM), c.tolist()):
        if v > 0:
            flags |= bit
    try:
        config._soar_cfg = (ver, config.data_ptr(), flags)
    except Exception:  # noqa: BLE001
        return 0
    return flags


def _view(P, M, degree, H, W, tan_fovx, tan_fovy, scale_modifier, prefiltered, render_front,
          sort_descending, debug, bg, means3D, sh, colors, opacity, scales, rotations, cov3D,
          viewmatrix, projmatrix, prcppoint, patchbbox, campos, config, config_flags=0):
    v = _capi.SoarRastView()
    v.P, v.sh_degree, v.M, v.width, v.height = P, int(degree), M, W, H
    v.tanfovx, v.tanfovy, v.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
    v.prefiltered, v.render_front = int(bool(prefiltered)), int(bool(render_front))
    v.sort_descending, v.debug = int(bool(sort_descending)), int(bool(debug))
    p = _capi.ptr
    v.bg, v.means3D, v.shs, v.colors_precomp = p(bg), p(means3D), p(sh), p(colors)
    v.opacities, v.scales, v.rotations, v.cov3D_precomp = p(opacity), p(scales), p(rotations), p(cov3D)
    v.viewmatrix, v.projmatrix, v.prcppoint = p(viewmatrix), p(projmatrix), p(prcppoint)
    v.patch_bbox, v.campos, v.config = p(patchbbox), p(campos), p(config)
    v.config_host_flags = config_flags
    return v


class _PinnedStatus:
    """Pool of page-locked 2-word host buffers for the asynchronous num_rendered read-back
    (shared by all threads, hence the lock)."""
    _free = []
    _lock = threading.Lock()

    @classmethod
    def get(cls):
        with cls._lock:
            if cls._free:
                return cls._free.pop()
        return torch.zeros(2, dtype=torch.int32).pin_memory()

    @classmethod
    def put(cls, t):
        with cls._lock:
            cls._free.append(t)


class PendingRaster:
    """Phase 1 of a forward (preprocess, depth order, scan) enqueued, count not yet read."""
    __slots__ = ("view", "keep", "P", "H", "W", "dev", "radii", "geom", "status", "event", "debug")


def rasterize_gaussians_begin(background, means3D, colors, opacity, scales, rotations,
                              scale_modifier, cov3D_precomp, viewmatrix, projmatrix, prcppoint,
                              patchbbox, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                              campos, prefiltered, render_front, sort_descending, debug, config):
    """Enqueue phase 1 of the forward on the current stream WITHOUT waiting for num_rendered
    (soar_rast_forward_bin_async). A multi-view step calls this for all its views, then
    rasterize_gaussians_finish() for each: the host wait of view k overlaps the device work of
    the views behind it. Same arguments as rasterize_gaussians."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    pend = PendingRaster()
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    pend.P, pend.H, pend.W, pend.dev, pend.debug = P, H, W, dev, bool(debug)
    pend.view = pend.keep = pend.radii = pend.geom = pend.status = pend.event = None
    if P == 0:
        return pend

    means3D = _f32(means3D, "means3D")
    sh, colors = _f32(sh, "sh"), _f32(colors, "colors_precomp")
    opacity, scales = _f32(opacity, "opacities"), _f32(scales, "scales")
    rotations, cov3D_precomp = _f32(rotations, "rotations"), _f32(cov3D_precomp, "cov3D_precomp")
    background, viewmatrix = _f32(background, "bg"), _f32(viewmatrix, "viewmatrix")
    projmatrix, prcppoint = _f32(projmatrix, "projmatrix"), _f32(prcppoint, "prcppoint")
    cfg_flags = _config_flags(config)
    patchbbox, campos, config = _f32(patchbbox, "patch_bbox"), _f32(campos, "campos"), _f32(config, "config")
    M = sh.size(1) if sh is not None else 0

    with torch.cuda.device(dev):
        pend.view = _view(P, M, degree, H, W, tan_fovx, tan_fovy, scale_modifier, prefiltered,
                          render_front, sort_descending, debug, background, means3D, sh, colors,
                          opacity, scales, rotations, cov3D_precomp, viewmatrix, projmatrix,
                          prcppoint, patchbbox, campos, config, cfg_flags)
        # the view struct holds raw pointers: keep the (possibly converted) tensors alive
        pend.keep = (background, means3D, sh, colors, opacity, scales, rotations, cov3D_precomp,
                     viewmatrix, projmatrix, prcppoint, patchbbox, campos, config)
        pend.radii = torch.empty(P, dtype=torch.int32, device=dev)
        pend.geom = torch.empty(_lib.soar_rast_geom_bytes(P), dtype=torch.uint8, device=dev)
        pend.status = _PinnedStatus.get()
        _capi.check(_lib.soar_rast_forward_bin_async(ctypes.byref(pend.view), pend.geom.data_ptr(),
                                                     pend.geom.numel(), pend.radii.data_ptr(),
                                                     pend.status.data_ptr(), _stream()),
                    "rasterize_gaussians (preprocess)")
        pend.event = torch.cuda.Event()
        pend.event.record(torch.cuda.current_stream())
    return pend


def rasterize_gaussians_finish(pend):
    """Wait for the count of a begun forward, size the binning buffer, enqueue phase 2 on the
    current stream (must be the stream phase 1 was enqueued on). Returns the 9-tuple of
    rasterize_gaussians."""
    P, H, W, dev = pend.P, pend.H, pend.W, pend.dev
    byte = dict(dtype=torch.uint8, device=dev)
    if P == 0:
        z = lambda c: torch.zeros((c, H, W), dtype=torch.float32, device=dev)
        e = torch.empty(0, **byte)
        return (0, z(3), z(3), z(1), z(1), torch.zeros(0, dtype=torch.int32, device=dev), e,
                e.clone(), e.clone())
    pend.event.synchronize()  # the one host wait of the forward (reference: rasterizer_impl.cu:250)
    R, flag = int(pend.status[0]) & 0xffffffff, int(pend.status[1])
    _PinnedStatus.put(pend.status)
    pend.status = None
    if flag:
        raise RuntimeError("rasterize_gaussians (preprocess) failed: Point is filtered although "
                           "prefiltered is set. This shouldn't happen!")
    with torch.cuda.device(dev):
        stream = _stream()
        img = torch.empty(_lib.soar_rast_image_bytes(W, H), **byte)
        binning = torch.empty(_lib.soar_rast_binning_bytes(R, W, H), **byte)
        f32 = dict(dtype=torch.float32, device=dev)
        color = torch.empty((3, H, W), **f32)
        normal = torch.empty((3, H, W), **f32)
        depth = torch.empty((1, H, W), **f32)
        opac = torch.empty((1, H, W), **f32)
        _capi.check(_lib.soar_rast_forward_render(ctypes.byref(pend.view), pend.geom.data_ptr(),
                                                  binning.data_ptr(), binning.numel(),
                                                  img.data_ptr(), img.numel(), R, color.data_ptr(),
                                                  normal.data_ptr(), depth.data_ptr(),
                                                  opac.data_ptr(), stream),
                    "rasterize_gaussians (render)")
        if pend.debug:
            torch.cuda.synchronize(dev)  # reference: CHECK_CUDA syncs after every launch
    radii, geom = pend.radii, pend.geom
    pend.view = pend.keep = None
    return R, color, normal, depth, opac, radii, geom, binning, img


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, prcppoint, patchbbox, tan_fovx,
                        tan_fovy, image_height, image_width, sh, degree, campos, prefiltered,
                        render_front, sort_descending, debug, config):
    """Forward. Returns (num_rendered, color[3,H,W], normal[3,H,W], depth[1,H,W],
    opac[1,H,W], radii[P] int32, geomBuffer, binningBuffer, imgBuffer) like
    RasterizeGaussiansCUDA (DGR/rasterize_points.cu:35-105)."""
    return rasterize_gaussians_finish(rasterize_gaussians_begin(
        background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
        viewmatrix, projmatrix, prcppoint, patchbbox, tan_fovx, tan_fovy, image_height, image_width,
        sh, degree, campos, prefiltered, render_front, sort_descending, debug, config))


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations,
                                 scale_modifier, cov3D_precomp, viewmatrix, projmatrix, prcppoint,
                                 patchbbox, tan_fovx, tan_fovy, dL_dout_color, dL_dout_normal,
                                 dL_dout_depth, dL_dout_opac, sh, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, debug, config, accumulate=None,
                                 accumulate_store=False, order=None):
    """Backward. Returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
    dL_dscales, dL_drotations, dL_dviewmat, dL_dprojmat, dL_dcampos) like
    RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:107-187).

    Extension: `accumulate` maps any of "means3D", "colors", "opacity", "sh", "scales",
    "rotations", "cov3D" to a caller-owned float32 buffer of the gradient's shape; this view's
    gradient is then ADDED to it in place (SoarRastGrads.accumulate_mask) and the buffer itself
    is returned in the tuple. With `accumulate_store=True` the buffers are OVERWRITTEN instead (every
    element, zeros for culled Gaussians): the first view of a step needs no zero-fill before it.
    `order` = (torch.cuda.Event or None, torch.cuda.Event): the launch that writes the gradient tensors
    waits for the first event and records the second (soar_rast_backward_ordered), for views of one
    step running on several streams."""
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    dev = means3D.device
    f32 = dict(dtype=torch.float32, device=dev)
    sh = _f32(sh, "sh")
    M = sh.size(1) if sh is not None else 0
    acc = accumulate or {}
    mask = 0

    def alloc(shape, name=None, bit=0):
        nonlocal mask
        buf = acc.get(name) if name else None
        if buf is not None and P > 0:
            if buf.dtype != torch.float32 or not buf.is_contiguous() or buf.numel() != P * (
                    shape[1] if len(shape) == 2 else shape[1] * shape[2]):
                raise RuntimeError(f"accumulate[{name!r}] must be a contiguous float32 buffer of shape {shape}")
            if not accumulate_store:
                mask |= bit
            return buf
        return (torch.empty if P > 0 else torch.zeros)(shape, **f32)

    g_means3D, g_means2D = alloc((P, 3), "means3D", _capi.ACC_MEANS3D), alloc((P, 3))
    g_colors, g_opacity = alloc((P, 3), "colors", _capi.ACC_COLORS), alloc((P, 1), "opacity", _capi.ACC_OPACITY)
    g_cov3D, g_sh = alloc((P, 6), "cov3D", _capi.ACC_COV3D), alloc((P, M, 3), "sh" if M else None, _capi.ACC_SH)
    g_scales, g_rot = alloc((P, 3), "scales", _capi.ACC_SCALES), alloc((P, 4), "rotations", _capi.ACC_ROTATIONS)
    g_view, g_proj, g_campos = torch.zeros((4, 4), **f32), torch.zeros((4, 4), **f32), torch.zeros(3, **f32)
    if P == 0:
        return (g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot, g_view,
                g_proj, g_campos)

    means3D, colors = _f32(means3D, "means3D"), _f32(colors, "colors_precomp")
    scales, rotations = _f32(scales, "scales"), _f32(rotations, "rotations")
    cov3D_precomp = _f32(cov3D_precomp, "cov3D_precomp")
    background, viewmatrix = _f32(background, "bg"), _f32(viewmatrix, "viewmatrix")
    projmatrix, prcppoint = _f32(projmatrix, "projmatrix"), _f32(prcppoint, "prcppoint")
    cfg_flags = _config_flags(config)
    patchbbox, campos, config = _f32(patchbbox, "patch_bbox"), _f32(campos, "campos"), _f32(config, "config")
    dL_dout_color, dL_dout_normal = _f32(dL_dout_color, "grad color"), _f32(dL_dout_normal, "grad normal")
    dL_dout_depth, dL_dout_opac = _f32(dL_dout_depth, "grad depth"), _f32(dL_dout_opac, "grad opac")
    radii = radii.contiguous()

    with torch.cuda.device(dev):
        # opacities are not needed by the backward (they live in the saved records)
        view = _view(P, M, degree, H, W, tan_fovx, tan_fovy, scale_modifier, False, False, False,
                     debug, background, means3D, sh, colors, None, scales, rotations,
                     cov3D_precomp, viewmatrix, projmatrix, prcppoint, patchbbox, campos, config,
                     cfg_flags)
        grads = _capi.SoarRastGrads()
        grads.dL_dmeans2D, grads.dL_dcolors = g_means2D.data_ptr(), g_colors.data_ptr()
        grads.dL_dopacity, grads.dL_dmeans3D = g_opacity.data_ptr(), g_means3D.data_ptr()
        grads.dL_dcov3D, grads.dL_dsh = g_cov3D.data_ptr(), _capi.ptr(g_sh)
        grads.dL_dscales, grads.dL_drotations = g_scales.data_ptr(), g_rot.data_ptr()
        grads.dL_dviewmat, grads.dL_dprojmat = g_view.data_ptr(), g_proj.data_ptr()
        grads.dL_dcampos = g_campos.data_ptr()
        grads.accumulate_mask = mask
        scratch = torch.empty(_lib.soar_rast_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
        after, done = order if order is not None else (None, None)
        _capi.check(_lib.soar_rast_backward_ordered(
            ctypes.byref(view), radii.data_ptr(), geomBuffer.data_ptr(), binningBuffer.data_ptr(),
            imageBuffer.data_ptr(), int(R), dL_dout_color.data_ptr(), dL_dout_normal.data_ptr(),
            dL_dout_depth.data_ptr(), dL_dout_opac.data_ptr(), scratch.data_ptr(), scratch.numel(),
            ctypes.byref(grads), ctypes.c_void_p(after.cuda_event) if after is not None else None,
            ctypes.c_void_p(done.cuda_event) if done is not None else None, _stream()),
            "rasterize_gaussians_backward")
        if debug:
            torch.cuda.synchronize(dev)
    return (g_means2D, g_colors, g_opacity, g_means3D, g_cov3D, g_sh, g_scales, g_rot, g_view,
            g_proj, g_campos)


def mark_visible(means3D, viewmatrix, projmatrix):
    """The reference kernel body is commented out (DGR/cuda_rasterizer/rasterizer_impl.cu:52-62):
    the all-False tensor allocated by the glue is what callers get (Q11)."""
    P = means3D.size(0)
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P != 0:
        _capi.check(_lib.soar_rast_mark_visible(P, means3D.data_ptr(), viewmatrix.data_ptr(),
                                                projmatrix.data_ptr(), present.data_ptr(), _stream()),
                    "mark_visible")
    return present
