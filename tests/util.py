"""Shared helpers of the test-suite: move synthetic scenes to the device and run the product
rasterizer through its public (reference-compatible) API."""
import torch

from soar_b200 import synthetic


def to_device(s, device="cuda"):
    return {k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in s.items()}


def settings_of(s):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=s["image_height"], image_width=s["image_width"], tanfovx=s["tanfovx"],
        tanfovy=s["tanfovy"], bg=s["bg"], scale_modifier=s["scale_modifier"],
        viewmatrix=s["viewmatrix"], projmatrix=s["projmatrix"], patch_bbox=s["patch_bbox"],
        prcppoint=s["prcppoint"], sh_degree=s["sh_degree"], campos=s["campos"],
        prefiltered=s["prefiltered"], render_front=s["render_front"],
        sort_descending=s["sort_descending"], debug=s["debug"], config=s["config"])


LEAVES = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations")


def run_ours(s, cot=None):
    """Forward (+ backward when cotangents are given) through GaussianRasterizer.
    Returns (outputs dict, grads dict or None)."""
    from diff_gaussian_rasterization import GaussianRasterizer
    leaves = {}
    for k in LEAVES:
        v = s.get(k)
        leaves[k] = None if v is None else v.detach().clone().requires_grad_(cot is not None)
    means2D = torch.zeros_like(leaves["means3D"], requires_grad=cot is not None)
    rast = GaussianRasterizer(settings_of(s))
    color, normal, depth, opac, radii = rast(
        means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
        shs=leaves["shs"], colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
        rotations=leaves["rotations"], cov3D_precomp=None)
    out = dict(color=color, normal=normal, depth=depth, opac=opac, radii=radii)
    if cot is None:
        return out, None
    loss = ((color * cot["color"]).sum() + (normal * cot["normal"]).sum()
            + (depth * cot["depth"]).sum() + (opac * cot["opac"]).sum())
    loss.backward()
    grads = dict(means2D=means2D.grad, means3D=leaves["means3D"].grad,
                 opacity=leaves["opacities"].grad, scales=leaves["scales"].grad,
                 rotations=leaves["rotations"].grad)
    if leaves["shs"] is not None:
        grads["sh"] = leaves["shs"].grad
    if leaves["colors_precomp"] is not None:
        grads["colors"] = leaves["colors_precomp"].grad
    return {k: v.detach() for k, v in out.items()}, grads


def rel_err(a, b):
    """max |a-b| / max(|b|max, tiny): relative to the tensor's scale (gradients near zero are
    compared against the tensor's max-abs)."""
    scale = b.abs().max().clamp_min(1e-30)
    return ((a - b).abs().max() / scale).item()
