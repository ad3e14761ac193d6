"""TEST INFRASTRUCTURE ONLY — CPU restatement (torch, autograd-differentiable) of the reference's
image-space post-ops. Only tests/, __graft_entry__.smoke() and bench.py's baseline legs may import it.

Pinned: tests/golden/postops_*.npz hold outputs AND input gradients of the reference's own
functions (TS/renderer/diff_gaussian_rasterizer.py), executed in the build container by
tests/golden/make_golden_postops.py; tests/test_oracle_golden.py checks this file against them.

Device-agnostic torch code: the GPU tests also evaluate it on the CUDA device at full image sizes.

Reference quirk (kept out of the comparisons, see tests): where a stencil is degenerate (all cross
products exactly 0 inside the mask), F.normalize's eps clamp turns the incoming gradient into
1e12-scale terms that cancel in whatever order autograd happens to add them, so the reference's
dL/ddepth in the 3x3 neighbourhood of such a pixel is cancellation noise (often exactly 0).

The restatement indexes neighbours with clamped coordinates (= the reference's replicate padding,
:391-394 / :437-440) instead of padding and slicing.
"""
import math

import torch


def _neighbours(img, H, W):
    """img [H,W,C] -> (up, left, below, right) with border replication."""
    ys = torch.arange(H, device=img.device)
    xs = torch.arange(W, device=img.device)
    up = img[(ys - 1).clamp(min=0)]
    below = img[(ys + 1).clamp(max=H - 1)]
    left = img[:, (xs - 1).clamp(min=0)]
    right = img[:, (xs + 1).clamp(max=W - 1)]
    return up, left, below, right


def camera_matrix_inverse(camera):
    """Kinv as the reference forms it (:378-381): float32 K = diag(focal(FoVy, H), focal(FoVx, W)),
    torch.inverse on the CPU."""
    k00 = camera.image_height / (2 * math.tan(camera.FoVy / 2))
    k11 = camera.image_width / (2 * math.tan(camera.FoVx / 2))
    return torch.inverse(torch.tensor([k00, 0, 0, k11]).reshape(2, 2))


def depth2normal(depth, mask, camera):
    """:359-430. depth [1,H,W] float32, mask [1,H,W] bool -> [3,H,W]."""
    _, H, W = depth.shape
    D = depth[0][..., None]                                   # [H,W,1]
    m = mask[0][..., None].to(torch.float32)
    dev = depth.device
    px = torch.arange(W, dtype=torch.float32, device=dev).view(1, W, 1).expand(H, W, 1)
    py = torch.arange(H, dtype=torch.float32, device=dev).view(H, 1, 1).expand(H, W, 1)
    pp = torch.as_tensor(camera.prcppoint, dtype=torch.float32).cpu()
    cx, cy = float(pp[0] * camera.image_width), float(pp[1] * camera.image_height)
    xy = torch.cat([px - cx, py - cy], -1) * D
    Kinv = camera_matrix_inverse(camera).to(D.device)
    cam = torch.cat([xy @ Kinv.t(), D], -1)                   # [H,W,3]  (:383-384)
    nb = _neighbours(cam, H, W)
    nm = _neighbours(m, H, W)
    c = cam * m
    u, l, b, r = ((p - c) * pm for p, pm in zip(nb, nm))      # :396-400
    n = torch.linalg.cross(u, l) + torch.linalg.cross(r, u) + torch.linalg.cross(b, r) \
        + torch.linalg.cross(l, b)                            # :402-420
    n = torch.nn.functional.normalize(n, dim=-1)              # :426
    return (n * m).permute(2, 0, 1)


def normal2curv(normal, mask):
    """:433-448. normal [3,H,W], mask [1,H,W] bool -> [1,H,W]."""
    _, H, W = normal.shape
    n = normal.permute(1, 2, 0)
    m = mask[0][..., None].to(torch.float32)
    nb = _neighbours(n, H, W)
    nm = _neighbours(m, H, W)
    c = n * m
    u, l, b, r = ((p - c) * pm for p, pm in zip(nb, nm))
    s = (u + l + b + r).permute(2, 0, 1) * mask
    return s.abs().sum(0, keepdim=True)


def surfel_postprocess(rendered_normal, rendered_depth, rendered_opac, camera):
    """The renderer's post block, :292-303 (functional: inputs are not modified)."""
    mask = rendered_opac.detach() > 1e-5
    keep = mask.expand(3, -1, -1)
    n = torch.where(keep, rendered_normal, rendered_normal.detach())
    n = n * torch.tensor([1.0, -1.0, -1.0], device=n.device).view(3, 1, 1)
    curv = normal2curv(n, mask)
    dn = depth2normal(rendered_depth, mask, camera)
    dn = dn * torch.tensor([1.0, -1.0, -1.0], device=n.device).view(3, 1, 1)
    return (n + 1) / 2, curv, (dn + 1) / 2, mask
