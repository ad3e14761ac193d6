"""Skinning of canonical Gaussians with SMPL-X bone transforms on the sm_100a
kernel (soar_b200/csrc/lbs.cu) — one fused op for the reference's torch chain:

    pt_mats = einsum("bnj,bjxy->bnxy", weights, cano2live_jnt_mats)   TS/utils/smpl.py:613
    points  = einsum("bnxy,bny->bnx", mat[..., :3, :3], points) + mat[..., :3, 3]
    points, T = transform_point_cloud(points, "+z,+x,+y")              (SDS views only)
    rot_mat = T.T @ (mat[..., :3, :3] @ quaternion_to_matrix(rot))
    rot     = normalize(matrix_to_quaternion(rot_mat))
                                   TS/renderer/diff_gaussian_rasterizer.py:103-114 / :138-149

Gradients flow to `points` and `rot` only: the reference detaches the skin
weights (TS/utils/smpl.py:611) and the pose tensors are plain buffers (:543-545).
"""
import ctypes

import torch

from . import _capi

_lib = _capi.load()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


class _SkinGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, quats, weights, bones, swizzle):
        N, J = weights.shape
        points_c = points.detach().float().contiguous()
        quats_c = quats.detach().float().contiguous()
        weights_c = weights.detach().float().contiguous()
        bones_c = bones.detach().float().reshape(J, 16).contiguous()
        dev = points.device
        out_p = torch.empty((N, 3), dtype=torch.float32, device=dev)
        out_q = torch.empty((N, 4), dtype=torch.float32, device=dev)
        mats = torch.empty((N, 12), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_lib.soar_lbs_forward(_capi.ptr(weights_c), bones_c.data_ptr(), J,
                                              _capi.ptr(points_c), _capi.ptr(quats_c), N,
                                              int(swizzle), _capi.ptr(out_p), _capi.ptr(out_q),
                                              _capi.ptr(mats), _stream()), "soar_lbs_forward")
        ctx.save_for_backward(mats, quats_c)
        ctx.swizzle = int(swizzle)
        ctx.mark_non_differentiable(mats)
        return out_p, out_q, mats

    @staticmethod
    def backward(ctx, g_points, g_quats, _g_mats):
        mats, quats = ctx.saved_tensors
        N = quats.shape[0]
        dev = quats.device
        if g_points is None:
            g_points = torch.zeros((N, 3), dtype=torch.float32, device=dev)
        if g_quats is None:
            g_quats = torch.zeros((N, 4), dtype=torch.float32, device=dev)
        g_points = g_points.float().contiguous()
        g_quats = g_quats.float().contiguous()
        d_points = torch.empty((N, 3), dtype=torch.float32, device=dev)
        d_quats = torch.empty((N, 4), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_lib.soar_lbs_backward(_capi.ptr(mats), _capi.ptr(quats), N, ctx.swizzle,
                                               _capi.ptr(g_points), _capi.ptr(g_quats),
                                               _capi.ptr(d_points), _capi.ptr(d_quats), _stream()),
                        "soar_lbs_backward")
        return d_points, d_quats, None, None, None


def skin_gaussians(points, quats, weights, bones, swizzle=False):
    """points [N,3], quats [N,4] (w,x,y,z), weights [N,J] (detached), bones [J,4,4] or [1,J,4,4]
    (cano->live) -> (points' [N,3], quats' [N,4] unit, pt_mats [N,12] = rows 0..2 of the blended 4x4).
    swizzle=True applies the SDS-view axis change (x,y,z)->(z,x,y)."""
    if not points.is_cuda:
        raise RuntimeError("skin_gaussians: CUDA tensors required (there is no CPU path)")
    bones = bones.reshape(-1, 4, 4)
    if weights.dim() == 3:
        weights = weights[0]
    if weights.shape[0] != points.shape[0] or weights.shape[1] != bones.shape[0]:
        raise RuntimeError(f"skin_gaussians: weights {tuple(weights.shape)} do not match "
                           f"{points.shape[0]} points x {bones.shape[0]} bones")
    return _SkinGaussians.apply(points, quats, weights, bones, swizzle)


class SkinWeightQuery:
    """SMPL_Guidance.query_weights_smpl (TS/utils/smpl.py:618-637) on the sm_100a kernel
    (soar_b200/csrc/lbs_query.cu): exact K nearest canonical vertices, distances clamped to
    [1e-4, 1], inverse-distance blend of their skin-weight rows. The result is detached, like
    the reference's (:611). `mats(points, bones)` fuses the bone blend of :613 and returns the
    per-point 3x4 transforms without materialising the [N,J] weights."""

    def __init__(self, cano_vertices, lbs_weights, K=30):
        if not cano_vertices.is_cuda:
            raise RuntimeError("SkinWeightQuery: CUDA tensors required (there is no CPU path)")
        dev = cano_vertices.device
        self.verts = cano_vertices.detach().to(device=dev, dtype=torch.float32).reshape(-1, 3).contiguous()
        self.weights = lbs_weights.detach().to(device=dev, dtype=torch.float32)
        if self.weights.dim() == 3:
            self.weights = self.weights[0]
        self.weights = self.weights.contiguous()
        if self.weights.shape[0] != self.verts.shape[0]:
            raise RuntimeError("SkinWeightQuery: lbs_weights must have one row per canonical vertex")
        self.K = int(K)
        self.brute_force = False   # True: the V-distances-per-point kernel (kept for ablation)
        self.per_point = False     # True: always the per-point grid walk (never the cell-batched kernel)
        V = self.verts.shape[0]
        self.grid = torch.empty(_lib.soar_lbs_query_grid_bytes(V), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _capi.check(_lib.soar_lbs_query_build_grid(self.verts.data_ptr(), V, self.grid.data_ptr(),
                                                       self.grid.numel(), _stream()),
                        "soar_lbs_query_build_grid")

    def _run(self, points, bones, want_weights, want_mats):
        if not points.is_cuda:
            raise RuntimeError("SkinWeightQuery: points must be a CUDA tensor")
        pts = points.detach().to(torch.float32).reshape(-1, 3).contiguous()
        N, (V, J) = pts.shape[0], self.weights.shape
        dev = pts.device
        w = torch.empty((N, J), dtype=torch.float32, device=dev) if want_weights else None
        m = torch.empty((N, 12), dtype=torch.float32, device=dev) if want_mats else None
        b = None
        if bones is not None:
            b = bones.detach().to(device=dev, dtype=torch.float32).reshape(-1, 16).contiguous()
            if b.shape[0] != J:
                raise RuntimeError(f"SkinWeightQuery: {b.shape[0]} bones for {J} skin-weight columns")
        with torch.cuda.device(dev):
            if self.brute_force:
                _capi.check(_lib.soar_lbs_query_weights(_capi.ptr(pts), N, self.verts.data_ptr(), V,
                                                        self.weights.data_ptr(), J, self.K, _capi.ptr(b),
                                                        _capi.ptr(w), _capi.ptr(m), _stream()),
                            "soar_lbs_query_weights")
            elif self.per_point:
                _capi.check(_lib.soar_lbs_query_weights_grid(_capi.ptr(pts), N, self.grid.data_ptr(), V,
                                                             self.weights.data_ptr(), J, self.K,
                                                             _capi.ptr(b), _capi.ptr(w), _capi.ptr(m),
                                                             _stream()),
                            "soar_lbs_query_weights_grid")
            else:
                # large N: queries sorted by grid cell, one ring walk per (warp, cell); the library falls
                # back to the per-point walk by itself for small N, K != 30 or very large V
                ws = torch.empty(_lib.soar_lbs_query_cells_workspace_bytes(N), dtype=torch.uint8, device=dev)
                _capi.check(_lib.soar_lbs_query_weights_cells(_capi.ptr(pts), N, self.grid.data_ptr(), V,
                                                              self.weights.data_ptr(), J, self.K,
                                                              _capi.ptr(b), _capi.ptr(w), _capi.ptr(m),
                                                              _capi.ptr(ws), ws.numel(), _stream()),
                            "soar_lbs_query_weights_cells")
        return w, m

    def __call__(self, points):
        """points [N,3] -> skin weights [N,J]."""
        return self._run(points, None, True, False)[0]

    def mats(self, points, bones):
        """points [N,3], bones [J,4,4] -> per-point transforms [N,12] (rows 0..2 of the blend)."""
        return self._run(points, bones, False, True)[1]
