"""TEST INFRASTRUCTURE ONLY — CPU (torch, autograd-capable) restatement of the reference's
skinning of canonical Gaussians.

Follows, line by line:
  * TS/utils/smpl.py:609-613   cano2live bones; pt_mats = einsum("bnj,bjxy->bnxy", weights, bones)
  * TS/utils/smpl.py:618-637   query_weights_smpl: K=30 nearest canonical vertices, d = sqrt(d2)
                               clamped to [1e-4, 1], inverse-distance blend of their skin weights
  * TS/renderer/diff_gaussian_rasterizer.py:103-114 (SDS views, with the "+z,+x,+y" swizzle of
    transform_point_cloud :321-352) and :138-149 (gt views, no swizzle)

Third-party arithmetic that is NOT under /root/reference: pytorch3d (unpinned git HEAD,
setup.py:146) — `quaternion_to_matrix` (a verbatim copy is vendored at
TS/data/uncond_multiview.py:2422-2451 and is what is restated here), `matrix_to_quaternion`
(not vendored; its published algorithm is restated: q_abs = sqrt(max(0, 1 +- m00 +- m11 +- m22)),
four candidates / (2*max(q_abs, 0.1)), one-hot selection of the candidate with the largest q_abs,
then standardize_quaternion makes the real part non-negative) and `knn_points` (exact K nearest,
squared distances ascending; restated with cdist+topk). PARITY UNPINNED for these three: the
reference tree holds no test or golden vector for them and pytorch3d cannot be installed here.
"""
import torch


def quaternion_to_matrix(q):
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack(
        (1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
         two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
         two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def _sqrt_positive_part(x):
    ret = torch.zeros_like(x)
    pos = x > 0
    return torch.where(pos, torch.sqrt(torch.where(pos, x, torch.ones_like(x))), ret)


def matrix_to_quaternion(matrix):
    batch = matrix.shape[:-2]
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(matrix.reshape(batch + (9,)), dim=-1)
    q_abs = _sqrt_positive_part(torch.stack(
        [1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22, 1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22],
        dim=-1))
    quat_by_rijk = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1),
    ], dim=-2)
    flr = torch.tensor(0.1, dtype=q_abs.dtype, device=q_abs.device)
    cand = quat_by_rijk / (2.0 * q_abs[..., None].max(flr))
    sel = torch.nn.functional.one_hot(q_abs.argmax(dim=-1), num_classes=4) > 0.5
    out = cand[sel, :].reshape(batch + (4,))
    return torch.where(out[..., 0:1] < 0, -out, out)  # standardize_quaternion


def blend_bones(weights, bones):
    """weights [N,J], bones [J,4,4] -> pt_mats [1,N,4,4] (TS/utils/smpl.py:613)."""
    return torch.einsum("bnj,bjxy->bnxy", weights[None], bones[None])


def skin(points, quats, weights, bones, swizzle):
    """-> (points' [N,3], quats' [N,4], pt_mats [N,4,4]); differentiable w.r.t. points, quats."""
    mat = blend_bones(weights.detach(), bones)
    pts = (torch.einsum("bnxy,bny->bnx", mat[..., :3, :3], points[None]) + mat[..., :3, 3])[0]
    rot_mat = quaternion_to_matrix(quats)
    rot_mat = torch.matmul(mat[..., :3, :3], rot_mat)
    if swizzle:
        T = torch.zeros(3, 3, dtype=points.dtype)
        T[2, 0] = 1.0  # "+z" -> column 0
        T[0, 1] = 1.0  # "+x" -> column 1
        T[1, 2] = 1.0  # "+y" -> column 2
        pts = torch.matmul(pts, T)
        rot_mat = torch.matmul(T.T, rot_mat)
    rot = matrix_to_quaternion(rot_mat)
    rot = torch.nn.functional.normalize(rot, p=2, dim=-1)[0]
    return pts, rot, mat[0]


def query_weights(points, cano_vertices, lbs_weights, K=30):
    """TS/utils/smpl.py:618-637 with pytorch3d.knn_points restated as cdist + topk."""
    d2 = torch.cdist(points.double(), cano_vertices.double()).pow(2).float()
    dist, idx = torch.topk(d2, K, dim=1, largest=False)
    dist = dist.sqrt().clamp_(0.0001, 1.0)
    w = lbs_weights[idx]
    ws = 1.0 / dist
    ws = ws / ws.sum(-1)[..., None]
    return (ws[..., None] * w).sum(-2)
