"""LBS skinning kernel vs. the torch-CPU restatement of the reference op chain
(oracle/lbs_oracle.py). Floating point: forward within 2e-6 abs on points (metres) and
2e-5 on unit quaternions; gradients within 1e-4 relative to the tensor's max-abs."""
import pytest
import torch

from oracle import lbs_oracle
from soar_b200 import synthetic
from util import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("N,swizzle", [(1, False), (127, True), (4097, False), (100_000, True)])
def test_forward_backward(N, swizzle):
    from soar_b200.lbs import skin_gaussians
    pb = synthetic.skinning_problem(N, seed=2)
    pts_c = pb["points"].clone().requires_grad_(True)
    q_c = pb["quats"].clone().requires_grad_(True)
    o_p, o_q, o_m = lbs_oracle.skin(pts_c, q_c, pb["weights"], pb["bones"], swizzle)
    gen = torch.Generator().manual_seed(5)
    gp, gq = torch.randn(N, 3, generator=gen), torch.randn(N, 4, generator=gen)
    ((o_p * gp).sum() + (o_q * gq).sum()).backward()

    pts_g = pb["points"].cuda().requires_grad_(True)
    q_g = pb["quats"].cuda().requires_grad_(True)
    p, q, m = skin_gaussians(pts_g, q_g, pb["weights"].cuda(), pb["bones"].cuda(), swizzle=swizzle)
    ((p * gp.cuda()).sum() + (q * gq.cuda()).sum()).backward()

    assert (p.detach().cpu() - o_p.detach()).abs().max().item() <= 2e-6
    assert (m.cpu().view(N, 3, 4) - o_m[:, :3, :].detach()).abs().max().item() <= 2e-6
    # q and -q are the same rotation, but the standardised sign must agree as well
    assert (q.detach().cpu() - o_q.detach()).abs().max().item() <= 2e-5
    assert rel_err(pts_g.grad.cpu(), pts_c.grad) <= 1e-4
    assert rel_err(q_g.grad.cpu(), q_c.grad) <= 1e-4


def test_rigid_bones_give_rigid_motion():
    """Property at BASELINE config-5 size (2M): with one-hot weights the blended transform is a
    rigid motion, so |x'_a - x'_b| = |x_a - x_b| for points bound to the same bone and the
    output quaternion is unit."""
    from soar_b200.lbs import skin_gaussians
    N = 2_000_000
    pb = synthetic.skinning_problem(4096, seed=1)
    reps = (N + 4095) // 4096
    pts = pb["points"].repeat(reps, 1)[:N].cuda()
    quats = pb["quats"].repeat(reps, 1)[:N].cuda()
    w = torch.zeros(N, synthetic.N_JOINTS, device="cuda")
    bone = torch.arange(N, device="cuda") % synthetic.N_JOINTS
    w[torch.arange(N, device="cuda"), bone] = 1.0
    p, q, _ = skin_gaussians(pts, quats, w, pb["bones"].cuda(), swizzle=False)
    assert (q.norm(dim=1) - 1).abs().max().item() <= 1e-5
    a = torch.arange(0, N - 55, 55 * 7, device="cuda")
    b = a + 55  # same bone (index congruent mod 55)
    d_in = (pts[a] - pts[b]).norm(dim=1)
    d_out = (p[a] - p[b]).norm(dim=1)
    assert (d_in - d_out).abs().max().item() <= 1e-5


# ---- SMPL-X forward (a17) and skin-weight query (a18) ----------------------------------------
import glob as _glob
import os as _os

import numpy as _np

_GOLDEN = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "golden")
_SMPLX_CASES = sorted(_os.path.basename(p)[:-4] for p in _glob.glob(_os.path.join(_GOLDEN, "smplx_*.npz")))


@pytest.mark.parametrize("name", _SMPLX_CASES)
def test_smplx_forward_matches_reference_golden(name):
    """soar_smplx_forward vs the outputs of the reference's own lbs() (float32; the kernels sum
    in a different order than torch's matmuls: 3e-6 absolute on metres-scale outputs)."""
    from soar_b200.smplx_fwd import SmplxPoser
    d = _np.load(_os.path.join(_GOLDEN, name + ".npz"))
    m = synthetic.smplx_template(V=int(d["V"]), NB=int(d["NB"]), seed=int(d["template_seed"]))
    poser = SmplxPoser(*(m[k].cuda() for k in ("v_template", "shapedirs", "posedirs", "J_regressor",
                                               "parents", "lbs_weights")))
    t = lambda k: torch.from_numpy(d[k]).cuda()
    verts, joints, A, bones = poser(t("betas"), t("full_pose"), t("transl"), t("inv_mats"))
    for got, k in ((verts, "verts"), (joints, "joints"), (A, "A"), (bones, "bones")):
        assert (got.cpu().numpy() - d[k]).reshape(-1).__abs__().max() <= 3e-6, k
    # twice on the same workspace: deterministic (no float atomics)
    v2, j2, A2, b2 = poser(t("betas"), t("full_pose"), t("transl"), t("inv_mats"))
    assert torch.equal(verts, v2) and torch.equal(A, A2) and torch.equal(bones, b2)


def test_smplx_forward_matches_oracle_other_pose():
    from oracle import smplx_oracle
    from soar_b200.smplx_fwd import SmplxPoser
    m = synthetic.smplx_template(V=3000, NB=16, seed=9)
    p = synthetic.smplx_pose(NB=16, seed=9, global_orient_zero=False)
    o = smplx_oracle.smplx_forward({k: v.numpy() for k, v in m.items()}, p["betas"].numpy(),
                                   p["full_pose"].numpy(), None)
    poser = SmplxPoser(*(m[k].cuda() for k in ("v_template", "shapedirs", "posedirs", "J_regressor",
                                               "parents", "lbs_weights")))
    verts, joints, A, bones = poser(p["betas"].cuda(), p["full_pose"].cuda(), None, None)
    assert bones is None
    assert _np.abs(verts.cpu().numpy() - o["verts"]).max() <= 3e-6
    assert _np.abs(joints.cpu().numpy() - o["joints"]).max() <= 3e-6
    assert _np.abs(A.cpu().numpy() - o["A"]).max() <= 3e-6


@pytest.mark.parametrize("N,V", [(1, 64), (1000, 500), (8_000, synthetic.N_VERTS)])
def test_skin_weight_query_matches_oracle(N, V):
    """Exact 30-NN + inverse-distance blend vs the torch restatement (cdist + topk). Weights are in
    [0,1]: 2e-5 absolute (a neighbour swap at a distance tie only reorders equal contributions)."""
    from soar_b200.lbs import SkinWeightQuery
    m = synthetic.smplx_template(V=V, NB=1, seed=6)
    gen = torch.Generator().manual_seed(N)
    src = m["v_template"][torch.randint(0, V, (N,), generator=gen)]
    pts = src + 0.02 * torch.randn(N, 3, generator=gen)
    pts[: N // 4] = src[: N // 4]  # exactly on a vertex: distance clamps at 1e-4 (Q18)
    want = lbs_oracle.query_weights(pts, m["v_template"], m["lbs_weights"], K=30)
    q = SkinWeightQuery(m["v_template"].cuda(), m["lbs_weights"].cuda(), K=30)
    got = q(pts.cuda())
    assert got.shape == (N, synthetic.N_JOINTS)
    assert (got.cpu() - want).abs().max().item() <= 2e-5
    assert (got.sum(1) - 1).abs().max().item() <= 1e-5
    # fused bone blend == weights @ bones
    bones = synthetic.skinning_problem(8, seed=3)["bones"].cuda()
    mats = q.mats(pts.cuda(), bones)
    ref = (got @ bones.reshape(-1, 16))[:, :12]
    assert (mats - ref).abs().max().item() <= 1e-5


def test_skin_weight_query_rejects_cpu_and_bad_sizes():
    from soar_b200.lbs import SkinWeightQuery
    m = synthetic.smplx_template(V=64, NB=1, seed=6)
    with pytest.raises(RuntimeError):
        SkinWeightQuery(m["v_template"], m["lbs_weights"])
    q = SkinWeightQuery(m["v_template"].cuda(), m["lbs_weights"].cuda(), K=64)
    with pytest.raises(RuntimeError):
        q(m["v_template"].cuda())


@pytest.mark.parametrize("V,K", [(synthetic.N_VERTS, 30), (500, 30), (20, 30), (1, 1), (3000, 7)])
def test_skin_weight_query_grid_is_bit_identical_to_brute_force(V, K):
    """The grid search must pick exactly the brute force's neighbours (same order, index
    tie-break) for points on the body, far outside the bounding box, on duplicated vertices and at
    non-finite coordinates — the blended weights are then bit-identical."""
    from soar_b200.lbs import SkinWeightQuery
    m = synthetic.smplx_template(V=max(V, 2), NB=1, seed=9)
    verts, lw = m["v_template"][:V].clone(), m["lbs_weights"][:V].clone()
    if V >= 20:
        verts[V // 2: V // 2 + 5] = verts[0]          # duplicated vertices: distance ties
    gen = torch.Generator().manual_seed(V + K)
    N = 20_000
    src = verts[torch.randint(0, V, (N,), generator=gen)]
    pts = src + 0.004 * torch.randn(N, 3, generator=gen)
    pts[:2000] = src[:2000]                            # exactly on vertices
    pts[2000:4000] += 0.3 * torch.randn(2000, 3, generator=gen)   # off the body
    pts[4000:4500] = 50.0 * torch.randn(500, 3, generator=gen)    # far outside the grid
    pts[4500] = torch.tensor([1e30, 0.0, 0.0])
    pts[4501] = torch.tensor([-3e38, 3e38, 0.0])
    q = SkinWeightQuery(verts.cuda(), lw.cuda(), K=K)
    fast = q(pts.cuda())
    q.brute_force = True
    slow = q(pts.cuda())
    same = (fast == slow) | (fast.isnan() & slow.isnan())
    assert same.all(), f"{(~same).any(1).sum().item()} of {N} points differ"
    bones = synthetic.skinning_problem(8, seed=3)["bones"].cuda()
    q.brute_force = False
    m_fast = q.mats(pts.cuda(), bones)
    q.brute_force = True
    m_slow = q.mats(pts.cuda(), bones)
    assert ((m_fast == m_slow) | (m_fast.isnan() & m_slow.isnan())).all()


@pytest.mark.parametrize("name", ("weights_full", "weights_dups", "weights_tiny"))
def test_skin_weight_query_matches_reference_golden(name):
    """Rows a18 / f1 against the reference's OWN lines: golden weights recorded by executing
    SMPL_Guidance.query_weights_smpl (TS/utils/smpl.py:618-637) unchanged with an exact knn_points
    stand-in (tests/golden/make_golden_weights.py) — on-vertex, near-surface, off-body, > 1 m away and
    duplicated-vertex queries. Both the grid kernel and the brute-force kernel."""
    import golden_util as G
    from soar_b200.lbs import SkinWeightQuery
    from test_lbs_oracle_golden import template_of
    d = G.load(name)
    verts, lw = template_of(d)
    want = torch.from_numpy(d["weights"])
    q = SkinWeightQuery(verts.cuda(), lw.cuda(), K=min(30, verts.shape[0]))
    pts = torch.from_numpy(d["points"]).cuda()
    for brute in (False, True):
        q.brute_force = brute
        got = q(pts).cpu()
        err = (got - want).abs().max().item()
        print(f"[{name}] {'brute' if brute else 'grid'}: max-abs {err:.2e}")
        assert err <= 2e-6, f"{name} ({'brute' if brute else 'grid'}): {err:.2e}"


def test_cell_batched_query_is_bit_identical_to_per_point_walk_at_2m():
    """BASELINE config-5 size: the cell-batched kernel (queries sorted by grid cell, a warp walks a
    cell's search cube once for all its lanes, per-lane register insertion) against the per-point
    grid walk — every weight bit for bit, including queries outside the grid (the sentinel tail)."""
    from soar_b200.lbs import SkinWeightQuery
    m = synthetic.smplx_template(NB=1, seed=0)
    N = 2_000_000
    pts = synthetic.avatar_gaussians(N, m["v_template"], seed=0)["points"].cuda()
    gen = torch.Generator(device="cuda").manual_seed(3)
    pts[:5000] += 0.5 * torch.randn(5000, 3, device="cuda", generator=gen)    # off the body, some outside the grid
    pts[5000:5100] = 30.0 * torch.randn(100, 3, device="cuda", generator=gen)  # far outside
    pts[5100] = float("nan")
    q = SkinWeightQuery(m["v_template"].cuda(), m["lbs_weights"].cuda(), K=30)
    fast = q(pts)
    q.per_point = True
    slow = q(pts)
    same = (fast == slow) | (fast.isnan() & slow.isnan())
    assert same.all(), f"{(~same).any(1).sum().item()} of {N} points differ"
    bones = synthetic.skinning_problem(8, seed=3)["bones"].cuda()
    m_slow = q.mats(pts, bones)
    q.per_point = False
    m_fast = q.mats(pts, bones)
    assert ((m_fast == m_slow) | (m_fast.isnan() & m_slow.isnan())).all()


def test_cell_batched_query_tie_overflow_goes_through_the_redo_list():
    """40 coincident template vertices: every query whose K nearest include that spot has 40 candidates
    at one and the same distance, more than the cell-batched kernel's per-lane list holds (32), so it
    must hand those queries to the per-point kernel through its redo list (lbs_query.cu, `overflow`).
    N is far above 32 x 24, so the cell-batched kernel is the one that runs. Bit-identical to the
    per-point walk, weights and fused transforms, and the overflow condition is shown to occur."""
    from soar_b200.lbs import SkinWeightQuery
    m = synthetic.smplx_template(NB=1, seed=0)
    verts = m["v_template"].clone()
    spot = verts[100].clone()
    verts[100:140] = spot                                   # 40 coincident vertices
    N = 200_000
    pts = synthetic.avatar_gaussians(N, m["v_template"], seed=1)["points"].cuda()
    gen = torch.Generator(device="cuda").manual_seed(5)
    pts[:4000] = spot.cuda() + 0.01 * torch.randn(4000, 3, device="cuda", generator=gen)
    # the condition itself, by brute force on a few of those queries: more than 32 vertices at or
    # below the 30th smallest distance
    d = ((pts[:64, None, :] - verts.cuda()[None, :, :]) ** 2).sum(-1)
    kth = d.kthvalue(30, dim=1).values
    assert ((d <= kth[:, None]).sum(1) > 32).any(), "the scene does not overflow the tie list"
    q = SkinWeightQuery(verts.cuda(), m["lbs_weights"].cuda(), K=30)
    fast = q(pts)
    bones = synthetic.skinning_problem(8, seed=3)["bones"].cuda()
    m_fast = q.mats(pts, bones)
    q.per_point = True
    slow = q(pts)
    m_slow = q.mats(pts, bones)
    assert torch.isfinite(fast).all()
    assert torch.equal(fast, slow), f"{(fast != slow).any(1).sum().item()} of {N} points differ"
    assert torch.equal(m_fast, m_slow)
    # and the weights are a convex blend (rows sum to 1 like the template's)
    assert (fast.sum(1) - 1).abs().max().item() < 1e-4


def _m2q_candidates(M):
    """The four candidate quaternions of pytorch3d's matrix_to_quaternion for M [N,3,3] (float64) and
    their q_abs, written out candidate by candidate (w-, x-, y-, z-dominant)."""
    m = M.double()
    m00, m01, m02 = m[:, 0, 0], m[:, 0, 1], m[:, 0, 2]
    m10, m11, m12 = m[:, 1, 0], m[:, 1, 1], m[:, 1, 2]
    m20, m21, m22 = m[:, 2, 0], m[:, 2, 1], m[:, 2, 2]
    q2 = torch.stack([1 + m00 + m11 + m22, 1 + m00 - m11 - m22, 1 - m00 + m11 - m22, 1 - m00 - m11 + m22], 1)
    q_abs = q2.clamp(min=0).sqrt()
    num = torch.stack([
        torch.stack([q2[:, 0].clamp(min=0), m21 - m12, m02 - m20, m10 - m01], 1),
        torch.stack([m21 - m12, q2[:, 1].clamp(min=0), m10 + m01, m02 + m20], 1),
        torch.stack([m02 - m20, m10 + m01, q2[:, 2].clamp(min=0), m12 + m21], 1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q2[:, 3].clamp(min=0)], 1)], 1)       # [N,4 cand,4]
    cand = num / (2 * q_abs.clamp(min=0.1))[:, :, None]
    cand = torch.where(cand[:, :, 0:1] < 0, -cand, cand)                                    # standardize
    return torch.nn.functional.normalize(cand, dim=-1), q_abs


def _skin_identity_quats(mats3):
    """Runs the product's skinning with one bone per point (one-hot weights) and identity input
    quaternions, so the matrix handed to its matrix_to_quaternion is exactly mats3[n]."""
    from soar_b200.lbs import skin_gaussians
    out = []
    for lo in range(0, mats3.shape[0], 48):
        blk = mats3[lo:lo + 48]
        n = blk.shape[0]
        bones = torch.zeros(n, 4, 4)
        bones[:, :3, :3] = blk
        bones[:, 3, 3] = 1
        quats = torch.tensor([1.0, 0, 0, 0]).repeat(n, 1)
        _, q, _ = skin_gaussians(torch.zeros(n, 3).cuda(), quats.cuda(), torch.eye(n).cuda(), bones.cuda())
        out.append(q.cpu())
    return torch.cat(out)


def test_matrix_to_quaternion_branch_choice_on_non_orthonormal_matrices():
    """a19's unpinned half (pytorch3d is not vendored by the reference and absent here): the branch
    choice, which is what matters once bone blends stop being rotations. For sheared /
    scaled matrices the four candidates are four DIFFERENT quaternions, so which one the kernel
    returned can be read off its output: it must be the candidate with the largest q_abs, the FIRST
    of them on an exact tie (torch.argmax). The 0.1 floor never binds on the selected candidate
    (the four q_abs^2 sum to 4, so the largest is >= 1); it is asserted rather than assumed."""
    gen = torch.Generator().manual_seed(11)
    # rotations whose dominant quaternion component is w, x, y, z in turn, then sheared and scaled
    q = torch.nn.functional.normalize(0.35 * torch.randn(4, 200, 4, generator=gen), dim=-1)
    for c in range(4):
        q[c, :, c] = q[c, :, c].abs() + 1.5
    q = torch.nn.functional.normalize(q.reshape(-1, 4), dim=-1)
    R = lbs_oracle.quaternion_to_matrix(q)
    A = torch.eye(3) + 0.25 * torch.randn(q.shape[0], 3, 3, generator=gen)
    mats = torch.matmul(A, R)
    # exact ties between the x and y candidates (m00 == m11) and between w and z (m00 == -m11), with
    # off-diagonal entries that make the tied candidates differ
    ties = torch.tensor([[[0.0, 1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.3, -1.0]],
                         [[0.5, 0.9, 0.1], [0.7, 0.5, 0.0], [0.2, 0.4, -1.0]],
                         [[0.25, -0.8, 0.2], [0.6, -0.25, 0.1], [0.0, 0.3, 0.5]]])
    mats = torch.cat([mats, ties])
    got = _skin_identity_quats(mats).double()
    cand, q_abs = _m2q_candidates(mats)
    want_sel = q_abs.argmax(1)                                # first maximal index on ties
    assert (q_abs.gather(1, want_sel[:, None]) >= 1.0 - 1e-12).all()
    for c in range(4):
        assert (want_sel[:800] == c).sum() >= 100, "the cases do not reach every branch"
    assert q_abs[800, 1] == q_abs[800, 2] and q_abs[801, 1] == q_abs[801, 2] and q_abs[802, 0] == q_abs[802, 3]
    err = (cand - got[:, None, :]).abs().amax(-1)             # [N,4]: distance to each candidate
    chosen = err.gather(1, want_sel[:, None])[:, 0]
    assert chosen.max().item() <= 2e-5, f"selected candidate off by {chosen.max().item():.2e}"
    # the identification is unambiguous wherever another candidate is far from the selected one
    other = err.clone()
    other.scatter_(1, want_sel[:, None], float("inf"))
    distinct = (cand - cand.gather(1, want_sel[:, None, None].expand(-1, 1, 4))).abs().amax(-1)
    distinct.scatter_(1, want_sel[:, None], float("inf"))
    clear = distinct.amin(1) > 1e-3
    assert clear[800:].all() and clear[:800].float().mean().item() > 0.9
    assert (other.amin(1)[clear] > 5e-4).all(), "the output matches a candidate other than the arg-max one"


def test_matrix_to_quaternion_against_pytorch3d_when_present():
    """Opportunistic pin: where a real pytorch3d is importable, the kernel's quaternions are compared
    with pytorch3d.transforms.matrix_to_quaternion itself on non-orthonormal bone blends. Skipped (and
    a19 stays 'parity unpinned') wherever pytorch3d is absent — it is in this image."""
    p3d = pytest.importorskip("pytorch3d.transforms")
    pb = synthetic.skinning_problem(50_000, seed=9)
    blended = torch.einsum("nj,jxy->nxy", pb["weights"], pb["bones"].reshape(-1, 4, 4))[:, :3, :3]
    sel = torch.arange(0, 50_000, 50)
    got = _skin_identity_quats(blended[sel])
    want = torch.nn.functional.normalize(p3d.matrix_to_quaternion(blended[sel]), dim=-1)
    want = torch.where(want[:, 0:1] < 0, -want, want)
    assert (got - want).abs().max().item() <= 2e-5
