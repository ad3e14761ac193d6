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
