"""distCUDA2 parity on the GPU: bit-exact against the reference simple-knn
(oracle/_ref/libref_knn.so) and against the CPU restatement (oracle/knn_oracle.py)."""
import numpy as np
import pytest
import torch

import refdgr
from oracle import knn_oracle

pytestmark = pytest.mark.gpu


def _pts(P, seed, kind="uniform"):
    rng = np.random.Generator(np.random.PCG64(seed))
    if kind == "uniform":
        p = rng.uniform(-1, 1, size=(P, 3))
    elif kind == "surface":  # points on a thin shell, like Gaussians on a body surface
        v = rng.normal(size=(P, 3))
        p = v / np.linalg.norm(v, axis=1, keepdims=True) * rng.normal(1.0, 0.002, size=(P, 1))
    elif kind == "dups":
        p = rng.uniform(-1, 1, size=(P, 3))
        p[: P // 4] = p[P // 4: 2 * (P // 4)]  # exact duplicates -> distance 0 (Q15)
    elif kind == "line":  # degenerate extent on two axes
        p = np.zeros((P, 3))
        p[:, 0] = rng.uniform(-1, 1, size=P)
    return torch.from_numpy(p.astype(np.float32)).cuda()


@pytest.mark.parametrize("P,kind", [(1, "uniform"), (3, "uniform"), (4, "uniform"), (33, "uniform"),
                                     (1025, "uniform"), (5000, "surface"), (4096, "dups"),
                                     (3000, "line"), (20000, "uniform")])
def test_matches_cpu_oracle_bit_exact(P, kind):
    from simple_knn._C import distCUDA2
    pts = _pts(P, 11, kind)
    got = distCUDA2(pts).cpu().numpy()
    with np.errstate(over="ignore", invalid="ignore"):
        want = knn_oracle.dist2(pts.cpu().numpy())
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), \
        f"max abs diff {np.nanmax(np.abs(got - want))}"


@pytest.mark.parametrize("P,kind", [(100_000, "uniform"), (200_000, "surface"), (2_000_000, "surface")])
def test_matches_reference_cuda_bit_exact(P, kind):
    if not refdgr.available():
        pytest.fail("oracle/_ref/libref_knn.so missing")
    from simple_knn._C import distCUDA2
    pts = _pts(P, 5, kind)
    got = distCUDA2(pts)
    want = refdgr.ref_dist2(pts)
    assert torch.equal(got.view(torch.int32), want.view(torch.int32)), \
        f"max abs diff {(got - want).abs().max().item()}"


def test_empty_and_errors():
    from simple_knn._C import distCUDA2
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).numel() == 0
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(4, 3))  # CPU tensor: no CPU path
