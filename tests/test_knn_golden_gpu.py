"""distCUDA2 on the sm_100a kernels against the golden outputs of the reference simple-knn
(tests/golden/knn_*.npz): bit-exact."""
import numpy as np
import pytest
import torch

import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", G.KNN_CASES)
def test_distCUDA2_matches_golden_bit_exact(name):
    from simple_knn._C import distCUDA2
    d = G.load(name)
    out = distCUDA2(torch.from_numpy(d["points"]).cuda()).cpu().numpy()
    assert np.array_equal(out, d["dist2"])
