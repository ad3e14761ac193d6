"""CPU suite: the numpy restatement of the reference's SMPL-X forward (oracle/smplx_oracle.py)
against golden vectors recorded from the reference's OWN `lbs()` (tests/golden/smplx_*.npz, made
by tests/golden/make_golden_smplx.py from /root/reference). Float32 throughout; the only freedom
is summation order inside the matmuls: 2e-6 absolute on metres-scale outputs."""
import glob
import os

import numpy as np
import pytest

from oracle import lbs_oracle, smplx_oracle
from soar_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "smplx_*.npz")))
TOL = 2e-6


def model_of(d):
    m = synthetic.smplx_template(V=int(d["V"]), NB=int(d["NB"]), seed=int(d["template_seed"]))
    return {k: v.numpy() for k, v in m.items()}


def test_fixtures_present():
    assert {"smplx_small", "smplx_full", "smplx_full_orient"} <= set(CASES)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_lbs(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"))
    m = model_of(d)
    o = smplx_oracle.smplx_forward(m, d["betas"], d["full_pose"], d["transl"], d["inv_mats"])
    for k in ("verts", "joints", "A", "bones"):
        assert np.abs(o[k] - d[k]).max() <= TOL, k


def test_generator_is_stable():
    """The big fixtures store no inputs: the generator must keep producing the stored small case."""
    d = np.load(os.path.join(GOLDEN, "smplx_small.npz"))
    m = model_of(d)
    for k, v in m.items():
        assert np.array_equal(v, d["in_" + k]), k
    p = synthetic.smplx_pose(NB=int(d["NB"]), seed=int(d["pose_seed"]))
    assert np.array_equal(p["betas"].numpy(), d["betas"])
    assert np.array_equal(p["full_pose"].numpy(), d["full_pose"])


def test_rest_pose_is_identity():
    """Zero pose: every bone transform is the identity (plus transl) and verts = v_shaped + transl."""
    m = {k: v.numpy() for k, v in synthetic.smplx_template(V=200, NB=5, seed=3).items()}
    betas = np.zeros(5, np.float32)
    o = smplx_oracle.smplx_forward(m, betas, np.zeros(165, np.float32), [0.0, 0.3, 0.0])
    eye = np.tile(np.eye(4, dtype=np.float32), (55, 1, 1))
    eye[:, 1, 3] = 0.3
    assert np.abs(o["A"] - eye).max() <= 1e-6
    assert np.abs(o["verts"] - (m["v_template"] + np.array([0, 0.3, 0], np.float32))).max() <= 1e-6


def test_query_weights_rows_sum_to_one():
    import torch
    m = synthetic.smplx_template(V=500, NB=2, seed=4)
    pts = m["v_template"][:50] + 0.01
    w = lbs_oracle.query_weights(pts, m["v_template"], m["lbs_weights"], K=30)
    assert w.shape == (50, 55)
    assert (w.sum(1) - 1).abs().max().item() <= 1e-5
    assert torch.all(w >= 0)
