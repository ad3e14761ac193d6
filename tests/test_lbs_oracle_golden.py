"""CPU: oracle/lbs_oracle.py against golden vectors made by the reference's OWN lines
(tests/golden/make_golden_weights.py: `SMPL_Guidance.query_weights_smpl`, TS/utils/smpl.py:618-637,
exec'd unchanged with an exact knn_points stand-in; the vendored `quaternion_to_matrix`,
TS/data/uncond_multiview.py:2422-2451, exec'd unchanged). Pins SURVEY §8 rows a18 / f1 and the first
half of a19; `matrix_to_quaternion` has no copy in the reference tree and stays unpinned."""
import hashlib

import numpy as np
import pytest
import torch

import golden_util as G
from oracle import lbs_oracle
from soar_b200 import synthetic

WEIGHT_CASES = ("weights_full", "weights_dups", "weights_tiny")


def template_of(d):
    """The synthetic template the golden file was made with (regenerated from its seed, checked by hash)."""
    V = int(d["V"])
    m = synthetic.smplx_template(V=V, NB=1, seed=int(d["template_seed"]))
    verts, lw = m["v_template"].clone(), m["lbs_weights"].clone()
    if int(d["duplicates"]):
        verts[V // 2: V // 2 + 5] = verts[0]
    digest = hashlib.sha256(verts.numpy().tobytes() + lw.numpy().tobytes()).hexdigest()
    assert digest == str(d["template_sha256"]), "synthetic template changed since the golden file was made"
    return verts, lw


@pytest.mark.parametrize("name", WEIGHT_CASES)
def test_query_weights_oracle_matches_reference_golden(name):
    d = G.load(name)
    verts, lw = template_of(d)
    got = lbs_oracle.query_weights(torch.from_numpy(d["points"]), verts, lw, K=30 if verts.shape[0] >= 30 else verts.shape[0])
    want = torch.from_numpy(d["weights"])
    # the oracle forms distances in float64 and rounds; the reference's float32 distances differ in the
    # last bit: 2e-6 on weights in [0, 1] (a swap between equidistant neighbours is exact)
    assert (got - want).abs().max().item() <= 2e-6
    assert (want.sum(1) - 1).abs().max().item() <= 1e-6


def test_quaternion_to_matrix_oracle_matches_vendored_reference_copy():
    d = G.load("quat2mat")
    got = lbs_oracle.quaternion_to_matrix(torch.from_numpy(d["quats"]))
    assert np.array_equal(got.numpy(), d["matrices"]), float(np.abs(got.numpy() - d["matrices"]).max())
