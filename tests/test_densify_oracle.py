"""CPU: oracle/densify_oracle.py replays the golden sequence recorded from the reference's own
methods + torch.optim.Adam (tests/golden/make_golden_densify.py) and must land on the same state
after every stage."""
import numpy as np
import pytest
import torch

import densify_replay as R
from oracle import densify_oracle as O


def test_cases_present():
    assert len(R.CASES) >= 2


@pytest.mark.parametrize("name", R.CASES)
def test_oracle_replays_reference_sequence(name):
    g, pb = R.load(name)
    st = O.new_state(pb["params"])
    get = lambda kind, k: {"p": st["p"], "m": st["m"], "v": st["v"], "s": st["stats"]}[kind][k]
    for gr in pb["adam_grads"]:
        O.adam_step(st, gr, pb["lrs"])
    R.compare(get, g, "adam")
    for vg, radii, sg in pb["views"]:
        O.add_densification_stats(st, vg, radii, sg)
    R.compare(get, g, "stats")
    O.adaptive_prune(st, 0.1, float(g["extent"]))
    assert st["p"]["xyz"].shape[0] == int(g["n_after_prune"])
    R.compare(get, g, "prune")
    O.adaptive_densify(st, float(g["max_grad"]), float(g["extent"]), torch.from_numpy(g["samples"]))
    R.compare(get, g, "densify")
    grads, (vg, radii, sg) = R.post_inputs(g)
    O.adam_step(st, grads, pb["lrs"])
    O.add_densification_stats(st, vg, radii, sg)
    R.compare(get, g, "post")
