"""CPU: oracle/postops_oracle.py against the golden vectors recorded from the reference's own
depth2normal / normal2curv / post block (tests/golden/make_golden_postops.py), forward and gradients."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import postops_oracle
from soar_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "postops_*.npz")))


def problem_of(g):
    pb = synthetic.postops_problem(int(g["W"]), int(g["H"]), seed=int(g["seed"]),
                                   prcppoint=tuple(g["prcppoint"]), touch_border=bool(g["touch_border"]))
    # the generator must still produce the recorded inputs
    assert np.array_equal(pb["normal"].numpy(), g["in_normal"])
    assert np.array_equal(pb["depth"].numpy(), g["in_depth"])
    assert np.array_equal(pb["opac"].numpy(), g["in_opac"])
    return pb


def close(a, b, atol, what):
    a = a.detach().cpu().numpy() if torch.is_tensor(a) else a
    err = np.abs(a.astype(np.float64) - b).max()
    assert err <= atol, f"{what}: max-abs {err:.3e} > {atol}"


def test_cases_present():
    assert len(CASES) >= 2


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_outputs(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    pb = problem_of(g)
    mask = pb["opac"] > 1e-5
    d = pb["depth"].clone().requires_grad_(True)
    dn = postops_oracle.depth2normal(d, mask, pb["camera"])
    (dn * pb["g_depth_normal"]).sum().backward()
    close(dn, g["d2n"], 1e-6, "depth2normal")
    close(d.grad, g["d2n_grad_depth"], 1e-3 * np.abs(g["d2n_grad_depth"]).max(), "depth2normal grad")
    n = pb["normal"].clone().requires_grad_(True)
    cv = postops_oracle.normal2curv(n, mask)
    (cv * pb["g_curv"]).sum().backward()
    close(cv, g["n2c"], 1e-6, "normal2curv")
    close(n.grad, g["n2c_grad_normal"], 1e-6 * max(1.0, np.abs(g["n2c_grad_normal"]).max()), "normal2curv grad")


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_post_block(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    pb = problem_of(g)
    n = pb["normal"].clone().requires_grad_(True)
    d = pb["depth"].clone().requires_grad_(True)
    on, oc, od, om = postops_oracle.surfel_postprocess(n, d, pb["opac"], pb["camera"])
    ((on * pb["g_normal"]).sum() + (oc * pb["g_curv"]).sum() + (od * pb["g_depth_normal"]).sum()).backward()
    assert np.array_equal(om.numpy(), g["post_mask"])
    close(on, g["post_normal"], 1e-6, "normal")
    close(oc, g["post_curv"], 1e-6, "curv")
    close(od, g["post_depth_normal"], 1e-6, "depth_normal")
    close(n.grad, g["post_grad_normal"], 1e-6 * max(1.0, np.abs(g["post_grad_normal"]).max()), "grad normal")
    close(d.grad, g["post_grad_depth"], 1e-3 * np.abs(g["post_grad_depth"]).max(), "grad depth")


def test_camera_constants_match_torch_inverse():
    # host logic of the product: Kinv without torch.inverse, principal point in float32
    from soar_b200 import postops
    cam = synthetic.ImageCamera(75, 53, 0.81, 0.66, (0.45, 0.58))
    cx, cy, kx, ky = postops.camera_constants(cam)
    Kinv = postops_oracle.camera_matrix_inverse(cam)
    assert np.float32(kx) == Kinv[0, 0].numpy() and np.float32(ky) == Kinv[1, 1].numpy()
    assert Kinv[0, 1] == 0 and Kinv[1, 0] == 0
    assert np.float32(cx) == (cam.prcppoint[0] * 75).numpy() and np.float32(cy) == (cam.prcppoint[1] * 53).numpy()
