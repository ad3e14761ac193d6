"""The packed two-pixels-per-thread blend kernels (blend_fwd2 / blend_bwd2) against their scalar twins
(forced with SOAR_BLEND_FWD_SCALAR / SOAR_BLEND_BWD_SCALAR = 0 / 1; unset, a launch picks by its size): the forward must be bit-identical
(images, n_contrib); the backward sums the pixel pair first, so gradients agree to rounding."""
import os

import pytest
import torch

from soar_b200 import synthetic
import test_rasterizer_gpu as T

pytestmark = pytest.mark.gpu

CASES = [c for c in T.CASES if c[0] in ("c2_sh0_512", "precomp_ragged", "surfel_soar", "surfel_front")]


def _run(s, cot, scalar):
    for k in ("SOAR_BLEND_FWD_SCALAR", "SOAR_BLEND_BWD_SCALAR"):
        os.environ[k] = "1" if scalar else "0"   # unset = the launch picks by its size
    try:
        out, g = T.run_ours(s, cot)
        dbg = T.debug_tensors(s)
        torch.cuda.synchronize()
    finally:
        for k in ("SOAR_BLEND_FWD_SCALAR", "SOAR_BLEND_BWD_SCALAR"):
            os.environ.pop(k, None)
    return ({k: v.clone() for k, v in out.items() if torch.is_tensor(v)}, dbg["n_contrib"].clone(),
            {k: v.clone() for k, v in g.items() if torch.is_tensor(v)})


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_packed_kernels_match_scalar_twins(case):
    name, P, W, H, deg, surfel, bg, desc, front, _ = case
    s = T.to_device(synthetic.scene(P, W, H, sh_degree=deg, seed=3, surfel=surfel, bg=bg,
                                    sort_descending=desc, render_front=front))
    cot = T.to_device(synthetic.cotangents(W, H, seed=3))
    o_s, n_s, g_s = _run(s, cot, scalar=True)
    o_p, n_p, g_p = _run(s, cot, scalar=False)
    assert torch.equal(n_s, n_p), "n_contrib differs between the packed and the scalar forward"
    for k in ("color", "normal", "depth", "opac"):
        assert torch.equal(o_s[k], o_p[k]), f"{k} is not bit-identical"
    for k in g_s:
        ref = g_s[k]
        tol = 2e-4 * float(ref.abs().max()) + 1e-12  # reduction-order noise, far inside the 1e-3 contract
        assert float((ref - g_p[k]).abs().max()) <= tol, k
