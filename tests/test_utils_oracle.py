"""CPU suite: the C restatement of gaussian2occgrid / contour_padding (oracle/utils_oracle.c) against
the golden outputs of the unmodified reference kernels (tests/golden/utils_*.npz), plus the
properties the kernels' definitions imply. SURVEY §8 row f4."""
import os

import numpy as np
import pytest

import golden_util as G
import utils_cases as U
from oracle import utils_oracle as O

pytestmark = pytest.mark.skipif(not O.available(), reason="make -C oracle")


def _golden_grid(d):
    g = np.zeros(int(np.prod(d["grid_shape"])), np.float32)
    g[d["grid_index"]] = d["grid_value"]
    return g.reshape(tuple(d["grid_shape"]))


def occ_inputs(d):
    return {k[3:]: d[k] for k in d.files if k.startswith("in_")}


@pytest.mark.parametrize("name", U.OCC_GOLDEN)
def test_occgrid_oracle_matches_reference_golden(name):
    d = G.load(f"utils_occgrid_{name}")
    c = occ_inputs(d)
    for k, v in U.occgrid_case(name).items():  # the fixture was made from these very inputs
        assert np.array_equal(c[k], v), k
    g, added = O.gaussian2occgrid(**c)
    ref = _golden_grid(d)
    assert g.shape == ref.shape and added > 0
    # same cells, and sums within float-atomic ordering noise: no accept decision differs
    assert np.array_equal(g != 0, ref != 0)
    tol = 1e-5 * np.abs(ref) + 1e-6 + float(d["spread"])
    assert np.all(np.abs(g - ref) <= tol), float(np.abs(g - ref).max())


@pytest.mark.parametrize("name", U.CONTOUR_GOLDEN)
def test_contour_oracle_matches_reference_golden_bit_exact(name):
    d = G.load(f"utils_contour_{name}")
    res = O.contour_padding(d["in_image"], d["in_mask"], int(d["in_size"]))
    assert np.array_equal(res, d["res"])


def test_contour_padding_properties():
    c = U.contour_case("mask")
    res = O.contour_padding(**c)
    m = c["mask"]
    assert np.all(res[0][m] == 0)  # masked pixels stay zero
    # image == mask: every value is 1 exactly where the window holds a masked pixel (a dilation ring)
    s = c["size"]
    H, W = m.shape
    pad = np.pad(m, s)
    dil = np.zeros_like(m)
    for dy in range(2 * s + 1):
        for dx in range(2 * s + 1):
            dil |= pad[dy:dy + H, dx:dx + W]
    assert np.array_equal(res[0] == 1.0, dil & ~m)
    assert np.all((res[0] == 0) | (res[0] == 1.0))


def test_occgrid_lexicographic_return_and_linearity():
    c = U.occgrid_case("cut")
    g_all, _ = O.gaussian2occgrid(**c)
    # the grid is a sum over Gaussians: two halves add up to the whole
    h = len(c["pos"]) // 2
    parts = []
    for sl in (slice(0, h), slice(h, None)):
        cc = dict(c, pos=c["pos"][sl], rot=c["rot"][sl], scale=c["scale"][sl], opac=c["opac"][sl])
        parts.append(O.gaussian2occgrid(**cc)[0].astype(np.float64))
    assert np.allclose(parts[0] + parts[1], g_all, rtol=1e-6, atol=1e-6)
    # a Gaussian whose lattice starts below pos_min contributes nothing (first sample returns, :809)
    far = dict(c, pos=(c["pos_min"][None] + np.float32(0.001)).astype(np.float32), rot=c["rot"][:1],
               scale=c["scale"][:1], opac=c["opac"][:1])
    g_far, added = O.gaussian2occgrid(**far)
    assert added == 0 and not g_far.any()
    # opacity below 1/255: the search radius is NaN, the loops never start
    dim = dict(c, opac=np.full_like(c["opac"], 0.003))
    assert O.gaussian2occgrid(**dim)[1] == 0
