"""GPU parity of SURVEY §8 row f4 through the C ABI (soar_occgrid_splat / soar_contour_padding behind
soar_b200.cuda_utils): against the golden outputs of the reference kernels, against the reference
kernels run live (oracle/_ref/libref_utils.so) at full sizes, and against the CPU oracle."""
import json
import time

import numpy as np
import pytest
import torch

import golden_util as G
import utils_cases as U

pytestmark = pytest.mark.gpu

OCC_KEYS = ("pos_min", "pos_max", "grid_len", "grid_dim", "pos", "rot", "scale", "opac", "cutoff")


def _cuda(c):
    return [torch.from_numpy(np.ascontiguousarray(c[k])).cuda() for k in OCC_KEYS]


def _close(a, ref, spread=0.0):
    """Cell sums differ only by the order of float atomics: 1e-5 relative (+ the reference's own spread)."""
    assert a.shape == ref.shape
    assert np.array_equal(a != 0, ref != 0), f"{int(((a != 0) != (ref != 0)).sum())} cells differ in occupancy"
    tol = 1e-5 * np.abs(ref) + 1e-6 + spread
    assert np.all(np.abs(a - ref) <= tol), float(np.abs(a - ref).max())


@pytest.mark.parametrize("name", U.OCC_GOLDEN)
def test_occgrid_matches_reference_golden(name):
    from soar_b200 import cuda_utils
    d = G.load(f"utils_occgrid_{name}")
    c = {k[3:]: d[k] for k in d.files if k.startswith("in_")}
    g = cuda_utils.gaussian2occgrid(*_cuda(c)).cpu().numpy()
    ref = np.zeros(int(np.prod(d["grid_shape"])), np.float32)
    ref[d["grid_index"]] = d["grid_value"]
    _close(g, ref.reshape(tuple(d["grid_shape"])), float(d["spread"]))


@pytest.mark.parametrize("name", ("avatar", "cut", "long_lattice", "large"))
def test_occgrid_matches_reference_kernel_live(name):
    import refutils
    from soar_b200 import cuda_utils
    if not refutils.available():
        pytest.skip("oracle/_ref/libref_utils.so not built")
    args = _cuda(U.occgrid_case(name))
    ref = refutils.gaussian2occgrid(*args)
    ref2 = refutils.gaussian2occgrid(*args)
    ours = cuda_utils.gaussian2occgrid(*args)
    _close(ours.cpu().numpy(), ref.cpu().numpy(), float((ref - ref2).abs().max()))
    if name == "large":  # stage timing, reported in the test log (pytest -rA)
        def timed(fn, reps):
            torch.cuda.synchronize()
            t = time.perf_counter()
            for _ in range(reps):
                fn()
            torch.cuda.synchronize()
            return (time.perf_counter() - t) / reps * 1e3
        ms_ours = timed(lambda: cuda_utils.gaussian2occgrid(*args), 10)
        ms_ref = timed(lambda: refutils.gaussian2occgrid(*args), 3)
        print("F4_STAGE " + json.dumps({"workload": "gaussian2occgrid, 100000 surfels, grid_dim_max 256",
                                         "grid": list(ours.shape), "soar_ms": ms_ours,
                                         "reference_kernel_same_gpu_ms": ms_ref, "speedup": ms_ref / ms_ours}))


@pytest.mark.parametrize("name", ("avatar", "cut"))
def test_occgrid_matches_cpu_oracle(name):
    from oracle import utils_oracle as O
    from soar_b200 import cuda_utils
    c = U.occgrid_case(name)
    _close(cuda_utils.gaussian2occgrid(*_cuda(c)).cpu().numpy(), O.gaussian2occgrid(**c)[0])


def test_occgrid_edge_cases():
    from soar_b200 import cuda_utils
    c = U.occgrid_case("avatar")
    args = _cuda(c)
    empty = list(args)
    for i in (4, 5, 6, 7):
        empty[i] = args[i][:0]
    g = cuda_utils.gaussian2occgrid(*empty)
    assert tuple(g.shape) == tuple(int(v) for v in c["grid_dim"]) and not g.any()
    dim = list(args)
    dim[7] = torch.full_like(args[7], 0.003)  # opacity < 1/255: NaN search radius, nothing splatted
    assert not cuda_utils.gaussian2occgrid(*dim).any()
    with pytest.raises(RuntimeError):
        cuda_utils.gaussian2occgrid(*[a.cpu() for a in args])
    huge = list(args)
    huge[6] = args[6] * 1e4  # lattice starts far below pos_min: the first sample returns (:809)
    assert not cuda_utils.gaussian2occgrid(*huge).any()
    tiny = list(args)
    tiny[2] = args[2] * 1e-7  # grid_len inconsistent with the bounds: > 65536 in-bounds steps, refused
    with pytest.raises(RuntimeError, match="65536"):
        cuda_utils.gaussian2occgrid(*tiny)


def test_to_occ_grid_mirrors_the_caller():
    """to_occ_grid = the bounds arithmetic of surfel_base.py:1151-1166 (torch ops on the device: its
    scalar division is a multiplication by the reciprocal, so grid_len is taken from torch, not numpy)
    + the splat."""
    from oracle import utils_oracle as O
    from soar_b200 import cuda_utils
    c = U.occgrid_case("avatar")
    xyz, rot = torch.from_numpy(c["pos"]).cuda(), torch.from_numpy(c["rot"]).cuda()
    grid, shift, inv_len, gd = cuda_utils.to_occ_grid(xyz, rot, torch.from_numpy(c["scale"][:, :1]).cuda(), 0.5,
                                                      grid_dim_max=160)
    lo, hi = xyz.min(0)[0], xyz.max(0)[0]
    ln = hi - lo
    lo, hi = lo - ln * 0.1, hi + ln * 0.1
    grid_len = (hi - lo).max() / 160
    assert torch.equal(-shift, lo) and torch.equal(inv_len, 1 / grid_len)
    assert torch.equal(gd, ((hi - lo) / grid_len + 0.5).to(torch.int32))
    cc = dict(c, pos_min=lo.cpu().numpy(), pos_max=hi.cpu().numpy(), grid_len=grid_len.reshape(1).cpu().numpy(),
              grid_dim=gd.cpu().numpy())
    _close(grid.cpu().numpy(), O.gaussian2occgrid(**cc)[0])


@pytest.mark.parametrize("name", U.CONTOUR_GOLDEN)
def test_contour_padding_matches_reference_golden_bit_exact(name):
    from soar_b200 import cuda_utils
    d = G.load(f"utils_contour_{name}")
    res = cuda_utils.contour_padding(torch.from_numpy(d["in_image"]).cuda(), torch.from_numpy(d["in_mask"]).cuda(),
                                     int(d["in_size"]))
    assert np.array_equal(res.cpu().numpy(), d["res"])


@pytest.mark.parametrize("name", ("mask", "rgb", "full_hd"))
def test_contour_padding_matches_reference_kernel_live_bit_exact(name):
    import refutils
    from soar_b200 import cuda_utils
    if not refutils.available():
        pytest.skip("oracle/_ref/libref_utils.so not built")
    c = U.contour_case(name)
    img, m = torch.from_numpy(c["image"]).cuda(), torch.from_numpy(c["mask"]).cuda()
    assert torch.equal(cuda_utils.contour_padding(img, m, c["size"]), refutils.contour_padding(img, m, c["size"]))


def test_contour_padding_rejects_bad_input():
    from soar_b200 import cuda_utils
    c = U.contour_case("mask")
    img, m = torch.from_numpy(c["image"]), torch.from_numpy(c["mask"])
    with pytest.raises(RuntimeError):
        cuda_utils.contour_padding(img, m, 3)  # CPU tensors
    with pytest.raises(RuntimeError):
        cuda_utils.contour_padding(img.cuda().double(), m.cuda(), 3)
