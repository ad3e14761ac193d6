"""Seeded inputs of the f4 tests (gaussian2occgrid / contour_padding), shared by the golden maker,
the CPU oracle tests and the GPU parity tests. numpy only."""
import numpy as np


def _surfels(rng, n, s_lo, s_hi):
    """Surfels on a stretched ellipsoid shell (avatar-sized, metres) with random tangent frames."""
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    pos = d * np.array([0.35, 0.85, 0.22]) + rng.normal(scale=0.004, size=(n, 3))
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    s = rng.uniform(s_lo, s_hi, size=(n, 1))
    return pos.astype(np.float32), q.astype(np.float32), np.repeat(s, 3, axis=1).astype(np.float32)


def _bounds_like_reference(pos, grid_dim_max, bound_overwrite=None):
    """TS/geometry/surfel_base.py:1151-1166 in float32."""
    if bound_overwrite is None:
        lo, hi = pos.min(0), pos.max(0)
        ln = hi - lo
        lo = (lo - ln * np.float32(0.1)).astype(np.float32)
        hi = (hi + ln * np.float32(0.1)).astype(np.float32)
    else:
        lo, hi = (np.asarray(b, dtype=np.float32) for b in bound_overwrite)
    ln = (hi - lo).astype(np.float32)
    grid_len = np.float32(ln.max() / np.float32(grid_dim_max))
    grid_dim = (ln / grid_len + np.float32(0.5)).astype(np.int32)
    return lo, hi, grid_len, grid_dim


def occgrid_case(name):
    """dict(pos_min, pos_max, grid_len, grid_dim, pos, rot, scale, opac, cutoff) as numpy."""
    if name == "avatar":  # the product's call: padded bounds, opacity 1 (surfel_base.py:1175-1185)
        rng = np.random.default_rng(11)
        pos, rot, scale = _surfels(rng, 3000, 0.004, 0.02)
        lo, hi, gl, gd = _bounds_like_reference(pos, 160)
        opac, cutoff = np.ones((len(pos), 1), np.float32), 0.5
    elif name == "cut":  # bound_overwrite through the cloud: both `return`s of the kernel fire
        rng = np.random.default_rng(12)
        pos, rot, scale = _surfels(rng, 1500, 0.006, 0.03)
        lo, hi, gl, gd = _bounds_like_reference(pos, 64, ((-0.30, -0.50, -0.20), (0.33, 0.62, 0.243)))
        opac, cutoff = rng.uniform(0.2, 1.0, size=(len(pos), 1)).astype(np.float32), 0.0
        rot = (rot * rng.uniform(0.8, 1.2, size=(len(pos), 1))).astype(np.float32)  # un-normalised quats
    elif name == "large":  # GPU-only size (ours vs the reference kernel live)
        rng = np.random.default_rng(13)
        pos, rot, scale = _surfels(rng, 100000, 0.003, 0.012)
        lo, hi, gl, gd = _bounds_like_reference(pos, 256)
        opac, cutoff = np.ones((len(pos), 1), np.float32), 0.3
    elif name == "long_lattice":  # lattices longer than the 256 staged coordinates per axis
        rng = np.random.default_rng(14)
        pos, rot, scale = _surfels(rng, 24, 0.15, 0.25)
        lo, hi, gl, gd = _bounds_like_reference(pos, 400, ((-2.0, -2.5, -2.0), (2.0, 2.5, 2.0)))
        opac, cutoff = np.ones((len(pos), 1), np.float32), 0.2
    else:
        raise KeyError(name)
    return dict(pos_min=lo, pos_max=hi, grid_len=np.array([gl], np.float32), grid_dim=gd, pos=pos, rot=rot,
                scale=scale, opac=opac, cutoff=np.array([cutoff], np.float32))


def contour_case(name):
    """dict(image [C,H,W] f32, mask [H,W] bool, size)."""
    if name == "mask":  # the product's call: image = mask as float (image_utils.py:316-317)
        rng = np.random.default_rng(21)
        H, W = 67, 93
        yy, xx = np.mgrid[0:H, 0:W]
        mask = ((yy - 30) ** 2 / 400.0 + (xx - 50) ** 2 / 900.0 < 1.0) | (rng.random((H, W)) < 0.02)
        mask[:3, :5] = True  # touches the border
        return dict(image=mask[None].astype(np.float32), mask=mask, size=3)
    if name == "rgb":  # 3 channels: the kernel's own indexing (image[C*c + q], last channel stays)
        rng = np.random.default_rng(22)
        H, W = 40, 56
        mask = rng.random((H, W)) < 0.3
        return dict(image=rng.random((3, H, W)).astype(np.float32), mask=mask, size=2)
    if name == "size0":
        rng = np.random.default_rng(23)
        mask = rng.random((16, 16)) < 0.5
        return dict(image=rng.random((1, 16, 16)).astype(np.float32), mask=mask, size=0)
    if name == "full_hd":  # GPU-only size
        rng = np.random.default_rng(24)
        H, W = 1080, 1920
        yy, xx = np.mgrid[0:H, 0:W]
        mask = ((yy - 540) ** 2 / 400.0 ** 2 + (xx - 960) ** 2 / 300.0 ** 2 < 1.0)
        return dict(image=mask[None].astype(np.float32), mask=mask, size=10)
    raise KeyError(name)


OCC_GOLDEN = ("avatar", "cut")
CONTOUR_GOLDEN = ("mask", "rgb", "size0")
