"""Drop-in for the two reachable functions of the reference's JIT-built `cuda_utils` extension
(`load(name="cuda_utils", sources=["utils/ext.cpp", "utils/cuda_utils.cu"])`,
TS/geometry/surfel_base.py:168-175, TS/utils/image_utils.py:312-314) — SURVEY.md §8 row f4.

    gaussian2occgrid(pos_min, pos_max, grid_len, grid_dim, pos, rot, scale, opac, cutoff)
        TS/utils/ext.cpp:201-255  ->  soar_occgrid_splat   (soar_b200/csrc/occgrid.cu)
    contour_padding(image, mask, size)
        TS/utils/ext.cpp:74-107   ->  soar_contour_padding

Same names, argument order and return values as the pybind functions, so
`self.utils_mod = soar_b200.cuda_utils` is the whole integration. `to_occ_grid` restates the caller
(TS/geometry/surfel_base.py:1150-1195) on plain tensors. CUDA only: there is no CPU path.
"""
import ctypes

import torch

from . import _capi

_lib = _capi.load()


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dev_f32(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")  # ext.cpp CHECK_CUDA
    return t.contiguous().to(torch.float32)


def gaussian2occgrid(pos_min, pos_max, grid_len, grid_dim, pos, rot, scale, opac, cutoff):
    """Occupancy grid [gx, gy, gz] (float32, CUDA) of surfels — TS/utils/ext.cpp:201-255.

    pos_min / pos_max [3], grid_len (0-dim or [1]), cutoff [1]: CUDA float tensors; grid_dim [3] int
    tensor (read on the host, like the reference's `grid_dim[i].item<int>()`); pos [N,3], rot [N,4]
    (w,x,y,z — not normalised by the kernel), scale [N,3] (columns 0 and 1 are used), opac [N,1]."""
    pos_min, pos_max = _dev_f32(pos_min, "pos_min"), _dev_f32(pos_max, "pos_max")
    grid_len, cutoff = _dev_f32(grid_len, "grid_len"), _dev_f32(cutoff, "cutoff")
    pos, rot = _dev_f32(pos, "pos"), _dev_f32(rot, "rot")
    scale, opac = _dev_f32(scale, "scale"), _dev_f32(opac, "opac")
    gd = [int(v) for v in grid_dim.detach().cpu().tolist()]
    n = int(pos.shape[0])
    if rot.shape[0] != n or scale.shape[0] != n or opac.numel() != n or scale.shape[-1] < 2:
        raise RuntimeError("gaussian2occgrid: pos [N,3], rot [N,4], scale [N,>=2], opac [N] expected")
    grid = torch.empty(gd, dtype=torch.float32, device=pos.device)
    flag = torch.empty(1, dtype=torch.int32, device=pos.device)
    gd_host = (ctypes.c_int32 * 3)(*gd)
    with torch.cuda.device(pos.device):
        _capi.check(_lib.soar_occgrid_splat(n, _capi.ptr(pos_min), _capi.ptr(pos_max), _capi.ptr(grid_len),
                                            _capi.ptr(cutoff), gd_host, _capi.ptr(pos), _capi.ptr(rot),
                                            _capi.ptr(scale), _capi.ptr(opac), grid.data_ptr() if grid.numel() else None,
                                            flag.data_ptr(), _stream()),
                    "soar_occgrid_splat")
    if grid.numel() and int(flag.item()) != 0:
        raise RuntimeError("gaussian2occgrid: a Gaussian's sample lattice exceeds 65536 in-bounds steps per "
                           "axis (grid_len does not match the bounds); the reference kernel would not "
                           "return in reasonable time on such input")
    return grid


def contour_padding(image, mask, size):
    """res [C,H,W]: zero inside the mask, mean of the masked window pixels outside it —
    TS/utils/ext.cpp:74-107. image [C,H,W] float32 CUDA, mask [H,W] (or [1,H,W]) bool CUDA."""
    if not image.is_cuda or not mask.is_cuda:
        raise RuntimeError("contour_padding: image and mask must be CUDA tensors")  # CHECK_INPUT
    if image.dtype != torch.float32 or mask.dtype != torch.bool:
        raise RuntimeError("contour_padding: image must be float32 and mask bool")
    if not image.is_contiguous() or not mask.is_contiguous():
        raise RuntimeError("contour_padding: image and mask must be contiguous")  # CHECK_CONTIGUOUS
    C, H, W = (int(v) for v in image.shape)
    if mask.numel() != H * W:
        raise RuntimeError("contour_padding: mask must hold H*W elements")
    res = torch.empty_like(image)
    with torch.cuda.device(image.device):
        _capi.check(_lib.soar_contour_padding(_capi.ptr(image), _capi.ptr(mask), C, H, W, int(size),
                                              _capi.ptr(res), _stream()),
                    "soar_contour_padding")
    return res


def to_occ_grid(xyz, rotation, scales, cutoff, grid_dim_max=512, bound_overwrite=None):
    """TS/geometry/surfel_base.py:1150-1195 on plain tensors: xyz [N,3], rotation [N,4] (normalised,
    `get_rotation`), scales [N,1] (the attribute field's activated scale; repeated to 3 columns like
    the reference), opacity 1. Returns (grid, -xyz_min, 1/grid_len, grid_dim)."""
    if bound_overwrite is None:
        xyz_min = xyz.min(0)[0]
        xyz_max = xyz.max(0)[0]
        xyz_len = xyz_max - xyz_min
        xyz_min = xyz_min - xyz_len * 0.1
        xyz_max = xyz_max + xyz_len * 0.1
    else:
        xyz_min, xyz_max = bound_overwrite
    xyz_len = xyz_max - xyz_min
    grid_len = xyz_len.max() / grid_dim_max
    grid_dim = (xyz_len / grid_len + 0.5).to(torch.int32)
    grid = gaussian2occgrid(xyz_min, xyz_max, grid_len, grid_dim, xyz, rotation,
                            scales.reshape(-1, 1).repeat(1, 3), torch.ones_like(xyz[:, :1]),
                            torch.tensor([cutoff]).to(torch.float32).to(xyz.device))
    return grid, -xyz_min, 1 / grid_len, grid_dim
