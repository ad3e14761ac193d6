"""TEST INFRASTRUCTURE ONLY — ctypes driver of oracle/_ref/libref_utils.so: the UNMODIFIED reference
cuda_utils kernels (TS/utils/cuda_utils.cu) behind oracle/ref_cuda/ref_utils_capi.cu. GPU box only."""
import ctypes
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libref_utils.so")
_lib = None


def available():
    return os.path.exists(LIB)


def _load():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB)
        _lib.refutils_gaussian2occgrid.restype = ctypes.c_int
        _lib.refutils_gaussian2occgrid.argtypes = [ctypes.c_int] + [ctypes.c_void_p] * 5 + [ctypes.c_size_t] + \
                                                  [ctypes.c_void_p] * 5
        _lib.refutils_contour_padding.restype = ctypes.c_int
        _lib.refutils_contour_padding.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                  ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    return _lib


def gaussian2occgrid(pos_min, pos_max, grid_len, grid_dim, pos, rot, scale, opac, cutoff):
    """torch CUDA tensors in the reference's argument order -> grid (torch, CUDA)."""
    lib = _load()
    f = lambda t: t.contiguous().float()
    pos_min, pos_max, grid_len, pos, rot, scale, opac, cutoff = map(f, (pos_min, pos_max, grid_len, pos, rot,
                                                                        scale, opac, cutoff))
    gd = grid_dim.contiguous().to(torch.int32).cuda()
    dims = [int(v) for v in gd.cpu().tolist()]
    grid = torch.empty(dims, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.refutils_gaussian2occgrid(int(pos.shape[0]), pos_min.data_ptr(), pos_max.data_ptr(),
                                       grid_len.data_ptr(), gd.data_ptr(), grid.data_ptr(), grid.numel(),
                                       pos.data_ptr(), rot.data_ptr(), scale.data_ptr(), opac.data_ptr(),
                                       cutoff.data_ptr())
    assert rc == 0, f"reference gaussian2occgrid failed: cuda error {rc}"
    return grid


def contour_padding(image, mask, size):
    lib = _load()
    image = image.contiguous().float()
    mask = mask.contiguous().to(torch.bool)
    C, H, W = image.shape
    res = torch.empty_like(image)
    torch.cuda.synchronize()
    rc = lib.refutils_contour_padding(image.data_ptr(), mask.data_ptr(), C, H, W, res.data_ptr(), int(size))
    assert rc == 0, f"reference contour_padding failed: cuda error {rc}"
    return res
