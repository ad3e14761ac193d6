"""Drop-in for `simple_knn._C` (KNN/ext.cpp:15-17 -> distCUDA2, KNN/spatial.cu:15-26):
mean squared distance from every point to its 3 nearest other points, computed
by the sm_100a kernels in soar_b200/csrc/knn.cu through the C-ABI
(soar_knn_dist2, include/soar_b200.h). No CPU fallback."""
import ctypes

import torch

from soar_b200 import _capi

_lib = _capi.load()


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """points [P,3] float32 CUDA -> [P] float32 (reference: KNN/spatial.cu:15-26)."""
    if not points.is_cuda:
        raise RuntimeError("distCUDA2: points must be a CUDA tensor")
    if points.dim() != 2 or points.size(1) != 3:
        raise RuntimeError("distCUDA2: points must have dimensions (num_points, 3)")
    pts = points.detach().to(torch.float32).contiguous()
    P = pts.size(0)
    out = torch.full((P,), 0.0, dtype=torch.float32, device=pts.device)
    if P == 0:
        return out
    with torch.cuda.device(pts.device):
        ws = torch.empty(_lib.soar_knn_workspace_bytes(P), dtype=torch.uint8, device=pts.device)
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        _capi.check(_lib.soar_knn_dist2(pts.data_ptr(), P, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                        stream), "distCUDA2")
    return out
