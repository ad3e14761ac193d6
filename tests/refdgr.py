"""TEST INFRASTRUCTURE ONLY. ctypes driver for the UNMODIFIED reference CUDA code built by
oracle/build_ref.sh into oracle/_ref/libref_dgr.so / libref_knn.so (C shim:
oracle/ref_cuda/ref_capi.cu). Used by the `-m gpu` parity tests, by
tests/golden/make_golden.py and by `bench.py --impl reference`; never by the product."""
import ctypes
import os
from ctypes import c_float, c_int, c_void_p

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DGR = os.path.join(ROOT, "oracle", "_ref", "libref_dgr.so")
REF_KNN = os.path.join(ROOT, "oracle", "_ref", "libref_knn.so")


def available():
    return os.path.exists(REF_DGR) and os.path.exists(REF_KNN)


_dgr = None
_knn = None


def _load_dgr():
    global _dgr
    if _dgr is None:
        lib = ctypes.CDLL(REF_DGR)
        lib.refdgr_create.restype = c_void_p
        lib.refdgr_destroy.argtypes = [c_void_p]
        lib.refdgr_forward.restype = c_int
        lib.refdgr_forward.argtypes = (
            [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int]
            + [c_void_p] * 5 + [c_float] + [c_void_p] * 7 + [c_float, c_float, c_int, c_int, c_int]
            + [c_void_p] * 6 + [c_int])
        lib.refdgr_backward.restype = c_int
        lib.refdgr_backward.argtypes = (
            [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int]
            + [c_void_p] * 4 + [c_float] + [c_void_p] * 7 + [c_float, c_float]
            + [c_void_p] * 19 + [c_int, c_void_p])
        for name in ("refdgr_binning_ptr", "refdgr_image_ptr", "refdgr_geom_ptr"):
            fn = getattr(lib, name)
            fn.restype = c_void_p
            fn.argtypes = [c_void_p, c_int]
        lib.refdgr_num_rendered.restype = c_int
        lib.refdgr_num_rendered.argtypes = [c_void_p]
        _dgr = lib
    return _dgr


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _from_ptr(ptr, n, dtype, device):
    """Copy n elements of dtype from a raw device pointer into a fresh tensor."""
    out = torch.empty(n, dtype=dtype, device=device)
    if n:
        rt = ctypes.CDLL("libcudart.so.12") if not hasattr(_from_ptr, "rt") else _from_ptr.rt
        _from_ptr.rt = rt
        rt.cudaMemcpy.argtypes = [c_void_p, c_void_p, ctypes.c_size_t, c_int]
        rc = rt.cudaMemcpy(out.data_ptr(), ptr, n * out.element_size(), 3)  # D2D
        assert rc == 0, f"cudaMemcpy failed: {rc}"
    return out


class RefRasterizer:
    """One reference view context (plays the role of the autograd ctx)."""

    def __init__(self):
        self.lib = _load_dgr()
        self.h = self.lib.refdgr_create()

    def close(self):
        if self.h:
            self.lib.refdgr_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def forward(self, s, sync=True):
        """s: scene dict of CUDA tensors (soar_b200.synthetic.scene moved to the device)."""
        dev = s["means3D"].device
        P, H, W = s["means3D"].shape[0], s["image_height"], s["image_width"]
        M = s["shs"].shape[1] if s.get("shs") is not None else 0
        z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)
        out = dict(color=z(3, H, W), normal=z(3, H, W), depth=z(1, H, W), opac=z(1, H, W),
                   radii=torch.zeros(P, dtype=torch.int32, device=dev))
        if sync:
            torch.cuda.synchronize()
        R = self.lib.refdgr_forward(
            self.h, P, s["sh_degree"], M, _p(s["bg"]), W, H, _p(s["means3D"]), _p(s.get("shs")),
            _p(s.get("colors_precomp")), _p(s["opacities"]), _p(s.get("scales")), s["scale_modifier"],
            _p(s.get("rotations")), _p(s.get("cov3D_precomp")), _p(s["viewmatrix"]),
            _p(s["projmatrix"]), _p(s["prcppoint"]), _p(s["patch_bbox"]), _p(s["campos"]),
            s["tanfovx"], s["tanfovy"], int(s["prefiltered"]), int(s["render_front"]),
            int(s["sort_descending"]), _p(s["config"]), _p(out["color"]), _p(out["normal"]),
            _p(out["depth"]), _p(out["opac"]), _p(out["radii"]), 0)
        if sync:
            torch.cuda.synchronize()
        assert R >= 0, "reference forward failed"
        out["num_rendered"] = R
        self.s, self.out, self.R = s, out, R
        return out

    def backward(self, cot, sync=True):
        s, dev = self.s, self.s["means3D"].device
        P, H, W = s["means3D"].shape[0], s["image_height"], s["image_width"]
        M = s["shs"].shape[1] if s.get("shs") is not None else 0
        z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), normal=z(P, 3),
                 depth=z(P, 1), means3D=z(P, 3), cov3D=z(P, 6), sh=z(P, M, 3), scales=z(P, 3),
                 rotations=z(P, 4), viewmat=z(4, 4), projmat=z(4, 4), campos=z(3))
        if sync:
            torch.cuda.synchronize()
        rc = self.lib.refdgr_backward(
            self.h, P, s["sh_degree"], M, self.R, _p(s["bg"]), W, H, _p(s["means3D"]),
            _p(s.get("shs")), _p(s.get("colors_precomp")), _p(s.get("scales")), s["scale_modifier"],
            _p(s.get("rotations")), _p(s.get("cov3D_precomp")), _p(s["viewmatrix"]),
            _p(s["projmatrix"]), _p(s["campos"]), _p(s["prcppoint"]), _p(s["patch_bbox"]),
            s["tanfovx"], s["tanfovy"], _p(self.out["radii"]), _p(cot["color"]), _p(cot["normal"]),
            _p(cot["depth"]), _p(cot["opac"]), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]),
            _p(g["colors"]), _p(g["normal"]), _p(g["depth"]), _p(g["means3D"]), _p(g["cov3D"]),
            _p(g["sh"]) if M else None, _p(g["scales"]), _p(g["rotations"]), _p(g["viewmat"]),
            _p(g["projmat"]), _p(g["campos"]), 0, _p(s["config"]))
        if sync:
            torch.cuda.synchronize()
        assert rc == 0, "reference backward failed"
        return g

    # ---- decoded private buffers -------------------------------------------------------
    def sorted_keys(self):
        p = self.lib.refdgr_binning_ptr(self.h, 0)
        return _from_ptr(p, self.R, torch.int64, self.s["means3D"].device)

    def point_list(self):
        p = self.lib.refdgr_binning_ptr(self.h, 1)
        return _from_ptr(p, self.R, torch.int32, self.s["means3D"].device)

    def ranges(self):
        H, W = self.s["image_height"], self.s["image_width"]
        T = ((W + 15) // 16) * ((H + 15) // 16)
        p = self.lib.refdgr_image_ptr(self.h, 0)
        return _from_ptr(p, 2 * T, torch.int32, self.s["means3D"].device).view(T, 2)

    def n_contrib(self):
        H, W = self.s["image_height"], self.s["image_width"]
        p = self.lib.refdgr_image_ptr(self.h, 1)
        return _from_ptr(p, H * W, torch.int32, self.s["means3D"].device)


def ref_dist2(points):
    """Reference distCUDA2 (KNN/simple_knn.cu:186-222) on a CUDA tensor [P,3]."""
    global _knn
    if _knn is None:
        _knn = ctypes.CDLL(REF_KNN)
        _knn.refknn_dist2.restype = c_int
        _knn.refknn_dist2.argtypes = [c_int, c_void_p, c_void_p]
    pts = points.float().contiguous()
    out = torch.zeros(pts.shape[0], dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    rc = _knn.refknn_dist2(pts.shape[0], pts.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    assert rc == 0, f"reference knn failed: {rc}"
    return out
