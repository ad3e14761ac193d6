"""ctypes binding of libsoar_b200.so — the C-ABI declared in include/soar_b200.h.

This is the only place that touches the shared library. There is no CPU
fallback: if the library is missing or was not built, importing the product
modules raises, and every entry point turns a non-zero return code into a
RuntimeError carrying soar_last_error() (the reference surfaces C++ exceptions
as RuntimeError the same way, DGR/rasterize_points.cu:50-52).
"""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int32, c_size_t, c_uint8, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsoar_b200.so")

c_float_p = c_void_p  # device pointers travel as integers


class SoarRastView(Structure):
    _fields_ = [
        ("P", c_int32),
        ("sh_degree", c_int32),
        ("M", c_int32),
        ("width", c_int32),
        ("height", c_int32),
        ("tanfovx", c_float),
        ("tanfovy", c_float),
        ("scale_modifier", c_float),
        ("prefiltered", c_int32),
        ("render_front", c_int32),
        ("sort_descending", c_int32),
        ("debug", c_int32),
        ("bg", c_void_p),
        ("means3D", c_void_p),
        ("shs", c_void_p),
        ("colors_precomp", c_void_p),
        ("opacities", c_void_p),
        ("scales", c_void_p),
        ("rotations", c_void_p),
        ("cov3D_precomp", c_void_p),
        ("viewmatrix", c_void_p),
        ("projmatrix", c_void_p),
        ("prcppoint", c_void_p),
        ("patch_bbox", c_void_p),
        ("campos", c_void_p),
        ("config", c_void_p),
        ("config_host_flags", c_uint32),
    ]


CFG_KNOWN, CFG_SURFACE, CFG_NORMALIZE_DEPTH, CFG_PERPIX_DEPTH, CFG_LRN_CAM = 1, 2, 4, 8, 16


class SoarRastGrads(Structure):
    _fields_ = [
        ("dL_dmeans2D", c_void_p),
        ("dL_dcolors", c_void_p),
        ("dL_dopacity", c_void_p),
        ("dL_dmeans3D", c_void_p),
        ("dL_dcov3D", c_void_p),
        ("dL_dsh", c_void_p),
        ("dL_dscales", c_void_p),
        ("dL_drotations", c_void_p),
        ("dL_dviewmat", c_void_p),
        ("dL_dprojmat", c_void_p),
        ("dL_dcampos", c_void_p),
        ("accumulate_mask", c_uint32),
    ]


ACC_MEANS2D, ACC_COLORS, ACC_OPACITY, ACC_MEANS3D = 1, 2, 4, 8
ACC_COV3D, ACC_SH, ACC_SCALES, ACC_ROTATIONS = 16, 32, 64, 128


class SoarSmplxModel(Structure):
    _fields_ = [
        ("V", c_int32),
        ("J", c_int32),
        ("NB", c_int32),
        ("v_template", c_void_p),
        ("shapedirs", c_void_p),
        ("posedirs", c_void_p),
        ("J_regressor", c_void_p),
        ("parents", c_void_p),
        ("lbs_weights", c_void_p),
    ]


class SoarPostArgs(Structure):
    _fields_ = [
        ("width", c_int32),
        ("height", c_int32),
        ("cx", c_float),
        ("cy", c_float),
        ("kinv_x", c_float),
        ("kinv_y", c_float),
        ("mask_threshold", c_float),
        ("surfel", c_int32),
        ("mask", c_void_p),
        ("opac", c_void_p),
        ("normal", c_void_p),
        ("depth", c_void_p),
    ]


class SoarAdamGroup(Structure):
    _fields_ = [
        ("param", c_void_p),
        ("grad", c_void_p),
        ("exp_avg", c_void_p),
        ("exp_avg_sq", c_void_p),
        ("numel", ctypes.c_int64),
        ("lr", ctypes.c_double),
        ("step", c_int32),
        ("reserved", c_int32),
    ]


class SoarDensifyStats(Structure):
    _fields_ = [(k, c_void_p) for k in ("xyz_gradient_accum", "scale_gradient_accum",
                                        "rot_gradient_accum", "opac_gradient_accum", "denom",
                                        "max_radii2D")]


class SoarRowTensor(Structure):
    _fields_ = [("data", c_void_p), ("width", c_int32), ("reserved", c_int32)]


ADAM_MAX_GROUPS = 16
ROWS_MAX_TENSORS = 32


# name -> (restype, argtypes); mirrors include/soar_b200.h one to one
SIGNATURES = {
    "soar_rast_geom_bytes": (c_size_t, [c_int32]),
    "soar_rast_image_bytes": (c_size_t, [c_int32, c_int32]),
    "soar_rast_binning_bytes": (c_size_t, [c_uint32, c_int32, c_int32]),
    "soar_rast_forward_bin": (
        c_int32,
        [POINTER(SoarRastView), c_void_p, c_size_t, c_void_p, POINTER(c_uint32), c_void_p],
    ),
    "soar_rast_forward_bin_async": (
        c_int32,
        [POINTER(SoarRastView), c_void_p, c_size_t, c_void_p, c_void_p, c_void_p],
    ),
    "soar_rast_forward_render": (
        c_int32,
        [POINTER(SoarRastView), c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_uint32,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_rast_backward_scratch_bytes": (c_size_t, [c_int32]),
    "soar_rast_backward": (
        c_int32,
        [POINTER(SoarRastView), c_void_p, c_void_p, c_void_p, c_void_p, c_uint32, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, POINTER(SoarRastGrads), c_void_p],
    ),
    "soar_rast_backward_ordered": (
        c_int32,
        [POINTER(SoarRastView), c_void_p, c_void_p, c_void_p, c_void_p, c_uint32, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, POINTER(SoarRastGrads), c_void_p, c_void_p,
         c_void_p],
    ),
    "soar_rast_mark_visible": (c_int32, [c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "soar_rast_debug_ptr": (
        c_void_p,
        [c_int32, c_int32, c_int32, c_int32, c_uint32, c_void_p, c_void_p, c_void_p],
    ),
    "soar_rast_debug_sorted_keys": (
        c_int32,
        [c_int32, c_int32, c_int32, c_uint32, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_knn_workspace_bytes": (c_size_t, [c_int32]),
    "soar_knn_dist2": (c_int32, [c_void_p, c_int32, c_void_p, c_void_p, c_size_t, c_void_p]),
    "soar_lbs_forward": (
        c_int32,
        [c_void_p, c_void_p, c_int32, c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_void_p,
         c_void_p, c_void_p],
    ),
    "soar_lbs_backward": (
        c_int32,
        [c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_lbs_query_weights": (
        c_int32,
        [c_void_p, c_int32, c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p, c_void_p,
         c_void_p, c_void_p],
    ),
    "soar_lbs_query_grid_bytes": (c_size_t, [c_int32]),
    "soar_lbs_query_build_grid": (c_int32, [c_void_p, c_int32, c_void_p, c_size_t, c_void_p]),
    "soar_lbs_query_weights_grid": (
        c_int32,
        [c_void_p, c_int32, c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p, c_void_p,
         c_void_p, c_void_p],
    ),
    "soar_lbs_query_cells_workspace_bytes": (c_size_t, [c_int32]),
    "soar_lbs_query_weights_cells": (
        c_int32,
        [c_void_p, c_int32, c_void_p, c_int32, c_void_p, c_int32, c_int32, c_void_p, c_void_p,
         c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "soar_smplx_workspace_bytes": (c_size_t, [c_int32, c_int32]),
    "soar_smplx_forward": (
        c_int32,
        [POINTER(SoarSmplxModel), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "soar_post_forward": (
        c_int32,
        [POINTER(SoarPostArgs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_post_backward": (
        c_int32,
        [POINTER(SoarPostArgs), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_occgrid_splat": (
        c_int32,
        [c_int32, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(c_int32), c_void_p, c_void_p, c_void_p,
         c_void_p, c_void_p, c_void_p, c_void_p],
    ),
    "soar_contour_padding": (
        c_int32,
        [c_void_p, c_void_p, c_int32, c_int32, c_int32, c_int32, c_void_p, c_void_p],
    ),
    "soar_adam_step": (c_int32, [POINTER(SoarAdamGroup), c_int32, ctypes.c_double, ctypes.c_double,
                                ctypes.c_double, c_void_p]),
    "soar_densify_add_stats": (
        c_int32,
        [c_int32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(SoarDensifyStats), c_void_p],
    ),
    "soar_densify_select": (
        c_int32,
        [c_int32, c_int32, POINTER(SoarDensifyStats), c_void_p, c_void_p, c_float, c_float, c_float,
         c_float, c_float, c_void_p, c_void_p, c_void_p],
    ),
    "soar_mask_index_workspace_bytes": (c_size_t, [c_int32]),
    "soar_mask_to_index": (
        c_int32,
        [c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "soar_mask_to_index_static": (
        c_int32,
        [c_void_p, c_int32, c_int32, c_void_p, c_void_p, c_void_p],
    ),
    "soar_rows_gather": (
        c_int32,
        [POINTER(SoarRowTensor), POINTER(SoarRowTensor), c_int32, c_void_p, c_int32, ctypes.c_int64,
         ctypes.c_int64, c_void_p],
    ),
    "soar_densify_split_rows": (
        c_int32,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int32, c_int32, c_void_p, c_int32, c_void_p,
         c_void_p, c_void_p],
    ),
    "soar_launch_count": (ctypes.c_ulonglong, []),
    "soar_prof_enable": (None, [c_int32]),
    "soar_prof_summary": (c_int32, [c_int32, POINTER(c_int32), POINTER(c_float)]),
    "soar_kernel_name": (c_char_p, [c_int32]),
    "soar_last_error": (c_char_p, []),
    "soar_version": (c_int32, []),
}

_lib = None


def load():
    """Load libsoar_b200.so (once). Raises ImportError with build instructions if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `bash soar_b200/csrc/build.sh` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). "
            "There is no CPU fallback for the render path."
        )
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    msg = load().soar_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {last_error()}")


def ptr(t):
    """Device pointer of a tensor, or None for an absent / empty one."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
