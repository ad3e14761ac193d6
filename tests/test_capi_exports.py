"""CPU suite, part 2: the C-ABI library loads and exports every symbol include/soar_b200.h
declares, the ctypes table in soar_b200/_capi.py covers exactly that set, and the pure size
queries (no device work) behave. No compute call is made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "soar_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(soar_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("soar_rast_forward_bin", "soar_rast_forward_render", "soar_rast_backward",
                 "soar_rast_mark_visible", "soar_knn_dist2", "soar_lbs_forward", "soar_lbs_backward",
                 "soar_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from soar_b200 import _capi
    assert os.path.exists(_capi.LIB_PATH), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(_capi.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/soar_b200.h but not exported"


def test_ctypes_table_matches_header():
    from soar_b200 import _capi
    assert sorted(_capi.SIGNATURES) == declared_functions()


def test_struct_layout_matches_header():
    from soar_b200 import _capi
    src = open(HEADER).read()
    body = re.search(r"typedef struct SoarRastView \{(.*?)\} SoarRastView;", src, flags=re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        names = decl.replace("*", " ").split()
        # "int32_t width, height" / "float tanfovx, tanfovy, scale_modifier" / "const float* bg"
        tail = decl.split(None, 2 if decl.startswith("const") else 1)[-1]
        fields += [n.strip(" *") for n in tail.split(",")]
    assert fields == [f[0] for f in _capi.SoarRastView._fields_]


def test_size_queries_need_no_device():
    from soar_b200 import _capi
    lib = _capi.load()
    assert lib.soar_version() >= 100
    g0, g1 = lib.soar_rast_geom_bytes(0), lib.soar_rast_geom_bytes(500_000)
    assert g1 > g0 and g1 >= 500_000 * 96
    assert lib.soar_rast_image_bytes(1024, 1024) >= 1024 * 1024 * 12
    b0, b1 = lib.soar_rast_binning_bytes(0, 1024, 1024), lib.soar_rast_binning_bytes(4_000_000, 1024, 1024)
    assert b1 > b0
    assert lib.soar_rast_backward_scratch_bytes(1000) >= 1000 * 13 * 4
    assert lib.soar_knn_workspace_bytes(1000) > 0
    assert lib.soar_kernel_name(0)


def test_null_view_is_an_error_not_a_crash():
    from soar_b200 import _capi
    lib = _capi.load()
    rendered = ctypes.c_uint32(7)
    rc = lib.soar_rast_forward_bin(None, None, 0, None, ctypes.byref(rendered), None)
    assert rc < 0 and "NULL" in _capi.last_error()
    with pytest.raises(RuntimeError):
        _capi.check(rc, "forward_bin")
