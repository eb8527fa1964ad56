"""CPU: libgrb_b200.so loads without a GPU and exports every symbol include/GraphBLAS_b200.h declares;
without a device it refuses to initialise (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "GraphBLAS_b200.h")
LIB = os.path.join(ROOT, "graph-python_b200", "libgrb_b200.so")


def _declared():
    text = open(HEADER).read()
    body = re.search(r"/\* cffi-begin \*/(.*)/\* cffi-end \*/", text, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    funcs = re.findall(r"\b(G[rx]B_\w+)\s*\(", body)
    globs = re.findall(r"^extern\s+[\w\s\*]+?\b(G[rx]B_\w+)\s*;", body, re.M)
    return sorted(set(funcs)), sorted(set(globs))


@pytest.fixture(scope="module")
def dll():
    if not os.path.exists(LIB):
        pytest.fail(f"{LIB} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
    return ctypes.CDLL(LIB)


def test_every_declared_symbol_is_exported(dll):
    funcs, globs = _declared()
    assert len(funcs) > 200 and len(globs) > 400
    missing = [s for s in funcs + globs if not hasattr(dll, s)]
    assert not missing, f"declared in the header but not exported: {missing[:10]}"


def test_hot_path_entry_points_present(dll):
    for name in ("GrB_mxm", "GrB_mxv", "GrB_vxm", "GrB_MIN_PLUS_SEMIRING_INT64", "GrB_PLUS_TIMES_SEMIRING_FP32",
                 "GxB_ANY_PAIR_BOOL", "GxB_PLUS_PAIR_INT64", "GrB_LOR_LAND_SEMIRING_BOOL", "GrB_MIN_FIRST_SEMIRING_UINT64",
                 "GrB_MIN_SECOND_SEMIRING_UINT32", "GxB_PLUS_SECOND_FP32", "GrB_DESC_RSC", "GrB_DESC_T0T1"):
        assert hasattr(dll, name), name


def test_binding_discovers_operators_by_name():
    import grb_b200
    from graphblas_b200 import binary, monoid, semiring

    assert semiring.min_plus["INT32"]._carg == grb_b200.lib.GrB_MIN_PLUS_SEMIRING_INT32
    assert semiring.any_pair["BOOL"].gb_name == "GxB_ANY_PAIR_BOOL"
    assert monoid.plus["FP64"].gb_name == "GrB_PLUS_MONOID_FP64"
    assert binary.plus["INT64"].gb_name == "GrB_PLUS_INT64"
    # INT8/INT16/UINT8/UINT16 semirings exist (computed in the 32-bit type, wrapped to the narrow one)
    assert semiring.plus_times["INT8"]._carg == grb_b200.lib.GrB_PLUS_TIMES_SEMIRING_INT8
    assert semiring.min_plus["UINT16"]._carg == grb_b200.lib.GrB_MIN_PLUS_SEMIRING_UINT16


def test_semiring_new_composes_builtins_and_udfs_are_refused(dll):
    """GrB_Semiring_new over built-in handles yields a handle (the reference's registry builds plus_pow & co at import,
    graphblas/core/operator/semiring.py:370-390); user-defined FUNCTIONS are refused: no JIT, no CPU fallback."""
    import grb_b200

    ffi, lib = grb_b200.ffi, grb_b200.lib
    s = ffi.new("GrB_Semiring*")
    assert lib.GrB_Semiring_new(s, lib.GrB_PLUS_MONOID_INT64, lib.GrB_TIMES_INT64) == lib.GrB_SUCCESS
    assert s[0] != ffi.NULL
    assert lib.GrB_Semiring_free(s) == lib.GrB_SUCCESS and s[0] == ffi.NULL
    b = ffi.new("GrB_BinaryOp*")
    assert lib.GrB_BinaryOp_new(b, ffi.NULL, lib.GrB_INT64, lib.GrB_INT64, lib.GrB_INT64) == lib.GrB_NOT_IMPLEMENTED
    u = ffi.new("GrB_UnaryOp*")
    assert lib.GrB_UnaryOp_new(u, ffi.NULL, lib.GrB_INT64, lib.GrB_INT64) == lib.GrB_NOT_IMPLEMENTED


def test_registry_symbols_the_reference_scans_for(dll):
    """Every name the reference's _parse_config regexes can match for non-complex types is exported
    (graphblas/core/operator/{unary,binary,monoid,semiring,indexunary}.py), so `import graphblas` does not stop at a
    missing attribute (round-1 probe: KeyError 'erf' at core/operator/unary.py:418)."""
    for name in ("GxB_ERF_FP32", "GxB_LGAMMA_FP64", "GxB_SQRT_FP32", "GxB_ISNAN_FP64", "GrB_VALUENE_INT64",
                 "GrB_VALUEEQ_BOOL", "GrB_ROWINDEX_INT64", "GrB_COLLE", "GrB_BOR_UINT8", "GxB_POW_FP64", "GxB_FIRSTI_INT64",
                 "GxB_BSHIFT_INT8", "GxB_BOR_UINT32_MONOID", "GxB_EQ_BOOL_MONOID", "GxB_PLUS_RMINUS_FP32",
                 "GxB_ANY_SECONDJ1_INT32", "GxB_LOR_GT_UINT8", "GxB_BOR_BAND_UINT64", "GxB_EQ_LXOR_BOOL",
                 "GrB_PLUS_TIMES_SEMIRING_INT8", "GrB_MIN_PLUS_SEMIRING_UINT16", "GxB_PLUS_TIMES_INT32",
                 "GxB_COMPRESSION_NONE", "GxB_SECURE_IMPORT", "GxB_Matrix_import_CSR", "GxB_Matrix_export_BitmapR",
                 "GxB_Vector_import_CSC", "GxB_Vector_export_Full", "GxB_Matrix_pack_FullC"):
        import grb_b200

        assert hasattr(dll, name) or hasattr(grb_b200.lib, name), name


def test_no_cpu_fallback_without_device():
    import torch

    if torch.cuda.is_available():
        pytest.skip("device present")
    import grb_b200

    if not grb_b200.is_initialized():
        with pytest.raises(RuntimeError, match="CUDA device"):
            grb_b200.initialize()
    A = grb_b200.ffi.new("GrB_Matrix*")
    assert grb_b200.lib.GrB_Matrix_new(A, grb_b200.lib.GrB_INT64, 3, 3) == grb_b200.lib.GrB_PANIC
