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
    with pytest.raises(KeyError):
        semiring.plus_times["INT8"]  # no sm_100a kernel instantiated: must raise, not fall back


def test_udf_semiring_is_refused(dll):
    import grb_b200

    s = grb_b200.ffi.new("GrB_Semiring*")
    assert grb_b200.lib.GrB_Semiring_new(s, grb_b200.lib.GrB_PLUS_MONOID_INT64, grb_b200.lib.GrB_TIMES_INT64) \
        == grb_b200.lib.GrB_NOT_IMPLEMENTED


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
