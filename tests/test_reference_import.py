"""CPU, this container only (skipped where /root/reference is absent, i.e. on the GPU box): the REAL python-graphblas
imports over grb_b200 and its expression machinery emits the C calls this repo's header declares.

Follows reference graphblas/tests/test_op.py:88-91 (handle identity of typed operators) and
graphblas/tests/test_recorder.py:15-39, 243-264 (textual form of the recorded calls)."""
import pytest

import refshim

pytestmark = pytest.mark.skipif(not refshim.available(), reason="/root/reference is not present on this machine")


@pytest.fixture(scope="module")
def ref():
    gb, dry = refshim.load()
    return gb, dry


def test_registry_imports_and_handles_are_the_exported_globals(ref):
    gb, dry = ref
    from graphblas import agg, binary, indexunary, monoid, op, select, semiring, unary  # noqa: F401

    lib = dry._lib
    # reference tests/test_op.py:88-91
    assert semiring.min_plus["INT32"].gb_obj == lib.GrB_MIN_PLUS_SEMIRING_INT32
    assert semiring.min_plus[gb.dtypes.UINT16].gb_obj == lib.GrB_MIN_PLUS_SEMIRING_UINT16
    assert monoid.max["INT32"].gb_obj == lib.GrB_MAX_MONOID_INT32
    assert binary.plus["UINT16"].gb_obj == lib.GrB_PLUS_UINT16
    assert unary.ainv["INT32"].gb_obj == lib.GrB_AINV_INT32
    # the north_star semirings, each on a wide and a narrow type
    for name, sym in [("plus_times", "GrB_PLUS_TIMES_SEMIRING_{}"), ("min_plus", "GrB_MIN_PLUS_SEMIRING_{}"),
                      ("any_pair", "GxB_ANY_PAIR_{}"), ("plus_pair", "GxB_PLUS_PAIR_{}"),
                      ("min_first", "GrB_MIN_FIRST_SEMIRING_{}"), ("min_second", "GrB_MIN_SECOND_SEMIRING_{}"),
                      ("plus_second", "GxB_PLUS_SECOND_{}")]:
        for t in ("INT64", "INT8", "UINT16", "FP32"):
            assert getattr(semiring, name)[t].gb_obj == getattr(lib, sym.format(t)), (name, t)
    assert semiring.lor_land["BOOL"].gb_obj == lib.GrB_LOR_LAND_SEMIRING_BOOL
    # names the reference imports at module level: core/mask.py:1-5, core/infix.py:4
    assert select.valuene["INT64"].gb_obj == lib.GrB_VALUENE_INT64
    assert unary.one["BOOL"].gb_obj == lib.GxB_ONE_BOOL
    assert semiring.any_pair["BOOL"].gb_obj == lib.GxB_ANY_PAIR_BOOL
    # the GxB math family the registry walks at import (core/operator/unary.py:395-425)
    assert unary.ss.erf["FP64"].gb_obj == lib.GxB_ERF_FP64


def test_recorder_mxm_as_in_reference(ref):
    gb, dry = ref
    A = gb.Matrix(int, 2, 2, name="A")
    B = gb.Matrix(int, 2, 2, name="B")
    with gb.Recorder() as rec:
        C = A.mxm(B).new(name="C")
        D = A.mxm(B.T, gb.semiring.min_plus).new(name="D")
        C(D.S) << A.T.ewise_mult(B)
    # reference tests/test_recorder.py:31-37, verbatim
    assert list(rec) == [
        "GrB_Matrix_new(&C, GrB_INT64, 2, 2);",
        "GrB_mxm(C, NULL, NULL, GrB_PLUS_TIMES_SEMIRING_INT64, A, B, NULL);",
        "GrB_Matrix_new(&D, GrB_INT64, 2, 2);",
        "GrB_mxm(D, NULL, NULL, GrB_MIN_PLUS_SEMIRING_INT64, A, B, GrB_DESC_T1);",
        "GrB_Matrix_eWiseMult_BinaryOp(C, D, NULL, GrB_TIMES_INT64, A, B, GrB_DESC_ST0);",
    ]


def test_hot_path_statements_assemble_the_seven_arguments(ref):
    """The north_star's own statements, through the reference's Updater / _update (core/base.py:438-483)."""
    gb, dry = ref
    from graphblas import binary, semiring

    n = 8
    A = gb.Matrix("INT64", n, n, name="A")
    M = gb.Matrix("INT64", n, n, name="M")
    w = gb.Vector("INT64", n, name="w")
    q = gb.Vector("BOOL", n, name="q")
    v = gb.Vector("INT32", n, name="v")
    G = gb.Matrix("BOOL", n, n, name="G")
    r = gb.Vector("FP32", n, name="r")
    F = gb.Matrix("FP32", n, n, name="F")
    x = gb.Vector("FP32", n, name="x")
    L = gb.Matrix("INT64", n, n, name="L")
    Cc = gb.Matrix("INT64", n, n, name="Cc")
    with gb.Recorder() as rec:
        M(M.V, accum=binary.plus) << A.mxm(A, semiring.min_plus)
        w(binary.min) << w.vxm(A, semiring.min_plus)                       # SSSP notebook cell 21
        q(~v.S, replace=True) << q.vxm(G, semiring.any_pair)               # level BFS
        r(binary.plus) << F.T.mxv(x, semiring.plus_times)                  # PageRank
        Cc(L.S) << L.mxm(L.T, semiring.plus_pair)                          # triangle counting
    assert list(rec) == [
        "GrB_mxm(M, M, GrB_PLUS_INT64, GrB_MIN_PLUS_SEMIRING_INT64, A, A, NULL);",
        "GrB_vxm(w, NULL, GrB_MIN_INT64, GrB_MIN_PLUS_SEMIRING_INT64, w, A, NULL);",
        "GrB_vxm(q, v, NULL, GxB_ANY_PAIR_INT64, q, G, GrB_DESC_RSC);",      # the reference types pair(BOOL, BOOL) as INT64
        "GrB_mxv(r, NULL, GrB_PLUS_FP32, GrB_PLUS_TIMES_SEMIRING_FP32, F, x, GrB_DESC_T0);",
        "GrB_mxm(Cc, L, NULL, GxB_PLUS_PAIR_INT64, L, L, GrB_DESC_ST1);",
    ]
    # and the cffi callbacks saw the exported global handles in the semiring slot
    lib = dry._lib
    mxv = [a for name, a in dry.calls if name == "GrB_mxv"][-1]
    assert mxv[3] == lib.GrB_PLUS_TIMES_SEMIRING_FP32 and mxv[2] == lib.GrB_PLUS_FP32 and mxv[6] == lib.GrB_DESC_T0


def test_ss_import_export_reach_the_gxb_entry_points(ref):
    """Matrix.ss.import_csr / Vector.ss.import_sparse / .ss.export call the GxB entry points with the argument lists
    declared in the header (reference core/ss/matrix.py:1282-1352, 744-1162; core/ss/vector.py:946-1016, 494-636)."""
    gb, dry = ref
    import numpy as np

    del dry.calls[:]
    gb.Matrix.ss.import_csr(nrows=2, ncols=2, indptr=np.array([0, 1, 2], np.uint64), col_indices=np.array([1, 0], np.uint64),
                            values=np.array([1.5, 2.5]), name="S")
    gb.Vector.ss.import_sparse(size=4, indices=np.array([1, 3], np.uint64), values=np.array([7, 9], np.int64), name="sv")
    gb.Vector.ss.import_bitmap(bitmap=np.array([1, 0, 1], bool), values=np.array([1, 2, 3], np.int32), name="bv")
    gb.Vector.ss.import_full(values=np.array([1.0, 2.0], np.float32), name="fv")
    names = [n for n, _ in dry.calls]
    for want in ("GxB_Matrix_import_CSR", "GxB_Vector_import_CSC", "GxB_Vector_import_Bitmap", "GxB_Vector_import_Full"):
        assert want in names, (want, names)


def test_operators_without_kernels_exist_but_are_marked(ref):
    """Registry-only handles (no sm_100a kernel) are distinct objects; the C side refuses them at call time
    (tests/test_gpu_* check the GrB_NOT_IMPLEMENTED on the device)."""
    gb, dry = ref
    lib = dry._lib
    assert lib.GxB_PLUS_ISEQ_INT64 != lib.GxB_PLUS_RMINUS_INT64
    assert lib.GxB_ERF_FP32 != lib.GxB_ERF_FP64
