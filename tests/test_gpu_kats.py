"""GPU: the reference's own golden vectors (tests/golden/reference_kats.json, transcribed from
graphblas/tests/test_matrix.py:307-392 and test_vector.py:297-366, 1479-1491) through the reference-shaped
API and the C-ABI of libgrb_b200.so, plus the reference's infix (`@`) and outer-product tests
(tests/test_infix.py:80-106, tests/test_vector.py:1520-1530) with their own fixtures.  Bit-exact (INT64)."""
import numpy as np
import pytest

from kat_util import load_kats, resolve

pytestmark = pytest.mark.gpu
FX, KATS = load_kats()


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    return gb


def _mat(gb, x, dtype="INT64"):
    x = resolve(FX, x)
    return gb.Matrix.from_coo(x["rows"], x["cols"], x["vals"], dtype=x.get("dtype", dtype),
                              nrows=x.get("nrows", 7), ncols=x.get("ncols", 7))


def _vec(gb, x):
    x = resolve(FX, x)
    return gb.Vector.from_coo(x["idx"], x["vals"], dtype=x.get("dtype", "INT64"), size=x.get("size", 7))


def _mask(m, k):
    mm = m.S if k.get("structure") else m.V
    return ~mm if k.get("comp") else mm


@pytest.mark.parametrize("k", KATS, ids=[k["name"] for k in KATS])
def test_reference_kat(gb, k):
    from graphblas_b200 import binary, semiring

    op = getattr(semiring, k["semiring"].lower())
    accum = getattr(binary, k["accum"].lower()) if k.get("accum") else None
    kw = {}
    if accum is not None:
        kw["accum"] = accum
    if k.get("replace"):
        kw["replace"] = True
    if k["op"] == "mxm":
        A, B = _mat(gb, k["A"]), _mat(gb, k["B"])
        Ae = A.T if k.get("tran0") else A
        Be = B.T if k.get("tran1") else B
        expr = Ae.mxm(Be, op)
        if k["C"] is None:
            C = expr.new()
        else:
            C = _mat(gb, k["C"])
            if "mask" in k:
                kw["mask"] = _mask(_mat(gb, k["mask"]), k)
            C(**kw) << expr
        r, c, v = C.to_coo()
        e = k["expect"]
        assert (r.tolist(), c.tolist(), v.tolist()) == (e["rows"], e["cols"], e["vals"])
    elif k["op"] in ("mxv", "vxm"):
        A, u = _mat(gb, k["A"]), _vec(gb, k["u"])
        if k["op"] == "mxv":
            expr = (A.T if k.get("tran0") else A).mxv(u, op)
        else:
            expr = u.vxm(A.T if k.get("tran1") else A, op)
        if k["w"] is None:
            w = expr.new()
        else:
            w = _vec(gb, k["w"])
            if "mask" in k:
                kw["mask"] = _mask(_vec(gb, k["mask"]), k)
            w(**kw) << expr
        i, v = w.to_coo()
        assert (i.tolist(), v.tolist()) == (k["expect"]["idx"], k["expect"]["vals"])
    else:
        u = _vec(gb, k["u"])
        s = u.inner(u, op).new()
        assert s.value == k["expect"]["value"]
        s(binary.plus) << u.inner(u, op)   # reference tests/test_vector.py:1489-1490
        assert s.value == 2 * k["expect"]["value"]


def test_dimension_mismatch_raises(gb):
    """reference tests/test_matrix.py:1920-1927, tests/test_vector.py:1114-1119"""
    A = _mat(gb, "A")
    v = gb.Vector.from_coo([0, 1], [1, 2], size=5)
    with pytest.raises(gb.DimensionMismatch):
        A.mxv(v)
    with pytest.raises(gb.DimensionMismatch):
        v.vxm(A)


def test_vxm_accum_spellings(gb):
    """reference tests/test_vector.py:349-366: accum as BinaryOp, Monoid, "+" and "plus"."""
    from graphblas_b200 import binary, monoid, semiring

    A, v = _mat(gb, "A"), _vec(gb, "v")
    expect = ([0, 1, 2, 3, 4, 5, 6], [3, 1, 3, 1, 10, 14, 4])
    for spelling in ((binary.plus,), (monoid.plus,), ("+",)):
        w = v.dup()
        w(*spelling) << v.vxm(A, semiring.plus_times)
        i, x = w.to_coo()
        assert (i.tolist(), x.tolist()) == expect
    w = v.dup()
    w(accum="plus") << v.vxm(A, semiring.plus_times)
    assert w.to_coo()[1].tolist() == expect[1]
    w = v.dup()
    w(~~gb.Vector.from_coo([0, 3, 4], [False, False, False], size=7).S) << v.vxm(A, semiring.plus_times)
    assert w.to_coo()[1].tolist() == [3, 1, 0, 8, 0]


def test_mask_must_be_marked(gb):
    """reference tests/test_matrix.py:373-374: a non-bool mask without .S/.V raises TypeError."""
    A, v = _mat(gb, "A"), _vec(gb, "v")
    with pytest.raises(TypeError, match="Mask must be"):
        v.vxm(A).new(mask=v)


def test_matmul_infix_as_in_reference(gb):
    """graphblas/tests/test_infix.py:80-106 (test_matmul) with its own fixtures (:11-29): `left @ right` is the method
    call with the default plus_times semiring, re-typable with `semiring.plus_times(left @ right)` and
    `semiring.plus_times[float](...)`, for vxm / mxv / mxm (plain and transposed operands) and inner."""
    from graphblas_b200 import Matrix, Vector, semiring

    v1 = Vector.from_coo([0, 2], [2.0, 5.0])
    v2 = Vector.from_coo([1, 2], [3.0, 7.0])
    A1 = Matrix.from_coo([0, 0], [0, 1], [0.0, 4.0], ncols=3)
    A2 = Matrix.from_coo([0, 2], [0, 0], [6.0, 8.0])
    for method, left, right in [("vxm", v2, A2), ("vxm", v2, A1.T), ("mxv", A1, v1), ("mxv", A2.T, v1), ("mxm", A1, A2),
                                ("mxm", A1.T, A2.T), ("mxm", A1, A1.T), ("mxm", A2.T, A2), ("inner", v1, v2),
                                ("inner", v1, v1)]:
        expected = getattr(left, method)(right, semiring.plus_times).new()
        assert expected.isequal(semiring.plus_times(left @ right).new()), method
        assert expected.isequal(semiring.plus_times[float](left @ right).new()), method
        assert expected.isequal((left @ right).new()), method
        if isinstance(left, Vector):
            if not isinstance(right, Vector):
                assert (left @ right).size == right.ncols
        elif isinstance(right, Vector):
            assert (left @ right).size == left.nrows
        else:
            assert (left @ right).nrows == left.nrows and (left @ right).ncols == right.ncols
    # the values the reference's fixtures give (worked by hand): v2 @ A2 = [7*8] at 0; A1 @ v1 = [0*2] at 0 (explicit zero)
    i, x = (v2 @ A2).new().to_coo()
    assert list(i) == [0] and list(x) == [56.0]
    i, x = (A1 @ v1).new().to_coo()
    assert list(i) == [0] and list(x) == [0.0]
    assert (v1 @ v2).new().value == 35.0 and (v1 @ v1).new().value == 29.0


def test_outer_as_in_reference(gb):
    """graphblas/tests/test_vector.py:1520-1530 (test_outer): v.outer(v) equals the column-by-row mxm, with the default
    `times` and with `monoid.times`; golden values from the fixture v = {1: 1, 3: 1, 4: 2, 6: 0} (test_vector.py:33-37)."""
    from graphblas_b200 import Matrix, Vector, monoid

    v = Vector.from_coo([1, 3, 4, 6], [1, 1, 2, 0], dtype="INT64", size=7)
    R = Matrix.from_coo([0] * 4, [1, 3, 4, 6], [1, 1, 2, 0], dtype="INT64", nrows=1, ncols=7)
    C = Matrix.from_coo([1, 3, 4, 6], [0] * 4, [1, 1, 2, 0], dtype="INT64", nrows=7, ncols=1)
    expected = C.mxm(R).new()
    assert v.outer(v).new().isequal(expected)
    assert v.outer(v, monoid.times).new().isequal(expected)
    r, c, x = expected.to_coo()
    assert expected.nvals == 16                                     # 4 x 4 entries, explicit zeros kept
    want = {(i, j): a * b for i, a in zip([1, 3, 4, 6], [1, 1, 2, 0]) for j, b in zip([1, 3, 4, 6], [1, 1, 2, 0])}
    assert {(int(i), int(j)): int(t) for i, j, t in zip(r, c, x)} == want


def test_power_as_in_reference(gb):
    """graphblas/tests/test_matrix.py:4367-4395 (test_power) on its fixture A (:33-41): A.power(i) for i < 50 against the
    chain `expected << A @ expected` (the 49 plus_times products overflow INT64: both sides must wrap alike), the
    transposed operand, min_plus, and the exceptional arguments.  The last link of the chain is also compared with the
    C oracle's chain, bit for bit."""
    from graphblas_b200 import semiring
    from oracle import oracle as O

    A = _mat(gb, "A")
    expected = A.dup()
    for i in range(1, 50):
        assert A.power(i).new().isequal(expected), i
        expected << A @ expected
    x = resolve(FX, "A")
    OA = O.Mat.from_coo(x["rows"], x["cols"], np.asarray(x["vals"], dtype=np.int64), 7, 7, "INT64")
    E = OA
    for _ in range(49):
        E = O.spgemm(OA, E, "PLUS", "TIMES", "INT64")
    r, c, v = expected.to_coo()
    er, ec, ev = E.to_coo()
    assert np.array_equal(r.astype(np.int64), er) and np.array_equal(c.astype(np.int64), ec) and np.array_equal(v, ev)
    expected = A.T.new()
    for i in range(1, 10):
        assert A.T.power(i).new().isequal(expected), i
        expected << A.T @ expected
    expected = A.dup()
    for i in range(1, 10):
        assert A.power(i, semiring.min_plus).new().isequal(expected), i
        expected << semiring.min_plus(A @ expected)
    with pytest.raises(TypeError, match="must be a positive integer"):
        A.power(1.5)
    with pytest.raises(ValueError, match="must be a positive integer"):
        A.power(-1)
    with pytest.raises(ValueError, match="must be a positive integer"):
        A.power(0)
    B = gb.Matrix("INT64", 2, 3)
    with pytest.raises(gb.DimensionMismatch):
        B.power(2)


def test_div_semirings_as_in_reference(gb):
    """graphblas/tests/test_op.py:950-965 (test_div_semirings): plus_div is not exposed; plus_cdiv is C division on INT64
    ((-1)/2 + (-3)/2 = 0 + -1), plus_truediv divides in FP64 (-0.5 + -1.5) and returns FP64."""
    from graphblas_b200 import Matrix, dtypes, semiring

    assert not hasattr(semiring, "plus_div")
    A1 = Matrix.from_coo([0, 1], [0, 0], [-1, -3])
    A2 = Matrix.from_coo([0, 1], [0, 0], [2, 2])
    result = A1.T.mxm(A2, semiring.plus_cdiv).new()
    assert result[0, 0].new() == -1
    assert result.dtype == dtypes.INT64
    result = A1.T.mxm(A2, semiring.plus_truediv).new()
    assert result[0, 0].new() == -2
    assert result.dtype == dtypes.FP64
    # the same through the vector products (operand order matters for a non-commutative multiply)
    from graphblas_b200 import Vector

    u = Vector.from_coo([0, 1], [12, 20])
    B = Matrix.from_coo([0, 1], [0, 0], [5, 3])
    i, x = u.vxm(B, semiring.plus_cdiv).new().to_coo()       # u(0)/B(0,0) + u(1)/B(1,0) = 2 + 6
    assert list(i) == [0] and list(x) == [8]
    i, x = B.T.mxv(u, semiring.plus_cdiv).new().to_coo()     # B(0,0)/u(0) + B(1,0)/u(1) = 0 + 0
    assert list(i) == [0] and list(x) == [0]


def test_set_element_in_place_drops_value_caches(gb):
    """GrB_Matrix_setElement on an existing entry edits the values in place; the typed copy of the
    values (INT64 matrix under an FP64 semiring), the cached transpose and the stream form must not survive it."""
    import numpy as np

    from graphblas_b200 import Matrix, Vector, semiring
    from oracle import oracle as O

    scale = 9
    n = 1 << scale
    A = O.rmat_matrix(scale, "INT64", value_mode=1)
    gA = Matrix.from_csr(A.indptr, A.indices, A.values, ncols=n)
    x = np.random.default_rng(3).random(n)
    gx = Vector.from_dense(x)
    ox = O.Vec(x, np.ones(n, np.uint8), "FP64")
    for tran in (False, True):
        for _ in range(3):                      # the third pull runs on the cached stream form
            (gA.T if tran else gA).mxv(gx, semiring.plus_times["FP64"]).new()
    i = int(np.flatnonzero(np.diff(A.indptr) > 0)[5])
    j = int(A.indices[A.indptr[i]])
    gA[i, j] = 1000
    A.values[A.indptr[i]] = 1000
    for tran in (False, True):
        w = O.Vec.empty(n, "FP64")
        O.mxv(w, A, ox, "PLUS", "TIMES", "FP64", tran0=tran)
        got = (gA.T if tran else gA).mxv(gx, semiring.plus_times["FP64"]).new()
        vals, pres = got.to_dense()
        p = w.present.astype(bool)
        assert np.array_equal(pres.astype(bool), p)
        assert np.allclose(vals[p], w.values[p], rtol=1e-12), tran


def test_registry_only_operators_are_refused_at_call_time(gb):
    """Handles that exist only so that python-graphblas's registry imports (tools/gen_abi.py registry_only_*) return
    GrB_NOT_IMPLEMENTED when used: no silent fallback (north_star)."""
    import grb_b200
    from graphblas_b200 import Matrix, Vector

    ffi, lib = grb_b200.ffi, grb_b200.lib
    A = Matrix.from_coo([0, 1], [1, 0], [1, 2], nrows=2, ncols=2, dtype="INT64")
    u = Vector.from_coo([0, 1], [3, 4], size=2, dtype="INT64")
    w = Vector("INT64", 2)
    assert lib.GrB_mxv(w._carg, ffi.NULL, ffi.NULL, lib.GxB_PLUS_RMINUS_INT64, A._carg, u._carg, ffi.NULL) == lib.GrB_NOT_IMPLEMENTED
    msg = ffi.new("char**")
    lib.GrB_Vector_error(msg, w._carg)
    assert b"no sm_100a kernel" in ffi.string(msg[0])
    assert lib.GrB_mxv(w._carg, ffi.NULL, lib.GxB_POW_INT64, lib.GrB_PLUS_TIMES_SEMIRING_INT64, A._carg, u._carg, ffi.NULL) \
        == lib.GrB_NOT_IMPLEMENTED
    assert lib.GrB_Vector_apply(w._carg, ffi.NULL, ffi.NULL, lib.GxB_SQRT_FP64, u._carg, ffi.NULL) == lib.GrB_NOT_IMPLEMENTED
    # a semiring composed from built-ins with a kernel runs; one without is refused when used
    s = ffi.new("GrB_Semiring*")
    assert lib.GrB_Semiring_new(s, lib.GrB_MIN_MONOID_INT64, lib.GrB_PLUS_INT64) == lib.GrB_SUCCESS
    assert lib.GrB_mxv(w._carg, ffi.NULL, ffi.NULL, s[0], A._carg, u._carg, ffi.NULL) == lib.GrB_SUCCESS
    idx, vals = w.to_coo()
    assert list(vals) == [5, 5]        # row 0: 1 + 4, row 1: 2 + 3
    lib.GrB_Semiring_free(s)
    assert lib.GrB_Semiring_new(s, lib.GrB_MIN_MONOID_INT64, lib.GrB_MINUS_INT64) == lib.GrB_SUCCESS
    assert lib.GrB_mxv(w._carg, ffi.NULL, ffi.NULL, s[0], A._carg, u._carg, ffi.NULL) == lib.GrB_NOT_IMPLEMENTED
    lib.GrB_Semiring_free(s)
