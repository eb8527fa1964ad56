"""GPU: the reference's own golden vectors (tests/golden/reference_kats.json, transcribed from
graphblas/tests/test_matrix.py:307-392 and test_vector.py:297-366, 1479-1491) through the reference-shaped
API and the C-ABI of libgrb_b200.so.  Bit-exact (INT64)."""
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
