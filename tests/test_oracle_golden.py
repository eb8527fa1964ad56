"""Pin both CPU restatements (oracle/pyref.py, oracle/grb_oracle.c) to the reference's own golden
vectors (graphblas/tests/test_matrix.py:307-392, test_vector.py:297-366, 1479-1491) and to each other."""
import numpy as np
import pytest

from kat_util import flags, load_kats, resolve
from oracle import oracle as O
from oracle import pyref

FX, KATS = load_kats()


def _omat(x, dtype="INT64"):
    if x is None:
        return None
    x = resolve(FX, x)
    nr, nc = x.get("nrows", 7), x.get("ncols", 7)
    return O.Mat.from_coo(x["rows"], x["cols"], x["vals"], nr, nc, x.get("dtype", dtype))


def _ovec(x, size=7):
    if x is None:
        return None
    x = resolve(FX, x)
    return O.Vec.from_coo(x["idx"], x["vals"], x.get("size", size), x.get("dtype", "INT64"))


def _pmat(x):
    if x is None:
        return {}
    x = resolve(FX, x)
    return {(i, j): np.int64(v) for i, j, v in zip(x["rows"], x["cols"], x["vals"])}


def _pvec(x):
    if x is None:
        return {}
    x = resolve(FX, x)
    return {i: np.int64(v) for i, v in zip(x["idx"], x["vals"])}


@pytest.mark.parametrize("k", KATS, ids=[k["name"] for k in KATS])
def test_c_oracle_matches_reference_kat(k):
    mon, mul = k["semiring"].split("_")
    f = flags(k)
    if k["op"] == "mxm":
        A, B = _omat(k["A"]), _omat(k["B"])
        nr = A.shape[1] if k.get("tran0") else A.shape[0]
        nc = B.shape[0] if k.get("tran1") else B.shape[1]
        C = _omat(k["C"]) or O.Mat(np.zeros(nr + 1), [], np.zeros(0, dtype=np.int64), (nr, nc), "INT64")
        for method in ("gustavson", "dot"):
            out = O.mxm(C, A, B, mon, mul, "INT64", mask=_omat(k.get("mask")), tran0=k.get("tran0", False),
                        tran1=k.get("tran1", False), method=method, **f)
            r, c, v = out.to_coo()
            e = k["expect"]
            assert (r.tolist(), c.tolist(), v.tolist()) == (e["rows"], e["cols"], e["vals"])
    elif k["op"] in ("mxv", "vxm"):
        A, u = _omat(k["A"]), _ovec(k["u"])
        n_out = (A.shape[1] if k.get("tran0") else A.shape[0]) if k["op"] == "mxv" else (
            A.shape[0] if k.get("tran1") else A.shape[1])
        w = _ovec(k["w"]) or O.Vec.empty(n_out, "INT64")
        mask = _ovec(k.get("mask"))
        if k["op"] == "mxv":
            O.mxv(w, A, u, mon, mul, "INT64", mask=mask, tran0=k.get("tran0", False), **f)
        else:
            O.vxm(w, u, A, mon, mul, "INT64", mask=mask, tran1=k.get("tran1", False), **f)
        i, v = w.to_coo()
        assert (i.tolist(), v.tolist()) == (k["expect"]["idx"], k["expect"]["vals"])
    else:  # inner: GrB_vxm on the vector cast to an n x 1 matrix (reference core/vector.py:1641-1648)
        u = _ovec(k["u"])
        i, v = u.to_coo()
        col = O.Mat.from_coo(i, np.zeros_like(i), v, u.size, 1, "INT64")
        w = O.Vec.empty(1, "INT64")
        O.vxm(w, u, col, mon, mul, "INT64")
        assert w.values[0] == k["expect"]["value"] and w.present[0] == 1


@pytest.mark.parametrize("k", KATS, ids=[k["name"] for k in KATS])
def test_pyref_matches_reference_kat(k):
    mon, mul = k["semiring"].split("_")
    f = flags(k)
    if k["op"] == "mxm":
        out = pyref.mxm(_pmat(k["C"]), _pmat(k["A"]), _pmat(k["B"]), mon, mul, "INT64",
                        M=_pmat(k["mask"]) if "mask" in k else None, tran0=k.get("tran0", False),
                        tran1=k.get("tran1", False), **f)
        assert out == _pmat(k["expect"])
    elif k["op"] == "mxv":
        out = pyref.mxv(_pvec(k["w"]), _pmat(k["A"]), _pvec(k["u"]), mon, mul, "INT64",
                        M=_pvec(k["mask"]) if "mask" in k else None, tran0=k.get("tran0", False), **f)
        assert out == _pvec(k["expect"])
    elif k["op"] == "vxm":
        out = pyref.vxm(_pvec(k["w"]), _pvec(k["u"]), _pmat(k["A"]), mon, mul, "INT64",
                        M=_pvec(k["mask"]) if "mask" in k else None, tran1=k.get("tran1", False), **f)
        assert out == _pvec(k["expect"])
    else:
        u = _pvec(k["u"])
        out = pyref.vxm({}, u, {(i, 0): x for i, x in u.items()}, mon, mul, "INT64")
        assert out == {0: np.int64(k["expect"]["value"])}


def test_power_chain_wraps_like_int64():
    """reference tests/test_matrix.py:4367-4395: 49 chained plus_times products overflow INT64; the C
    oracle and the Python restatement must wrap identically (bit-exact integer parity)."""
    A = _omat("A")
    E = A
    P = _pmat("A")
    PE = dict(P)
    for _ in range(49):
        E = O.spgemm(A, E, "PLUS", "TIMES", "INT64")
        PE = pyref.semiring_product(P, PE, "PLUS", "TIMES", "INT64")
    r, c, v = E.to_coo()
    assert {(int(i), int(j)): int(x) for i, j, x in zip(r, c, v)} == {k: int(x) for k, x in PE.items()}


SEMIRINGS = [("PLUS", "TIMES"), ("MIN", "PLUS"), ("MAX", "PLUS"), ("ANY", "PAIR"), ("PLUS", "PAIR"),
             ("MIN", "FIRST"), ("MIN", "SECOND"), ("PLUS", "SECOND"), ("MAX", "MIN")]


@pytest.mark.parametrize("mon,mul", SEMIRINGS)
@pytest.mark.parametrize("seed", [0, 1])
def test_c_oracle_vs_pyref_random(mon, mul, seed):
    """Cross-check the two restatements on random small inputs with every mask/accum/replace combination."""
    rng = np.random.default_rng(seed)
    n, m = 13, 9
    def rmat(nr, nc, dens):
        msk = rng.random((nr, nc)) < dens
        r, c = np.nonzero(msk)
        return r, c, rng.integers(-3, 4, size=r.size)
    ar, ac, av = rmat(n, m, 0.3)
    br, bc, bv = rmat(m, n, 0.3)
    cr, cc, cv = rmat(n, n, 0.2)
    mr, mc, mv = rmat(n, n, 0.4)
    A, B = O.Mat.from_coo(ar, ac, av, n, m, "INT64"), O.Mat.from_coo(br, bc, bv, m, n, "INT64")
    C, M = O.Mat.from_coo(cr, cc, cv, n, n, "INT64"), O.Mat.from_coo(mr, mc, mv, n, n, "INT64")
    pd = lambda r, c, v: {(int(i), int(j)): np.int64(x) for i, j, x in zip(r, c, v)}
    for accum in (None, "PLUS", "MIN"):
        for replace in (False, True):
            for comp in (False, True):
                for structure in (False, True):
                    out = O.mxm(C, A, B, mon, mul, "INT64", mask=M, accum=accum, replace=replace,
                                comp=comp, structure=structure)
                    exp = pyref.mxm(pd(cr, cc, cv), pd(ar, ac, av), pd(br, bc, bv), mon, mul, "INT64",
                                    M=pd(mr, mc, mv), accum=accum, replace=replace, comp=comp,
                                    structure=structure)
                    got = pd(*out.to_coo())
                    if mon == "ANY" and mul != "PAIR":
                        assert set(got) == set(exp)
                    else:
                        assert got == exp


def test_rmat_generator_is_deterministic_and_skewed():
    r1, c1 = O.rmat_pattern(12, 16)
    r2, c2 = O.rmat_pattern(12, 16)
    assert np.array_equal(r1, r2) and np.array_equal(c1, c2)
    assert (r1 != c1).all()
    deg = np.bincount(r1, minlength=1 << 12)
    assert deg.max() > 20 * max(1, int(np.median(deg)))  # heavy tail
    assert r1.size > 0.7 * (16 << 12)


def test_div_semirings_reference_values():
    """reference tests/test_op.py:950-965: A1' (+).(/) A2 with A1 = [-1, -3]', A2 = [2, 2]' is -1 under C division on INT64
    and -2 in FP64; plus the operand order of vxm / mxv and the integer-division corner cases SuiteSparse defines
    (x / 0 saturates by the sign of x, 0 / 0 = 0, INT64_MIN / -1 wraps)."""
    A1 = O.Mat.from_coo([0, 1], [0, 0], np.array([-1, -3], dtype=np.int64), 2, 1, "INT64")
    A2 = O.Mat.from_coo([0, 1], [0, 0], np.array([2, 2], dtype=np.int64), 2, 1, "INT64")
    assert O.spgemm(A1.transpose(), A2, "PLUS", "DIV", "INT64").values.tolist() == [-1]
    F1 = O.Mat.from_coo([0, 1], [0, 0], np.array([-1.0, -3.0]), 2, 1, "FP64")
    F2 = O.Mat.from_coo([0, 1], [0, 0], np.array([2.0, 2.0]), 2, 1, "FP64")
    assert O.spgemm(F1.transpose(), F2, "PLUS", "DIV", "FP64").values.tolist() == [-2.0]
    u = O.Vec.from_coo([0, 1], [12, 20], 2, "INT64")
    B = O.Mat.from_coo([0, 1], [0, 0], np.array([5, 3], dtype=np.int64), 2, 1, "INT64")
    w = O.Vec.from_coo([], [], 1, "INT64")
    O.vxm(w, u, B, "PLUS", "DIV", "INT64")                     # u(i) / B(i,0): 2 + 6
    assert w.to_coo()[1].tolist() == [8]
    w = O.Vec.from_coo([], [], 1, "INT64")
    O.mxv(w, B, u, "PLUS", "DIV", "INT64", tran0=True)         # B(i,0) / u(i): 0 + 0
    assert w.to_coo()[1].tolist() == [0]
    imin, imax = np.iinfo(np.int64).min, np.iinfo(np.int64).max
    X = O.Mat.from_coo([0, 1, 2, 3], [0, 1, 2, 3], np.array([7, -7, 0, imin], dtype=np.int64), 4, 4, "INT64")
    Y = O.Mat.from_coo([0, 1, 2, 3], [0, 1, 2, 3], np.array([0, 0, 0, -1], dtype=np.int64), 4, 4, "INT64")
    assert O.spgemm(X, Y, "PLUS", "DIV", "INT64").values.tolist() == [imax, imin, 0, imin]
