"""GPU: loop-body vector operations around the products: unary apply, extract u[I],
indexed assign w[I] << u, row/column-wise reduce of a matrix, and the option getters the reference queries.
Checked against numpy restatements of the C API semantics (mask / accum / replace as in oracle.vec_writeback)."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    return gb


def _vec(gb, vals, pres, dt):
    idx = np.nonzero(pres)[0]
    return gb.Vector.from_coo(idx, vals[idx], size=vals.size, dtype=dt)


def _rand(rng, n, dt, density):
    pres = rng.random(n) < density
    if dt == "BOOL":
        vals = rng.random(n) < 0.5
    elif dt.startswith("FP"):
        vals = (rng.random(n) * 20 - 10).astype(O.NP_TYPES[dt])
    elif dt.startswith("U"):
        vals = rng.integers(0, 50, size=n).astype(O.NP_TYPES[dt])
    else:
        vals = rng.integers(-50, 50, size=n).astype(O.NP_TYPES[dt])
    return vals, pres


@pytest.mark.parametrize("dt", ["INT8", "INT32", "INT64", "UINT32", "FP32", "FP64"])
def test_apply_unary(gb, dt):
    from graphblas_b200 import unary

    rng = np.random.default_rng(3)
    n = 5000
    vals, pres = _rand(rng, n, dt, 0.6)
    u = _vec(gb, vals, pres, dt)
    npdt = O.NP_TYPES[dt]
    with np.errstate(all="ignore"):
        expect = {
            "identity": vals,
            "ainv": (np.zeros(n, dtype=npdt) - vals).astype(npdt),
            "abs": np.abs(vals).astype(npdt) if not dt.startswith("U") else vals,
            "one": np.ones(n, dtype=npdt),
        }
        if dt.startswith("FP"):
            expect["minv"] = (npdt(1) / vals).astype(npdt)
        if "INT" in dt:
            expect["bnot"] = (~vals).astype(npdt)
    for name, ref in expect.items():
        w = u.apply(getattr(unary, name)).new()
        gv, gp = w.to_dense()
        assert np.array_equal(gp.astype(bool), pres), name
        if dt.startswith("FP"):
            assert np.allclose(gv[pres], ref[pres], rtol=1e-6, equal_nan=True), name
        else:
            assert np.array_equal(gv[pres], ref[pres]), name
    # masked, accumulated: w(m.S, plus) << apply(abs)
    from graphblas_b200 import binary

    wv, wp = _rand(rng, n, dt, 0.5)
    mv, mp = _rand(rng, n, "BOOL", 0.5)
    w, m = _vec(gb, wv, wp, dt), _vec(gb, mv, mp, "BOOL")
    w(m.S, binary.plus) << u.apply(unary.abs)
    gv, gp = w.to_dense()
    t = expect["abs"]
    both, only_t = wp & pres & mp, ~wp & pres & mp
    ref_v, ref_p = wv.copy(), wp.copy()
    ref_v[both] = (wv[both] + t[both]).astype(npdt)
    ref_v[only_t] = t[only_t]
    ref_p[only_t] = True
    assert np.array_equal(gp.astype(bool), ref_p)
    assert np.array_equal(gv[ref_p], ref_v[ref_p]) if not dt.startswith("FP") else np.allclose(gv[ref_p], ref_v[ref_p])


def test_apply_lnot_and_integer_minv(gb):
    from graphblas_b200 import unary

    b = gb.Vector.from_coo([0, 2, 3], [True, False, True], size=5, dtype="BOOL")
    idx, vals = b.apply(unary.lnot).new().to_coo()
    assert idx.tolist() == [0, 2, 3] and vals.tolist() == [False, True, False]
    v = gb.Vector.from_coo([0, 1, 2, 3], [1, -1, 2, -7], size=4, dtype="INT32")
    idx, vals = v.apply(unary.minv).new().to_coo()
    assert vals.tolist() == [1, -1, 0, 0]


@pytest.mark.parametrize("dt", ["INT64", "FP32", "BOOL"])
def test_extract_index_list(gb, dt):
    rng = np.random.default_rng(5)
    n, ni = 4000, 2500
    vals, pres = _rand(rng, n, dt, 0.7)
    u = _vec(gb, vals, pres, dt)
    I = rng.integers(0, n, size=ni)           # duplicates are legal for extract
    w = u[I].new()
    gv, gp = w.to_dense()
    assert w.size == ni
    assert np.array_equal(gp.astype(bool), pres[I])
    assert np.array_equal(gv[pres[I]], vals[I][pres[I]])
    # into an existing vector under a complemented structural mask, replace
    wv, wp = _rand(rng, ni, dt, 0.5)
    mv, mp = _rand(rng, ni, "BOOL", 0.4)
    w2, m = _vec(gb, wv, wp, dt), _vec(gb, mv, mp, "BOOL")
    w2(~m.S, replace=True) << u[I]
    gv, gp = w2.to_dense()
    ref_p = pres[I] & ~mp
    assert np.array_equal(gp.astype(bool), ref_p)
    assert np.array_equal(gv[ref_p], vals[I][ref_p])
    with pytest.raises(gb.IndexOutOfBound):
        u[[0, n]].new()


@pytest.mark.parametrize("dt", ["INT32", "FP64"])
def test_assign_index_list(gb, dt):
    from graphblas_b200 import binary

    rng = np.random.default_rng(8)
    n, ni = 3000, 1200
    wv, wp = _rand(rng, n, dt, 0.6)
    uv, up = _rand(rng, ni, dt, 0.7)
    I = rng.permutation(n)[:ni]               # no duplicates: the C API leaves those undefined
    # plain: w[I] << u  (absent u(k) deletes w(I[k]))
    w, u = _vec(gb, wv, wp, dt), _vec(gb, uv, up, dt)
    w[I] << u
    ref_v, ref_p = wv.copy(), wp.copy()
    ref_v[I[up]] = uv[up]
    ref_p[I] = up
    gv, gp = w.to_dense()
    assert np.array_equal(gp.astype(bool), ref_p)
    assert np.array_equal(gv[ref_p], ref_v[ref_p])
    # accumulate: w(min)[I] << u  (absent u(k) leaves w alone)
    w = _vec(gb, wv, wp, dt)
    w(binary.min)[I] << u
    ref_v, ref_p = wv.copy(), wp.copy()
    tgt = I[up]
    had = wp[tgt]
    ref_v[tgt[had]] = np.minimum(wv[tgt[had]], uv[up][had])
    ref_v[tgt[~had]] = uv[up][~had]
    ref_p[tgt] = True
    gv, gp = w.to_dense()
    assert np.array_equal(gp.astype(bool), ref_p)
    assert np.array_equal(gv[ref_p], ref_v[ref_p])
    # scalar into an index list
    w = _vec(gb, wv, wp, dt)
    w[I] << 7
    gv, gp = w.to_dense()
    ref_p = wp.copy(); ref_p[I] = True
    ref_v = wv.copy(); ref_v[I] = 7
    assert np.array_equal(gp.astype(bool), ref_p) and np.array_equal(gv[ref_p], ref_v[ref_p])
    with pytest.raises(gb.IndexOutOfBound):
        w[[n + 3]] << 1


@pytest.mark.parametrize("dt", ["UINT32", "INT64", "INT32"])
@pytest.mark.parametrize("acc", ["min", "max", "plus"])
def test_assign_duplicate_indices_fold_under_the_accumulator(gb, dt, acc):
    """`f(binary.min)[I] << mngp` with DUPLICATE indices: FastSV's hooking step (reference notebooks/Connected Components --
    FastSV.ipynb cell 30) relies on every duplicate being folded with the accumulator (SuiteSparse does; the C API leaves it
    undefined).  k_assign_atomic against a numpy restatement (ufunc.at), bit-exact on integers, heavy duplication."""
    from graphblas_b200 import binary

    rng = np.random.default_rng(21)
    n, ni = 2000, 9000                        # 4.5 writes per target on average, some targets untouched
    wv, wp = _rand(rng, n, dt, 0.7)
    uv, up = _rand(rng, ni, dt, 0.8)
    I = rng.integers(0, n // 2, size=ni)      # only the lower half is hit
    w, u = _vec(gb, wv, wp, dt), _vec(gb, uv, up, dt)
    w(getattr(binary, acc))[I] << u
    ufunc = {"min": np.minimum, "max": np.maximum, "plus": np.add}[acc]
    tgt, src = I[up], uv[up]
    ref_v, ref_p = wv.copy(), wp.copy()
    fresh = ~wp
    if acc == "plus":
        ref_v[fresh] = 0
    else:
        info = np.iinfo(ref_v.dtype)
        ref_v[fresh] = info.max if acc == "min" else info.min       # identity where w had no entry
    ufunc.at(ref_v, tgt, src)
    ref_p[tgt] = True
    gv, gp = w.to_dense()
    assert np.array_equal(gp.astype(bool), ref_p)
    assert np.array_equal(gv[ref_p], ref_v[ref_p])
    assert w.nvals == int(ref_p.sum())


@pytest.mark.parametrize("dt,mon", [("INT64", "plus"), ("FP32", "plus"), ("INT32", "min"), ("UINT32", "max"), ("BOOL", "lor"),
                                    ("BOOL", "land")])
def test_reduce_rowwise_columnwise(gb, dt, mon):
    from graphblas_b200 import monoid

    scale = 11
    A = O.rmat_matrix(scale, dt, value_mode=0 if dt == "BOOL" else 1)
    gA = gb.Matrix.rmat(scale, dt, value_mode=0 if dt == "BOOL" else 1)
    n = A.shape[0]
    rows = np.repeat(np.arange(n), np.diff(A.indptr))
    npdt = O.NP_TYPES[dt]
    ufunc = {"plus": np.add, "min": np.minimum, "max": np.maximum, "lor": np.logical_or, "land": np.logical_and}[mon]
    ident = {"plus": 0, "min": np.iinfo(npdt).max if dt != "BOOL" and not dt.startswith("FP") else 0,
             "max": 0, "lor": False, "land": True}[mon]
    for axis, idx, expr in ((0, rows, gA.reduce_rowwise(getattr(monoid, mon))),
                            (1, A.indices, gA.reduce_columnwise(getattr(monoid, mon))),
                            (1, A.indices, gA.T.reduce_rowwise(getattr(monoid, mon)))):
        ref = np.full(n, ident, dtype=npdt)
        ufunc.at(ref, idx, A.values)
        has = np.zeros(n, dtype=bool)
        has[idx] = True
        w = expr.new()
        gv, gp = w.to_dense()
        assert np.array_equal(gp.astype(bool), has), axis
        if dt.startswith("FP"):
            assert np.allclose(gv[has], ref[has], rtol=1e-5), axis
        else:
            assert np.array_equal(gv[has], ref[has]), axis


def test_option_getters(gb):
    lib, ffi = gb.lib, gb.ffi
    v = ffi.new("int32_t*")
    assert lib.GxB_Global_Option_get_INT32(lib.GxB_MODE, v) == lib.GrB_SUCCESS
    assert v[0] in (lib.GrB_BLOCKING, lib.GrB_NONBLOCKING)
    A = gb.Matrix.from_coo([0, 1], [1, 0], [1.0, 2.0], nrows=2, ncols=2)
    assert lib.GxB_Matrix_Option_get_INT32(A._carg, lib.GxB_FORMAT, v) == lib.GrB_SUCCESS and v[0] == lib.GxB_BY_ROW
    assert lib.GxB_Matrix_Option_get_INT32(A._carg, lib.GxB_SPARSITY_STATUS, v) == lib.GrB_SUCCESS and v[0] == lib.GxB_SPARSE
    u = gb.Vector.from_coo([0], [1.0], size=3)
    assert lib.GxB_Vector_Option_get_INT32(u._carg, lib.GxB_SPARSITY_STATUS, v) == lib.GrB_SUCCESS and v[0] == lib.GxB_BITMAP


def test_fused_isequal_matches_reference_semantics(gb):
    """GxB_b200_Vector_isequal (one pass, one read-back) against what the reference's sequence eWiseMult(eq) + nvals +
    reduce(land) decides: same pattern and `eq` on every common entry (so -0.0 == +0.0 and NaN != NaN)."""
    from graphblas_b200 import Vector

    n = 100003
    rng = np.random.default_rng(3)
    vals = rng.integers(-5, 5, size=n).astype(np.int64)
    pres = rng.random(n) < 0.4
    a = Vector.from_dense(vals, present=pres)
    assert a.isequal(a.dup())
    v2 = vals.copy(); k = int(np.flatnonzero(pres)[-1]); v2[k] += 1
    assert not a.isequal(Vector.from_dense(v2, present=pres))          # one value differs (last present entry)
    v3 = vals.copy(); k0 = int(np.flatnonzero(~pres)[0]); v3[k0] += 7
    assert a.isequal(Vector.from_dense(v3, present=pres))              # a value under an absent entry does not count
    p4 = pres.copy(); p4[k0] = True
    assert not a.isequal(Vector.from_dense(vals, present=p4))          # pattern differs, values equal
    assert not a.isequal(Vector.from_dense(vals[:-1], present=pres[:-1]))
    f = Vector.from_dense(np.array([0.0, 1.5, np.nan]), present=np.array([True, True, False]))
    g = Vector.from_dense(np.array([-0.0, 1.5, 3.0]), present=np.array([True, True, False]))
    assert f.isequal(g)
    h = Vector.from_dense(np.array([0.0, np.nan, 0.0]), present=np.array([True, True, False]))
    assert not h.isequal(h.dup())                                      # NaN != NaN under eq, as in the reference
    # different dtypes: the reference-shaped path (eWiseMult(eq) with type promotion)
    i32 = Vector.from_dense(vals.astype(np.int32), present=pres)
    assert a.isequal(i32) and not a.isequal(i32, check_dtype=True)
    e1, e2 = Vector("FP32", 10), Vector("FP32", 10)
    assert e1.isequal(e2)


# ---- the deferred element-wise pair of non-blocking mode (csrc/grb_internal.cuh PendingOp): `w << u.ewise_mult(v)` followed by a
# ---- plain fill `r[:] << x` runs as ONE kernel at the next entry point; results are what two separate statements give
@pytest.mark.parametrize("dt", ["FP32", "FP64", "INT32"])
@pytest.mark.parametrize("n", [1000, 4099])
def test_deferred_ewise_fill_pair(gb, dt, n):
    from graphblas_b200 import binary

    rng = np.random.default_rng(n)
    npdt = O.NP_TYPES[dt]
    a = (rng.random(n) * 9 + 1).astype(npdt)
    d = (rng.random(n) * 3 + 1).astype(npdt)
    op, ref = (binary.truediv, a / d) if dt.startswith("FP") else (binary.times, a * d)
    r, dv = gb.Vector.from_dense(a), gb.Vector.from_dense(d)
    w = gb.Vector.from_dense(np.zeros(n, dtype=npdt))
    l0 = gb.launch_count()
    w << r.ewise_mult(dv, op)
    r[:] << 7
    assert gb.launch_count() == l0                      # both wait for the next entry point ...
    wv, wp = w.to_dense()                               # ... which is this one
    assert gb.launch_count() - l0 >= 1
    rv, rp = r.to_dense()
    assert wp.all() and rp.all() and r.nvals == n
    np.testing.assert_allclose(wv, ref.astype(npdt), rtol=1e-6)
    assert (rv == 7).all()
    # one fused launch: check on a fresh pair with nothing else in between
    w << r.ewise_mult(dv, op)
    r[:] << 3
    l1 = gb.launch_count()
    gb.sync()
    assert gb.launch_count() - l1 == 1 and gb.last_kernel() == "ewise_fill_pair"
    # aliases: the product in place, then the fill over the same vector
    w << w.ewise_mult(dv, op)
    w[:] << 5
    assert (w.to_dense()[0] == 5).all()
    # a product alone is flushed by whatever comes next; a fill into a vector that was not full sets its presence
    w << r.ewise_mult(dv, op)
    np.testing.assert_allclose(w.to_dense()[0], (np.full(n, 3, dtype=npdt) / d if dt.startswith("FP") else np.full(n, 3, dtype=npdt) * d).astype(npdt), rtol=1e-6)
    e = gb.Vector(dt, n)
    w << r.ewise_mult(dv, op)
    e[:] << 2
    ev, ep = e.to_dense()
    assert ep.all() and (ev == 2).all() and e.nvals == n
    # an object named by deferred work may be freed before anything else happens: the free runs the work first
    tmp = gb.Vector.from_dense(np.zeros(n, dtype=npdt))
    tmp << r.ewise_mult(dv, op)
    del tmp
    import gc

    gc.collect()
    gb.sync()
    assert (r.to_dense()[0] == 3).all()
    # switched off: two launches, same results
    gb.set_option("lazy", 0)
    try:
        l2 = gb.launch_count()
        w << r.ewise_mult(dv, op)
        r[:] << 9
        assert gb.launch_count() - l2 == 2
        assert (r.to_dense()[0] == 9).all()
    finally:
        gb.set_option("lazy", 1)
