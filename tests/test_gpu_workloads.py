"""GPU: the five BASELINE workloads in the reference's notebook syntax (graphblas_b200/workloads.py) against
independent CPU implementations (scipy.sparse / scipy.sparse.csgraph), the cross-library checks SURVEY 8c calls for."""
import numpy as np
import pytest
import scipy.sparse as sp
from scipy.sparse import csgraph

from oracle import oracle as O

pytestmark = pytest.mark.gpu
SCALE = 13


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    return gb


def _scipy(A):
    return sp.csr_matrix((A.values, A.indices, A.indptr), shape=A.shape)


def test_level_bfs_vs_scipy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(SCALE, "BOOL", symmetrize=True)
    gA = gb.Matrix.rmat(SCALE, "BOOL", symmetrize=True)
    S = _scipy(A).astype(np.int8)
    deg = np.diff(A.indptr)
    for src in (int(np.argmax(deg)), int(np.nonzero(deg)[0][17])):
        dist = csgraph.shortest_path(S, unweighted=True, indices=src)
        v, st = W.bfs_levels(gA, src)
        vals, pres = v.to_dense()
        reach = np.isfinite(dist)
        assert np.array_equal(pres, reach)
        assert np.array_equal(vals[reach], (dist[reach] + 1).astype(np.int32))
        assert st["levels"] == int(dist[reach].max()) + 1
    # lor_land gives the same levels as any_pair (the notebook's original semiring)
    from graphblas_b200 import semiring
    v2, _ = W.bfs_levels(gA, src, semiring.lor_land)
    assert v2.isequal(v)


def test_sssp_vs_scipy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(SCALE, "INT64", symmetrize=True, value_mode=1)
    gA = gb.Matrix.rmat(SCALE, "INT64", symmetrize=True, value_mode=1)
    S = _scipy(A).astype(np.float64)
    src = int(np.argmax(np.diff(A.indptr)))
    dist = csgraph.dijkstra(S, indices=src)
    w, st = W.sssp_bellman_ford(gA, src)
    vals, pres = w.to_dense()
    reach = np.isfinite(dist)
    assert np.array_equal(pres, reach)
    assert np.array_equal(vals[reach], dist[reach].astype(np.int64))
    assert st["iterations"] < 64


def test_triangle_count_vs_scipy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(11, "INT64", symmetrize=True)
    gA = gb.Matrix.rmat(11, "INT64", symmetrize=True)
    S = _scipy(A).astype(np.int64)
    expect = int((S @ S @ S).diagonal().sum() // 6)
    ntri, st = W.triangle_count(gA)
    assert ntri == expect
    assert st["nnz_L"] == A.nnz // 2


def test_spgemm_vs_scipy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(12, "FP64", value_mode=2)
    gA = gb.Matrix.rmat(12, "FP64", value_mode=2)
    S = _scipy(A)
    C = (S @ S).tocsr()
    C.sort_indices()
    gC, st = W.spgemm(gA)
    ap, ai, ax = gC.to_csr()
    # scipy drops nothing here (no cancellation with positive values): identical pattern, values within 1e-5 relative
    assert st["nnz_C"] == C.nnz
    assert np.array_equal(ap.astype(np.int64), C.indptr.astype(np.int64)) and np.array_equal(ai.astype(np.int64), C.indices.astype(np.int64))
    assert np.allclose(ax, C.data, rtol=1e-5, atol=0)


def test_connected_components_vs_scipy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(SCALE, "BOOL", symmetrize=True)
    gA = gb.Matrix.rmat(SCALE, "BOOL", symmetrize=True)
    ncomp, labels = csgraph.connected_components(_scipy(A).astype(np.int8), directed=False)
    f, st = W.cc_min_label(gA)
    got, _ = f.to_dense()
    n = A.shape[0]
    # scipy's labels are arbitrary: compare as partitions through the minimum vertex id of each component
    mins = np.full(ncomp, n, dtype=np.int64)
    np.minimum.at(mins, labels, np.arange(n))
    assert np.array_equal(got.astype(np.int64), mins[labels])


def test_pagerank_vs_numpy(gb):
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(SCALE, "FP32", value_mode=2)
    gA = gb.Matrix.rmat(SCALE, "FP32", value_mode=2)
    n = A.shape[0]
    d = np.diff(A.indptr).astype(np.float32) / np.float32(0.85)
    d[d == 0] = 1
    r, _ = W.pagerank(gA, gb.Vector.from_dense(d), iters=10)
    P = _scipy(A).T.tocsr().astype(np.float64)
    x = np.full(n, 1.0 / n)
    for _ in range(10):
        x = (1 - 0.85) / n + P @ (x / d)
    got, _ = r.to_dense()
    assert np.allclose(got, x, rtol=2e-5, atol=1e-9)


def test_fastsv_vs_scipy_and_permutation(gb):
    """The notebook's own FastSV loop (cell 30) against scipy's components, and its permutation-invariance check
    (cells 54-55, 63): relabelling the vertices must give the same partition."""
    from graphblas_b200 import workloads as W

    A = O.rmat_matrix(SCALE, "BOOL", symmetrize=True)
    S = _scipy(A).astype(np.int8)
    n = A.shape[0]
    ncomp, labels = csgraph.connected_components(S, directed=False)
    ref = np.full(ncomp, n, dtype=np.int64)
    np.minimum.at(ref, labels, np.arange(n))
    gA = gb.Matrix.rmat(SCALE, "BOOL", symmetrize=True)
    f, st = W.fastsv(gA)
    vals, pres = f.to_dense()
    assert pres.all()
    assert np.array_equal(vals.astype(np.int64), ref[labels])
    assert st["iterations"] <= 12
    # permuted copy of the graph
    rng = np.random.default_rng(4)
    p = rng.permutation(n)
    coo = S.tocoo()
    gP = gb.Matrix.from_coo(p[coo.row], p[coo.col], np.ones(coo.nnz, dtype=bool), nrows=n, ncols=n, dup_op="lor")
    fp, _ = W.fastsv(gP)
    vp, _ = fp.to_dense()
    same = vals[coo.row] == vals[coo.col]
    assert same.all()
    # two vertices share a label before the permutation iff they share one after it
    a, b = rng.integers(0, n, size=20000), rng.integers(0, n, size=20000)
    assert np.array_equal(vals[a] == vals[b], vp[p[a]] == vp[p[b]])


# ---- BASELINE's full sizes through size-independent properties; oracle-vs-CUDA parity on the same graphs at scale 16-18
# (SpGEMM, whole CSR), 20 (triangle counting), 22 (Bellman-Ford) and 18 (BFS) is tests/test_gpu_baseline_parity.py ----
@pytest.mark.parametrize("scale,dt", [(18, "INT64"), (20, "FP64")])
def test_spgemm_full_size_row_sums(gb, scale, dt):
    """(A A) 1 = A (A 1): the row sums of C = A.mxm(A, plus_times) against two mxv's of the ones vector.  INT64 wraps
    like C, so the identity is bit-exact; FP64 values lie in (0, 1] (no cancellation), tolerance 1e-5 relative as
    north_star states.  Scale 20 FP64 is BASELINE configs[4] itself (9.7 G entries out)."""
    from graphblas_b200 import Vector, monoid, semiring

    A = gb.Matrix.rmat(scale, dt, value_mode=1 if dt == "INT64" else 2)
    n = A.nrows
    ones = Vector.from_dense(np.ones(n, dtype=np.int64 if dt == "INT64" else np.float64))
    t = A.mxv(ones, semiring.plus_times).new()
    # rows without entries give no entry in t; A (A 1) only reads t where A has columns, absent = 0 under plus
    expect = A.mxv(t, semiring.plus_times).new()
    C = A.mxm(A, semiring.plus_times).new()
    got = C.reduce_rowwise(monoid.plus).new()
    assert got.nvals == expect.nvals
    gv, gp = got.to_dense()
    ev, ep = expect.to_dense()
    assert np.array_equal(gp, ep)
    if dt == "INT64":
        assert np.array_equal(gv[gp.astype(bool)], ev[ep.astype(bool)])
    else:
        assert np.allclose(gv[gp.astype(bool)], ev[ep.astype(bool)], rtol=1e-5, atol=0)
    del C, got, expect


def test_triangle_count_scale20_two_algorithms(gb):
    """BASELINE configs[3] at scale 20: the hash-set membership path and the binary-search path of the masked dot
    product are independent implementations; they must agree entry for entry, and with trace-free counting
    sum(C) = sum over edges (i > j) of |N<(i) & N<(j)|."""
    from graphblas_b200 import Matrix, monoid, semiring

    A = gb.Matrix.rmat(20, "INT64", symmetrize=True)
    L = A.select("tril", -1).new()
    out = []
    for h in (1, 0):
        gb.set_option("dot_hash", h)
        try:
            C = Matrix("INT64", A.nrows, A.ncols)
            C(L.S) << L.mxm(L.T, semiring.plus_pair)
            out.append(C)
        finally:
            gb.set_option("dot_hash", 1)
    assert out[0].nvals == out[1].nvals
    assert out[0].isequal(out[1])
    # the count, pattern and values themselves are checked against the CPU oracle's masked dot product in
    # tests/test_gpu_baseline_parity.py::test_triangle_count_scale20_pattern_values_and_count; here the two device paths only
    # have to agree with each other (both alike wrong would be caught there)
    assert int(out[0].reduce_scalar(monoid.plus).value) == int(out[1].reduce_scalar(monoid.plus).value)


def test_fastsv_device_index_vectors_match_the_notebook_form(gb):
    """FastSV with I kept in HBM (GxB_b200_Vector_assign_Vector / _extract_Vector, I = the parent vector) gives the labels of the
    notebook's own statements (host index lists) and of scipy's connected_components, iteration count included."""
    from graphblas_b200 import workloads as W

    A = gb.Matrix.rmat(13, "BOOL", symmetrize=True)
    f_host, st_host = W.fastsv(A)
    f_dev, st_dev = W.fastsv(A, device_indices=True)
    assert st_host["iterations"] == st_dev["iterations"]
    assert f_host.isequal(f_dev)
    f_cc, _ = W.cc_min_label(A)
    assert f_cc.isequal(f_dev)


def test_vector_index_vector_assign_and_extract(gb):
    """w(min)[I] << u and w << u[I] with I a device vector == the host-index-list forms."""
    from graphblas_b200 import Vector, binary

    rng = np.random.default_rng(4)
    n, ni = 500, 2000
    idx = rng.integers(0, n, size=ni)
    u = Vector.from_dense(rng.integers(0, 1000, size=ni).astype(np.uint32))
    w1 = Vector.from_dense(rng.integers(500, 1000, size=n).astype(np.uint32))
    w2 = w1.dup()
    Iv = Vector.from_dense(idx.astype(np.uint32))
    w1(binary.min)[idx] << u
    w2(binary.min)[Iv] << u
    assert w1.isequal(w2)
    src = Vector.from_dense(rng.integers(0, 99, size=n).astype(np.int64))
    a = src[idx].new()
    b = Vector("INT64", ni)
    b << src[Iv]
    assert a.isequal(b)
    bad = Vector.from_dense(np.array([n + 5] * ni, dtype=np.uint32))
    with pytest.raises(gb.IndexOutOfBound):
        b << src[bad]
