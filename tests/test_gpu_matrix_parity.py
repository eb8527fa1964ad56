"""GPU: GrB_mxm (hash Gustavson + masked dot product), GrB_transpose, matrix eWise, select and the device R-MAT
generator through the C-ABI vs the CPU oracle on seeded inputs.  Bit-exact patterns always; bit-exact values for
integer / boolean semirings; 1e-5 relative for fp plus_times (north_star tolerance)."""
import numpy as np
import pytest

from gb_util import mat_equal, to_gpu_matrix
from oracle import oracle as O

pytestmark = pytest.mark.gpu
FP_TOL = 1e-5


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    return gb


def _sr(mon, mul):
    from graphblas_b200 import semiring

    return getattr(semiring, f"{mon.lower()}_{mul.lower()}")


def _rand_mat(rng, nrows, ncols, nnz, dtype):
    r = rng.integers(0, nrows, size=nnz)
    c = rng.integers(0, ncols, size=nnz)
    keys = np.unique(r * ncols + c)
    r, c = keys // ncols, keys % ncols
    if dtype.startswith("FP"):
        v = rng.random(r.size)
    elif dtype == "BOOL":
        v = rng.random(r.size) < 0.7
    else:
        v = rng.integers(0 if dtype.startswith("U") else -20, 50, size=r.size)
    return O.Mat.from_coo(r, c, v.astype(O.NP_TYPES[dtype]), nrows, ncols, dtype)


@pytest.mark.parametrize("scale,dt,vm,sym", [(9, "INT64", 1, False), (11, "FP64", 2, False), (10, "BOOL", 0, True),
                                             (12, "FP32", 2, True), (8, "UINT32", 1, False)])
def test_device_rmat_equals_oracle_rmat(gb, scale, dt, vm, sym):
    """The graph generated in HBM is bit-identical to the host oracle's (same integer hash arithmetic)."""
    A = O.rmat_matrix(scale, dt, symmetrize=sym, value_mode=vm)
    gA = gb.Matrix.rmat(scale, dt, symmetrize=sym, value_mode=vm)
    assert gA.nvals == A.nnz
    assert mat_equal(gA, A)


SEMIRINGS = [("PLUS", "TIMES", "INT64"), ("PLUS", "TIMES", "FP64"), ("PLUS", "TIMES", "FP32"), ("MIN", "PLUS", "INT64"),
             ("PLUS", "PAIR", "INT64"), ("ANY", "PAIR", "BOOL"), ("LOR", "LAND", "BOOL"), ("MIN", "FIRST", "UINT64"),
             ("MIN", "SECOND", "INT32"), ("MAX", "PLUS", "INT32"), ("PLUS", "SECOND", "FP32"), ("PLUS", "FIRST", "UINT32"),
             # narrow integer types: 32-bit kernels, wrapped (see test_gpu_vector_parity.CASES)
             ("PLUS", "TIMES", "INT8"), ("MIN", "PLUS", "UINT8"), ("MAX", "TIMES", "INT16"), ("PLUS", "PAIR", "UINT16")]


@pytest.mark.parametrize("mon,mul,dt", SEMIRINGS)
def test_spgemm_rmat_vs_oracle(gb, mon, mul, dt):
    """A.mxm(A): every bin of the hash Gustavson (warp / CTA tables / dense hub rows) on a skewed graph."""
    scale = 10
    vm = 2 if dt.startswith("FP") else (0 if dt == "BOOL" else 1)
    A = O.rmat_matrix(scale, dt, value_mode=vm)
    gA = to_gpu_matrix(gb, A)
    tol = FP_TOL if dt.startswith("FP") else None
    for t0, t1 in ((False, False), (True, False), (False, True)):
        Ae, Be = (A.transpose() if t0 else A), (A.transpose() if t1 else A)
        T = O.spgemm(Ae, Be, mon, mul, dt)
        gC = ((gA.T if t0 else gA).mxm(gA.T if t1 else gA, _sr(mon, mul))).new()
        assert gC.nvals == T.nnz
        assert mat_equal(gC, T, tol), (t0, t1, gb.last_kernel())


def test_spgemm_hub_rows_dense_path(gb):
    """Rows whose product count exceeds every shared-memory table go through the bitmap + dense accumulator."""
    rng = np.random.default_rng(5)
    n = 20000
    A = _rand_mat(rng, 64, n, 64 * 600, "INT64")      # 64 rows x ~600 entries
    B = _rand_mat(rng, n, n, n * 40, "INT64")         # ~24000 products per row of A
    T = O.spgemm(A, B, "PLUS", "TIMES", "INT64")
    gC = to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, B), _sr("PLUS", "TIMES")).new()
    assert mat_equal(gC, T)
    T = O.spgemm(A, B, "MIN", "PLUS", "INT64")
    gC = to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, B), _sr("MIN", "PLUS")).new()
    assert mat_equal(gC, T)


@pytest.mark.parametrize("panels", [1, 2, 0])
def test_spgemm_heavy_rows_hash_ranges(gb, panels):
    """Heavy rows whose columns are cut into several shared-memory hash ranges or dense base panels (25 base panels
    of 4 K columns), next to the variant that keeps the values in an L2-resident buffer (2) and the HBM dense-accumulator fallback
    (spgemm_panels = 0): same bits every way."""
    rng = np.random.default_rng(7)
    n = 100000
    B = _rand_mat(rng, n, n, n * 100, "FP64")
    gB = to_gpu_matrix(gb, B)
    gb.set_option("spgemm_panels", panels)
    try:
        # ~11 000 products per row (sparse ranges: hash tables over several base panels) and ~30 000 (dense base panels)
        for per_row in (110, 300):
            A = _rand_mat(rng, 48, n, 48 * per_row, "FP64")
            gA = to_gpu_matrix(gb, A)
            T = O.spgemm(A, B, "PLUS", "TIMES", "FP64")
            gC = gA.mxm(gB, _sr("PLUS", "TIMES")).new()
            assert mat_equal(gC, T, FP_TOL), (per_row, gb.last_kernel())
            T = O.spgemm(A, B, "MIN", "SECOND", "FP64")
            gC = gA.mxm(gB, _sr("MIN", "SECOND")).new()
            assert mat_equal(gC, T), (per_row, gb.last_kernel())
    finally:
        gb.set_option("spgemm_panels", 1)


def test_spgemm_clustered_columns_ordered_emission(gb):
    """Output rows whose columns sit in three narrow clusters (at 0, 100 K and 1.4 M): the order-preserving buckets of the
    sort-free emission then hold hundreds of entries each (ranked by counting inside the bucket) -- rows must still come
    out sorted and exact, in the hash-row kernel (~7 K entries per row) and in the heavy-row kernel (~10 K entries, the
    first two clusters sharing one hash range)."""
    rng = np.random.default_rng(13)
    n = 1_500_000
    for a_per_row, width in ((40, 4000), (80, 6000)):
        m = 6
        ar = np.repeat(np.arange(m), a_per_row)
        ac = np.concatenate([rng.choice(2000, size=a_per_row, replace=False) for _ in range(m)])   # among the first 2000 rows of B
        A = O.Mat.from_coo(ar, ac, rng.integers(1, 9, size=ar.size).astype(np.int64), m, n, "INT64")
        br = np.repeat(np.arange(2000), 200)
        lo = np.array([0, 100_000, 1_400_000])[rng.integers(0, 3, size=br.size)]
        bc = lo + rng.integers(0, width, size=br.size)
        keys = np.unique(br.astype(np.int64) * n + bc)
        br, bc = keys // n, keys % n
        B = O.Mat.from_coo(br, bc, rng.integers(1, 9, size=br.size).astype(np.int64), n, n, "INT64")
        T = O.spgemm(A, B, "PLUS", "TIMES", "INT64")
        gC = to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, B), _sr("PLUS", "TIMES")).new()
        assert gC.nvals == T.nnz
        assert mat_equal(gC, T), (a_per_row, gb.last_kernel())


def test_spgemm_rmat_scale14_fp64(gb):
    """BASELINE config 5a in small: A.mxm(A, plus_times) FP64 on a directed R-MAT graph, where > 80 % of the
    products sit in heavy rows (hash ranges and dense base panels); pattern exact, values within 1e-5."""
    A = O.rmat_matrix(14, "FP64", value_mode=2)
    gA = to_gpu_matrix(gb, A)
    T = O.spgemm(A, A, "PLUS", "TIMES", "FP64")
    gC = gA.mxm(gA, _sr("PLUS", "TIMES")).new()
    assert gC.nvals == T.nnz
    assert mat_equal(gC, T, FP_TOL), gb.last_kernel()
    # integer semiring on the same pattern: bit-exact
    Ai = O.rmat_matrix(14, "INT64", value_mode=1)
    Ti = O.spgemm(Ai, Ai, "PLUS", "TIMES", "INT64")
    gCi = to_gpu_matrix(gb, Ai).mxm(to_gpu_matrix(gb, Ai), _sr("PLUS", "TIMES")).new()
    assert mat_equal(gCi, Ti), gb.last_kernel()


def test_spgemm_nonsquare_and_empty(gb):
    rng = np.random.default_rng(6)
    A = _rand_mat(rng, 37, 91, 400, "FP64")
    B = _rand_mat(rng, 91, 13, 300, "FP64")
    T = O.spgemm(A, B, "PLUS", "TIMES", "FP64")
    gC = to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, B), _sr("PLUS", "TIMES")).new()
    assert mat_equal(gC, T, FP_TOL)
    E = gb.Matrix("FP64", 91, 13)
    gC = to_gpu_matrix(gb, A).mxm(E, _sr("PLUS", "TIMES")).new()
    assert gC.nvals == 0 and gC.shape == (37, 13)
    with pytest.raises(gb.DimensionMismatch):
        to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, A))


MASKS = [dict(structure=True, comp=False), dict(structure=False, comp=False), dict(structure=True, comp=True),
         dict(structure=False, comp=True)]


@pytest.mark.parametrize("mk", MASKS, ids=lambda m: f"{'S' if m['structure'] else 'V'}{'c' if m['comp'] else ''}")
@pytest.mark.parametrize("accum", [None, "PLUS", "MIN"])
@pytest.mark.parametrize("replace", [False, True])
def test_mxm_mask_accum_replace_vs_oracle(gb, mk, accum, replace):
    """C<M,replace> = accum(C, A (+).(x) B) for every mask kind; covers the dot path (non-complemented masks),
    the Gustavson path (complemented) and the three-way merge write-back."""
    from graphblas_b200 import binary

    rng = np.random.default_rng(11)
    n = 300
    A = _rand_mat(rng, n, n, 3000, "INT64")
    B = _rand_mat(rng, n, n, 3000, "INT64")
    C0 = _rand_mat(rng, n, n, 2500, "INT64")
    M = _rand_mat(rng, n, n, 4000, "INT32")     # has explicit zeros: value masks differ from structural ones
    for mon, mul in (("PLUS", "TIMES"), ("MIN", "PLUS")):
        C = O.mxm(C0, A, B, mon, mul, "INT64", mask=M, accum=accum, replace=replace, **mk)
        gC, gM = to_gpu_matrix(gb, C0), to_gpu_matrix(gb, M)
        m = gM.S if mk["structure"] else gM.V
        m = ~m if mk["comp"] else m
        kw = dict(mask=m, replace=replace)
        if accum:
            kw["accum"] = getattr(binary, accum.lower())
        gC(**kw) << to_gpu_matrix(gb, A).mxm(to_gpu_matrix(gb, B), _sr(mon, mul))
        assert mat_equal(gC, C), (mon, mul, gb.last_kernel())


def test_mxm_axb_method_hints_agree(gb):
    """reference tests/test_matrix.py:4277-4319: axb_method hints change the kernel, never the result."""
    rng = np.random.default_rng(12)
    n = 200
    A = _rand_mat(rng, n, n, 2000, "INT64")
    M = _rand_mat(rng, n, n, 3000, "BOOL")
    gA, gM = to_gpu_matrix(gb, A), to_gpu_matrix(gb, M)
    ref = gA.mxm(gA.T, _sr("PLUS", "TIMES")).new(mask=gM.S)
    for method in ("dot", "gustavson", "hash", "saxpy"):
        C = gA.mxm(gA.T, _sr("PLUS", "TIMES")).new(mask=gM.S, axb_method=method)
        assert C.isequal(ref), method


def test_triangle_count_masked_dot(gb):
    """C(L.S) << L.mxm(L.T, plus_pair) (BASELINE config 4) against the oracle and a dense count."""
    scale = 9
    A = O.rmat_matrix(scale, "INT64", symmetrize=True, value_mode=0)
    r, c, _ = A.to_coo()
    keep = c < r
    n = 1 << scale
    L = O.Mat.from_coo(r[keep], c[keep], np.ones(keep.sum(), dtype=np.int64), n, n, "INT64")
    gA = gb.Matrix.rmat(scale, "INT64", symmetrize=True)
    gL = gA.select("tril", -1).new()
    assert mat_equal(gL, L, structure_only=True)
    C = O.mxm(O.Mat.from_coo([], [], np.zeros(0, dtype=np.int64), n, n, "INT64"), L, L, "PLUS", "PAIR", "INT64",
              mask=L, structure=True, tran1=True, method="dot")
    gC = gb.Matrix("INT64", n, n)
    gC(gL.S) << gL.mxm(gL.T, _sr("PLUS", "PAIR"))
    assert mat_equal(gC, C)
    dense = np.zeros((n, n), dtype=np.int64)
    dense[r, c] = 1
    assert int(gC.reduce_scalar().value) == int(np.trace(dense @ dense @ dense) // 6)


@pytest.mark.parametrize("dot_hash", [1, 0])
def test_masked_dot_pair_row_hash_sets(gb, dot_hash):
    """Masked dot product with a pair multiply on two DIFFERENT operands with long rows: membership through per-row hash
    sets (dot_hash = 1) and through binary searches (0) give the oracle's bits, for plus_pair and any_pair."""
    rng = np.random.default_rng(11)
    m, k, n = 300, 5000, 280
    A = _rand_mat(rng, m, k, m * 200, "INT64")         # rows of ~200 entries (> 32: hashed)
    B = _rand_mat(rng, k, n, k * 12, "INT64")          # columns of ~210 entries
    M = _rand_mat(rng, m, n, 20000, "INT64")
    gA, gB, gM = to_gpu_matrix(gb, A), to_gpu_matrix(gb, B), to_gpu_matrix(gb, M)
    empty = O.Mat.from_coo([], [], np.zeros(0, dtype=np.int64), m, n, "INT64")
    gb.set_option("dot_hash", dot_hash)
    try:
        for mon in ("PLUS", "ANY"):
            ref = O.mxm(empty, A, B, mon, "PAIR", "INT64", mask=M, structure=True, method="dot")
            gC = gb.Matrix("INT64", m, n)
            gC(gM.S) << gA.mxm(gB, _sr(mon, "PAIR"))
            assert mat_equal(gC, ref), (mon, gb.last_kernel())
    finally:
        gb.set_option("dot_hash", 1)


def test_transpose_ewise_select_vs_oracle(gb):
    from graphblas_b200 import binary

    rng = np.random.default_rng(13)
    A = _rand_mat(rng, 120, 80, 1500, "INT64")
    B = _rand_mat(rng, 120, 80, 1500, "INT64")
    gA, gB = to_gpu_matrix(gb, A), to_gpu_matrix(gb, B)
    assert mat_equal(gA.T.new(), A.transpose())
    ak, bk = A.keys(), B.keys()
    av, bv = A.to_coo()[2], B.to_coo()[2]
    both = np.intersect1d(ak, bk)
    gm = gA.ewise_mult(gB, binary.times).new()
    r, c, v = gm.to_coo()
    assert np.array_equal(r.astype(np.int64) * 80 + c.astype(np.int64), both)
    assert np.array_equal(v, av[np.searchsorted(ak, both)] * bv[np.searchsorted(bk, both)])
    ga = gA.ewise_add(gB, binary.plus).new()
    r, c, v = ga.to_coo()
    uni = np.union1d(ak, bk)
    assert np.array_equal(r.astype(np.int64) * 80 + c.astype(np.int64), uni)
    exp = np.zeros(uni.size, dtype=np.int64)
    exp[np.searchsorted(uni, ak)] += av
    exp[np.searchsorted(uni, bk)] += bv
    assert np.array_equal(v, exp)
    for which, thunk, pred in (("tril", 0, lambda d: d <= 0), ("triu", 1, lambda d: d >= 1), ("diag", 0, lambda d: d == 0),
                               ("offdiag", 0, lambda d: d != 0)):
        gs = gA.select(which, thunk).new()
        ar, ac, axv = A.to_coo()
        k = pred(ac - ar)
        r, c, v = gs.to_coo()
        assert np.array_equal(r.astype(np.int64), ar[k]) and np.array_equal(c.astype(np.int64), ac[k])
        assert np.array_equal(v, axv[k])


def test_rowblock_shards_reassemble(gb):
    A = O.rmat_matrix(9, "INT64", value_mode=1)
    gA = to_gpu_matrix(gb, A)
    n = 512
    cuts = [0, 100, 101, 400, n]
    rows, cols, vals = [], [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        blk = gA.rowblock(a, b)
        assert blk.shape == (b - a, n)
        r, c, v = blk.to_coo()
        rows.append(r.astype(np.int64) + a), cols.append(c.astype(np.int64)), vals.append(v)
    ar, ac, av = A.to_coo()
    assert np.array_equal(np.concatenate(rows), ar) and np.array_equal(np.concatenate(cols), ac)
    assert np.array_equal(np.concatenate(vals), av)


@pytest.mark.parametrize("fmt", ["csr", "csc", "coo", "lil"])
def test_from_scipy_sparse_as_in_reference(gb, fmt):
    """io.from_scipy_sparse / to_scipy_sparse (reference graphblas/io/_scipy.py:8-119): every scipy format lands as the same
    matrix, unsorted column indices included, duplicates of a COO input folded by dup_op; the product of the ingested matrix
    equals scipy's."""
    import scipy.sparse as sp
    from graphblas_b200 import binary, io, semiring

    rng = np.random.default_rng(17)
    S = sp.random(300, 200, density=0.05, random_state=3, format="csr", dtype=np.float64)
    S.data = np.round(S.data * 100) + 1
    A = io.from_scipy_sparse(S.asformat(fmt))
    assert A.nvals == S.nnz and (A.nrows, A.ncols) == S.shape
    back = io.to_scipy_sparse(A, "csr")
    assert (back != S).nnz == 0
    x = rng.integers(0, 5, size=200).astype(np.float64)
    w = A.mxv(gb.Vector.from_dense(x), semiring.plus_times).new()
    vals, pres = w.to_dense()
    ref = S @ x
    has = np.diff(S.indptr) > 0
    assert np.array_equal(pres.astype(bool), has) and np.array_equal(vals[has], ref[has])
    if fmt == "coo":                      # duplicates: summed by dup_op, as the reference's from_coo does
        C = S.tocoo()
        D = sp.coo_array((np.concatenate([C.data, C.data]), (np.concatenate([C.row, C.row]), np.concatenate([C.col, C.col]))), shape=S.shape)
        A2 = io.from_scipy_sparse(D, dup_op=binary.plus)
        assert (io.to_scipy_sparse(A2) != 2 * S).nnz == 0
    if fmt == "csr":                      # unsorted indices inside rows
        U = S.copy()
        for i in range(U.shape[0]):
            s_, e_ = U.indptr[i], U.indptr[i + 1]
            o = rng.permutation(e_ - s_)
            U.indices[s_:e_], U.data[s_:e_] = U.indices[s_:e_][o], U.data[s_:e_][o]
        U.has_sorted_indices = False
        assert (io.to_scipy_sparse(io.from_scipy_sparse(U)) != S).nnz == 0
