"""GPU: GrB_mxv / GrB_vxm through the C-ABI vs the CPU oracle on seeded inputs — every kernel path
(streaming pull, early-exit pull, push), masks, accumulators, replace, transposes, types.
Bit-exact for integer / boolean semirings; 1e-5 relative for fp plus_times (north_star tolerance)."""
import numpy as np
import pytest

from gb_util import to_gpu_matrix, to_gpu_vector, vec_equal
from oracle import oracle as O

pytestmark = pytest.mark.gpu
FP_TOL = 1e-5


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    yield gb
    gb.set_option("force_direction", 0)


def _rand_vec(rng, n, dtype, density):
    p = (rng.random(n) < density).astype(np.uint8)
    if dtype.startswith("FP"):
        v = rng.random(n).astype(O.NP_TYPES[dtype])
    elif dtype == "BOOL":
        v = rng.random(n) < 0.7
    else:
        v = rng.integers(0 if dtype.startswith("U") else -50, 100, size=n).astype(O.NP_TYPES[dtype])
    return O.Vec(v, p, dtype)


def _sr(gb, mon, mul):
    from graphblas_b200 import semiring

    return getattr(semiring, f"{mon.lower()}_{mul.lower()}")


CASES = [
    ("PLUS", "TIMES", "INT64"), ("PLUS", "TIMES", "FP32"), ("PLUS", "TIMES", "FP64"), ("PLUS", "SECOND", "FP32"),
    ("PLUS", "PAIR", "INT64"), ("PLUS", "FIRST", "INT32"), ("MIN", "PLUS", "INT64"), ("MIN", "PLUS", "FP64"),
    ("MIN", "FIRST", "UINT64"), ("MIN", "SECOND", "UINT32"), ("MAX", "PLUS", "INT32"), ("MAX", "MIN", "FP32"),
    ("ANY", "PAIR", "BOOL"), ("ANY", "PAIR", "INT32"), ("LOR", "LAND", "BOOL"), ("LXOR", "LAND", "BOOL"),
    ("MIN", "TIMES", "UINT32"), ("PLUS", "MIN", "INT64"),
    # INT8 / INT16 / UINT8 / UINT16 (reference graphblas/core/operator/semiring.py:170-204): computed in 32 bits and wrapped;
    # sums overflow the narrow type on purpose (255-valued weights), min_plus / max_times wrap inside the multiply
    ("PLUS", "TIMES", "INT8"), ("PLUS", "TIMES", "UINT16"), ("MIN", "PLUS", "INT8"), ("MIN", "PLUS", "UINT16"),
    ("MAX", "TIMES", "INT16"), ("MIN", "TIMES", "UINT8"), ("ANY", "PAIR", "UINT8"), ("MIN", "SECOND", "INT16"),
    ("PLUS", "PAIR", "INT8"),
]


@pytest.mark.parametrize("mon,mul,dt", CASES)
@pytest.mark.parametrize("direction", [1, 2])   # 1 = forced pull, 2 = forced push
def test_products_vs_oracle_rmat(gb, mon, mul, dt, direction):
    rng = np.random.default_rng(hash((mon, mul, dt)) % 2**31)
    scale = 11
    n = 1 << scale
    vm = 2 if dt.startswith("FP") else (0 if dt == "BOOL" else 1)
    A = O.rmat_matrix(scale, dt, value_mode=vm)
    gA = to_gpu_matrix(gb, A)
    At = A.transpose()
    tol = FP_TOL if (dt.startswith("FP") and mon == "PLUS") else None
    gb.set_option("force_direction", direction)
    sr = _sr(gb, mon, mul)
    for density in (1.0, 0.3, 0.004):
        u = _rand_vec(rng, n, dt, density)
        gu = to_gpu_vector(gb, u)
        # mxv, vxm, and their transposed forms
        for kind, tran in (("mxv", False), ("mxv", True), ("vxm", False), ("vxm", True)):
            w = O.Vec.empty(n, dt)
            if kind == "mxv":
                O.mxv(w, A, u, mon, mul, dt, tran0=tran, At=At)
                gw = ((gA.T if tran else gA).mxv(gu, sr)).new()
            else:
                O.vxm(w, u, A, mon, mul, dt, tran1=tran, At=At)
                gw = (gu.vxm(gA.T if tran else gA, sr)).new()
            if mon == "ANY" and mul != "PAIR":
                assert np.array_equal(gw.to_dense()[1], w.present.astype(bool))
            else:
                assert vec_equal(gw, w, tol), (kind, tran, density, gb.last_kernel())
            assert gw.nvals == w.nvals


@pytest.mark.parametrize("direction", [1, 2])
@pytest.mark.parametrize("accum", [None, "PLUS", "MIN"])
@pytest.mark.parametrize("replace", [False, True])
@pytest.mark.parametrize("comp", [False, True])
@pytest.mark.parametrize("structure", [False, True])
def test_writeback_semantics_vs_oracle(gb, direction, accum, replace, comp, structure):
    """C API 2.0 write-back: every mask kind x accum x replace, fused epilogue (pull) and dense write-back (push)."""
    from graphblas_b200 import binary, semiring

    rng = np.random.default_rng(7)
    n = 1500
    dt = "INT64"
    A = O.rmat_matrix(11, dt, value_mode=1)
    A = O.Mat(A.indptr[: n + 1], A.indices[: A.indptr[n]] % n, A.values[: A.indptr[n]], (n, n), dt)
    A = O.Mat.from_coo(*_dedup(A), n, n, dt)
    gA = to_gpu_matrix(gb, A)
    gb.set_option("force_direction", direction)
    u = _rand_vec(rng, n, dt, 0.2)
    w0 = _rand_vec(rng, n, dt, 0.5)
    m = _rand_vec(rng, n, "INT32", 0.5)
    m.values[rng.random(n) < 0.5] = 0          # stored zeros: false under a value mask, true structurally
    w = w0.dup()
    O.mxv(w, A, u, "MIN", "PLUS", dt, mask=m, accum=accum, replace=replace, comp=comp, structure=structure)
    gw, gm = to_gpu_vector(gb, w0), to_gpu_vector(gb, m)
    mask = gm.S if structure else gm.V
    mask = ~mask if comp else mask
    kw = dict(mask=mask, replace=replace)
    if accum:
        kw["accum"] = getattr(binary, accum.lower())
    gw(**kw) << gA.mxv(to_gpu_vector(gb, u), semiring.min_plus)
    assert vec_equal(gw, w)
    assert gw.nvals == w.nvals


def _dedup(A):
    r, c, v = A.to_coo()
    keys = r * A.shape[1] + c
    _, first = np.unique(keys, return_index=True)
    return r[first], c[first], v[first]


def test_aliased_output_pingpong(gb):
    """w(min) << w.vxm(A, min_plus): the output aliases the input; Bellman-Ford to a fixed point vs scipy."""
    import scipy.sparse as sp
    from scipy.sparse.csgraph import bellman_ford

    from graphblas_b200 import binary, semiring

    scale = 10
    n = 1 << scale
    A = O.rmat_matrix(scale, "INT64", value_mode=1)
    gA = to_gpu_matrix(gb, A)
    gb.set_option("force_direction", 0)
    src = int(np.argmax(np.diff(A.indptr)))
    gw = gb.Vector.from_coo([src], [0], dtype="INT64", size=n)
    for _ in range(n):
        before = gw.dup()
        gw(binary.min) << gw.vxm(gA, semiring.min_plus)
        if gw.isequal(before):
            break
    dist = bellman_ford(sp.csr_matrix((A.values.astype(np.float64), A.indices, A.indptr), shape=A.shape),
                        indices=src, directed=True)
    vals, pres = gw.to_dense()
    assert np.array_equal(pres, np.isfinite(dist))
    assert np.array_equal(vals[pres], dist[pres].astype(np.int64))


def test_level_bfs_direction_optimising(gb):
    """Level BFS (reference notebooks/Example B.1 cell 8): q(~v.S, replace) << q.vxm(A, any_pair); levels
    must match scipy's BFS order distances, with the push / early-exit pull switch left on auto."""
    import scipy.sparse as sp
    from scipy.sparse.csgraph import shortest_path

    from graphblas_b200 import semiring

    scale = 12
    n = 1 << scale
    A = O.rmat_matrix(scale, "BOOL", symmetrize=True)
    gA = to_gpu_matrix(gb, A)
    gb.set_option("force_direction", 0)
    src = int(np.argmax(np.diff(A.indptr)))
    q = gb.Vector.from_coo([src], [True], dtype="BOOL", size=n)
    v = gb.Vector("INT32", n)
    level, kernels = 0, set()
    while q.nvals:
        v(q.S)[:] = level
        q(~v.S, replace=True) << q.vxm(gA, semiring.any_pair)
        kernels.add(gb.last_kernel())
        level += 1
    d = shortest_path(sp.csr_matrix((np.ones(A.nnz), A.indices, A.indptr), shape=A.shape), method="D",
                      unweighted=True, indices=src)
    vals, pres = v.to_dense()
    assert np.array_equal(pres, np.isfinite(d))
    assert np.array_equal(vals[pres], d[pres].astype(np.int32))


def test_type_mismatch_is_cast_like_c(gb):
    """Operands of other types are typecast to the semiring's type; output of another type is cast on write-back."""
    from graphblas_b200 import semiring

    rng = np.random.default_rng(3)
    n = 500
    A = O.rmat_matrix(9, "INT32", value_mode=1)
    A = O.Mat(A.indptr[: n + 1], A.indices[: A.indptr[n]], A.values[: A.indptr[n]], (n, 512), "INT32")
    u = _rand_vec(rng, 512, "FP64", 0.5)
    gw = gb.Vector("FP32", n)
    gw << to_gpu_matrix(gb, A).mxv(to_gpu_vector(gb, u), semiring.plus_times["FP64"])
    w = O.Vec.empty(n, "FP32")
    t = O.spmv(A, u, "PLUS", "TIMES", "FP64")
    O.vec_writeback(w, t)
    assert vec_equal(gw, w, 1e-6)


# ---------------------------------------------------------------------------------------------- flagged-stream pull
def _skewed_matrix(rng, n, dt):
    """Rows of very different lengths: empty rows, length-1 rows, and hub rows spanning many 256-entry tiles."""
    lens = rng.choice([0, 0, 1, 2, 5, 16, 40, 300, 2000], size=n, p=[.1, .1, .2, .2, .2, .1, .05, .04, .01])
    lens = np.minimum(lens, n)
    rows = np.repeat(np.arange(n), lens)
    cols = np.concatenate([rng.choice(n, size=k, replace=False) for k in lens if k]) if rows.size else np.zeros(0, int)
    if dt.startswith("FP"):
        v = rng.random(rows.size)
    elif dt == "BOOL":
        v = np.ones(rows.size, dtype=bool)
    else:
        v = rng.integers(1, 50, size=rows.size)
    return O.Mat.from_coo(rows, cols, v.astype(O.NP_TYPES[dt]), n, n, dt)


@pytest.mark.parametrize("mon,mul,dt", [("PLUS", "TIMES", "FP32"), ("PLUS", "TIMES", "INT64"), ("MIN", "PLUS", "INT64"),
                                        ("MIN", "SECOND", "UINT32"), ("PLUS", "PAIR", "INT64"), ("LOR", "LAND", "BOOL"),
                                        ("MAX", "PLUS", "FP64"), ("PLUS", "SECOND", "FP32")])
@pytest.mark.parametrize("hub_mode", [0, 2, 3])
def test_stream_pull_full_vector(gb, mon, mul, dt, hub_mode):
    """u full -> k_spmv_stream (hub_mode 2 builds the stream form at once) vs the staged kernel (hub_mode 0) vs oracle;
    masks, accumulators and replace through the fused epilogue; hub rows crossing tiles; empty rows."""
    from graphblas_b200 import binary

    rng = np.random.default_rng(21)
    n = 3000
    A = _skewed_matrix(rng, n, dt)
    gA = to_gpu_matrix(gb, A)
    At = A.transpose()
    tol = FP_TOL if dt.startswith("FP") and mon == "PLUS" else None
    gb.set_option("force_direction", 1)
    gb.set_option("hub_mode", hub_mode)
    try:
        u = _rand_vec(rng, n, dt, 1.0)
        gu = to_gpu_vector(gb, u)
        sr = _sr(gb, mon, mul)
        for tran in (False, True):
            w = O.Vec.empty(n, dt)
            O.mxv(w, A, u, mon, mul, dt, tran0=tran, At=At)
            gw = ((gA.T if tran else gA).mxv(gu, sr)).new()
            assert vec_equal(gw, w, tol), (tran, gb.last_kernel())
            assert gw.nvals == w.nvals
        m = _rand_vec(rng, n, "INT32", 0.5)
        m.values[rng.random(n) < 0.5] = 0
        w0 = _rand_vec(rng, n, dt, 0.6)
        acc_name = None if dt == "BOOL" else ("MIN" if mon == "MIN" else "PLUS")
        # no mask, accumulator == monoid (atomic-segment kernel) and != monoid (flagged-stream kernel); w partly / fully present
        for accum in ([mon] if mon in ("PLUS", "MIN", "MAX") and dt != "BOOL" else []) + ([acc_name] if acc_name else []):
            for dens in (0.6, 1.0):
                wa = _rand_vec(rng, n, dt, dens)
                w = wa.dup()
                O.mxv(w, A, u, mon, mul, dt, accum=accum)
                gw = to_gpu_vector(gb, wa)
                gw(getattr(binary, accum.lower())) << gA.mxv(gu, sr)
                assert vec_equal(gw, w, tol), (accum, dens, gb.last_kernel())
                assert gw.nvals == w.nvals
        for comp, structure, replace, accum in ((False, False, False, None), (True, True, True, acc_name),
                                                (False, True, False, acc_name), (True, False, True, None)):
            w = w0.dup()
            O.mxv(w, A, u, mon, mul, dt, mask=m, accum=accum, replace=replace, comp=comp, structure=structure)
            gw, gm = to_gpu_vector(gb, w0), to_gpu_vector(gb, m)
            mask = gm.S if structure else gm.V
            mask = ~mask if comp else mask
            kw = dict(mask=mask, replace=replace)
            if accum:
                kw["accum"] = getattr(binary, accum.lower())
            gw(**kw) << gA.mxv(gu, sr)
            assert vec_equal(gw, w, tol), (comp, structure, replace, accum, gb.last_kernel())
            assert gw.nvals == w.nvals
        # w aliasing u: w(min) << w.vxm(A) ping-pong through the stream kernel
        if dt == "INT64" and mon == "MIN":
            w = u.dup()
            O.vxm(w, w.dup(), A, mon, mul, dt, accum="MIN", At=At)
            gw = to_gpu_vector(gb, u)
            gw(binary.min) << gw.vxm(gA, sr)
            assert vec_equal(gw, w)
    finally:
        gb.set_option("hub_mode", 1)
        gb.set_option("force_direction", 0)


def test_stream_pull_tile_edges(gb):
    """Row boundaries on and around the 256-entry tile borders and the 32-entry look-ahead window."""
    from graphblas_b200 import semiring

    rng = np.random.default_rng(22)
    gb.set_option("hub_mode", 2)
    gb.set_option("force_direction", 1)
    try:
        for lens in ([256, 256, 1], [255, 1, 256, 31, 32, 33, 1], [200, 56 + 31, 300], [200, 56 + 32, 5], [200, 56 + 33, 5],
                     [1] * 300, [700, 1, 1, 254, 290], [256 * 3], [3, 0, 0, 253, 0, 32], [250, 6 + 256 + 10, 20]):
            n = max(max(lens), len(lens)) + 5
            rows = np.repeat(np.arange(len(lens)), lens)
            cols = np.concatenate([np.sort(rng.choice(n, size=k, replace=False)) for k in lens if k])
            vals = rng.integers(1, 9, size=rows.size).astype(np.int64)
            A = O.Mat.from_coo(rows, cols, vals, n, n, "INT64")
            u = O.Vec(rng.integers(1, 9, size=n).astype(np.int64), np.ones(n, dtype=np.uint8), "INT64")
            w = O.Vec.empty(n, "INT64")
            O.mxv(w, A, u, "PLUS", "TIMES", "INT64")
            gw = to_gpu_matrix(gb, A).mxv(to_gpu_vector(gb, u), semiring.plus_times).new()
            assert vec_equal(gw, w), lens
    finally:
        gb.set_option("hub_mode", 1)
        gb.set_option("force_direction", 0)


@pytest.mark.parametrize("mon,mul,dt,vm", [("PLUS", "TIMES", "INT64", 1), ("PLUS", "TIMES", "FP32", 2), ("MIN", "PLUS", "INT64", 1),
                                           ("MIN", "SECOND", "UINT32", 1)])
@pytest.mark.parametrize("hub_mode", [2, 3])
def test_stream_pull_hubs_and_non_hubs(gb, mon, mul, dt, vm, hub_mode):
    """A graph with more columns than the shared-memory hub table holds (scale 17: 131072 columns), so both the hub
    and the global-gather lanes of k_spmv_red / k_spmv_stream run; three chained products like the PageRank loop."""
    from graphblas_b200 import binary

    scale = 17
    n = 1 << scale
    A = O.rmat_matrix(scale, dt, value_mode=vm)
    gA = gb.Matrix.rmat(scale, dt, value_mode=vm)
    At = A.transpose()
    rng = np.random.default_rng(33)
    tol = FP_TOL if dt.startswith("FP") else None
    gb.set_option("force_direction", 1)
    gb.set_option("hub_mode", hub_mode)
    try:
        u = _rand_vec(rng, n, dt, 1.0)
        if dt.startswith("FP"):
            u.values[:] = u.values / n
        gu = to_gpu_vector(gb, u)
        sr = _sr(gb, mon, mul)
        accum = mon
        # overwrite, then accumulate (w full), with A and with A' (cached columns); masked variant through k_spmv_stream
        w = O.Vec.empty(n, dt)
        O.mxv(w, A, u, mon, mul, dt, tran0=True, At=At)
        gw = gA.T.mxv(gu, sr).new()
        assert vec_equal(gw, w, tol) and gw.nvals == w.nvals
        for _ in range(2):
            O.mxv(w, A, u, mon, mul, dt, accum=accum)
            gw(getattr(binary, accum.lower())) << gA.mxv(gu, sr)
            assert vec_equal(gw, w, tol), gb.last_kernel()
        m = _rand_vec(rng, n, "BOOL", 0.5)
        gm = to_gpu_vector(gb, m)
        O.mxv(w, A, u, mon, mul, dt, mask=m, accum=accum, structure=True, tran0=True, At=At)
        gw(gm.S, getattr(binary, accum.lower())) << gA.T.mxv(gu, sr)
        assert vec_equal(gw, w, tol), gb.last_kernel()
        assert gw.nvals == w.nvals
    finally:
        gb.set_option("hub_mode", 1)
        gb.set_option("force_direction", 0)


@pytest.mark.parametrize("mon,mul,dt,vm", [("MIN", "PLUS", "INT64", 1), ("PLUS", "TIMES", "FP32", 2), ("ANY", "PAIR", "BOOL", 0),
                                           ("MAX", "FIRST", "INT32", 1)])
@pytest.mark.parametrize("density", [0.001, 0.3, 0.97])
def test_red_pull_sparse_vector(gb, mon, mul, dt, vm, density):
    """k_spmv_red<SPARSE>: u has holes, so hub presence bits (shared memory) and the presence gather of non-hub
    columns decide which products exist; rows whose products are all absent must stay absent.  Overwrite,
    accumulate into a sparse w, and the aliased Bellman-Ford form w(accum) << w.vxm(A)."""
    from graphblas_b200 import binary

    scale = 17
    n = 1 << scale
    A = O.rmat_matrix(scale, dt, value_mode=vm)
    gA = gb.Matrix.rmat(scale, dt, value_mode=vm)
    At = A.transpose()
    rng = np.random.default_rng(int(density * 1000) + 5)
    tol = FP_TOL if dt.startswith("FP") else None
    gb.set_option("force_direction", 1)
    gb.set_option("hub_mode", 2)
    try:
        u = _rand_vec(rng, n, dt, density)
        if dt.startswith("FP"):
            u.values[:] = u.values / n
        gu = to_gpu_vector(gb, u)
        sr = _sr(gb, mon, mul)
        w = O.Vec.empty(n, dt)
        O.mxv(w, A, u, mon, mul, dt)
        gw = gA.mxv(gu, sr).new()
        assert vec_equal(gw, w, tol) and gw.nvals == w.nvals
        if mon != "ANY":
            acc = getattr(binary, mon.lower())
            O.mxv(w, A, u, mon, mul, dt, accum=mon, tran0=True, At=At)
            gw(acc) << gA.T.mxv(gu, sr)
            assert vec_equal(gw, w, tol) and gw.nvals == w.nvals
            # aliased: u(accum) << u.vxm(A)   (vxm(A) == mxv with A')
            for _ in range(2):
                O.vxm(u, u.dup(), A, mon, mul, dt, accum=mon, At=At)
                gu(acc) << gu.vxm(gA, sr)
                assert vec_equal(gu, u, tol) and gu.nvals == u.nvals
    finally:
        gb.set_option("hub_mode", 1)
        gb.set_option("force_direction", 0)


@pytest.mark.parametrize("mon,mul,dt,mdt,vm,density", [("PLUS", "TIMES", "FP32", "FP32", 2, 1.0), ("MIN", "PLUS", "INT64", "INT64", 1, 0.4),
                                                       ("MIN", "SECOND", "UINT32", "BOOL", 0, 1.0), ("PLUS", "TIMES", "INT64", "INT32", 1, 1.0)])
@pytest.mark.parametrize("panel_kb", [64, 200])
def test_red_pull_column_panels(gb, mon, mul, dt, mdt, vm, density, panel_kb):
    """Column panels (option panel_kb): the pull matrix is split by column ranges so that one panel's slice of the
    vector stays in L2; every panel accumulates into the same output with the monoid's atomic.  Results must not
    depend on the split (8 and 3 panels here), with matrix values in the semiring's type, in another type (cast
    once per matrix) or ignored by the multiply (pattern-only panels)."""
    from graphblas_b200 import binary

    scale = 17
    n = 1 << scale
    A = O.rmat_matrix(scale, mdt, value_mode=vm)
    gA = gb.Matrix.rmat(scale, mdt, value_mode=vm)
    At = A.transpose()
    rng = np.random.default_rng(panel_kb)
    tol = FP_TOL if dt.startswith("FP") else None
    gb.set_option("force_direction", 1)
    gb.set_option("hub_mode", 2)
    gb.set_option("panel_kb", panel_kb)
    try:
        u = _rand_vec(rng, n, dt, density)
        if dt.startswith("FP"):
            u.values[:] = u.values / n
        gu = to_gpu_vector(gb, u)
        sr = _sr(gb, mon, mul)
        Ac, Atc = (A, At) if mdt == dt else (A.astype(dt), At.astype(dt))
        w = O.Vec.empty(n, dt)
        O.mxv(w, Ac, u, mon, mul, dt)
        gw = gA.mxv(gu, sr).new()
        assert vec_equal(gw, w, tol) and gw.nvals == w.nvals
        acc = getattr(binary, mon.lower())
        for _ in range(2):
            O.mxv(w, Ac, u, mon, mul, dt, accum=mon, tran0=True, At=Atc)
            gw(acc) << gA.T.mxv(gu, sr)
            assert vec_equal(gw, w, tol) and gw.nvals == w.nvals
    finally:
        gb.set_option("panel_kb", 24576)
        gb.set_option("hub_mode", 1)
        gb.set_option("force_direction", 0)
