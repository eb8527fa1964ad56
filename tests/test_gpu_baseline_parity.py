"""GPU: CUDA vs the CPU oracle at (or next to) the BASELINE.json sizes, on the real R-MAT skew -- where round 1 only had
self-consistency checks ("What's weak" 1).  Everything goes through the C-ABI (graphblas_b200 -> libgrb_b200.so):
  * SpGEMM A.mxm(A) plus_times, FP64 and INT64, R-MAT scale 16-18: the whole CSR against oracle.spgemm -- pattern bit-exact,
    INT64 values bit-exact, FP64 within the north_star's 1e-5 relative.  Scale 18 puts 93 % of the products in rows that go
    through the column-panel kernels (k_panel_symbolic / k_panel_row);
  * triangle counting C(L.S) << L.mxm(L.T, plus_pair) at scale 20: the count AND C's pattern and values against the oracle's
    masked dot product (replaces the literal this test used to hold);
  * Bellman-Ford SSSP min_plus INT64 at scale 22 and level BFS any_pair at scale 18 (BASELINE configs[2], [0]): bit-exact
    against the same loop on the oracle.
The host graphs come from oracle.rmat_csr (parallel C, == the numpy path, tests/test_oracle_golden.py), the device graphs
from the generator proved bit-identical in test_gpu_matrix_parity.py."""
import numpy as np
import pytest

from gb_util import mat_equal
from oracle import oracle as O

pytestmark = pytest.mark.gpu
FP_TOL = 1e-5


@pytest.fixture(scope="module")
def gb():
    import os

    import graphblas_b200 as gb

    gb.init("b200")
    O.build()
    O.set_num_threads(len(os.sched_getaffinity(0)))
    return gb


def _host_ram_gb():
    try:
        import psutil

        return psutil.virtual_memory().available / 2**30
    except Exception:
        return 0.0


@pytest.mark.parametrize("scale,dt", [(16, "FP64"), (16, "INT64"), (17, "FP64"), (18, "FP64"), (18, "INT64")])
def test_spgemm_full_csr_vs_oracle(gb, scale, dt):
    from graphblas_b200 import Matrix, semiring

    if scale >= 18 and _host_ram_gb() < 96:
        pytest.skip("the scale-18 product (1.28 G entries) needs ~60 GB of host memory for the two CSR copies")
    vm = 2 if dt == "FP64" else 1
    A = O.rmat_csr(scale, dt, value_mode=vm)
    gA = Matrix.rmat(scale, dt, value_mode=vm)
    assert gA.nvals == A.nnz
    T = O.spgemm(A, A, "PLUS", "TIMES", dt)
    gC = gA.mxm(gA, semiring.plus_times).new()
    assert gC.nvals == T.nnz
    ap, ai, ax = gC.to_csr()
    del gC
    assert np.array_equal(ap.astype(np.int64), T.indptr), "row pointers differ"
    assert np.array_equal(ai.view(np.int64), T.indices), "pattern differs"
    if dt == "INT64":
        assert np.array_equal(ax, T.values)
    else:
        # relative error of the sums (all products are positive): different summation orders only
        err = np.abs(ax - T.values) / np.abs(T.values)
        assert float(err.max()) <= FP_TOL, float(err.max())


def test_triangle_count_scale20_pattern_values_and_count(gb):
    from graphblas_b200 import Matrix, monoid, semiring

    scale = 20
    G = O.rmat_csr(scale, "INT64", symmetrize=True)
    rows = np.repeat(np.arange(G.shape[0], dtype=np.int64), np.diff(G.indptr))
    keep = G.indices < rows                                   # strict lower triangle
    ip = np.zeros(G.shape[0] + 1, dtype=np.int64)
    ip[1:] = np.cumsum(np.bincount(rows[keep], minlength=G.shape[0]))
    L = O.Mat(ip, G.indices[keep], np.ones(int(keep.sum()), dtype=np.int64), G.shape, "INT64")
    del G, rows, keep
    T = O.masked_dot(L, L, L, "PLUS", "PAIR", "INT64")        # C<L> = L (+).pair L'  (Bt = L by rows)
    gG = Matrix.rmat(scale, "INT64", symmetrize=True)
    gL = gG.select("tril", -1).new()
    del gG
    assert gL.nvals == L.nnz
    C = Matrix("INT64", 1 << scale, 1 << scale)
    C(gL.S) << gL.mxm(gL.T, semiring.plus_pair)
    assert C.nvals == T.nnz
    assert mat_equal(C, T)                                    # pattern and per-entry counts, bit-exact
    assert int(C.reduce_scalar(monoid.plus).value) == int(T.values.sum())


def test_bellman_ford_scale22_vs_oracle(gb):
    from graphblas_b200 import Matrix, workloads as W

    scale = 22
    n = 1 << scale
    A = O.rmat_csr(scale, "INT64", symmetrize=True, value_mode=1)
    At = O.rmat_csr(scale, "INT64", symmetrize=True, value_mode=1, transpose=True)   # the pattern is symmetric, the weights are not
    src = int(np.argmax(np.diff(A.indptr)))
    w = O.Vec.from_coo([src], [0], n, "INT64")
    iters = 0
    while True:                                               # notebooks/Intro to GraphBLAS + SSSP example.ipynb cell 21
        prev_v, prev_p = w.values.copy(), w.present.copy()
        O.vxm(w, w.dup(), A, "MIN", "PLUS", "INT64", accum="MIN", At=At)
        iters += 1
        if np.array_equal(prev_p, w.present) and np.array_equal(prev_v[prev_p.astype(bool)], w.values[prev_p.astype(bool)]):
            break
    gA = Matrix.rmat(scale, "INT64", symmetrize=True, value_mode=1)
    gw, st = W.sssp_bellman_ford(gA, src)
    assert st["iterations"] == iters
    gv, gp = gw.to_dense()
    p = w.present.astype(bool)
    assert np.array_equal(gp.astype(bool), p)
    assert np.array_equal(gv[p], w.values[p])


def test_level_bfs_scale18_vs_oracle(gb):
    from graphblas_b200 import Matrix, workloads as W

    scale = 18
    n = 1 << scale
    A = O.rmat_csr(scale, "BOOL", symmetrize=True)
    src = int(np.argmax(np.diff(A.indptr)))
    v = O.Vec.empty(n, "INT32")
    q = O.Vec.from_coo([src], [True], n, "BOOL")
    level = 0
    while q.nvals:                                            # notebooks/Example B.1 -- Level BFS.ipynb cell 8
        level += 1
        v.values[q.present.astype(bool)] = level
        v.present[q.present.astype(bool)] = 1
        nq = O.Vec.empty(n, "BOOL")
        O.vxm(nq, q, A, "ANY", "PAIR", "BOOL", mask=v, comp=True, structure=True, replace=True, At=A)
        q = nq
    gA = Matrix.rmat(scale, "BOOL", symmetrize=True)
    gv, st = W.bfs_levels(gA, src)
    assert st["levels"] == level
    vals, pres = gv.to_dense()
    p = v.present.astype(bool)
    assert np.array_equal(pres.astype(bool), p)
    assert np.array_equal(vals[p], v.values[p])
