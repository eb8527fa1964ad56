"""The five BASELINE.json workloads, written the way the reference's notebooks write them (same statements, same
operators), on top of graphblas_b200.  bench.py times them; tests/test_gpu_workloads.py checks them against
scipy.sparse.csgraph / dense arithmetic.  Each returns (result, stats) with the work units the metric needs."""
import numpy as np

from . import binary, monoid, semiring
from .core import Matrix, Scalar, Vector


def bfs_levels(A, source, op=None):
    """Level BFS: `q(~v.S, replace=True) << q.vxm(A, any_pair)` (reference notebooks/Example B.1 -- Level BFS.ipynb
    cell 8, with the north-star semiring any_pair in place of lor_land).  v[i] = 1 + distance from source."""
    op = op or semiring.any_pair
    n = A.nrows
    v = Vector("INT32", n)
    q = Vector("BOOL", n)
    q[source] = True
    level = 0
    frontier_sizes = []
    while True:
        level += 1
        nq = q.nvals
        if nq == 0:
            break
        frontier_sizes.append(nq)
        v(q.V)[:] << level
        q(~v.S, replace=True) << q.vxm(A, op)
    return v, {"levels": level - 1, "frontier_sizes": frontier_sizes}


def pagerank(A, d_out, damping=0.85, iters=20):
    """notebooks/Pagerank Demo.ipynb cell 9 with a fixed iteration count: w = r / d; r[:] = teleport;
    r(plus) << A.T.mxv(w, plus_times)."""
    n = A.nrows
    teleport = float(np.float32((1 - damping) / n))
    r = Vector.from_dense(np.full(n, 1.0 / n, dtype=np.float32))
    w = Vector("FP32", n)
    for _ in range(iters):
        w << r.ewise_mult(d_out, binary.truediv)
        r[:] << teleport
        r(binary.plus) << A.T.mxv(w, semiring.plus_times)
    return r, {"iterations": iters}


def sssp_bellman_ford(A, source, max_iters=None):
    """`w(min) << w.vxm(A, min_plus)` to the fixed point (notebooks/Intro to GraphBLAS + SSSP example.ipynb cell 21)."""
    n = A.nrows
    w = Vector(A.dtype, n)
    w[source] = 0
    iters = 0
    max_iters = max_iters or n
    while iters < max_iters:
        prev_nvals = w.nvals
        prev = w.dup()
        w(binary.min) << w.vxm(A, semiring.min_plus)
        iters += 1
        if w.nvals == prev_nvals and w.isequal(prev):
            break
    return w, {"iterations": iters}


def triangle_count(A):
    """Sandia-LL: L = tril(A, -1); C(L.S) << L.mxm(L.T, plus_pair); sum(C) (BASELINE config 4)."""
    L = A.select("tril", -1).new()
    C = Matrix("INT64", A.nrows, A.ncols)
    C(L.S) << L.mxm(L.T, semiring.plus_pair)
    ntri = C.reduce_scalar(monoid.plus).value
    return int(ntri or 0), {"nnz_L": L.nvals, "nnz_C": C.nvals}


def spgemm(A, op=None):
    """C = A (+).(x) A (BASELINE config 5a)."""
    C = A.mxm(A, op or semiring.plus_times).new()
    return C, {"nnz_C": C.nvals}


def cc_min_label(A, max_iters=None, f0=None):
    """Connected components by min-label propagation with the FastSV product `mngp << A.mxv(gp, min_second)`
    (notebooks/Connected Components -- FastSV.ipynb cell 30).  FastSV's hooking / shortcutting steps need
    assign-by-index and extract (SURVEY 8f "next"); this loop keeps the product and the element-wise min, which is the
    part BASELINE config 5b measures (GTEPS = iterations x nnz / time)."""
    n = A.nrows
    f = f0.dup() if f0 is not None else Vector.from_dense(np.arange(n, dtype=np.uint32))   # f0: device-resident 0..n-1
    iters = 0
    max_iters = max_iters or n
    while iters < max_iters:
        prev = f.dup()
        f(binary.min) << A.mxv(f, semiring.min_second)
        iters += 1
        if f.isequal(prev):
            break
    return f, {"iterations": iters}


def fastsv(A, max_iters=None, device_indices=False, f0=None):
    """Connected components by FastSV, statement for statement the reference's
    notebooks/Connected Components -- FastSV.ipynb cell 30 (hooking through an indexed min-assign with duplicate
    indices, shortcutting, grandparents by extract).  f[i] = smallest vertex id of i's component.
    device_indices: the notebook refreshes the index list with `_, I = f.to_coo()` every iteration (f is full, so I is
    just f's values pulled to the host and pushed back); with True, I is the Vector f itself and stays in HBM
    (GxB_b200_Vector_assign_Vector / _extract_Vector) -- same statements, same result, no PCIe round trip."""
    n = A.nrows
    I = np.arange(n)
    dt = "UINT32" if n < (1 << 32) else "UINT64"
    # the parent of each vertex is the vertex itself (f0: the labels 0..n-1 already resident in HBM)
    f = f0.dup() if f0 is not None else Vector.from_coo(I, I.astype(dt.lower()), size=n, dtype=dt)
    gp = f.dup()              # grandparents
    gp_dup = gp.dup()
    mngp = f.dup()            # minimum grandparent of each star vertex
    mod = Vector("BOOL", n)
    change = True
    iters = 0
    if device_indices:
        I = f.dup()
    while change and (max_iters is None or iters < max_iters):
        # Step 1: hooking
        mngp << A.mxv(gp, semiring.min_second)
        f(binary.min)[I] << mngp
        f << f.ewise_add(mngp, binary.min)
        # Step 2: shortcutting
        f << f.ewise_add(gp, binary.min)
        # Step 3: grandparents
        if device_indices:
            I << f
        else:
            _, I = f.to_coo()
        gp << f[I]
        # termination
        mod << gp_dup.ewise_mult(gp, binary.ne)
        change = bool(mod.reduce(monoid.lor).value)
        gp_dup << gp
        iters += 1
    return f, {"iterations": iters}
