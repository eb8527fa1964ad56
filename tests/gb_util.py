"""Conversions between the oracle's host containers and graphblas_b200 objects (GPU tests only)."""
import numpy as np


def to_gpu_matrix(gb, M):
    vals = np.broadcast_to(M.values[:1], M.indices.shape).copy() if M.iso else M.values
    return gb.Matrix.from_csr(M.indptr, M.indices, vals, dtype=M.dtype, ncols=M.shape[1], nrows=M.shape[0]) \
        if M.nnz else gb.Matrix(M.dtype, *M.shape)


def to_gpu_vector(gb, v):
    if v.nvals == v.size:
        return gb.Vector.from_dense(v.values, dtype=v.dtype)
    return gb.Vector.from_dense(v.values, dtype=v.dtype, present=v.present)


def vec_equal(gv, ov, tol=None):
    vals, pres = gv.to_dense()
    p = ov.present.astype(bool)
    if not np.array_equal(pres.astype(bool), p):
        return False
    if tol is None:
        return bool(np.array_equal(vals[p], ov.values[p]))
    return bool(np.allclose(vals[p].astype(np.float64), ov.values[p].astype(np.float64), rtol=tol, atol=tol))


def mat_equal(gm, om, tol=None, structure_only=False):
    ap, ai, ax = gm.to_csr()
    if not (np.array_equal(ap.astype(np.int64), om.indptr) and np.array_equal(ai.astype(np.int64), om.indices)):
        return False
    if structure_only:
        return True
    ov = np.broadcast_to(om.values[:1], om.indices.shape) if om.iso else om.values
    if tol is None:
        return bool(np.array_equal(ax, ov))
    return bool(np.allclose(ax.astype(np.float64), ov.astype(np.float64), rtol=tol, atol=tol))
