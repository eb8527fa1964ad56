"""Ingest from / export to scipy.sparse, with the reference's entry-point names and argument meaning
(reference graphblas/io/_scipy.py:8-64 `from_scipy_sparse`, :67-119 `to_scipy_sparse`).

csr / csc inputs go across the C-ABI as they are (GrB_Matrix_import_<T>: uint64 index arrays, typed values; a CSC input keeps
only its columns in HBM until somebody needs rows); every other format goes through COO and the device build
(radix sort + dedup with `dup_op`, grb_storage.cu).  Unsorted indices inside rows are sorted on the device."""
import numpy as np

from .core import Matrix, Vector
from .dtypes import lookup_dtype


def from_scipy_sparse(A, *, dup_op=None, name=None):
    """Create a Matrix from a scipy.sparse array or matrix."""
    nrows, ncols = A.shape
    dtype = lookup_dtype(A.dtype)
    if A.nnz == 0:
        return Matrix(dtype, nrows, ncols, name=name)
    if A.format == "csr":
        return Matrix.from_csr(A.indptr, A.indices, A.data, dtype=dtype, nrows=nrows, ncols=ncols, name=name)
    if A.format == "csc":
        return Matrix.from_csc(A.indptr, A.indices, A.data, dtype=dtype, nrows=nrows, ncols=ncols, name=name)
    if A.format != "coo":
        A = A.tocoo()
    return Matrix.from_coo(A.row, A.col, A.data, nrows=nrows, ncols=ncols, dtype=dtype, dup_op=dup_op, name=name)


def to_scipy_sparse(A, format="csr"):
    """Create a scipy.sparse array from a Matrix or Vector (a Vector becomes an n x 1 array, as in the reference)."""
    import scipy.sparse as ss

    format = format.lower()
    if isinstance(A, Vector):
        idx, vals = A.to_coo()
        out = ss.coo_array((vals, (idx.astype(np.int64), np.zeros(idx.size, dtype=np.int64))), shape=(A.size, 1))
        return out.asformat(format)
    if format == "csc":
        indptr, indices, data = A.to_csc()
        return ss.csc_array((data, indices.astype(np.int64), indptr.astype(np.int64)), shape=(A.nrows, A.ncols))
    indptr, indices, data = A.to_csr()
    out = ss.csr_array((data, indices.astype(np.int64), indptr.astype(np.int64)), shape=(A.nrows, A.ncols))
    return out if format == "csr" else out.asformat(format)
