"""oracle/oracle.py — TEST INFRASTRUCTURE ONLY (checker; never on the product path).

numpy + ctypes front end of oracle/grb_oracle.c: GraphBLAS mxm / mxv / vxm with mask, accumulator,
replace and transpose descriptors on host CSR arrays.  Semantics follow the GraphBLAS C API 2.0 as
used by python-graphblas (reference graphblas/core/base.py:438-483 argument assembly,
core/descriptor.py:51-89 descriptor bits, core/mask.py:131-200 mask kinds).  Pinned to the
reference's golden vectors by tests/test_oracle_golden.py.

Data model
  Mat: CSR, indptr int64[nrows+1], indices int64[nnz] sorted in rows, values dtype[nnz], shape
  Vec: dense values dtype[n] + present uint8[n]
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TYPE_CODES = {"BOOL": 0, "INT8": 1, "INT16": 2, "INT32": 3, "INT64": 4, "UINT8": 5, "UINT16": 6,
              "UINT32": 7, "UINT64": 8, "FP32": 9, "FP64": 10}
NP_TYPES = {"BOOL": np.bool_, "INT8": np.int8, "INT16": np.int16, "INT32": np.int32,
            "INT64": np.int64, "UINT8": np.uint8, "UINT16": np.uint16, "UINT32": np.uint32,
            "UINT64": np.uint64, "FP32": np.float32, "FP64": np.float64}
OPS = {"FIRST": 0, "SECOND": 1, "ONEB": 2, "PAIR": 2, "ANY": 3, "PLUS": 4, "MINUS": 5, "TIMES": 6,
       "DIV": 7, "MIN": 8, "MAX": 9, "LOR": 10, "LAND": 11, "LXOR": 12, "LXNOR": 13, "EQ": 14}
# on BOOL the arithmetic names are the logical ones (C API 2.0 table of built-in ops)
_BOOL_ALIAS = {"PLUS": "LOR", "TIMES": "LAND", "MIN": "LAND", "MAX": "LOR"}


def build(force=False):
    """Compile oracle/grb_oracle.c -> oracle/liboracle.so (gcc, OpenMP)."""
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "grb_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O3", "-march=native", "-fopenmp", "-fPIC", "-shared",
                               "-o", so, src, "-lm"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.orc_num_threads.restype = ctypes.c_int
        L.orc_free.argtypes = [ctypes.c_void_p]
        _LIB = L
    return _LIB


def num_threads():
    return lib().orc_num_threads()


def _p(a):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def _opcode(op, dt):
    if dt == "BOOL":
        op = _BOOL_ALIAS.get(op, op)
    return OPS[op]


class Mat:
    def __init__(self, indptr, indices, values, shape, dtype=None, iso=False):
        self.indptr = np.ascontiguousarray(indptr, dtype=np.int64)
        self.indices = np.ascontiguousarray(indices, dtype=np.int64)
        self.dtype = dtype or _dtype_name(values.dtype)
        self.values = np.ascontiguousarray(values, dtype=NP_TYPES[self.dtype])
        self.shape = (int(shape[0]), int(shape[1]))
        self.iso = iso

    @property
    def nnz(self):
        return int(self.indptr[-1])

    @classmethod
    def from_coo(cls, rows, cols, vals, nrows, ncols, dtype):
        rows = np.asarray(rows, dtype=np.int64)
        cols = np.asarray(cols, dtype=np.int64)
        vals = np.asarray(vals, dtype=NP_TYPES[dtype])
        order = np.lexsort((cols, rows))
        rows, cols, vals = rows[order], cols[order], vals[order]
        indptr = np.zeros(nrows + 1, dtype=np.int64)
        np.add.at(indptr, rows + 1, 1)
        return cls(np.cumsum(indptr), cols, vals, (nrows, ncols), dtype)

    def to_coo(self):
        rows = np.repeat(np.arange(self.shape[0], dtype=np.int64), np.diff(self.indptr))
        vals = np.broadcast_to(self.values[:1], rows.shape).copy() if self.iso else self.values
        return rows, self.indices, vals

    def transpose(self):
        rows, cols, vals = self.to_coo()
        order = np.lexsort((rows, cols))
        indptr = np.zeros(self.shape[1] + 1, dtype=np.int64)
        np.add.at(indptr, cols + 1, 1)
        return Mat(np.cumsum(indptr), rows[order], vals[order], (self.shape[1], self.shape[0]),
                   self.dtype)

    def astype(self, dtype):
        if dtype == self.dtype:
            return self
        return Mat(self.indptr, self.indices, _cast(self.values, dtype), self.shape, dtype, self.iso)

    def keys(self):
        rows, cols, _ = self.to_coo()
        return rows * np.int64(max(self.shape[1], 1)) + cols


class Vec:
    def __init__(self, values, present, dtype=None):
        self.dtype = dtype or _dtype_name(np.asarray(values).dtype)
        self.values = np.ascontiguousarray(values, dtype=NP_TYPES[self.dtype])
        self.present = np.ascontiguousarray(present, dtype=np.uint8)

    @property
    def size(self):
        return self.values.shape[0]

    @property
    def nvals(self):
        return int(self.present.sum())

    @classmethod
    def from_coo(cls, idx, vals, size, dtype):
        v = np.zeros(size, dtype=NP_TYPES[dtype])
        p = np.zeros(size, dtype=np.uint8)
        idx = np.asarray(idx, dtype=np.int64)
        v[idx] = np.asarray(vals, dtype=NP_TYPES[dtype])
        p[idx] = 1
        return cls(v, p, dtype)

    @classmethod
    def empty(cls, size, dtype):
        return cls(np.zeros(size, dtype=NP_TYPES[dtype]), np.zeros(size, dtype=np.uint8), dtype)

    def to_coo(self):
        idx = np.nonzero(self.present)[0].astype(np.int64)
        return idx, self.values[idx]

    def dup(self):
        return Vec(self.values.copy(), self.present.copy(), self.dtype)


def _dtype_name(dt):
    dt = np.dtype(dt)
    for k, v in NP_TYPES.items():
        if np.dtype(v) == dt:
            return k
    raise TypeError(dt)


def _cast(a, dtype):
    """C-style typecast between built-in types (C API 2.0: as in ANSI C; bool(x) = x != 0)."""
    if dtype == "BOOL":
        return a != 0
    with np.errstate(invalid="ignore", over="ignore"):
        return a.astype(NP_TYPES[dtype])


# --------------------------------------------------------------------------------------------------
# vector products
# --------------------------------------------------------------------------------------------------
def _vec_mask(mask, comp, structure, n):
    """Evaluate a vector mask to a truth byte per position (None = no mask, everything allowed)."""
    if mask is None:
        return np.zeros(n, dtype=np.uint8) if comp else None
    m = mask.present.astype(bool)
    if not structure:
        m = m & (mask.values != 0)
    if comp:
        m = ~m
    return m.astype(np.uint8)


def spmv(A, u, mon, mul, dt, flip=False, row_mask=None):
    """t = A (+).(x) u over CSR rows of A; returns Vec of type dt."""
    L = lib()
    A = A.astype(dt)
    uv = _cast(u.values, dt) if u.dtype != dt else u.values
    uv = np.ascontiguousarray(uv, dtype=NP_TYPES[dt])
    up = None if u.nvals == u.size else u.present
    t = Vec.empty(A.shape[0], dt)
    rc = L.orc_spmv(TYPE_CODES[dt], _opcode(mon, dt), _opcode(mul, dt), int(flip),
                    ctypes.c_int64(A.shape[0]), _p(A.indptr), _p(A.indices), _p(A.values),
                    int(A.iso), _p(uv), _p(up), _p(row_mask), _p(t.values), _p(t.present))
    assert rc == 0
    return t


def vec_writeback(w, t, mask=None, accum=None, replace=False, comp=False, structure=False):
    """w<m,replace> = accum(w, t) in place; t is cast to w's type (accum computed in w's type)."""
    L = lib()
    tv = t.values if t.dtype == w.dtype else np.ascontiguousarray(_cast(t.values, w.dtype))
    m = _vec_mask(mask, comp, structure, w.size)
    acc = -1 if accum is None else _opcode(accum, w.dtype)
    rc = L.orc_vec_writeback(TYPE_CODES[w.dtype], ctypes.c_int64(w.size), _p(w.values), _p(w.present),
                             _p(tv), _p(t.present), _p(m), acc, int(replace))
    assert rc == 0
    return w


def mxv(w, A, u, mon, mul, dt, mask=None, accum=None, replace=False, comp=False, structure=False,
        tran0=False, At=None):
    """GrB_mxv: w<m> = accum(w, A (+).(x) u).  At = optional cached transpose of A."""
    if tran0:
        A = At if At is not None else A.transpose()
    assert A.shape[1] == u.size and A.shape[0] == w.size, "dimension mismatch"
    rm = None
    if mon == "ANY" and mask is not None:
        rm = _vec_mask(mask, comp, structure, w.size)
    t = spmv(A, u, mon, mul, dt, flip=False, row_mask=rm)
    return vec_writeback(w, t, mask, accum, replace, comp, structure)


def vxm(w, u, A, mon, mul, dt, mask=None, accum=None, replace=False, comp=False, structure=False,
        tran1=False, At=None):
    """GrB_vxm: w'<m'> = accum(w', u' (+).(x) A); evaluated as a pull over the rows of A'."""
    if not tran1:
        A = At if At is not None else A.transpose()
    assert A.shape[1] == u.size and A.shape[0] == w.size, "dimension mismatch"
    rm = None
    if mon == "ANY" and mask is not None:
        rm = _vec_mask(mask, comp, structure, w.size)
    t = spmv(A, u, mon, mul, dt, flip=True, row_mask=rm)
    return vec_writeback(w, t, mask, accum, replace, comp, structure)


# --------------------------------------------------------------------------------------------------
# matrix products
# --------------------------------------------------------------------------------------------------
def _take_csr(nrows, ncols, Cp, Cj, Cx, dt):
    L = lib()
    indptr = np.ctypeslib.as_array(Cp, shape=(nrows + 1,)).copy()
    nnz = int(indptr[-1])
    indices = np.ctypeslib.as_array(Cj, shape=(max(nnz, 1),))[:nnz].copy()
    nb = np.dtype(NP_TYPES[dt]).itemsize
    if nnz:      # (ctypes.string_at takes a C int: results beyond 2 GB, e.g. the scale-17 SpGEMM, need the array route)
        raw = np.ctypeslib.as_array(ctypes.cast(Cx, ctypes.POINTER(ctypes.c_ubyte)), shape=(nnz * nb,))
        values = raw.view(NP_TYPES[dt]).copy()
    else:
        values = np.zeros(0, dtype=NP_TYPES[dt])
    L.orc_free(Cp), L.orc_free(Cj), L.orc_free(Cx)
    return Mat(indptr, indices, values, (nrows, ncols), dt)


def spgemm(A, B, mon, mul, dt, mask=None, comp=False):
    """T = A (+).(x) B (Gustavson); mask = structural pattern Mat restricting T (optional)."""
    L = lib()
    A, B = A.astype(dt), B.astype(dt)
    assert A.shape[1] == B.shape[0], "dimension mismatch"
    Cp, Cj = ctypes.POINTER(ctypes.c_int64)(), ctypes.POINTER(ctypes.c_int64)()
    Cx = ctypes.c_void_p()
    rc = L.orc_spgemm(TYPE_CODES[dt], _opcode(mon, dt), _opcode(mul, dt), ctypes.c_int64(A.shape[0]),
                      ctypes.c_int64(B.shape[1]), _p(A.indptr), _p(A.indices), _p(A.values), int(A.iso),
                      _p(B.indptr), _p(B.indices), _p(B.values), int(B.iso),
                      _p(mask.indptr) if mask is not None else None,
                      _p(mask.indices) if mask is not None else None, int(comp),
                      ctypes.byref(Cp), ctypes.byref(Cj), ctypes.byref(Cx))
    assert rc == 0
    return _take_csr(A.shape[0], B.shape[1], Cp, Cj, Cx, dt)


def masked_dot(M, A, Bt, mon, mul, dt):
    """T<M.S> = A (+).(x) Bt' by per-mask-entry sorted intersections (Bt given by rows)."""
    L = lib()
    A, Bt = A.astype(dt), Bt.astype(dt)
    val = np.zeros(max(M.nnz, 1), dtype=NP_TYPES[dt])
    has = np.zeros(max(M.nnz, 1), dtype=np.uint8)
    rc = L.orc_masked_dot(TYPE_CODES[dt], _opcode(mon, dt), _opcode(mul, dt), ctypes.c_int64(M.shape[0]),
                          _p(M.indptr), _p(M.indices), _p(A.indptr), _p(A.indices), _p(A.values),
                          int(A.iso), _p(Bt.indptr), _p(Bt.indices), _p(Bt.values), int(Bt.iso),
                          _p(val), _p(has))
    assert rc == 0
    keep = has[:M.nnz].astype(bool)
    rows = np.repeat(np.arange(M.shape[0], dtype=np.int64), np.diff(M.indptr))[keep]
    indptr = np.zeros(M.shape[0] + 1, dtype=np.int64)
    np.add.at(indptr, rows + 1, 1)
    return Mat(np.cumsum(indptr), M.indices[keep], val[:M.nnz][keep], M.shape, dt)


def _apply_np(op, x, y, dt):
    if dt == "BOOL":
        op = _BOOL_ALIAS.get(op, op)
    with np.errstate(over="ignore", invalid="ignore", divide="ignore"):
        if op == "PLUS":
            return x + y
        if op == "MINUS":
            return x - y
        if op == "TIMES":
            return x * y
        if op == "MIN":
            return np.minimum(x, y) if dt not in ("FP32", "FP64") else np.fmin(x, y)
        if op == "MAX":
            return np.maximum(x, y) if dt not in ("FP32", "FP64") else np.fmax(x, y)
        if op == "FIRST":
            return x
        if op in ("SECOND", "ANY"):
            return y
        if op in ("ONEB", "PAIR"):
            return np.ones_like(x)
        if op == "LOR":
            return (x != 0) | (y != 0)
        if op == "LAND":
            return (x != 0) & (y != 0)
        if op == "LXOR":
            return (x != 0) ^ (y != 0)
        if op == "LXNOR":
            return ~((x != 0) ^ (y != 0))
    raise ValueError(op)


def mat_writeback(C, T, mask=None, accum=None, replace=False, comp=False, structure=False):
    """C<M,replace> = accum(C, T) on CSR matrices; returns a new Mat of C's type."""
    ct = C.dtype
    nc = np.int64(max(C.shape[1], 1))
    ck, tk = C.keys(), T.keys()
    cv = C.to_coo()[2]
    tv = _cast(T.to_coo()[2], ct)
    if accum is None:
        zk, zv = tk, tv
    else:
        zk = np.union1d(ck, tk)
        zv = np.zeros(zk.shape, dtype=NP_TYPES[ct])
        ci, ti = np.searchsorted(zk, ck), np.searchsorted(zk, tk)
        inC = np.zeros(zk.shape, dtype=bool)
        inC[ci] = True
        zv[ci] = cv
        both = inC[ti]
        zv[ti[~both]] = tv[~both]
        zv[ti[both]] = _cast(np.asarray(_apply_np(accum, zv[ti[both]], tv[both], ct)), ct)
    if mask is None and not comp:
        ok, ov = zk, zv
    else:
        if mask is None:
            mk = np.zeros(0, dtype=np.int64)
        else:
            mk = mask.keys()
            if not structure:
                mk = mk[mask.to_coo()[2] != 0]
        z_in = np.isin(zk, mk) != comp
        c_keep = (np.isin(ck, mk) == comp) & (not replace)
        ok = np.concatenate([zk[z_in], ck[c_keep]])
        ov = np.concatenate([zv[z_in], cv[c_keep]])
        order = np.argsort(ok, kind="stable")
        ok, ov = ok[order], ov[order]
    rows, cols = ok // nc, ok % nc
    indptr = np.zeros(C.shape[0] + 1, dtype=np.int64)
    np.add.at(indptr, rows + 1, 1)
    return Mat(np.cumsum(indptr), cols, ov.astype(NP_TYPES[ct]), C.shape, ct)


def mxm(C, A, B, mon, mul, dt, mask=None, accum=None, replace=False, comp=False, structure=False,
        tran0=False, tran1=False, method="gustavson"):
    """GrB_mxm: C<M> = accum(C, A (+).(x) B)."""
    if tran0:
        A = A.transpose()
    if method == "dot" and mask is not None and not comp and structure:
        Bt = B if tran1 else B.transpose()
        T = masked_dot(mask, A, Bt, mon, mul, dt)
    else:
        if tran1:
            B = B.transpose()
        T = spgemm(A, B, mon, mul, dt)
    assert C.shape == (A.shape[0], T.shape[1]), "dimension mismatch"
    return mat_writeback(C, T, mask, accum, replace, comp, structure)


# --------------------------------------------------------------------------------------------------
# synthetic graphs
# --------------------------------------------------------------------------------------------------
def rmat_edges(scale, edgefactor=16, seed=None):
    L = lib()
    seed = 42 + scale if seed is None else seed
    ne = edgefactor << scale
    src = np.empty(ne, dtype=np.uint64)
    dst = np.empty(ne, dtype=np.uint64)
    L.orc_rmat_edges(int(scale), ctypes.c_uint64(ne), ctypes.c_uint64(seed), _p(src), _p(dst))
    return src, dst


def rmat_pattern(scale, edgefactor=16, seed=None, symmetrize=False):
    """Deduplicated, self-loop-free edge keys (row*n+col, sorted) of the R-MAT graph."""
    src, dst = rmat_edges(scale, edgefactor, seed)
    n = np.uint64(1) << np.uint64(scale)
    keep = src != dst
    src, dst = src[keep], dst[keep]
    keys = src * n + dst
    if symmetrize:
        keys = np.concatenate([keys, dst * n + src])
    keys = np.unique(keys)
    return (keys // n).astype(np.int64), (keys % n).astype(np.int64)


def set_num_threads(n):
    """OpenMP threads of the C restatement (bench.py's CPU arm: torchrun exports OMP_NUM_THREADS=1)."""
    lib().orc_set_num_threads(int(n))


def rmat_csr(scale, dtype, edgefactor=16, seed=None, symmetrize=False, value_mode=0, transpose=False):
    """rmat_matrix(...) (transpose=False) or its transpose, built by the parallel C path: the same Mat, bit for bit, as
    the numpy path above (tests/test_oracle_golden.py), at the sizes bench.py's CPU arm needs (scale 23-26)."""
    L = lib()
    seed = 42 + scale if seed is None else seed
    n = 1 << scale
    Ap, Aj, nnz = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_int64()
    rc = L.orc_rmat_csr(int(scale), ctypes.c_uint64(edgefactor << scale), ctypes.c_uint64(seed), int(transpose),
                        int(symmetrize), ctypes.byref(Ap), ctypes.byref(Aj), ctypes.byref(nnz))
    assert rc == 0, f"orc_rmat_csr failed ({rc})"
    nnz = nnz.value
    indptr = np.ctypeslib.as_array(ctypes.cast(Ap, ctypes.POINTER(ctypes.c_int64)), shape=(n + 1,)).copy()
    indices = np.ctypeslib.as_array(ctypes.cast(Aj, ctypes.POINTER(ctypes.c_int64)), shape=(max(nnz, 1),))[:nnz].copy()
    L.orc_free(Ap)
    L.orc_free(Aj)
    if value_mode == 0:
        vals = np.ones(nnz, dtype=NP_TYPES[dtype])
    else:
        vals = np.empty(nnz, dtype=NP_TYPES[dtype])
        rc = L.orc_csr_edge_values(TYPE_CODES[dtype], int(value_mode), int(transpose), ctypes.c_int64(n), _p(indptr),
                                   _p(indices), _p(vals))
        assert rc == 0, f"orc_csr_edge_values: unsupported type {dtype}"
    return Mat(indptr, indices, vals, (n, n), dtype)


def edge_value(rows, cols, mode, dtype):
    """Deterministic per-edge value shared with the CUDA generator (grb_rmat.cu):
    mode 0: iso 1; mode 1: integer weight in [1,255]; mode 2: float in (0,1]."""
    t = NP_TYPES[dtype]
    if mode == 0:
        return np.ones(rows.shape, dtype=t)
    h = (rows.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)) ^ \
        (cols.astype(np.uint64) * np.uint64(0xC2B2AE3D27D4EB4F))
    h ^= h >> np.uint64(29)
    h *= np.uint64(0xBF58476D1CE4E5B9)
    h ^= h >> np.uint64(32)
    if mode == 1:
        return (np.uint64(1) + h % np.uint64(255)).astype(t)
    return ((h >> np.uint64(40)).astype(np.float64) + 1.0).astype(t) / t(16777216.0)


def rmat_matrix(scale, dtype, edgefactor=16, seed=None, symmetrize=False, value_mode=0):
    rows, cols = rmat_pattern(scale, edgefactor, seed, symmetrize)
    n = 1 << scale
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.add.at(indptr, rows + 1, 1)
    return Mat(np.cumsum(indptr), cols, edge_value(rows, cols, value_mode, dtype), (n, n), dtype)
