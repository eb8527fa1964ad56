"""oracle/pyref.py — TEST INFRASTRUCTURE ONLY (never imported by the product path).

A dict-of-keys restatement of GraphBLAS C-API-2.0 semantics for the semiring products

    C<M, replace> = accum(C, A (+).(x) B)           (GrB_mxm / GrB_mxv / GrB_vxm)

as python-graphblas uses them.  The arithmetic itself lives in SuiteSparse:GraphBLAS 7.4.x (PyPI
`suitesparse-graphblas >=7.4.0.0,<7.5`, reference pyproject.toml:65), which is not in /root/reference;
this file therefore follows the published C API 2.0 semantics as *used* at the reference call sites:
  - argument assembly [C, mask, accum, op, A, B, desc]      reference graphblas/core/base.py:438-483
  - descriptor bits (replace, comp, structure, tran0, tran1) reference graphblas/core/descriptor.py:51-89
  - mask kinds (structural / value / complemented)           reference graphblas/core/mask.py:131-200
  - user-level statement of the same rules                   reference docs/user_guide/fundamentals.rst:16-38
Pinned against the literal golden vectors of the reference's own tests (tests/golden/reference_kats.json,
transcribed from graphblas/tests/test_matrix.py:307-392 and test_vector.py:297-366) by
tests/test_oracle_golden.py.  Pure-Python loops: small cases only.

Matrices are dicts {(i, j): value}; vectors are dicts {i: value}; values are numpy scalars of the
semiring's dtype so integer overflow wraps exactly as in C.
"""
import numpy as np

_NP = {
    "BOOL": np.bool_, "INT8": np.int8, "INT16": np.int16, "INT32": np.int32, "INT64": np.int64,
    "UINT8": np.uint8, "UINT16": np.uint16, "UINT32": np.uint32, "UINT64": np.uint64,
    "FP32": np.float32, "FP64": np.float64,
}


def np_dtype(name):
    return _NP[name]


def _binop(op, dt):
    """z = op(x, y) in dtype dt, C semantics (wraparound for ints)."""
    t = _NP[dt]
    with np.errstate(over="ignore", invalid="ignore", divide="ignore"):
        if op == "PLUS":
            return lambda x, y: t(x) + t(y) if dt != "BOOL" else t(bool(x) or bool(y))
        if op == "TIMES":
            return lambda x, y: t(x) * t(y) if dt != "BOOL" else t(bool(x) and bool(y))
        if op == "MINUS":
            return lambda x, y: t(x) - t(y)
        if op == "MIN":
            return lambda x, y: t(min(t(x), t(y)))
        if op == "MAX":
            return lambda x, y: t(max(t(x), t(y)))
        if op == "FIRST":
            return lambda x, y: t(x)
        if op in ("SECOND", "ANY"):
            # ANY may return either argument; the oracle picks the later one.  Only compare ANY
            # results bit-exactly when both candidates are equal (any_pair) or compare structure.
            return lambda x, y: t(y)
        if op in ("ONEB", "PAIR"):
            return lambda x, y: t(1)
        if op == "LOR":
            return lambda x, y: t(bool(x) or bool(y))
        if op == "LAND":
            return lambda x, y: t(bool(x) and bool(y))
        if op == "LXOR":
            return lambda x, y: t(bool(x) != bool(y))
        if op in ("LXNOR", "EQ"):
            return lambda x, y: t(bool(x) == bool(y)) if dt == "BOOL" else np.bool_(t(x) == t(y))
    raise ValueError(op)


def _wrap(f):
    def g(x, y):
        with np.errstate(over="ignore", invalid="ignore"):
            return f(x, y)
    return g


def semiring_product(A, B, mon, mul, dt):
    """T = A (+).(x) B over the structural intersection (explicit zeros are entries)."""
    add = _wrap(_binop(mon, dt))
    mult = _wrap(_binop(mul, dt))
    t = _NP[dt]
    Brows = {}
    for (k, j), b in B.items():
        Brows.setdefault(k, []).append((j, b))
    T = {}
    for (i, k), a in sorted(A.items()):
        for j, b in sorted(Brows.get(k, ())):
            z = mult(t(a), t(b))
            T[(i, j)] = add(T[(i, j)], z) if (i, j) in T else z
    return T


def transpose(A):
    return {(j, i): v for (i, j), v in A.items()}


def writeback(C, T, ctype, M=None, accum=None, accum_type=None, replace=False, comp=False,
              structure=False):
    """C<M,replace> = accum(C, T).  C API 2.0 section 'standard method for masks and accumulators':
    Z = accum ? (C union T, accum on the intersection) : T;  inside the mask C takes Z (entries absent
    from Z are deleted); outside the mask C is kept, unless replace, then deleted."""
    ct = _NP[ctype]
    if accum is None:
        Z = {k: ct(v) for k, v in T.items()}
    else:
        f = _wrap(_binop(accum, accum_type or ctype))
        Z = {k: ct(v) for k, v in C.items()}
        for k, v in T.items():
            Z[k] = ct(f(C[k], v)) if k in C else ct(v)

    def allowed(k):
        if M is None:
            m = True
            return not comp  # complemented empty mask allows nothing
        m = (k in M) and (structure or bool(M[k]))
        return m != comp

    out = {}
    for k in set(C) | set(Z):
        if allowed(k):
            if k in Z:
                out[k] = Z[k]
        elif not replace and k in C:
            out[k] = ct(C[k])
    return out


def mxm(C, A, B, mon, mul, dt, ctype=None, M=None, accum=None, replace=False, comp=False,
        structure=False, tran0=False, tran1=False):
    if tran0:
        A = transpose(A)
    if tran1:
        B = transpose(B)
    T = semiring_product(A, B, mon, mul, dt)
    return writeback(C, T, ctype or dt, M, accum, ctype or dt, replace, comp, structure)


def _col(v):
    return {(i, 0): x for i, x in v.items()}


def _uncol(Cm):
    return {i: x for (i, _), x in Cm.items()}


def mxv(w, A, u, mon, mul, dt, **kw):
    """w<m> = accum(w, A (+).(x) u); tran0 transposes A (GrB_DESC_T0)."""
    kw.pop("tran1", None)
    M = kw.pop("M", None)
    return _uncol(mxm(_col(w), A, _col(u), mon, mul, dt, M=None if M is None else _col(M), **kw))


def vxm(w, u, A, mon, mul, dt, **kw):
    """w'<m'> = accum(w', u' (+).(x) A); tran1 transposes A (GrB_DESC_T1).  Note the multiply
    operand order: z = mult(u(i), A(i,j)) — first/second are not symmetric with mxv."""
    tran1 = kw.pop("tran1", False)
    kw.pop("tran0", None)
    M = kw.pop("M", None)
    row = lambda v: {(0, i): x for i, x in v.items()}
    Cm = mxm(row(w), row(u), A, mon, mul, dt, M=None if M is None else row(M), tran1=tran1, **kw)
    return {j: x for (_, j), x in Cm.items()}
