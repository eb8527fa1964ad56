"""Host-side mirror of python-graphblas's collections and expression/update machinery for the semiring
hot path.  Same names, argument meaning and error behaviour as the reference, re-implemented over
grb_b200's cffi `lib`:

    C(mask.V, accum=binary.plus, replace=True) << A.mxm(B.T, semiring.min_plus)
    w = A.mxv(v, semiring.plus_times).new()
    q(~v.S, replace=True) << q.vxm(A, semiring.any_pair)

  - `call()`                       reference graphblas/core/base.py:23-54
  - BaseType.__call__ / Updater    reference graphblas/core/base.py:192-263, core/expr.py:402-471
  - _update -> [C, mask, accum, op, A, B, desc]   reference graphblas/core/base.py:438-483
  - descriptor lookup              reference graphblas/core/descriptor.py:51-156
  - masks .S / .V / ~              reference graphblas/core/mask.py:9-200
  - Matrix.mxv / mxm, Vector.vxm / inner / outer  reference core/matrix.py:2163-2251, core/vector.py:1259-1306, 1609-1693
"""
import numpy as np

from grb_b200 import errors, ffi, lib

from . import dtypes
from .dtypes import lookup_dtype, unify
from .operators import Op, TypedOp, binary, binary_from_string, get_typed_op, monoid, semiring, unary

NULL = ffi.NULL
replace = object()  # sentinel usable positionally: C(mask, replace) << ...


def call(cfunc_name, args):
    """Invoke one C entry point: each argument contributes its `_carg`, None is NULL; a non-zero
    GrB_Info raises the matching exception with the library's error string."""
    cargs = [NULL if a is None else getattr(a, "_carg", a) for a in args]
    info = getattr(lib, cfunc_name)(*cargs)
    if info == lib.GrB_SUCCESS:
        return info
    if info == lib.GrB_NO_VALUE:
        return info
    msg = ""
    first = args[0]
    if isinstance(first, BaseType):
        s = ffi.new("char**")
        getattr(lib, f"GrB_{first._ckind}_error")(s, first._carg)
        if s[0] != NULL:
            msg = ffi.string(s[0]).decode()
    raise errors.error_for(info)(f"{cfunc_name}: {msg}" if msg else cfunc_name)


# ------------------------------------------------------------------------------------------ descriptors
def _descriptor(replace_=False, comp=False, structure=False, t0=False, t1=False, **opts):
    """One of the 32 predefined GrB_DESC_* handles (NULL for the default), or a fresh descriptor when
    SuiteSparse-style options (axb_method=, nthreads=, sort=) are given
    (reference core/descriptor.py:92-156, core/ss/descriptor.py:123-197)."""
    if opts:
        return _custom_descriptor(replace_, comp, structure, t0, t1, opts)
    name = ("R" if replace_ else "") + ("S" if structure else "") + ("C" if comp else "")
    name += ("T0" if t0 else "") + ("T1" if t1 else "")
    if not name:
        return None
    return _Handle(getattr(lib, f"GrB_DESC_{name}"))


_AXB = {"default": "GxB_DEFAULT", "gustavson": "GxB_AxB_GUSTAVSON", "dot": "GxB_AxB_DOT",
        "hash": "GxB_AxB_HASH", "saxpy": "GxB_AxB_SAXPY"}
_VALID_OPTS = ("axb_method", "nthreads", "chunk", "sort", "secure_import", "compression", "gpu_id")


class _Handle:
    __slots__ = ("_carg", "_keep", "return_type")

    def __init__(self, carg, keep=None, return_type=None):
        self._carg = carg
        self._keep = keep
        self.return_type = return_type


def _custom_descriptor(replace_, comp, structure, t0, t1, opts):
    low = {k.lower(): v for k, v in opts.items()}
    if len(low) != len(opts):
        raise ValueError("Duplicate descriptor options given (descriptor options are case-insensitive)")
    bad = [k for k in low if k not in _VALID_OPTS]
    if bad:
        raise ValueError(f"Unknown descriptor option(s) {bad}; valid options: {', '.join(_VALID_OPTS)}")
    d = ffi.new("GrB_Descriptor*")
    call("GrB_Descriptor_new", [d])
    h = _Handle(d[0], keep=ffi.gc(d, lib.GrB_Descriptor_free))
    for cond, field, val in ((replace_, lib.GrB_OUTP, lib.GrB_REPLACE), (comp, lib.GrB_MASK, lib.GrB_COMP),
                             (structure, lib.GrB_MASK, lib.GrB_STRUCTURE), (t0, lib.GrB_INP0, lib.GrB_TRAN),
                             (t1, lib.GrB_INP1, lib.GrB_TRAN)):
        if cond:
            lib.GrB_Descriptor_set(d[0], field, val)
    if "axb_method" in low and low["axb_method"] is not None:
        m = str(low["axb_method"]).lower()
        if m not in _AXB:
            raise ValueError(f"axb_method must be one of {', '.join(_AXB)}; got {low['axb_method']!r}")
        lib.GrB_Descriptor_set(d[0], lib.GxB_AxB_METHOD, getattr(lib, _AXB[m]))
    if "sort" in low and low["sort"] not in (None, False, True):
        raise ValueError("sort must be one of False, True")
    return h


# ------------------------------------------------------------------------------------------ masks
class Mask:
    complement = False
    structure = False
    value = False

    def __init__(self, parent):
        self.parent = parent

    @property
    def _carg(self):
        return self.parent._carg

    def __repr__(self):
        return f"{type(self).__name__}({self.parent!r})"


class StructuralMask(Mask):
    structure = True

    def __invert__(self):
        return ComplementedStructuralMask(self.parent)


class ValueMask(Mask):
    value = True

    def __invert__(self):
        return ComplementedValueMask(self.parent)


class ComplementedStructuralMask(Mask):
    complement = True
    structure = True

    def __invert__(self):
        return StructuralMask(self.parent)


class ComplementedValueMask(Mask):
    complement = True
    value = True

    def __invert__(self):
        return ValueMask(self.parent)


def _check_mask(mask, output):
    if isinstance(mask, BaseType):
        if type(mask) is not type(output) and not (isinstance(mask, Matrix) and isinstance(output, Matrix)):
            raise TypeError(f"Mask object must be type {type(output).__name__}; got {type(mask).__name__}")
        if mask.dtype is dtypes.BOOL:
            return ValueMask(mask)
        raise TypeError("Mask must be indicated with one of the following methods:\n"
                        "    M.S, M.V, ~M.S, ~M.V (bool collections may be used as-is)")
    if not isinstance(mask, Mask):
        raise TypeError(f"Invalid mask: {type(mask)}")
    return mask


# ------------------------------------------------------------------------------------------ expressions
class Expression:
    """A deferred C call: nothing is computed until `.new()` or `out << expr`."""

    def __init__(self, method, cfunc_name, args, op, out_cls, shape, at=False, bt=False, scalar_args=()):
        self.method = method
        self.cfunc_name = cfunc_name
        self.args = args
        self.op = op
        self.out_cls = out_cls
        self.shape = shape
        self.at = at
        self.bt = bt
        self.scalar_args = scalar_args

    @property
    def dtype(self):
        return self.op.return_type if self.op is not None else self.args[0].dtype

    # shape of the result, as the reference's expression objects expose it (core/expr.py; tests/test_infix.py:96-106)
    @property
    def size(self):
        if self.out_cls is not Vector:
            raise AttributeError(f"{self.out_cls.__name__} expression has no attribute 'size'")
        return self.shape[0]

    @property
    def nrows(self):
        if self.out_cls is not Matrix:
            raise AttributeError(f"{self.out_cls.__name__} expression has no attribute 'nrows'")
        return self.shape[0]

    @property
    def ncols(self):
        if self.out_cls is not Matrix:
            raise AttributeError(f"{self.out_cls.__name__} expression has no attribute 'ncols'")
        return self.shape[1]

    def _new_output(self, dtype):
        dt = lookup_dtype(dtype) if dtype is not None else self.dtype
        if self.out_cls is Vector:
            return Vector(dt, self.shape[0])
        if self.out_cls is Scalar:
            return Scalar(dt)
        return Matrix(dt, *self.shape)

    def new(self, dtype=None, *, mask=None, name=None, **opts):
        out = self._new_output(dtype)
        if mask is not None:
            mask = _check_mask(mask, out)
        out._update(self, mask=mask, opts=opts)
        return out

    def __repr__(self):
        return f"<{self.out_cls.__name__}Expression {self.method} op={self.op!r}>"


class MatMulInfix(Expression):
    """`A @ B`: an mxm / mxv / vxm / inner expression whose semiring defaults to plus_times and can be
    re-typed with `semiring.min_plus(A @ B)` (reference core/infix.py:444-497)."""

    def __init__(self, left, right):
        self.left, self.right = left, right
        e = _matmul_expr(left, right, semiring.plus_times)
        super().__init__(e.method, e.cfunc_name, e.args, e.op, e.out_cls, e.shape, e.at, e.bt)

    def _with_op(self, op):
        return _matmul_expr(self.left, self.right, op)


def _matmul_expr(left, right, op):
    if isinstance(left, (Matrix, TransposedMatrix)):
        if isinstance(right, (Matrix, TransposedMatrix)):
            return left.mxm(right, op)
        if isinstance(right, Vector):
            return left.mxv(right, op)
    elif isinstance(left, Vector):
        if isinstance(right, (Matrix, TransposedMatrix)):
            return left.vxm(right, op)
        if isinstance(right, Vector):
            return left.inner(right, op)
    raise TypeError(f"Bad types for @: {type(left).__name__} and {type(right).__name__}")


class Updater:
    def __init__(self, parent, mask=None, accum=None, replace_=False, opts=None):
        self.parent = parent
        self.kwargs = dict(mask=mask, accum=accum, replace_=replace_)
        self.opts = opts or {}

    def __lshift__(self, expr):
        self.parent._update(expr, opts=self.opts, **self.kwargs)

    def update(self, expr):
        self.parent._update(expr, opts=self.opts, **self.kwargs)

    def __setitem__(self, keys, value):
        self.parent._assign(keys, value, opts=self.opts, **self.kwargs)

    def __getitem__(self, keys):
        return _Assigner(self, keys)


class _Assigner:
    """w(mask)[:] << value   /   w[:](mask) << value"""

    def __init__(self, updater, keys):
        self.updater, self.keys = updater, keys

    def __lshift__(self, value):
        self.updater[self.keys] = value

    def update(self, value):
        self.updater[self.keys] = value

    def __call__(self, *args, **kwargs):
        return _Assigner(self.updater.parent(*args, **kwargs), self.keys)


# ------------------------------------------------------------------------------------------ base collection
class BaseType:
    _ckind = "Matrix"
    _is_scalar = False

    def __init__(self, gb_obj, dtype, name=None):
        self.gb_obj = gb_obj
        self.dtype = dtype
        self.name = name

    @property
    def _carg(self):
        return self.gb_obj[0]

    def __call__(self, *positional, mask=None, accum=None, replace=False, **opts):
        """out(mask, accum, replace=) -> Updater (argument sorting as reference core/base.py:192-263)."""
        replace_ = replace
        mask_arg = accum_arg = None
        for arg in positional:
            if arg is globals()["replace"]:
                replace_ = True
            elif isinstance(arg, (BaseType, Mask, TransposedMatrix)):
                if self._is_scalar:
                    raise TypeError("Mask not allowed for Scalars")
                if mask_arg is not None:
                    raise TypeError("Got multiple values for argument 'mask'")
                mask_arg = arg
            else:
                if accum_arg is not None:
                    raise TypeError("Got multiple values for argument 'accum'")
                if isinstance(arg, str):
                    arg = binary_from_string(arg)
                elif not isinstance(arg, (Op, TypedOp)):
                    raise TypeError(f"Invalid item found in output params: {type(arg)}")
                accum_arg = arg
        if mask_arg is not None and mask is not None:
            raise TypeError("Got multiple values for argument 'mask'")
        mask = mask_arg if mask_arg is not None else mask
        if mask is None:
            if replace_:
                raise TypeError("'replace' argument may only be True if a mask is provided")
        elif self._is_scalar:
            raise TypeError("Mask not allowed for Scalars")
        else:
            mask = _check_mask(mask, self)
        if accum_arg is not None and accum is not None:
            raise TypeError("Got multiple values for argument 'accum'")
        accum = accum_arg if accum_arg is not None else accum
        if accum is not None:
            if isinstance(accum, str):
                accum = binary_from_string(accum)
            accum = get_typed_op(accum, self.dtype, kind="binary")
            if accum.opclass == "Monoid":
                accum = accum.parent.binaryop[accum.dtype]
            elif accum.opclass != "BinaryOp":
                raise TypeError(f"accum must be a BinaryOp or Monoid; got {accum!r}")
        return Updater(self, mask=mask, accum=accum, replace_=replace_, opts=opts)

    def __lshift__(self, expr):
        self._update(expr, opts={})

    def update(self, expr, **opts):
        self._update(expr, opts=opts)

    def __bool__(self):
        raise TypeError(f"__bool__ not defined for objects of type {type(self)}.  Perhaps use .nvals attribute instead.")

    def _update(self, expr, mask=None, accum=None, replace_=False, *, opts):
        if isinstance(expr, BaseType) and type(expr) is type(self):
            expr = expr._as_expression()
        if not isinstance(expr, Expression):
            raise TypeError(f"update expected an expression; got {type(expr).__name__}. "
                            "Hint: did you forget a method such as .mxm/.mxv/.vxm?")
        if expr.out_cls is not type(self) and not (expr.out_cls is Scalar and self._is_scalar):
            if expr.out_cls is Scalar:
                raise TypeError(f"Bad type for output: scalar expression into {type(self).__name__} "
                                "(enable autocompute or use .new() first)")
            raise TypeError(f"Bad output type: expression yields {expr.out_cls.__name__}, output is {type(self).__name__}")
        comp = structure = False
        if mask is not None:
            comp, structure = mask.complement, mask.structure
        desc = _descriptor(replace_, comp, structure, expr.at, expr.bt, **opts)
        args = [self, mask, accum]
        if expr.op is not None:
            args.append(expr.op)
        args.extend(expr.args)
        args.extend(expr.scalar_args)
        args.append(desc)
        call(expr.cfunc_name, args)

    @property
    def nvals(self):
        n = ffi.new("GrB_Index*")
        call(f"GrB_{self._ckind}_nvals", [n, self])
        return int(n[0])

    def wait(self):
        call(f"GrB_{self._ckind}_wait", [self, lib.GrB_MATERIALIZE])
        return self

    def clear(self):
        call(f"GrB_{self._ckind}_clear", [self])

    @property
    def S(self):
        return StructuralMask(self)

    @property
    def V(self):
        return ValueMask(self)

    def __matmul__(self, other):
        return MatMulInfix(self, other)


def _free_later(gb_obj, free_name):
    return ffi.gc(gb_obj, getattr(lib, free_name))


def _semiring_for(op, a, b):
    if isinstance(op, TypedOp):
        if op.opclass != "Semiring":
            raise TypeError(f"Expected a Semiring; got {op!r}")
        return op
    if isinstance(op, str):
        op = getattr(semiring, op)
    if not isinstance(op, Op) or op.opclass != "Semiring":
        raise TypeError(f"Expected a Semiring; got {op!r}")
    return op[unify(a.dtype, b.dtype)]


class _Cast:
    """A handle viewed as another handle type (reference Vector._as_matrix, core/vector.py:194-202)."""

    def __init__(self, parent, ctype, ckind):
        self.parent, self._ctype, self._ckind = parent, ctype, ckind
        self.dtype = parent.dtype

    @property
    def _carg(self):
        return ffi.cast(self._ctype, self.parent._carg)


def _np_values(values, dtype):
    return np.ascontiguousarray(values, dtype=dtype.np_type)


def _np_index(idx):
    return np.ascontiguousarray(idx, dtype=np.uint64)


def _slice_index(key, size):
    """The GrB_Index list a Python slice stands for on a vector of `size` (the reference resolves slices to GrB_ALL, a
    (lo, hi, stride) triple or a list, core/utils.py / core/expr.py AmbiguousAssignOrExtract; this library takes GrB_ALL or a list).
    Negative steps give the indices in the order the slice visits them, as numpy does."""
    return np.arange(*key.indices(int(size)), dtype=np.int64).astype(np.uint64)


def _buf(arr, ctype):
    return ffi.cast(ctype, ffi.from_buffer(arr)) if arr.size else ffi.cast(ctype, 0)


# ------------------------------------------------------------------------------------------ Scalar
class Scalar(BaseType):
    _ckind = "Scalar"
    _is_scalar = True

    def __init__(self, dtype=dtypes.FP64, name=None):
        dtype = lookup_dtype(dtype)
        gb_obj = ffi.new("GrB_Scalar*")
        call("GrB_Scalar_new", [gb_obj, dtype])
        super().__init__(_free_later(gb_obj, "GrB_Scalar_free"), dtype, name)

    @classmethod
    def from_value(cls, value, dtype=None):
        dt = lookup_dtype(dtype) if dtype is not None else lookup_dtype(type(value) if not isinstance(value, np.generic) else value.dtype)
        s = cls(dt)
        s.value = value
        return s

    @property
    def value(self):
        out = ffi.new(f"{_CTYPE[self.dtype.name]}*")
        info = call(f"GrB_Scalar_extractElement_{self.dtype.name}", [out, self])
        if info == lib.GrB_NO_VALUE:
            return None
        return self.dtype.np_type.type(out[0]).item()

    @value.setter
    def value(self, val):
        if val is None:
            self.clear()
        else:
            call(f"GrB_Scalar_setElement_{self.dtype.name}", [self, self.dtype.np_type.type(val).item()])

    @property
    def is_empty(self):
        return self.nvals == 0

    def __eq__(self, other):
        if isinstance(other, Scalar):
            other = other.value
        return self.value == other

    __hash__ = None

    def isequal(self, other, *, check_dtype=False):
        if isinstance(other, Scalar):
            if check_dtype and other.dtype is not self.dtype:
                return False
            other = other.value
        return self.value == other

    def _update(self, expr, mask=None, accum=None, replace_=False, *, opts):
        if isinstance(expr, Expression) and expr.cfunc_name == "GrB_vxm":
            # inner product: the scalar is handed to GrB_vxm as a length-1 vector
            desc = _descriptor(False, False, False, expr.at, expr.bt, **opts)
            call("GrB_vxm", [_Cast(self, "GrB_Vector", "Vector"), None, accum, expr.op, *expr.args, desc])
            return
        if isinstance(expr, Expression) and expr.out_cls is Scalar and expr.method == "reduce":
            val = expr.new().value
            if accum is not None and self.value is not None and val is not None:
                tmp = Vector.from_coo([0, 1], [self.value, val], dtype=self.dtype, size=2)
                val = tmp.reduce(getattr(monoid, accum.parent.name)).new().value
            self.value = val
            return
        raise TypeError(f"Cannot update a Scalar from {expr!r}")

    def __repr__(self):
        return f"Scalar({self.value}, dtype={self.dtype})"


_CTYPE = {"BOOL": "bool", "INT8": "int8_t", "INT16": "int16_t", "INT32": "int32_t", "INT64": "int64_t",
          "UINT8": "uint8_t", "UINT16": "uint16_t", "UINT32": "uint32_t", "UINT64": "uint64_t",
          "FP32": "float", "FP64": "double"}


# ------------------------------------------------------------------------------------------ Vector
class Vector(BaseType):
    _ckind = "Vector"

    def __init__(self, dtype=dtypes.FP64, size=0, *, name=None):
        dtype = lookup_dtype(dtype)
        gb_obj = ffi.new("GrB_Vector*")
        call("GrB_Vector_new", [gb_obj, dtype, int(size)])
        super().__init__(_free_later(gb_obj, "GrB_Vector_free"), dtype, name)
        self._size = int(size)

    @classmethod
    def _from_handle(cls, gb_obj, dtype, size, name=None):
        """Adopt a GrB_Vector* made by a library call (GxB_b200_Vector_wrap_device); freed with GrB_Vector_free."""
        out = object.__new__(cls)
        BaseType.__init__(out, _free_later(gb_obj, "GrB_Vector_free"), lookup_dtype(dtype), name)
        out._size = int(size)
        return out

    @property
    def size(self):
        return self._size

    @property
    def shape(self):
        return (self._size,)

    # ---- construction / extraction ----
    @classmethod
    def from_coo(cls, indices, values=1.0, dtype=None, *, size=None, dup_op=None, name=None):
        indices = _np_index(indices)
        if np.ndim(values) == 0:
            values = np.full(indices.shape, values)
        values = np.asarray(values)
        dt = lookup_dtype(dtype) if dtype is not None else lookup_dtype(values.dtype)
        if size is None:
            if indices.size == 0:
                raise ValueError("No indices provided. Unable to infer size.")
            size = int(indices.max()) + 1
        v = cls(dt, size, name=name)
        v.build(indices, values, dup_op=dup_op)
        return v

    def build(self, indices, values, *, dup_op=None, clear=False):
        if clear:
            self.clear()
        indices = _np_index(indices)
        values = _np_values(values, self.dtype)
        if indices.size != values.size:
            raise ValueError(f"`indices` and `values` lengths must match: {indices.size} != {values.size}")
        if dup_op is not None:
            dup_op = get_typed_op(dup_op, self.dtype, kind="binary")
        call(f"GrB_Vector_build_{self.dtype.name}",
             [self, _buf(indices, "GrB_Index*"), _buf(values, f"{_CTYPE[self.dtype.name]}*"), indices.size, dup_op])

    @classmethod
    def from_dense(cls, values, dtype=None, *, present=None, name=None):
        """Full (or bitmap, with `present`) vector from a host array: one H2D copy, no sort."""
        values = np.asarray(values)
        dt = lookup_dtype(dtype) if dtype is not None else lookup_dtype(values.dtype)
        v = cls(dt, values.shape[0], name=name)
        v.assign_dense(values, present)
        return v

    def assign_dense(self, values, present=None):
        values = _np_values(values, self.dtype)
        pres_arr = None if present is None else np.ascontiguousarray(present, dtype=np.int8)  # keep alive
        pres = NULL if pres_arr is None else _buf(pres_arr, "int8_t*")
        call("GxB_b200_Vector_import_dense", [self, _buf(values, "void*"), pres, self._size])

    def to_dense(self, out=None, present_out=None):
        vals = out if out is not None else np.empty(self._size, dtype=self.dtype.np_type)
        pres = present_out if present_out is not None else np.empty(self._size, dtype=np.int8)
        call("GxB_b200_Vector_export_dense", [_buf(vals, "void*"), _buf(pres, "int8_t*"), self._size, self])
        return vals, pres.view(np.bool_) if pres.dtype == np.int8 else pres

    def to_coo(self, dtype=None, *, indices=True, values=True, sort=True):
        n = self.nvals
        dt = lookup_dtype(dtype) if dtype is not None else self.dtype
        idx = np.empty(n, dtype=np.uint64)
        vals = np.empty(n, dtype=dt.np_type)
        nv = ffi.new("GrB_Index*", n)
        call(f"GrB_Vector_extractTuples_{dt.name}", [_buf(idx, "GrB_Index*"), _buf(vals, f"{_CTYPE[dt.name]}*"), nv, self])
        return (idx if indices else None), (vals if values else None)

    def dup(self, dtype=None, *, name=None):
        if dtype is not None and lookup_dtype(dtype) is not self.dtype:
            i, v = self.to_coo()
            return Vector.from_coo(i, v.astype(lookup_dtype(dtype).np_type), size=self._size)
        gb_obj = ffi.new("GrB_Vector*")
        call("GrB_Vector_dup", [gb_obj, self])
        out = object.__new__(Vector)
        BaseType.__init__(out, _free_later(gb_obj, "GrB_Vector_free"), self.dtype, name)
        out._size = self._size
        return out

    def _as_expression(self):
        return _IdentityExpr(self)

    def isequal(self, other, *, check_dtype=False):
        if not isinstance(other, Vector):
            raise TypeError(f"Expected Vector; got {type(other).__name__}")
        if check_dtype and self.dtype is not other.dtype:
            return False
        if self._size != other._size:
            return False
        if self.dtype is other.dtype:
            # one fused device pass + one read-back (GxB_b200_Vector_isequal) instead of the reference's sequence below
            # (nvals, nvals, eWiseMult(eq), nvals, reduce(land): four host synchronisations per call); same answer
            res = ffi.new("bool*")
            call("GxB_b200_Vector_isequal", [_Handle(res), self, other])
            return bool(res[0])
        if self.nvals != other.nvals:
            return False
        if self.nvals == 0:
            return True
        # on the device, as the reference does it (core/vector.py isequal -> eWiseMult(eq) + reduce(land)):
        # equal patterns <=> the intersection has as many entries as either operand
        matches = Vector(dtypes.BOOL, self._size)
        matches << self.ewise_mult(other, binary.eq)
        if matches.nvals != self.nvals:
            return False
        return bool(matches.reduce(monoid.land).value)

    def isclose(self, other, *, rel_tol=1e-7, abs_tol=0.0, check_dtype=False):
        if self._size != other._size or self.nvals != other.nvals:
            return False
        i1, v1 = self.to_coo()
        i2, v2 = other.to_coo()
        return bool(np.array_equal(i1, i2) and np.allclose(v1, v2, rtol=rel_tol, atol=abs_tol))

    def __getitem__(self, key):
        if isinstance(key, slice) and key == slice(None):
            return _Assigner(Updater(self), key)
        if isinstance(key, slice):
            return _IndexedVector(self, _slice_index(key, self._size))
        if isinstance(key, (list, tuple, np.ndarray)):
            return _IndexedVector(self, _np_index(key))
        if isinstance(key, Vector):
            return _VectorIndexedVector(self, key)
        out = ffi.new(f"{_CTYPE[self.dtype.name]}*")
        info = call(f"GrB_Vector_extractElement_{self.dtype.name}", [out, self, int(key)])
        s = Scalar(self.dtype)
        if info != lib.GrB_NO_VALUE:
            s.value = out[0]
        return _ScalarResult(s)

    def __setitem__(self, key, value):
        if isinstance(key, (slice, list, tuple, np.ndarray, Vector)):
            self._assign(key, value, opts={})
        else:
            call(f"GrB_Vector_setElement_{self.dtype.name}", [self, self.dtype.np_type.type(value).item(), int(key)])

    def __delitem__(self, key):
        call("GrB_Vector_removeElement", [self, int(key)])

    # ---- the hot path ----
    def vxm(self, other, op=semiring.plus_times):
        """w' = u' (+).(x) A  ->  GrB_vxm (reference core/vector.py:1259-1306)."""
        if not isinstance(other, (Matrix, TransposedMatrix)):
            raise TypeError(f"Expected Matrix or TransposedMatrix for `other`; got {type(other).__name__}")
        op = _semiring_for(op, self, other)
        expr = Expression("vxm", "GrB_vxm", [self, other], op, Vector, (other.ncols,), bt=other._is_transposed)
        if self._size != other.nrows:
            expr.new()  # let the C library raise DimensionMismatch, like the reference (:1304-1305)
        return expr

    def inner(self, other, op=semiring.plus_times):
        """Scalar s = u' (+).(x) v: GrB_vxm on v cast to an n x 1 matrix (reference core/vector.py:1609-1648)."""
        if not isinstance(other, Vector):
            raise TypeError(f"Expected Vector for `other`; got {type(other).__name__}")
        op = _semiring_for(op, self, other)
        expr = Expression("inner", "GrB_vxm", [self, _Cast(other, "GrB_Matrix", "Matrix")], op, Scalar, ())
        if self._size != other._size:
            expr.new()
        return expr

    def outer(self, other, op=binary.times):
        """C = u (x) v': GrB_mxm with any_<op> on two n x 1 matrices, second transposed (core/vector.py:1650-1693)."""
        if not isinstance(other, Vector):
            raise TypeError(f"Expected Vector for `other`; got {type(other).__name__}")
        if isinstance(op, (Op, TypedOp)) and op.opclass in ("BinaryOp", "Monoid"):
            name = op.name
            op = getattr(semiring, f"any_{name}")
        op = _semiring_for(op, self, other)
        return Expression("outer", "GrB_mxm", [_Cast(self, "GrB_Matrix", "Matrix"), _Cast(other, "GrB_Matrix", "Matrix")],
                          op, Matrix, (self._size, other._size), bt=True)

    # ---- loop-body operations around the products ----
    def _assign(self, key, value, mask=None, accum=None, replace_=False, *, opts):
        """w(mask, accum)[I] << value  ->  GrB_Vector_assign / GrB_Vector_assign_<T> (reference core/vector.py:1772-1950);
        I is `:` (GrB_ALL) or an index list."""
        if isinstance(key, slice) and key == slice(None):
            I, ni, keep = lib.GrB_ALL, self._size, None
        elif isinstance(key, slice):
            keep = _slice_index(key, self._size)          # w[lo:hi:step]: the index list the slice stands for
            I, ni = _buf(keep, "GrB_Index*"), keep.size
        elif isinstance(key, (list, tuple, np.ndarray)):
            keep = _np_index(key)
            I, ni = _buf(keep, "GrB_Index*"), keep.size
        elif isinstance(key, Vector):
            # the index list is the value array of a full integer vector that already lives in HBM
            # (GxB_b200_Vector_assign_Vector, the counterpart of SuiteSparse's GxB_Vector_assign_Vector)
            comp = structure = False
            if mask is not None:
                comp, structure = mask.complement, mask.structure
            if isinstance(value, Expression):
                value = value.new()
            if not isinstance(value, Vector):
                raise TypeError("w[index vector] << value needs a Vector value")
            call("GxB_b200_Vector_assign_Vector", [self, mask, accum, value, key, _descriptor(replace_, comp, structure, **opts)])
            return
        else:
            raise TypeError("assignment takes w[:], w[slice], w[index list] or w[index vector]")
        comp = structure = False
        if mask is not None:
            comp, structure = mask.complement, mask.structure
        desc = _descriptor(replace_, comp, structure, **opts)
        if isinstance(value, Scalar):
            value = value.value
        if isinstance(value, Expression):
            value = value.new()
        if isinstance(value, Vector):
            call("GrB_Vector_assign", [self, mask, accum, value, I, ni, desc])
        else:
            call(f"GrB_Vector_assign_{self.dtype.name}",
                 [self, mask, accum, self.dtype.np_type.type(value).item(), I, ni, desc])

    def ewise_mult(self, other, op=binary.times):
        op = _binary_for(op, self, other)
        return Expression("ewise_mult", "GrB_Vector_eWiseMult_BinaryOp", [self, other], op, Vector, (self._size,))

    def ewise_add(self, other, op=binary.plus):
        op = _binary_for(op, self, other)
        return Expression("ewise_add", "GrB_Vector_eWiseAdd_BinaryOp", [self, other], op, Vector, (self._size,))

    def apply(self, op, right=None):
        """apply(unary) -> GrB_Vector_apply; apply(binary, right=scalar) -> GrB_Vector_apply_BinaryOp2nd_<T>
        (reference core/vector.py:1308-1420)."""
        if right is None:
            op = get_typed_op(op, self.dtype, kind="unary")
            if op.opclass != "UnaryOp":
                raise TypeError("apply with a binary operator needs right=<scalar>")
            return Expression("apply", "GrB_Vector_apply", [self], op, Vector, (self._size,))
        op = get_typed_op(op, self.dtype, kind="binary")
        if op.opclass == "Monoid":
            op = op.parent.binaryop[op.dtype]
        val = op.dtype.np_type.type(right).item()
        return Expression("apply", f"GrB_Vector_apply_BinaryOp2nd_{op.dtype.name}", [self], op, Vector, (self._size,),
                          scalar_args=(val,))

    def reduce(self, op=monoid.plus):
        op = get_typed_op(op, self.dtype, kind="monoid")
        return _ReduceExpr(self, op, "Vector")

    def __repr__(self):
        return f"Vector(dtype={self.dtype}, size={self._size}, nvals={self.nvals})"


class _IndexedVector(Expression):
    """u[I]: extract as an expression (`w << u[I]`, `u[I].new()`), or the target of an indexed assignment
    (`u[I] << v`, `u(mask, accum)[I] << v`)."""

    def __init__(self, src, index):
        super().__init__("extract", "GrB_Vector_extract", [src], None, Vector, (index.size,),
                         scalar_args=(_buf(index, "GrB_Index*"), index.size))
        self._src, self._index = src, index

    def __lshift__(self, value):
        self._src._assign(self._index, value, opts={})

    def update(self, value):
        self._src._assign(self._index, value, opts={})


class _VectorIndexedVector(Expression):
    """u[I] with I a full integer Vector in HBM: `w << u[I]` -> GxB_b200_Vector_extract_Vector; `u(accum)[I] << v` ->
    GxB_b200_Vector_assign_Vector.  No host round trip of the index list (FastSV: I is the parent vector itself)."""

    def __init__(self, src, index):
        super().__init__("extract", "GxB_b200_Vector_extract_Vector", [src, index], None, Vector, (index.size,))
        self._src, self._index = src, index

    def __lshift__(self, value):
        self._src._assign(self._index, value, opts={})

    def update(self, value):
        self._src._assign(self._index, value, opts={})


class _ScalarResult:
    """`v[i]` -> object with .new() / .value, like the reference's extract-element expression."""

    def __init__(self, scalar):
        self._s = scalar

    def new(self, dtype=None, **kw):
        return self._s

    @property
    def value(self):
        return self._s.value


class _IdentityExpr(Expression):
    """`w << u`: copy through apply(first, 0)-free path — a masked/accumulated assign of a whole vector."""

    def __init__(self, src):
        if not isinstance(src, Vector):
            raise TypeError("only Vector << Vector copies are supported; use .dup() for matrices")
        super().__init__("assign", "GrB_Vector_assign", [src], None, Vector, src.shape,
                         scalar_args=(lib.GrB_ALL, src.size))
        self.src = src

    @property
    def dtype(self):
        return self.src.dtype


class _ReduceExpr(Expression):
    def __init__(self, src, op, kind):
        super().__init__("reduce", f"GrB_{kind}_reduce_{op.dtype.name}", [src], op, Scalar, ())
        self.src, self.kind = src, kind

    def new(self, dtype=None, **kw):
        dt = self.op.dtype
        out = ffi.new(f"{_CTYPE[dt.name]}*")
        # identity goes in first so an empty collection reduces to "no value"
        if self.src.nvals == 0:
            return Scalar(lookup_dtype(dtype) if dtype else dt)
        call(self.cfunc_name, [out, None, self.op, self.src, None])
        s = Scalar(lookup_dtype(dtype) if dtype else dt)
        s.value = out[0]
        return s

    @property
    def value(self):
        return self.new().value


def _binary_for(op, a, b):
    if isinstance(op, str):
        op = binary_from_string(op)
    if isinstance(op, TypedOp):
        typed = op
    else:
        typed = op[unify(a.dtype, b.dtype)]
    if typed.opclass == "Monoid":
        typed = typed.parent.binaryop[typed.dtype]
    if typed.opclass != "BinaryOp":
        raise TypeError(f"Expected a BinaryOp or Monoid; got {op!r}")
    return typed


# ------------------------------------------------------------------------------------------ Matrix
class Matrix(BaseType):
    _ckind = "Matrix"
    _is_transposed = False

    def __init__(self, dtype=dtypes.FP64, nrows=0, ncols=0, *, name=None):
        dtype = lookup_dtype(dtype)
        gb_obj = ffi.new("GrB_Matrix*")
        call("GrB_Matrix_new", [gb_obj, dtype, int(nrows), int(ncols)])
        super().__init__(_free_later(gb_obj, "GrB_Matrix_free"), dtype, name)
        self._nrows, self._ncols = int(nrows), int(ncols)

    @classmethod
    def _wrap(cls, gb_obj, dtype, nrows, ncols, name=None):
        out = object.__new__(cls)
        BaseType.__init__(out, _free_later(gb_obj, "GrB_Matrix_free"), dtype, name)
        out._nrows, out._ncols = int(nrows), int(ncols)
        return out

    @property
    def nrows(self):
        return self._nrows

    @property
    def ncols(self):
        return self._ncols

    @property
    def shape(self):
        return (self._nrows, self._ncols)

    @property
    def T(self):
        return TransposedMatrix(self)

    # ---- construction / extraction ----
    @classmethod
    def from_coo(cls, rows, columns, values=1.0, dtype=None, *, nrows=None, ncols=None, dup_op=None, name=None):
        rows, columns = _np_index(rows), _np_index(columns)
        if np.ndim(values) == 0:
            values = np.full(rows.shape, values)
        values = np.asarray(values)
        dt = lookup_dtype(dtype) if dtype is not None else lookup_dtype(values.dtype)
        if nrows is None:
            if rows.size == 0:
                raise ValueError("No row indices provided. Unable to infer nrows.")
            nrows = int(rows.max()) + 1
        if ncols is None:
            if columns.size == 0:
                raise ValueError("No column indices provided. Unable to infer ncols.")
            ncols = int(columns.max()) + 1
        A = cls(dt, nrows, ncols, name=name)
        A.build(rows, columns, values, dup_op=dup_op)
        return A

    def build(self, rows, columns, values, *, dup_op=None, clear=False):
        if clear:
            self.clear()
        rows, columns = _np_index(rows), _np_index(columns)
        values = _np_values(values, self.dtype)
        if not (rows.size == columns.size == values.size):
            raise ValueError("`rows`, `columns` and `values` lengths must match")
        if dup_op is not None:
            dup_op = get_typed_op(dup_op, self.dtype, kind="binary")
        call(f"GrB_Matrix_build_{self.dtype.name}",
             [self, _buf(rows, "GrB_Index*"), _buf(columns, "GrB_Index*"), _buf(values, f"{_CTYPE[self.dtype.name]}*"),
              rows.size, dup_op])

    @classmethod
    def from_csr(cls, indptr, col_indices, values=1.0, dtype=None, *, ncols=None, nrows=None, name=None):
        return cls._from_csx("GrB_CSR_FORMAT", indptr, col_indices, values, dtype, ncols, nrows, name)

    @classmethod
    def from_csc(cls, indptr, row_indices, values=1.0, dtype=None, *, nrows=None, ncols=None, name=None):
        return cls._from_csx("GrB_CSC_FORMAT", indptr, row_indices, values, dtype, nrows, ncols, name)

    @classmethod
    def _from_csx(cls, fmt, indptr, indices, values, dtype, minor, major, name):
        """GrB_Matrix_import_<T> with copy semantics (reference core/matrix.py:1057-1133)."""
        indptr, indices = _np_index(indptr), _np_index(indices)
        if np.ndim(values) == 0:
            values = np.full(indices.shape, values)
        values = np.asarray(values)
        dt = lookup_dtype(dtype) if dtype is not None else lookup_dtype(values.dtype)
        values = _np_values(values, dt)
        if major is None:
            major = indptr.size - 1
        if minor is None:
            if indices.size == 0:
                raise ValueError("No indices provided. Unable to infer the minor dimension.")
            minor = int(indices.max()) + 1
        nrows, ncols = (major, minor) if fmt == "GrB_CSR_FORMAT" else (minor, major)
        gb_obj = ffi.new("GrB_Matrix*")
        info = getattr(lib, f"GrB_Matrix_import_{dt.name}")(
            gb_obj, dt._carg, nrows, ncols, _buf(indptr, "GrB_Index*"), _buf(indices, "GrB_Index*"),
            _buf(values, f"{_CTYPE[dt.name]}*"), indptr.size, indices.size, values.size, getattr(lib, fmt))
        if info != lib.GrB_SUCCESS:
            raise errors.error_for(info)("GrB_Matrix_import")
        return cls._wrap(gb_obj, dt, nrows, ncols, name)

    @classmethod
    def rmat(cls, scale, dtype=dtypes.BOOL, *, edgefactor=16, seed=None, symmetrize=False, value_mode=0, name=None):
        """Graph500-style R-MAT graph generated in HBM (bench utility): bit-identical to
        oracle.rmat_matrix(scale, dtype, edgefactor, seed, symmetrize, value_mode)."""
        dt = lookup_dtype(dtype)
        seed = 42 + scale if seed is None else seed
        gb_obj = ffi.new("GrB_Matrix*")
        info = lib.GxB_b200_Matrix_rmat(gb_obj, dt._carg, int(scale), int(edgefactor), int(seed), int(bool(symmetrize)),
                                        int(value_mode))
        if info != lib.GrB_SUCCESS:
            raise errors.error_for(info)("GxB_b200_Matrix_rmat")
        n = 1 << scale
        return cls._wrap(gb_obj, dt, n, n, name)

    def rowblock(self, row_begin, row_end, *, name=None):
        """Rows [row_begin, row_end) as a new matrix: the 1-D row-block shard of SURVEY 8e."""
        gb_obj = ffi.new("GrB_Matrix*")
        call("GxB_b200_Matrix_rowblock", [gb_obj, self, int(row_begin), int(row_end)])
        return Matrix._wrap(gb_obj, self.dtype, int(row_end) - int(row_begin), self._ncols, name)

    def to_coo(self, dtype=None, *, rows=True, columns=True, values=True, sort=True):
        n = self.nvals
        dt = lookup_dtype(dtype) if dtype is not None else self.dtype
        r = np.empty(n, dtype=np.uint64)
        c = np.empty(n, dtype=np.uint64)
        v = np.empty(n, dtype=dt.np_type)
        nv = ffi.new("GrB_Index*", n)
        call(f"GrB_Matrix_extractTuples_{dt.name}",
             [_buf(r, "GrB_Index*"), _buf(c, "GrB_Index*"), _buf(v, f"{_CTYPE[dt.name]}*"), nv, self])
        return (r if rows else None), (c if columns else None), (v if values else None)

    def to_csr(self, dtype=None):
        return self._to_csx("GrB_CSR_FORMAT", dtype)

    def to_csc(self, dtype=None):
        return self._to_csx("GrB_CSC_FORMAT", dtype)

    def _to_csx(self, fmt, dtype):
        """GrB_Matrix_exportSize + GrB_Matrix_export_<T> (reference core/matrix.py:1658-1702)."""
        dt = lookup_dtype(dtype) if dtype is not None else self.dtype
        f = getattr(lib, fmt)
        lp, li, lx = ffi.new("GrB_Index*"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
        call("GrB_Matrix_exportSize", [lp, li, lx, f, self])
        ap = np.empty(lp[0], dtype=np.uint64)
        ai = np.empty(li[0], dtype=np.uint64)
        ax = np.empty(lx[0], dtype=dt.np_type)
        call(f"GrB_Matrix_export_{dt.name}",
             [_buf(ap, "GrB_Index*"), _buf(ai, "GrB_Index*"), _buf(ax, f"{_CTYPE[dt.name]}*"), lp, li, lx, f, self])
        return ap, ai, ax

    def dup(self, dtype=None, *, name=None):
        if dtype is not None and lookup_dtype(dtype) is not self.dtype:
            r, c, v = self.to_coo()
            return Matrix.from_coo(r, c, v.astype(lookup_dtype(dtype).np_type), nrows=self._nrows, ncols=self._ncols)
        gb_obj = ffi.new("GrB_Matrix*")
        call("GrB_Matrix_dup", [gb_obj, self])
        return Matrix._wrap(gb_obj, self.dtype, self._nrows, self._ncols, name)

    def _as_expression(self):
        raise TypeError("Matrix << Matrix is not supported by this backend; use .dup()")

    def isequal(self, other, *, check_dtype=False):
        if isinstance(other, TransposedMatrix):
            other = other.new()
        if not isinstance(other, Matrix):
            raise TypeError(f"Expected Matrix; got {type(other).__name__}")
        if check_dtype and self.dtype is not other.dtype:
            return False
        if self.shape != other.shape or self.nvals != other.nvals:
            return False
        if self.nvals == 0:
            return True
        # reference core/matrix.py:387-398: eWiseMult(eq) + reduce(land), all on the device
        matches = Matrix(dtypes.BOOL, self._nrows, self._ncols)
        matches << self.ewise_mult(other, binary.eq)
        if matches.nvals != self.nvals:
            return False
        return bool(matches.reduce_scalar(monoid.land).value)

    def isclose(self, other, *, rel_tol=1e-7, abs_tol=0.0, check_dtype=False):
        if self.shape != other.shape or self.nvals != other.nvals:
            return False
        r1, c1, v1 = self.to_coo()
        r2, c2, v2 = other.to_coo()
        return bool(np.array_equal(r1, r2) and np.array_equal(c1, c2) and np.allclose(v1, v2, rtol=rel_tol, atol=abs_tol))

    def __getitem__(self, key):
        i, j = key
        out = ffi.new(f"{_CTYPE[self.dtype.name]}*")
        info = call(f"GrB_Matrix_extractElement_{self.dtype.name}", [out, self, int(i), int(j)])
        s = Scalar(self.dtype)
        if info != lib.GrB_NO_VALUE:
            s.value = out[0]
        return _ScalarResult(s)

    def __setitem__(self, key, value):
        i, j = key
        call(f"GrB_Matrix_setElement_{self.dtype.name}", [self, self.dtype.np_type.type(value).item(), int(i), int(j)])

    def __delitem__(self, key):
        i, j = key
        call("GrB_Matrix_removeElement", [self, int(i), int(j)])

    # ---- the hot path ----
    def mxv(self, other, op=semiring.plus_times):
        """w = A (+).(x) u  ->  GrB_mxv (reference core/matrix.py:2163-2204)."""
        if not isinstance(other, Vector):
            raise TypeError(f"Expected Vector for `other`; got {type(other).__name__}")
        op = _semiring_for(op, self, other)
        expr = Expression("mxv", "GrB_mxv", [self, other], op, Vector, (self.nrows,), at=self._is_transposed)
        if self.ncols != other.size:
            expr.new()  # DimensionMismatch comes from the C library, as in the reference (:2202-2203)
        return expr

    def mxm(self, other, op=semiring.plus_times):
        """C = A (+).(x) B  ->  GrB_mxm (reference core/matrix.py:2206-2251)."""
        if not isinstance(other, (Matrix, TransposedMatrix)):
            raise TypeError(f"Expected Matrix or TransposedMatrix for `other`; got {type(other).__name__}")
        op = _semiring_for(op, self, other)
        expr = Expression("mxm", "GrB_mxm", [self, other], op, Matrix, (self.nrows, other.ncols),
                          at=self._is_transposed, bt=other._is_transposed)
        if self.ncols != other.nrows:
            expr.new()
        return expr

    def power(self, n, op=semiring.plus_times):
        """A^n by repeated squaring over mxm (reference core/matrix.py:95-150, 2754)."""
        if not isinstance(n, (int, np.integer)):
            raise TypeError(f"n must be a positive integer; got {n!r}")
        if n <= 0:
            raise ValueError(f"n must be a positive integer; got {n}")
        if self.nrows != self.ncols:
            raise errors.DimensionMismatch(f"power only works for square Matrix; shape is {self.shape}")
        result, base = None, (self.new() if self._is_transposed else self)
        while n:
            if n & 1:
                result = base if result is None else result.mxm(base, op).new()
            n >>= 1
            if n:
                base = base.mxm(base, op).new()
        return _Materialized(result)

    # ---- helpers around the products ----
    def reduce_rowwise(self, op=monoid.plus):
        """w(i) = (+)_j A(i,j) -> GrB_Matrix_reduce_Monoid (reference core/matrix.py:2754+, aggregators agg.py:303-321)."""
        op = get_typed_op(op, self.dtype, kind="monoid")
        return Expression("reduce_rowwise", "GrB_Matrix_reduce_Monoid", [self], op, Vector, (self.nrows,),
                          at=self._is_transposed)

    def reduce_columnwise(self, op=monoid.plus):
        op = get_typed_op(op, self.dtype, kind="monoid")
        return Expression("reduce_columnwise", "GrB_Matrix_reduce_Monoid", [self], op, Vector, (self.ncols,),
                          at=not self._is_transposed)

    def reduce_scalar(self, op=monoid.plus):
        op = get_typed_op(op, self.dtype, kind="monoid")
        return _ReduceExpr(self, op, "Matrix")

    def select(self, which, thunk=0):
        """tril / triu / diag / offdiag (builds the triangle-count operand L on the device)."""
        opname = {"tril": "GrB_TRIL", "triu": "GrB_TRIU", "diag": "GrB_DIAG", "offdiag": "GrB_OFFDIAG"}[which]
        return Expression("select", "GrB_Matrix_select_INT64", [self], _Handle(getattr(lib, opname), return_type=self.dtype), Matrix, self.shape,
                          scalar_args=(int(thunk),), at=self._is_transposed)

    def ewise_mult(self, other, op=binary.times):
        op = _binary_for(op, self, other)
        return Expression("ewise_mult", "GrB_Matrix_eWiseMult_BinaryOp", [self, other], op, Matrix, self.shape)

    def ewise_add(self, other, op=binary.plus):
        op = _binary_for(op, self, other)
        return Expression("ewise_add", "GrB_Matrix_eWiseAdd_BinaryOp", [self, other], op, Matrix, self.shape)

    def __repr__(self):
        return f"Matrix(dtype={self.dtype}, nrows={self._nrows}, ncols={self._ncols}, nvals={self.nvals})"


class _Materialized:
    def __init__(self, obj):
        self._obj = obj

    def new(self, dtype=None, **kw):
        return self._obj.dup(dtype) if dtype is not None else self._obj


class TransposedMatrix:
    """`A.T`: same handle, transposed flag folded into the descriptor (reference core/matrix.py:3614-3700)."""
    _is_transposed = True
    _is_scalar = False

    def __init__(self, matrix):
        self._matrix = matrix
        self.dtype = matrix.dtype

    @property
    def _carg(self):
        return self._matrix._carg

    @property
    def nrows(self):
        return self._matrix._ncols

    @property
    def ncols(self):
        return self._matrix._nrows

    @property
    def shape(self):
        return (self.nrows, self.ncols)

    @property
    def T(self):
        return self._matrix

    def new(self, dtype=None, *, mask=None, name=None, **opts):
        out = Matrix(lookup_dtype(dtype) if dtype else self.dtype, self.nrows, self.ncols, name=name)
        comp = structure = False
        if mask is not None:
            mask = _check_mask(mask, out)
            comp, structure = mask.complement, mask.structure
        call("GrB_transpose", [out, mask, None, self._matrix, _descriptor(False, comp, structure, **opts)])
        return out

    mxv = Matrix.mxv
    mxm = Matrix.mxm
    power = Matrix.power
    select = Matrix.select
    reduce_rowwise = Matrix.reduce_rowwise
    reduce_columnwise = Matrix.reduce_columnwise
    __matmul__ = BaseType.__matmul__

    def __repr__(self):
        return f"TransposedMatrix({self._matrix!r})"
