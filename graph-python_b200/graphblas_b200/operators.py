"""Semiring / Monoid / BinaryOp namespaces discovered from the symbols libgrb_b200 exports.

Like the reference (graphblas/core/operator/base.py:291, semiring.py:170-204) the registry is built by
scanning dir(lib) with name patterns, so ``semiring.min_plus[INT64]._carg`` is the exported global
``GrB_MIN_PLUS_SEMIRING_INT64``.  A semiring without an sm_100a kernel simply has no symbol: asking for it
raises (north_star: unsupported / user-defined semirings raise, there is no CPU fallback)."""
import re
import types

from grb_b200 import lib

from . import dtypes

_TYPES = "BOOL|INT8|INT16|INT32|INT64|UINT8|UINT16|UINT32|UINT64|FP32|FP64"
_SEMIRING_RES = [
    re.compile(rf"^GrB_([A-Z]+)_([A-Z]+)_SEMIRING_({_TYPES})$"),
    re.compile(rf"^GxB_(PLUS|MIN|MAX|ANY|TIMES|LOR|LAND|LXOR|LXNOR)_([A-Z]+)_({_TYPES})$"),
]
_MONOID_RES = [
    re.compile(rf"^GrB_([A-Z]+)_MONOID_({_TYPES})$"),
    re.compile(rf"^GxB_(ANY)_({_TYPES})_MONOID$"),
]
_BINARY_RES = [
    re.compile(rf"^G[rx]B_(FIRST|SECOND|ONEB|PAIR|ANY|PLUS|MINUS|TIMES|DIV|MIN|MAX|EQ|NE|GT|LT|GE|LE)_({_TYPES})$"),
]
_UNARY_RES = [re.compile(rf"^G[rx]B_(IDENTITY|AINV|MINV|ABS|ONE|BNOT)_({_TYPES})$")]
_BOOL_BINARY = {"GrB_LOR": "lor", "GrB_LAND": "land", "GrB_LXOR": "lxor", "GrB_LXNOR": "lxnor"}


class TypedOp:
    __slots__ = ("parent", "dtype", "gb_name", "_carg", "return_type")

    def __init__(self, parent, dtype, gb_name, return_type):
        self.parent = parent
        self.dtype = dtype
        self.gb_name = gb_name
        self._carg = getattr(lib, gb_name)
        self.return_type = return_type

    @property
    def name(self):
        return self.parent.name

    @property
    def opclass(self):
        return self.parent.opclass

    def __repr__(self):
        return f"{self.parent.opclass.lower()}.{self.parent.name}[{self.dtype}]"

    def __call__(self, expr):
        return self.parent(expr, _typed=self)


class Op:
    opclass = "Op"

    def __init__(self, name):
        self.name = name
        self._typed = {}

    def _add(self, dtype, gb_name, return_type=None):
        # the library exports both spellings of the official semirings (GrB_MIN_PLUS_SEMIRING_T and GxB_MIN_PLUS_T), as
        # SuiteSparse does; the GrB_ handle wins, as in the reference (tests/test_op.py:88-91)
        have = self._typed.get(dtype.name)
        if have is not None and have.gb_name.startswith("GrB_") and gb_name.startswith("GxB_"):
            return
        self._typed[dtype.name] = TypedOp(self, dtype, gb_name, return_type or dtype)

    @property
    def types(self):
        return sorted(self._typed)

    def __getitem__(self, dtype):
        dt = dtypes.lookup_dtype(dtype)
        try:
            return self._typed[dt.name]
        except KeyError:
            raise KeyError(
                f"{self.opclass.lower()}.{self.name} has no kernel for {dt} in libgrb_b200 "
                f"(available: {', '.join(self.types)})"
            ) from None

    def __contains__(self, dtype):
        try:
            return dtypes.lookup_dtype(dtype).name in self._typed
        except ValueError:
            return False

    def __repr__(self):
        return f"{self.opclass.lower()}.{self.name}"


class BinaryOp(Op):
    opclass = "BinaryOp"


class UnaryOp(Op):
    opclass = "UnaryOp"


class Monoid(Op):
    opclass = "Monoid"

    @property
    def binaryop(self):
        return getattr(binary, self.name)


class Semiring(Op):
    opclass = "Semiring"

    def __call__(self, expr, _typed=None):
        """semiring.min_plus(A @ B): re-type an infix matmul expression."""
        from .core import MatMulInfix

        if not isinstance(expr, MatMulInfix):
            raise TypeError(f"{self!r} can only be applied to an `A @ B` expression")
        return expr._with_op(_typed or self)


unary = types.SimpleNamespace()
binary = types.SimpleNamespace()
monoid = types.SimpleNamespace()
semiring = types.SimpleNamespace()


def _get(ns, cls, name):
    op = getattr(ns, name, None)
    if op is None:
        op = cls(name)
        setattr(ns, name, op)
    return op


def _initialize():
    for sym in dir(lib):
        for r in _SEMIRING_RES:
            m = r.match(sym)
            if m:
                mon, mul, t = m.groups()
                _get(semiring, Semiring, f"{mon.lower()}_{mul.lower()}")._add(dtypes.lookup_dtype(t), sym)
                break
        else:
            for r in _MONOID_RES:
                m = r.match(sym)
                if m:
                    mon, t = m.groups()
                    _get(monoid, Monoid, mon.lower())._add(dtypes.lookup_dtype(t), sym)
                    break
            else:
                for r in _BINARY_RES:
                    m = r.match(sym)
                    if m:
                        op, t = m.groups()
                        ret = dtypes.BOOL if op in ("EQ", "NE", "GT", "LT", "GE", "LE") else None
                        name = "pair" if op == "ONEB" else op.lower()
                        if sym.startswith("GrB_ONEB"):
                            _get(binary, BinaryOp, "oneb")._add(dtypes.lookup_dtype(t), sym, ret)
                        _get(binary, BinaryOp, name)._add(dtypes.lookup_dtype(t), sym, ret)
                        break
                else:
                    if sym in _BOOL_BINARY:
                        _get(binary, BinaryOp, _BOOL_BINARY[sym])._add(dtypes.BOOL, sym)
                    elif sym == "GrB_LNOT":
                        _get(unary, UnaryOp, "lnot")._add(dtypes.BOOL, sym)
                    else:
                        m = _UNARY_RES[0].match(sym)
                        if m:
                            op, t = m.groups()
                            _get(unary, UnaryOp, op.lower())._add(dtypes.lookup_dtype(t), sym)
    # on BOOL, plus/times/min/max are lor/land/land/lor (C API table of built-in operators); alias them so
    # semiring.plus_times[bool] style lookups resolve the way the reference's coercion tables do
    # (reference graphblas/core/operator/semiring.py:391-510)
    for name, alias in (("lor_land", ("plus_times", "max_min", "max_times")), ("land_lor", ("min_max", "times_plus")),
                        ("any_land", ("any_times", "any_min")), ("lor_pair", ("plus_pair", "max_pair"))):
        base = getattr(semiring, name, None)
        if base is None or "BOOL" not in base._typed:
            continue
        for a in alias:
            op = _get(semiring, Semiring, a)
            if "BOOL" not in op._typed:
                op._typed["BOOL"] = base._typed["BOOL"]
    for name, alias in (("lor", ("plus", "max")), ("land", ("times", "min"))):
        for ns, cls in ((monoid, Monoid), (binary, BinaryOp)):
            base = getattr(ns, name, None)
            if base is None or "BOOL" not in base._typed:
                continue
            for a in alias:
                op = _get(ns, cls, a)
                if "BOOL" not in op._typed:
                    op._typed["BOOL"] = base._typed["BOOL"]
    # binary.truediv: true division, defined on the floating-point types where it equals GrB_DIV (the reference
    # builds it for integers too by casting to FP64, graphblas/core/operator/binary.py; not needed on this path)
    div = getattr(binary, "div", None)
    if div is not None:
        td = _get(binary, BinaryOp, "truediv")
        for t in ("FP32", "FP64"):
            if t in div._typed:
                td._typed[t] = div._typed[t]


    # Division semirings as the reference names them (graphblas/core/operator/semiring.py: `plus_div` is hidden; `plus_cdiv`
    # is C division on the integer types, `plus_truediv` divides in floating point -- integer operands are cast to FP64
    # by the library, the result is FP64; tests/test_op.py:950-965)
    pdiv = getattr(semiring, "plus_div", None)
    if pdiv is not None:
        cdiv_, tdiv = _get(semiring, Semiring, "plus_cdiv"), _get(semiring, Semiring, "plus_truediv")
        for t, typed in pdiv._typed.items():
            if t.startswith("FP"):
                tdiv._typed[t] = typed
            else:
                cdiv_._typed[t] = typed
        if "FP64" in pdiv._typed:
            for t in ("INT8", "INT16", "INT32", "INT64", "UINT8", "UINT16", "UINT32", "UINT64", "BOOL"):
                tdiv._typed.setdefault(t, pdiv._typed["FP64"])
        delattr(semiring, "plus_div")


_initialize()

_STRING_BINARY = {"+": "plus", "-": "minus", "*": "times", "/": "div", "min": "min", "max": "max",
                  "|": "lor", "&": "land", "^": "lxor", "==": "eq", "!=": "ne", "<": "lt", ">": "gt"}


def binary_from_string(s):
    name = _STRING_BINARY.get(s, s)
    op = getattr(binary, name, None)
    if op is None:
        raise ValueError(f"Unknown binary operator string: {s!r}")
    return op


def get_typed_op(op, dtype, dtype2=None, *, kind):
    """Resolve an (un)typed operator for the operand types (reference core/operator/utils.py:38-60)."""
    if isinstance(op, TypedOp):
        return op
    if isinstance(op, str):
        ns = {"unary": unary, "binary": binary, "monoid": monoid, "semiring": semiring}[kind]
        op = binary_from_string(op) if kind == "binary" else getattr(ns, op)
    if not isinstance(op, Op):
        raise TypeError(f"Invalid operator: {op!r}")
    dt = dtypes.unify(dtype, dtype2) if dtype2 is not None else dtypes.lookup_dtype(dtype)
    return op[dt]
