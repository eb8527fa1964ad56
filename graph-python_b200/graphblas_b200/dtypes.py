"""Built-in GraphBLAS data types bound to libgrb_b200's GrB_<TYPE> handles.
Mirrors the lookup/unify behaviour of reference graphblas/core/dtypes.py:154-282, 377-394
(numpy promotion rules decide the type a semiring runs in)."""
import numpy as np

from grb_b200 import lib


class DataType:
    __slots__ = ("name", "np_type", "gb_name", "_carg")

    def __init__(self, name, np_type):
        self.name = name
        self.np_type = np.dtype(np_type)
        self.gb_name = f"GrB_{name}"
        self._carg = getattr(lib, self.gb_name)

    def __repr__(self):
        return self.name

    def __eq__(self, other):
        try:
            return lookup_dtype(other) is self
        except (TypeError, ValueError):
            return False

    def __hash__(self):
        return hash(self.name)


BOOL = DataType("BOOL", np.bool_)
INT8 = DataType("INT8", np.int8)
INT16 = DataType("INT16", np.int16)
INT32 = DataType("INT32", np.int32)
INT64 = DataType("INT64", np.int64)
UINT8 = DataType("UINT8", np.uint8)
UINT16 = DataType("UINT16", np.uint16)
UINT32 = DataType("UINT32", np.uint32)
UINT64 = DataType("UINT64", np.uint64)
FP32 = DataType("FP32", np.float32)
FP64 = DataType("FP64", np.float64)

_ALL = [BOOL, INT8, INT16, INT32, INT64, UINT8, UINT16, UINT32, UINT64, FP32, FP64]
_BY_NAME = {t.name: t for t in _ALL}
_BY_NP = {t.np_type: t for t in _ALL}
_BY_PY = {bool: BOOL, int: INT64, float: FP64}


def lookup_dtype(x):
    if isinstance(x, DataType):
        return x
    if isinstance(x, str):
        key = x.upper()
        if key in _BY_NAME:
            return _BY_NAME[key]
        key = {"FLOAT": "FP64", "INT": "INT64", "DOUBLE": "FP64"}.get(key)
        if key:
            return _BY_NAME[key]
    if x in _BY_PY:
        return _BY_PY[x]
    try:
        return _BY_NP[np.dtype(x)]
    except (TypeError, KeyError):
        raise ValueError(f"Unknown dtype: {x!r}") from None


def unify(a, b):
    """Type two operands are promoted to (numpy rules), as reference core/dtypes.py:377-394."""
    a, b = lookup_dtype(a), lookup_dtype(b)
    if a is b:
        return a
    return lookup_dtype(np.promote_types(a.np_type, b.np_type))
