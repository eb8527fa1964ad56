"""GrB_Info -> exception classes, same names and codes as reference graphblas/exceptions.py:10-117."""


class GraphblasException(Exception):
    pass


class NoValue(GraphblasException):
    pass


class UninitializedObject(GraphblasException):
    pass


class InvalidObject(GraphblasException):
    pass


class NullPointer(GraphblasException):
    pass


class InvalidValue(GraphblasException):
    pass


class InvalidIndex(GraphblasException):
    pass


class DomainMismatch(GraphblasException):
    pass


class DimensionMismatch(GraphblasException):
    pass


class OutputNotEmpty(GraphblasException):
    pass


class EmptyObject(GraphblasException):
    pass


class OutOfMemory(GraphblasException):
    pass


class IndexOutOfBound(GraphblasException):
    pass


class Panic(GraphblasException):
    pass


class InsufficientSpace(GraphblasException):
    pass


class NotImplementedException(GraphblasException):
    pass


_BY_CODE = {
    1: NoValue, -1: UninitializedObject, -2: NullPointer, -3: InvalidValue, -4: InvalidIndex,
    -5: DomainMismatch, -6: DimensionMismatch, -7: OutputNotEmpty, -8: NotImplementedException,
    -101: Panic, -102: OutOfMemory, -103: InsufficientSpace, -104: InvalidObject, -105: IndexOutOfBound,
    -106: EmptyObject,
}


def error_for(info):
    return _BY_CODE.get(int(info), GraphblasException)
