"""graphblas_b200 — the host-side mirror of python-graphblas's API for the semiring hot path, running on
libgrb_b200.so (hand-written CUDA for sm_100a).  Same spelling as the reference:

    import graphblas_b200 as gb
    from graphblas_b200 import Matrix, Vector, semiring, binary, monoid, dtypes
    gb.init("b200")
    C(M.S, accum=binary.plus) << A.mxm(B.T, semiring.plus_pair)

`init()` mirrors reference graphblas/__init__.py:105-197 (backend selection + initialize(blocking,
memory_manager="numpy")); the only accepted backend is "b200", and there is no CPU fallback.
"""
import grb_b200 as _binding

backend = "b200"
_init_params = None


def init(backend_name="b200", blocking=False):
    global _init_params
    if backend_name != "b200":
        raise ValueError(f'Bad backend name. Must be "b200"; got {backend_name!r} (there is no CPU fallback)')
    params = dict(backend=backend_name, blocking=blocking)
    if _init_params is not None:
        if _init_params != params:
            raise RuntimeError(f"graphblas_b200 is already initialized with {_init_params}")
        return
    if not _binding.is_initialized():
        _binding.initialize(blocking=blocking, memory_manager="numpy")
    _init_params = params


def _autoinit():
    if _init_params is None:
        init()


from grb_b200 import ffi, lib  # noqa: E402
from grb_b200.errors import (  # noqa: E402,F401
    DimensionMismatch, DomainMismatch, EmptyObject, GraphblasException, IndexOutOfBound, InsufficientSpace,
    InvalidIndex, InvalidObject, InvalidValue, NoValue, NotImplementedException, NullPointer, OutOfMemory,
    OutputNotEmpty, Panic, UninitializedObject)

from . import dtypes  # noqa: E402
from .operators import binary, monoid, semiring, unary  # noqa: E402
from .core import Matrix, Scalar, TransposedMatrix, Vector, replace  # noqa: E402
from . import io  # noqa: E402,F401  (from_scipy_sparse / to_scipy_sparse, reference graphblas/io/_scipy.py)


def use_torch_stream():
    """Run libgrb_b200 on torch's current CUDA stream so torch.cuda.Event timing brackets its kernels."""
    import torch

    lib.GxB_b200_set_stream(ffi.cast("void*", torch.cuda.current_stream().cuda_stream))


def launch_count():
    return int(lib.GxB_b200_launch_count())


def last_kernel():
    return ffi.string(lib.GxB_b200_last_kernel()).decode()


def set_option(name, value):
    info = lib.GxB_b200_set_option(name.encode(), int(value))
    if info != lib.GrB_SUCCESS:
        raise ValueError(f"unknown option {name!r}")


def sync():
    lib.GxB_b200_sync()
