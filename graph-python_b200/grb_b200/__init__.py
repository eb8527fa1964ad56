"""grb_b200 — cffi binding of libgrb_b200.so with the surface python-graphblas imports from
``suitesparse_graphblas``:  ``from grb_b200 import ffi, lib, initialize, is_initialized``
(reference graphblas/__init__.py:141, :156, :171) plus ``grb_b200.utils.claim_buffer/unclaim_buffer``
(reference graphblas/core/ss/matrix.py:4) and ``__version__`` (reference graphblas/core/ss/__init__.py:1-3).

ABI mode: the declarations are read from include/GraphBLAS_b200.h (the region between the cffi markers)
and the shared library is dlopen'ed in-tree, so the loaded native code is visible as
graph-python_b200/libgrb_b200.so.  There is no CPU fallback: without the library, or without a CUDA device,
loading / initialize() raises.
"""
import os
import re
import threading

import cffi

__version__ = "7.4.0.0+b200.1"  # major 7: python-graphblas takes its SuiteSparse-7 code paths

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
HEADER = os.path.join(_ROOT, "include", "GraphBLAS_b200.h")
LIBRARY = os.path.join(os.path.dirname(_HERE), "libgrb_b200.so")


def _cdef_text():
    with open(HEADER) as f:
        text = f.read()
    m = re.search(r"/\* cffi-begin \*/(.*)/\* cffi-end \*/", text, re.S)
    if m is None:
        raise RuntimeError(f"{HEADER}: cffi markers not found")
    return m.group(1)


ffi = cffi.FFI()
ffi.cdef(_cdef_text())
if not os.path.exists(LIBRARY):
    raise ImportError(
        f"{LIBRARY} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C graph-python_b200/csrc`). grb_b200 has no CPU fallback."
    )
lib = ffi.dlopen(LIBRARY)

_lock = threading.Lock()
_initialized = False
_keepalive = []


def is_initialized():
    return _initialized


def initialize(*, blocking=False, memory_manager="numpy"):
    """GrB_init / GxB_init.  memory_manager="numpy" registers numpy's allocator so that buffers handed
    over with GxB_*_pack_* / received from GxB_*_unpack_* can be claimed by numpy
    (reference graphblas/__init__.py:171)."""
    global _initialized
    with _lock:
        if _initialized:
            raise RuntimeError("GraphBLAS is already initialized")
        mode = lib.GrB_BLOCKING if blocking else lib.GrB_NONBLOCKING
        if memory_manager == "numpy":
            from . import utils

            info = lib.GxB_init(mode, *utils.numpy_allocators(ffi))
        elif memory_manager == "c":
            info = lib.GrB_init(mode)
        else:
            raise ValueError(f'memory_manager must be "numpy" or "c"; got {memory_manager!r}')
        if info != lib.GrB_SUCCESS:
            raise RuntimeError(
                f"GrB_init failed with GrB_Info {info}: libgrb_b200 needs a CUDA device (sm_100a) "
                "and does not fall back to the CPU"
            )
        _initialized = True


def check(info, obj=None, kind="Matrix"):
    """Raise on a GrB_Info error code (mirrors reference graphblas/exceptions.py:119-155)."""
    if info == lib.GrB_SUCCESS or info == lib.GrB_NO_VALUE:
        return info
    msg = ""
    if obj is not None:
        s = ffi.new("char**")
        getattr(lib, f"GrB_{kind}_error")(s, obj)
        if s[0] != ffi.NULL:
            msg = ffi.string(s[0]).decode()
    from .errors import error_for

    raise error_for(info)(msg)
