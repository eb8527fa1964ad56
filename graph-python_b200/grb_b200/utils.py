"""numpy allocator plumbing: the counterpart of suitesparse_graphblas.utils (claim_buffer & co,
reference graphblas/core/ss/matrix.py:4, :895-897, :1349-1351)."""
import ctypes

import numpy as np

_libc = ctypes.CDLL(None)
_libc.malloc.restype = ctypes.c_void_p
_libc.malloc.argtypes = [ctypes.c_size_t]
_libc.free.argtypes = [ctypes.c_void_p]


def numpy_allocators(ffi):
    """Function pointers (malloc, calloc, realloc, free) compatible with numpy's default data allocator
    (which is the C allocator unless NUMPY's tracemalloc domain hooks are replaced)."""
    c = ffi.dlopen(None)
    ffi_local = ffi
    try:
        ffi_local.cdef("void *malloc(size_t); void *calloc(size_t, size_t); void *realloc(void *, size_t); void free(void *);")
    except Exception:  # already declared
        pass
    return (ffi_local.addressof(c, "malloc"), ffi_local.addressof(c, "calloc"),
            ffi_local.addressof(c, "realloc"), ffi_local.addressof(c, "free"))


class _Owner:
    """Frees a malloc'd buffer when the last numpy view of it dies."""

    def __init__(self, addr):
        self.addr = addr

    def __del__(self):
        if self.addr:
            _libc.free(ctypes.c_void_p(self.addr))
            self.addr = 0


def claim_buffer(ffi, cdata, size, dtype):
    """Wrap a library-allocated buffer (from GxB_*_unpack_*) in a numpy array that owns it."""
    dtype = np.dtype(dtype)
    addr = int(ffi.cast("uintptr_t", cdata))
    nbytes = int(size) * dtype.itemsize
    buf = (ctypes.c_char * nbytes).from_address(addr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(size))
    owner = _Owner(addr)
    # keep the owner alive as long as any view of the buffer exists
    buf._grb_owner = owner
    return arr


def claim_buffer_2d(ffi, cdata, cdata_size, nrows, ncols, dtype, is_c_order):
    arr = claim_buffer(ffi, cdata, cdata_size, dtype)[: nrows * ncols]
    return arr.reshape((nrows, ncols) if is_c_order else (ncols, nrows)) if is_c_order else \
        arr.reshape((ncols, nrows)).T


def unclaim_buffer(arr):
    """The library took ownership of arr's memory (GxB_*_pack_*): numpy must not free it."""
    arr.flags.writeable = False  # best effort; OWNDATA cannot be cleared from Python without the C API


def malloc_copy(ffi, arr):
    """Copy a numpy array into a malloc'd buffer the library may take ownership of (pack)."""
    arr = np.ascontiguousarray(arr)
    nbytes = max(arr.nbytes, 1)
    addr = _libc.malloc(nbytes)
    ctypes.memmove(addr, arr.ctypes.data, arr.nbytes)
    return ffi.cast("void *", addr), nbytes
