"""Test harness: import the REAL python-graphblas (from /root/reference, read-only, this container only) over grb_b200.

What it proves without a GPU:
  * the reference's operator registry (regex scan of dir(lib), graphblas/core/operator/base.py:291,410-486) and every
    module-level `lib.X` access find what they need in include/GraphBLAS_b200.h / libgrb_b200.so;
  * the argument lists the reference's expression machinery assembles (graphblas/core/base.py:438-483) convert against
    the C signatures this repo declares: every function of `lib` is replaced by a cffi callback OF THE SAME C TYPE that
    records the call and reports success, so cffi still type-checks each argument against the header.

The real functions need a CUDA device (GrB_init fails without one: there is no CPU fallback), and /root/reference does not
travel to the GPU box, so the kernels are exercised through the mirror (graphblas_b200) there and the reference's Python
is exercised here.  Environment shims only: `donfig` (absent from the image) and `np.cast` (removed in NumPy 2).
"""
import contextlib
import os
import sys
import types

REFERENCE = "/root/reference"


def available():
    return os.path.isdir(os.path.join(REFERENCE, "graphblas"))


class DryLib:
    """cffi `lib` stand-in: data symbols and enum constants are the real ones (so handles compare equal to the exported
    globals of libgrb_b200.so); functions are cffi callbacks with the declared signature that log (name, args)."""

    def __init__(self, ffi, lib):
        object.__setattr__(self, "_ffi", ffi)
        object.__setattr__(self, "_lib", lib)
        object.__setattr__(self, "_cache", {})
        object.__setattr__(self, "calls", [])
        object.__setattr__(self, "_fake", [0x10000])

    def __dir__(self):
        return dir(self._lib)

    def __getattr__(self, name):
        cache = self._cache
        if name in cache:
            return cache[name]
        val = getattr(self._lib, name)          # AttributeError passes through, as with the real lib
        ffi = self._ffi
        if isinstance(val, ffi.CData) and ffi.typeof(val).kind == "function":
            val = ffi.callback(ffi.typeof(val), self._make(name))
        cache[name] = val
        return val

    def _make(self, name):
        ffi, calls, fake = self._ffi, self.calls, self._fake

        def impl(*args):
            calls.append((name, args))
            first = args[0] if args else None
            if isinstance(first, ffi.CData) and ffi.typeof(first).kind == "pointer" and first != ffi.NULL:
                item = ffi.typeof(first).item
                if name.endswith("_new") and item.kind == "pointer":
                    fake[0] += 0x100
                    first[0] = ffi.cast(item, fake[0])            # a distinct non-NULL handle
                elif name.endswith(("_nvals", "_nrows", "_ncols", "_size")):
                    first[0] = 0
                elif name.endswith("_error"):
                    first[0] = ffi.NULL
            return 0   # GrB_SUCCESS

        return impl


def load():
    """Import graphblas from /root/reference over a dry grb_b200; returns (graphblas module, DryLib)."""
    import numpy as np

    if not hasattr(np, "cast"):        # graphblas/core/utils.py:46 (NumPy < 2 API)

        class _Cast(dict):
            def __missing__(self, key):
                dt = np.dtype(key)
                return lambda x: np.asarray(x).astype(dt)

        np.cast = _Cast()
    if "donfig" not in sys.modules:    # graphblas/__init__.py:22-33 uses Config(name), update_defaults, get, set
        donfig = types.ModuleType("donfig")

        class Config:
            def __init__(self, name):
                self.name, self.config = name, {}

            def update_defaults(self, d):
                for k, v in d.items():
                    self.config.setdefault(k, v)

            def get(self, key, default=None):
                return self.config.get(key, default)

            @contextlib.contextmanager
            def _restore(self, old):
                try:
                    yield self
                finally:
                    self.config = old

            def set(self, arg=None, **kwargs):
                old = dict(self.config)
                self.config.update(arg or {})
                self.config.update(kwargs)
                return self._restore(old)

        donfig.Config = Config
        sys.modules["donfig"] = donfig
    import grb_b200
    import grb_b200.utils

    dry = DryLib(grb_b200.ffi, grb_b200.lib)
    shim = types.ModuleType("suitesparse_graphblas")
    shim.ffi, shim.lib = grb_b200.ffi, dry
    shim.__version__ = grb_b200.__version__
    shim.is_initialized = lambda: False
    shim.initialize = lambda **kwargs: None      # the real one needs a CUDA device
    shim.utils = grb_b200.utils
    sys.modules["suitesparse_graphblas"] = shim
    sys.modules["suitesparse_graphblas.utils"] = grb_b200.utils
    if REFERENCE not in sys.path:
        sys.path.insert(0, REFERENCE)
    import graphblas as gb

    gb.init("suitesparse")
    return gb, dry
