"""GPU: GxB_b200_mxv_plus_times_FP32_noalloc (the allocation-free product; scratch in a static __device__ array) through the
C-ABI against the existing entry point GrB_mxv on the same matrix, two calls in a row, within the FP32 plus_times tolerance
the other parity tests use (1e-5 relative).  Collected last on purpose (file name): the round's end runs `pytest -x`."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FP_TOL = 1e-5


@pytest.fixture(scope="module")
def gb():
    import graphblas_b200 as gb

    gb.init("b200")
    return gb


def _device_ptrs_vec(v):
    import grb_b200

    ffi, lib = grb_b200.ffi, grb_b200.lib
    vals, pres = ffi.new("void**"), ffi.new("void**")
    assert lib.GxB_b200_Vector_device_ptrs(vals, pres, v._carg) == lib.GrB_SUCCESS
    return vals[0], pres[0]


def _device_ptrs_mat(A):
    import grb_b200

    ffi, lib = grb_b200.ffi, grb_b200.lib
    rp, ci, vx = ffi.new("void**"), ffi.new("void**"), ffi.new("void**")
    assert lib.GxB_b200_Matrix_device_ptrs(rp, ci, vx, A._carg) == lib.GrB_SUCCESS
    return rp[0], ci[0], vx[0]


def _noalloc(A, u, out, n, nnz, want_count=True):
    """out: an FP32 GrB_Vector whose device buffers serve as the caller-provided output arrays."""
    import grb_b200

    ffi, lib = grb_b200.ffi, grb_b200.lib
    rp, ci, vx = _device_ptrs_mat(A)
    uv, _ = _device_ptrs_vec(u)
    wv, wp = _device_ptrs_vec(out)
    cnt = ffi.new("GrB_Index*") if want_count else ffi.NULL
    err = ffi.new("char**")
    info = lib.GxB_b200_mxv_plus_times_FP32_noalloc(ffi.cast("float*", wv), ffi.cast("uint8_t*", wp), cnt,
                                                   ffi.cast("int64_t*", rp), ffi.cast("uint32_t*", ci),
                                                   ffi.cast("float*", vx), ffi.cast("float*", uv), n, n, nnz, err)
    assert info == lib.GrB_SUCCESS, ffi.string(err[0])
    # the library did not write `out` itself: tell it to recount the presence bytes on the device
    assert lib.GxB_b200_Vector_set_dense_nvals(out._carg, -1) == lib.GrB_SUCCESS
    return int(cnt[0]) if want_count else None


@pytest.mark.parametrize("scale", [10, 16])
def test_noalloc_matches_mxv_twice(gb, scale):
    from graphblas_b200 import Matrix, Vector, semiring

    n = 1 << scale
    A = Matrix.rmat(scale, "FP32", value_mode=2)        # directed R-MAT: has empty rows, hub rows spanning many tiles
    nnz = A.nvals
    rng = np.random.default_rng(scale)
    out = Vector.from_dense(np.zeros(n, np.float32))    # caller-provided output buffers (contents are overwritten)
    for call in range(2):                               # two calls in a row share the static scratch
        x = rng.random(n, dtype=np.float32)
        u = Vector.from_dense(x)
        ref = A.mxv(u, semiring.plus_times).new()       # the existing entry point: GrB_mxv
        count = _noalloc(A, u, out, n, nnz)
        rv, rp = ref.to_dense()
        ov, op = out.to_dense()
        assert np.array_equal(op.astype(bool), rp.astype(bool)), call
        p = rp.astype(bool)
        assert count == int(p.sum()) == ref.nvals
        assert np.allclose(ov[p].astype(np.float64), rv[p].astype(np.float64), rtol=FP_TOL, atol=FP_TOL), call


def test_noalloc_rejects_sizes_beyond_the_static_scratch(gb):
    import grb_b200

    ffi, lib = grb_b200.ffi, grb_b200.lib
    assert lib.GXB_B200_NOALLOC_MAX_NVALS == 1 << 26
    err = ffi.new("char**")
    one = ffi.cast("float*", 256)          # never dereferenced: the size check comes first
    info = lib.GxB_b200_mxv_plus_times_FP32_noalloc(one, ffi.cast("uint8_t*", 256), ffi.NULL, ffi.cast("int64_t*", 256),
                                                   ffi.cast("uint32_t*", 256), one, one, 10, 10,
                                                   lib.GXB_B200_NOALLOC_MAX_NVALS + 1, err)
    assert info == lib.GrB_INSUFFICIENT_SPACE
    assert b"GXB_B200_NOALLOC_MAX_NVALS" in ffi.string(err[0])
    info = lib.GxB_b200_mxv_plus_times_FP32_noalloc(ffi.NULL, ffi.NULL, ffi.NULL, ffi.NULL, ffi.NULL, ffi.NULL, ffi.NULL,
                                                   1, 1, 0, err)
    assert info == lib.GrB_NULL_POINTER
