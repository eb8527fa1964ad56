"""GPU: the SuiteSparse-style storage hand-off of python-graphblas's `.ss` namespace through the C-ABI:
GxB_{Matrix,Vector}_{pack,unpack,import,export}_{CSR,CSC,BitmapR,BitmapC,FullR,FullC,Bitmap,Full} with the numpy-allocator
ownership protocol (reference graphblas/core/ss/matrix.py:744-1162, 1282-1352, 1472-1543; core/ss/vector.py:494-636,
946-1016, 1128-1202, 1303-1360): pack/import free the caller's malloc'd buffers and NULL the pointers, unpack/export
return malloc'd buffers that numpy claims (grb_b200.utils.claim_buffer).  Values and patterns are compared with the
numpy data they came from, bit for bit; iso and jumbled inputs included."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    import graphblas_b200 as gb
    import grb_b200
    from grb_b200 import utils

    gb.init("b200")
    return gb, grb_b200.ffi, grb_b200.lib, utils


def _rand_csr(rng, nrows, ncols, density, dtype, jumble=False):
    dense = rng.random((nrows, ncols)) < density
    indptr = np.zeros(nrows + 1, dtype=np.uint64)
    indptr[1:] = np.cumsum(dense.sum(axis=1))
    cols = [np.flatnonzero(dense[i]).astype(np.uint64) for i in range(nrows)]
    if jumble:
        cols = [rng.permutation(c) for c in cols]
    indices = np.concatenate(cols) if cols else np.zeros(0, np.uint64)
    if np.issubdtype(dtype, np.floating):
        values = rng.random(indices.size).astype(dtype)
    else:
        values = rng.integers(1, 100, size=indices.size).astype(dtype)
    return indptr, indices, values


def _pp(ffi, utils, arr, ctype):
    """pointer-to-pointer on a malloc'd copy the library may take ownership of"""
    p, nbytes = utils.malloc_copy(ffi, arr)
    return ffi.new(ctype + "*", ffi.cast(ctype, p)), nbytes


def _sorted_rows(indptr, indices, values):
    idx, val = indices.copy(), values.copy()
    for i in range(indptr.size - 1):
        s, e = int(indptr[i]), int(indptr[i + 1])
        o = np.argsort(idx[s:e], kind="stable")
        idx[s:e], val[s:e] = idx[s:e][o], val[s:e][o]
    return idx, val


TYPES = [("FP64", np.float64, "GrB_FP64"), ("INT32", np.int32, "GrB_INT32"), ("UINT8", np.uint8, "GrB_UINT8")]


@pytest.mark.parametrize("tname,npdt,sym", TYPES)
@pytest.mark.parametrize("by_col", [False, True])
@pytest.mark.parametrize("iso,jumbled", [(False, False), (True, False), (False, True)])
def test_matrix_pack_unpack_csx_round_trip(G, tname, npdt, sym, by_col, iso, jumbled):
    gb, ffi, lib, utils = G
    rng = np.random.default_rng(11)
    nrows, ncols = 37, 53
    # CSC of an nrows x ncols matrix is the CSR of its transpose: generate in the orientation handed over
    outer, inner = (ncols, nrows) if by_col else (nrows, ncols)
    indptr, indices, values = _rand_csr(rng, outer, inner, 0.2, npdt, jumble=jumbled)
    xvals = values[:1].copy() if iso else values
    A = ffi.new("GrB_Matrix*")
    assert lib.GrB_Matrix_new(A, getattr(lib, sym), nrows, ncols) == lib.GrB_SUCCESS
    Ap, ap_n = _pp(ffi, utils, indptr, "GrB_Index*")
    Ai, ai_n = _pp(ffi, utils, indices, "GrB_Index*")
    Ax, ax_n = _pp(ffi, utils, xvals, "void*")
    fmt = "CSC" if by_col else "CSR"
    info = getattr(lib, f"GxB_Matrix_pack_{fmt}")(A[0], Ap, Ai, Ax, ap_n, ai_n, ax_n, iso, jumbled, ffi.NULL)
    assert info == lib.GrB_SUCCESS
    assert Ap[0] == ffi.NULL and Ai[0] == ffi.NULL and Ax[0] == ffi.NULL        # ownership moved to the library
    nv = ffi.new("GrB_Index*")
    assert lib.GrB_Matrix_nvals(nv, A[0]) == lib.GrB_SUCCESS and nv[0] == indices.size
    is_iso = ffi.new("bool*")
    assert lib.GxB_Matrix_iso(is_iso, A[0]) == lib.GrB_SUCCESS and bool(is_iso[0]) == (iso and indices.size > 0)
    # ... and back
    Bp, Bi, Bx = ffi.new("GrB_Index**"), ffi.new("GrB_Index**"), ffi.new("void**")
    sp, si, sx = ffi.new("GrB_Index*"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
    o_iso, o_j = ffi.new("bool*"), ffi.new("bool*")
    info = getattr(lib, f"GxB_Matrix_unpack_{fmt}")(A[0], Bp, Bi, Bx, sp, si, sx, o_iso, o_j, ffi.NULL)
    assert info == lib.GrB_SUCCESS
    got_p = utils.claim_buffer(ffi, Bp[0], sp[0] // 8, np.uint64)[: outer + 1]
    got_i = utils.claim_buffer(ffi, Bi[0], si[0] // 8, np.uint64)[: indices.size]
    got_x = utils.claim_buffer(ffi, Bx[0], sx[0] // np.dtype(npdt).itemsize, npdt)
    want_i, want_x = _sorted_rows(indptr, indices, values if not iso else np.broadcast_to(values[:1], values.shape).copy())
    assert not o_j[0]                                                           # rows come back sorted
    assert np.array_equal(got_p, indptr) and np.array_equal(got_i, want_i)
    if o_iso[0]:
        assert iso and got_x[0] == values[0]
    else:
        assert np.array_equal(got_x[: indices.size], want_x)
    assert lib.GrB_Matrix_nvals(nv, A[0]) == lib.GrB_SUCCESS and nv[0] == 0     # unpack empties the matrix
    assert lib.GrB_Matrix_free(A) == lib.GrB_SUCCESS
    del got_p, got_i, got_x                                                     # numpy frees the exported buffers


@pytest.mark.parametrize("by_col", [False, True])
def test_matrix_import_export_csx_and_product(G, by_col):
    """GxB_Matrix_import_CSR/CSC as Matrix.ss.import_csr calls it, used in a product, then GxB_Matrix_export_*."""
    gb, ffi, lib, utils = G
    from graphblas_b200 import Matrix, Vector, semiring

    rng = np.random.default_rng(5)
    n = 64
    indptr, indices, values = _rand_csr(rng, n, n, 0.1, np.int64)
    A = ffi.new("GrB_Matrix*")
    Ap, ap_n = _pp(ffi, utils, indptr, "GrB_Index*")
    Ai, ai_n = _pp(ffi, utils, indices, "GrB_Index*")
    Ax, ax_n = _pp(ffi, utils, values, "void*")
    fmt = "CSC" if by_col else "CSR"
    info = getattr(lib, f"GxB_Matrix_import_{fmt}")(A, lib.GrB_INT64, n, n, Ap, Ai, Ax, ap_n, ai_n, ax_n, False, False, ffi.NULL)
    assert info == lib.GrB_SUCCESS and Ap[0] == ffi.NULL
    ref = Matrix.from_csc(indptr, indices, values, nrows=n, ncols=n) if by_col else Matrix.from_csr(indptr, indices, values, ncols=n)
    x = Vector.from_dense(rng.integers(0, 9, size=n).astype(np.int64))
    w1, w2 = Vector("INT64", n), Vector("INT64", n)
    assert lib.GrB_mxv(w1._carg, ffi.NULL, ffi.NULL, lib.GrB_PLUS_TIMES_SEMIRING_INT64, A[0], x._carg, ffi.NULL) == lib.GrB_SUCCESS
    w2 << ref.mxv(x, semiring.plus_times)
    assert w1.isequal(w2)
    typ, nr, nc = ffi.new("GrB_Type*"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
    Bp, Bi, Bx = ffi.new("GrB_Index**"), ffi.new("GrB_Index**"), ffi.new("void**")
    sp, si, sx = ffi.new("GrB_Index*"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
    o_iso = ffi.new("bool*")
    info = getattr(lib, f"GxB_Matrix_export_{fmt}")(A, typ, nr, nc, Bp, Bi, Bx, sp, si, sx, o_iso, ffi.NULL, ffi.NULL)
    assert info == lib.GrB_SUCCESS and A[0] == ffi.NULL and typ[0] == lib.GrB_INT64 and (nr[0], nc[0]) == (n, n)
    assert np.array_equal(utils.claim_buffer(ffi, Bp[0], sp[0] // 8, np.uint64)[: n + 1], indptr)
    assert np.array_equal(utils.claim_buffer(ffi, Bi[0], si[0] // 8, np.uint64)[: indices.size], indices)
    assert np.array_equal(utils.claim_buffer(ffi, Bx[0], sx[0] // 8, np.int64)[: indices.size], values)


@pytest.mark.parametrize("rowmajor", [True, False])
@pytest.mark.parametrize("iso", [False, True])
def test_matrix_bitmap_and_full_formats(G, rowmajor, iso):
    gb, ffi, lib, utils = G
    rng = np.random.default_rng(3)
    nrows, ncols = 13, 21
    bitmap = rng.random((nrows, ncols)) < 0.4
    values = rng.integers(1, 50, size=(nrows, ncols)).astype(np.int32)
    if iso:
        values[:] = 7
    order = "C" if rowmajor else "F"
    sfx = "R" if rowmajor else "C"
    A = ffi.new("GrB_Matrix*")
    assert lib.GrB_Matrix_new(A, lib.GrB_INT32, nrows, ncols) == lib.GrB_SUCCESS
    Ab, ab_n = _pp(ffi, utils, np.asarray(bitmap, order=order).ravel(order=order).astype(np.int8), "int8_t*")
    xv = np.array([7], np.int32) if iso else np.asarray(values, order=order).ravel(order=order)
    Ax, ax_n = _pp(ffi, utils, xv, "void*")
    info = getattr(lib, f"GxB_Matrix_pack_Bitmap{sfx}")(A[0], Ab, Ax, ab_n, ax_n, iso, int(bitmap.sum()), ffi.NULL)
    assert info == lib.GrB_SUCCESS and Ab[0] == ffi.NULL and Ax[0] == ffi.NULL
    nv = ffi.new("GrB_Index*")
    assert lib.GrB_Matrix_nvals(nv, A[0]) == lib.GrB_SUCCESS and nv[0] == int(bitmap.sum())
    # entries as tuples against the numpy data
    I, J, X = np.empty(nv[0], np.uint64), np.empty(nv[0], np.uint64), np.empty(nv[0], np.int32)
    cnt = ffi.new("GrB_Index*", nv[0])
    assert lib.GrB_Matrix_extractTuples_INT32(ffi.from_buffer("GrB_Index*", I), ffi.from_buffer("GrB_Index*", J),
                                              ffi.from_buffer("int32_t*", X), cnt, A[0]) == lib.GrB_SUCCESS
    ri, ci = np.nonzero(bitmap)
    assert np.array_equal(I, ri.astype(np.uint64)) and np.array_equal(J, ci.astype(np.uint64)) and np.array_equal(X, values[ri, ci])
    # bitmap out again
    Bb, Bx, sb, sx = ffi.new("int8_t**"), ffi.new("void**"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
    o_iso, o_nv = ffi.new("bool*"), ffi.new("GrB_Index*")
    assert getattr(lib, f"GxB_Matrix_unpack_Bitmap{sfx}")(A[0], Bb, Bx, sb, sx, o_iso, o_nv, ffi.NULL) == lib.GrB_SUCCESS
    gotb = utils.claim_buffer_2d(ffi, Bb[0], sb[0], nrows, ncols, np.int8, rowmajor)
    assert o_nv[0] == int(bitmap.sum()) and np.array_equal(gotb.astype(bool), bitmap)
    if o_iso[0]:
        assert utils.claim_buffer(ffi, Bx[0], sx[0] // 4, np.int32)[0] == 7
    else:
        gotx = utils.claim_buffer_2d(ffi, Bx[0], sx[0] // 4, nrows, ncols, np.int32, rowmajor)
        assert np.array_equal(gotx[bitmap], values[bitmap])
    # full: every cell present
    Fx, fx_n = _pp(ffi, utils, xv if iso else np.asarray(values, order=order).ravel(order=order), "void*")
    assert getattr(lib, f"GxB_Matrix_pack_Full{sfx}")(A[0], Fx, fx_n, iso, ffi.NULL) == lib.GrB_SUCCESS and Fx[0] == ffi.NULL
    assert lib.GrB_Matrix_nvals(nv, A[0]) == lib.GrB_SUCCESS and nv[0] == nrows * ncols
    Gx, gs, g_iso = ffi.new("void**"), ffi.new("GrB_Index*"), ffi.new("bool*")
    assert getattr(lib, f"GxB_Matrix_unpack_Full{sfx}")(A[0], Gx, gs, g_iso, ffi.NULL) == lib.GrB_SUCCESS
    if g_iso[0]:
        assert utils.claim_buffer(ffi, Gx[0], gs[0] // 4, np.int32)[0] == 7
    else:
        assert np.array_equal(utils.claim_buffer_2d(ffi, Gx[0], gs[0] // 4, nrows, ncols, np.int32, rowmajor), values)
    # a matrix with a hole is not full
    assert lib.GrB_Matrix_setElement_INT32(A[0], 1, 0, 0) == lib.GrB_SUCCESS
    assert getattr(lib, f"GxB_Matrix_unpack_Full{sfx}")(A[0], Gx, gs, g_iso, ffi.NULL) == lib.GrB_INVALID_VALUE
    lib.GrB_Matrix_free(A)


@pytest.mark.parametrize("iso", [False, True])
def test_vector_sparse_bitmap_full_hand_off(G, iso):
    gb, ffi, lib, utils = G
    rng = np.random.default_rng(8)
    n = 1000
    idx = np.sort(rng.choice(n, size=120, replace=False)).astype(np.uint64)
    vals = np.full(idx.size, 2.5) if iso else rng.random(idx.size)
    jumbled = rng.permutation(idx.size)
    v = ffi.new("GrB_Vector*")
    vi, vi_n = _pp(ffi, utils, idx[jumbled], "GrB_Index*")
    vx, vx_n = _pp(ffi, utils, vals[:1] if iso else vals[jumbled], "void*")
    assert lib.GxB_Vector_import_CSC(v, lib.GrB_FP64, n, vi, vx, vi_n, vx_n, iso, idx.size, True, ffi.NULL) == lib.GrB_SUCCESS
    assert vi[0] == ffi.NULL and vx[0] == ffi.NULL
    nv = ffi.new("GrB_Index*")
    assert lib.GrB_Vector_nvals(nv, v[0]) == lib.GrB_SUCCESS and nv[0] == idx.size
    oi, ox, si, sx = ffi.new("GrB_Index**"), ffi.new("void**"), ffi.new("GrB_Index*"), ffi.new("GrB_Index*")
    o_iso, o_nv, o_j = ffi.new("bool*"), ffi.new("GrB_Index*"), ffi.new("bool*")
    assert lib.GxB_Vector_unpack_CSC(v[0], oi, ox, si, sx, o_iso, o_nv, o_j, ffi.NULL) == lib.GrB_SUCCESS
    assert o_nv[0] == idx.size and not o_j[0]
    assert np.array_equal(utils.claim_buffer(ffi, oi[0], si[0] // 8, np.uint64)[: idx.size], idx)
    assert np.array_equal(utils.claim_buffer(ffi, ox[0], sx[0] // 8, np.float64)[: idx.size], vals)
    # bitmap form
    present = np.zeros(n, np.int8)
    present[idx.astype(np.int64)] = 1
    dense = np.zeros(n)
    dense[idx.astype(np.int64)] = vals
    vb, vb_n = _pp(ffi, utils, present, "int8_t*")
    vx, vx_n = _pp(ffi, utils, dense, "void*")
    assert lib.GxB_Vector_pack_Bitmap(v[0], vb, vx, vb_n, vx_n, False, idx.size, ffi.NULL) == lib.GrB_SUCCESS and vb[0] == ffi.NULL
    typ, nn = ffi.new("GrB_Type*"), ffi.new("GrB_Index*")
    ob, sb = ffi.new("int8_t**"), ffi.new("GrB_Index*")
    assert lib.GxB_Vector_export_Bitmap(v, typ, nn, ob, ox, sb, sx, o_iso, o_nv, ffi.NULL) == lib.GrB_SUCCESS
    assert v[0] == ffi.NULL and typ[0] == lib.GrB_FP64 and nn[0] == n and o_nv[0] == idx.size
    gb_ = utils.claim_buffer(ffi, ob[0], sb[0], np.int8)[:n]
    gx_ = utils.claim_buffer(ffi, ox[0], sx[0] // 8, np.float64)[:n]
    assert np.array_equal(gb_, present) and np.array_equal(gx_[present.astype(bool)], dense[present.astype(bool)])
    # full form
    fx, fx_n = _pp(ffi, utils, np.array([4.0]) if iso else dense, "void*")
    assert lib.GxB_Vector_import_Full(v, lib.GrB_FP64, n, fx, fx_n, iso, ffi.NULL) == lib.GrB_SUCCESS and fx[0] == ffi.NULL
    assert lib.GrB_Vector_nvals(nv, v[0]) == lib.GrB_SUCCESS and nv[0] == n
    assert lib.GxB_Vector_export_Full(v, typ, nn, ox, sx, o_iso, ffi.NULL) == lib.GrB_SUCCESS and v[0] == ffi.NULL
    gx_ = utils.claim_buffer(ffi, ox[0], sx[0] // 8, np.float64)[:n]
    assert np.array_equal(gx_, np.full(n, 4.0) if iso else dense)


def test_pack_rejects_out_of_range_indices(G):
    gb, ffi, lib, utils = G
    v = ffi.new("GrB_Vector*")
    assert lib.GrB_Vector_new(v, lib.GrB_INT64, 10) == lib.GrB_SUCCESS
    vi, vi_n = _pp(ffi, utils, np.array([1, 99], np.uint64), "GrB_Index*")
    vx, vx_n = _pp(ffi, utils, np.array([5, 6], np.int64), "void*")
    assert lib.GxB_Vector_pack_CSC(v[0], vi, vx, vi_n, vx_n, False, 2, False, ffi.NULL) == lib.GrB_INDEX_OUT_OF_BOUNDS
    lib.GrB_Vector_free(v)
