// grb_containers.cu — GrB_Matrix / GrB_Vector / GrB_Scalar objects and the storage hand-off across the ABI.
// Reference call sites: graphblas/core/matrix.py:185 (new), :213 (free), :643-697 (build), :1057-1133
// (import), :1658-1702 (export); graphblas/core/ss/matrix.py:1282-1352 (pack CSR), :744-1162 (unpack);
// graphblas/core/ss/vector.py:946-1016, 494-636; error strings graphblas/exceptions.py:143-154.
#include "grb_internal.cuh"
#include "grb_ops.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

// ---------------------------------------------------------------- object helpers
GB_Object *obj_check_lazy(const void *h, const char *what) {
    if (!h) grb_throw(GrB_NULL_POINTER, std::string(what) + " is NULL");
    GB_Object *o = (GB_Object *)h;
    if (o->magic != GRB_MAGIC) grb_throw(o->magic == GRB_FREED ? GrB_UNINITIALIZED_OBJECT : GrB_INVALID_OBJECT,
                                         std::string(what) + " is not a valid object");
    return o;
}
GB_Object *obj_check(const void *h, const char *what) {
    GB_Object *o = obj_check_lazy(h, what);
    if (o->csr_stale) mat_ensure_csr(o);
    return o;
}
void mat_ensure_csr(GB_Object *o) {
    if (!o->csr_stale) return;
    csr_free(o->csr);
    csr_transpose(o->csc, o->csr, o->type->size);
    o->csr_stale = false;
}

void obj_invalidate_caches(GB_Object *o) {
    o->csr_stale = false;
    if (o->csc_valid) { csr_free(o->csc); o->csc_valid = false; }
    csr_drop_caches(o->csr);
}

void obj_release_storage(GB_Object *o) {
    csr_free(o->csr);
    if (o->csc_valid) csr_free(o->csc);
    o->csc_valid = false;
    if (!o->external_vals) dfree(o->vals);
    dfree(o->present); dfree(o->vals_alt); dfree(o->present_alt); dfree(o->list);
    dfree(o->d_nvals);
    o->vals = o->vals_alt = nullptr; o->present = o->present_alt = nullptr; o->list = nullptr;
    o->d_nvals = nullptr; o->list_cap = 0; o->list_valid = false;
}

void vec_alloc(GB_Object *o) {
    int64_t n = o->nrows;
    o->vals = dmalloc((size_t)(n > 0 ? n : 1) * o->type->size);
    o->present = dalloc<uint8_t>(n);
    o->d_nvals = dalloc<int64_t>(1);
    dzero(o->vals, (size_t)n * o->type->size);
    dzero(o->present, (size_t)n);
    dzero(o->d_nvals, sizeof(int64_t));
    o->nvals = 0; o->nvals_known = true; o->list_valid = false;
}
void vec_ensure_alt(GB_Object *o) {
    if (o->external_vals)
        grb_throw(GrB_NOT_IMPLEMENTED, "a vector that wraps caller-owned device memory cannot be the aliased output of a product "
                                       "(its buffer would have to be swapped)");
    if (o->vals_alt) return;
    int64_t n = o->nrows;
    o->vals_alt = dmalloc((size_t)(n > 0 ? n : 1) * o->type->size);
    o->present_alt = dalloc<uint8_t>(n);
}
int64_t vec_nvals(GB_Object *o) {
    if (!o->nvals_known) { o->nvals = read_i64(o->d_nvals); o->nvals_known = true; }
    return o->nvals;
}
void vec_set_unknown_nvals(GB_Object *o) { o->nvals_known = false; o->list_valid = false; }

static void csr_empty(Csr &c, int64_t nrows, int64_t ncols, size_t tsize) {
    c = Csr();
    c.nrows = nrows; c.ncols = ncols;
    c.rowptr = dalloc<int64_t>(nrows + 1);
    dzero(c.rowptr, (size_t)(nrows + 1) * sizeof(int64_t));
    c.colidx = dalloc<uint32_t>(1);
    c.vals = dmalloc(tsize);
}

void mat_set_csr(GB_Object *o, Csr &c) {
    csr_free(o->csr);
    obj_invalidate_caches(o);
    o->csr = c;
    c = Csr();
}

void mat_ensure_csc(GB_Object *o) {
    if (o->csc_valid) return;
    csr_transpose(o->csr, o->csc, o->type->size);
    o->csc_valid = true;
}

GB_Object *obj_new(int kind, GrB_Type type, uint64_t nrows, uint64_t ncols) {
    require_init();
    if (!type) grb_throw(GrB_NULL_POINTER, "type is NULL");
    if (nrows >= (1ull << 32) || ncols >= (1ull << 32))
        grb_throw(GrB_NOT_IMPLEMENTED, "libgrb_b200 keeps 32-bit indices in HBM: dimensions must be < 2^32");
    if (kind != K_MATRIX && nrows > (1ull << 31))
        grb_throw(GrB_OUT_OF_MEMORY, "dense vector storage: size must be <= 2^31");
    GB_Object *o = (kind == K_MATRIX) ? (GB_Object *)new GB_Matrix_opaque()
                 : (kind == K_VECTOR) ? (GB_Object *)new GB_Vector_opaque() : (GB_Object *)new GB_Scalar_opaque();
    o->kind = kind; o->type = type; o->nrows = (int64_t)nrows; o->ncols = (int64_t)ncols;
    try {
        if (kind == K_MATRIX) csr_empty(o->csr, o->nrows, o->ncols, type->size);
        else vec_alloc(o);
    } catch (...) { delete o; throw; }
    return o;
}
void obj_free(GB_Object *o) {
    obj_release_storage(o);
    o->magic = GRB_FREED;
    delete o;
}
static GB_Object *obj_dup(GB_Object *a) {
    GB_Object *o = (a->kind == K_MATRIX) ? (GB_Object *)new GB_Matrix_opaque()
                 : (a->kind == K_VECTOR) ? (GB_Object *)new GB_Vector_opaque() : (GB_Object *)new GB_Scalar_opaque();
    o->kind = a->kind; o->type = a->type; o->nrows = a->nrows; o->ncols = a->ncols;
    if (a->kind == K_MATRIX) o->csr = csr_dup(a->csr, a->type->size);
    else {
        vec_alloc(o);
        d2d(o->vals, a->vals, (size_t)a->nrows * a->type->size);
        d2d(o->present, a->present, (size_t)a->nrows);
        d2d(o->d_nvals, a->d_nvals, sizeof(int64_t));
        o->nvals = a->nvals; o->nvals_known = a->nvals_known;
    }
    return o;
}
void obj_clear(GB_Object *o) {
    if (o->kind == K_MATRIX) {
        Csr c; csr_empty(c, o->nrows, o->ncols, o->type->size);
        mat_set_csr(o, c);
    } else {
        dzero(o->present, (size_t)o->nrows);
        dzero(o->d_nvals, sizeof(int64_t));
        o->nvals = 0; o->nvals_known = true; o->list_valid = false;
    }
}
static int64_t obj_nvals(GB_Object *o) { return o->kind == K_MATRIX ? o->csr.nnz : vec_nvals(o); }

// A matrix-typed handle may really be a vector (reference Vector._as_matrix): give its dimensions as n x 1
static inline int64_t obj_ncols(GB_Object *o) { return o->kind == K_MATRIX ? o->ncols : 1; }

extern "C" {

// ---------------------------------------------------------------- Matrix basics
GrB_Info GrB_Matrix_new(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols) {
    if (!A) return GrB_NULL_POINTER;
    *A = nullptr;
    API_BEGIN
    *A = (GrB_Matrix)obj_new(K_MATRIX, type, nrows, ncols);
    API_END(nullptr)
}
GrB_Info GrB_Matrix_dup(GrB_Matrix *C, const GrB_Matrix A) {
    if (!C) return GrB_NULL_POINTER;
    API_BEGIN
    *C = (GrB_Matrix)obj_dup(obj_check(A, "A"));
    API_END(A)
}
GrB_Info GrB_Matrix_free(GrB_Matrix *A) {
    if (!A || !*A) return GrB_SUCCESS;
    GB_Object *o = (GB_Object *)*A;
    // deferred element-wise work may name this object (grb_internal.cuh PendingOp): run it before the storage goes away
    if (g.pending_n) { try { pending_flush(); } catch (const GrbError &) { g.pending_n = 0; } }
    if (o->magic == GRB_MAGIC) obj_free(o);
    *A = nullptr;
    return GrB_SUCCESS;
}
GrB_Info GrB_Matrix_clear(GrB_Matrix A) {
    API_BEGIN
    obj_clear(obj_check(A, "A"));
    API_END(A)
}
GrB_Info GrB_Matrix_nrows(GrB_Index *nrows, const GrB_Matrix A) {
    if (!nrows) return GrB_NULL_POINTER;
    API_BEGIN
    *nrows = (GrB_Index)obj_check(A, "A")->nrows;
    API_END(A)
}
GrB_Info GrB_Matrix_ncols(GrB_Index *ncols, const GrB_Matrix A) {
    if (!ncols) return GrB_NULL_POINTER;
    API_BEGIN
    *ncols = (GrB_Index)obj_ncols(obj_check(A, "A"));
    API_END(A)
}
GrB_Info GrB_Matrix_nvals(GrB_Index *nvals, const GrB_Matrix A) {
    if (!nvals) return GrB_NULL_POINTER;
    API_BEGIN
    GB_Object *o = obj_check_lazy(A, "A");
    *nvals = (GrB_Index)(o->csr_stale ? o->csc.nnz : obj_nvals(o));
    API_END(A)
}
GrB_Info GrB_Matrix_wait(GrB_Matrix A, GrB_WaitMode mode) {
    (void)mode;
    API_BEGIN
    obj_check_lazy(A, "A");
    stream_sync();
    API_END(A)
}
GrB_Info GrB_Matrix_error(const char **error, const GrB_Matrix A) {
    if (!error) return GrB_NULL_POINTER;
    GB_Object *o = (GB_Object *)A;
    *error = (o && o->magic == GRB_MAGIC) ? o->err.c_str() : "";
    return GrB_SUCCESS;
}
GrB_Info GxB_Matrix_type(GrB_Type *type, const GrB_Matrix A) {
    if (!type) return GrB_NULL_POINTER;
    API_BEGIN
    *type = obj_check(A, "A")->type;
    API_END(A)
}
GrB_Info GxB_Matrix_iso(bool *iso, const GrB_Matrix A) {
    if (!iso) return GrB_NULL_POINTER;
    API_BEGIN
    *iso = obj_check(A, "A")->csr.iso;
    API_END(A)
}

// ---------------------------------------------------------------- Vector basics
GrB_Info GrB_Vector_new(GrB_Vector *v, GrB_Type type, GrB_Index n) {
    if (!v) return GrB_NULL_POINTER;
    *v = nullptr;
    API_BEGIN
    *v = (GrB_Vector)obj_new(K_VECTOR, type, n, 1);
    API_END(nullptr)
}
GrB_Info GrB_Vector_dup(GrB_Vector *w, const GrB_Vector u) {
    if (!w) return GrB_NULL_POINTER;
    API_BEGIN
    *w = (GrB_Vector)obj_dup(obj_check(u, "u"));
    API_END(u)
}
GrB_Info GrB_Vector_free(GrB_Vector *v) { return GrB_Matrix_free((GrB_Matrix *)v); }
GrB_Info GrB_Vector_clear(GrB_Vector v) { return GrB_Matrix_clear((GrB_Matrix)v); }
GrB_Info GrB_Vector_size(GrB_Index *n, const GrB_Vector v) { return GrB_Matrix_nrows(n, (GrB_Matrix)v); }
GrB_Info GrB_Vector_nvals(GrB_Index *nvals, const GrB_Vector v) { return GrB_Matrix_nvals(nvals, (GrB_Matrix)v); }
GrB_Info GrB_Vector_wait(GrB_Vector v, GrB_WaitMode mode) { return GrB_Matrix_wait((GrB_Matrix)v, mode); }
GrB_Info GrB_Vector_error(const char **error, const GrB_Vector v) { return GrB_Matrix_error(error, (GrB_Matrix)v); }
GrB_Info GxB_Vector_type(GrB_Type *type, const GrB_Vector v) { return GxB_Matrix_type(type, (GrB_Matrix)v); }
GrB_Info GxB_Vector_iso(bool *iso, const GrB_Vector v) {
    if (!iso) return GrB_NULL_POINTER;
    *iso = false;
    return GrB_SUCCESS;
}

// ---------------------------------------------------------------- Scalar basics (a size-1 vector)
GrB_Info GrB_Scalar_new(GrB_Scalar *s, GrB_Type type) {
    if (!s) return GrB_NULL_POINTER;
    *s = nullptr;
    API_BEGIN
    *s = (GrB_Scalar)obj_new(K_SCALAR, type, 1, 1);
    API_END(nullptr)
}
GrB_Info GrB_Scalar_dup(GrB_Scalar *s, const GrB_Scalar t) { return GrB_Vector_dup((GrB_Vector *)s, (GrB_Vector)t); }
GrB_Info GrB_Scalar_free(GrB_Scalar *s) { return GrB_Matrix_free((GrB_Matrix *)s); }
GrB_Info GrB_Scalar_clear(GrB_Scalar s) { return GrB_Matrix_clear((GrB_Matrix)s); }
GrB_Info GrB_Scalar_nvals(GrB_Index *nvals, const GrB_Scalar s) { return GrB_Matrix_nvals(nvals, (GrB_Matrix)s); }
GrB_Info GrB_Scalar_wait(GrB_Scalar s, GrB_WaitMode mode) { return GrB_Matrix_wait((GrB_Matrix)s, mode); }
GrB_Info GrB_Scalar_error(const char **error, const GrB_Scalar s) { return GrB_Matrix_error(error, (GrB_Matrix)s); }

}  // extern "C"

// ================================================================ typed entry points (templates)
__global__ void k_make_keys2(const uint64_t *I, const uint64_t *J, uint64_t nrows, uint64_t ncols, int64_t n,
                             uint64_t *keys, int *bad) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint64_t i = I[k], j = J ? J[k] : 0;
    if (i >= nrows || j >= ncols) { *bad = 1; keys[k] = 0; return; }
    keys[k] = i * ncols + j;
}

static int dup_code(const GrB_BinaryOp dup, GB_Object *o) {
    if (!dup) return -1;
    op_need_kernel(dup->op, dup->name);
    if (op_is_cmp(dup->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot be used as dup");
    (void)o;
    return dup->op;
}

// host COO (uint64 indices, typed values) -> device CSR of the object's type
static Csr build_from_host(int64_t nrows, int64_t ncols, const uint64_t *I, const uint64_t *J, const void *X,
                           int xcode, int out_code, int64_t n, int dup_op) {
    uint64_t *dI = dalloc<uint64_t>(n), *dJ = J ? dalloc<uint64_t>(n) : nullptr, *keys = dalloc<uint64_t>(n);
    void *dX = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[xcode]);
    h2d(dI, I, (size_t)n * 8);
    if (J) h2d(dJ, J, (size_t)n * 8);
    h2d(dX, X, (size_t)n * kTypeSize[xcode]);
    int *dflag = (int *)g.d_scratch;
    dzero(dflag, sizeof(int));
    if (n) GRB_LAUNCH("make_keys", k_make_keys2, nblk(n), 256, 0, dI, dJ, (uint64_t)nrows, (uint64_t)ncols, n, keys, dflag);
    int bad = 0;
    d2h(&bad, dflag, sizeof(int));
    stream_sync();
    dfree(dI); dfree(dJ);
    if (bad) { dfree(keys); dfree(dX); grb_throw(GrB_INDEX_OUT_OF_BOUNDS, "index out of bounds in build"); }
    if (xcode != out_code) {
        void *cx = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[out_code]);
        cast_array(dX, xcode, cx, out_code, n);
        dfree(dX); dX = cx;
    }
    Csr c;
    try { c = coo_to_csr(nrows, ncols, keys, dX, false, out_code, n, dup_op, false); }
    catch (...) { dfree(keys); dfree(dX); throw; }
    dfree(keys); dfree(dX);
    return c;
}

template <class T>
static GrB_Info matrix_build(GrB_Matrix C_, const GrB_Index *I, const GrB_Index *J, const T *X, GrB_Index nvals,
                             const GrB_BinaryOp dup, int xcode) {
    API_BEGIN
    GB_Object *C = obj_check(C_, "C");
    if (nvals && (!I || !J || !X)) grb_throw(GrB_NULL_POINTER, "I, J or X is NULL");
    if (obj_nvals(C) != 0) grb_throw(GrB_OUTPUT_NOT_EMPTY, "build requires an empty output");
    if (C->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "matrix build on a non-matrix");
    Csr c = build_from_host(C->nrows, C->ncols, I, J, X, xcode, C->type->code, (int64_t)nvals, dup_code(dup, C));
    mat_set_csr(C, c);
    API_END(C_)
}

template <class T>
static GrB_Info vector_build(GrB_Vector w_, const GrB_Index *I, const T *X, GrB_Index nvals,
                             const GrB_BinaryOp dup, int xcode) {
    API_BEGIN
    GB_Object *w = obj_check(w_, "w");
    if (nvals && (!I || !X)) grb_throw(GrB_NULL_POINTER, "I or X is NULL");
    if (obj_nvals(w) != 0) grb_throw(GrB_OUTPUT_NOT_EMPTY, "build requires an empty output");
    Csr c = build_from_host(w->nrows, 1, I, nullptr, X, xcode, w->type->code, (int64_t)nvals, dup_code(dup, w));
    csr_to_vec(c, w, true);
    csr_free(c);
    API_END(w_)
}

// device CSR -> host tuples; X may be of another type (cast on the device)
static void csr_extract_host(const Csr &c, int code, uint64_t *I, uint64_t *J, void *X, int xcode, bool by_col) {
    int64_t nnz = c.nnz;
    if (nnz == 0) return;
    if (I || J) {
        uint64_t *tmp = dalloc<uint64_t>(nnz);
        uint32_t *rows = csr_row_ids(c);
        uint64_t *rowdst = by_col ? J : I, *coldst = by_col ? I : J;
        if (rowdst) { widen_indices(rows, nnz, tmp); d2h(rowdst, tmp, (size_t)nnz * 8); }
        if (coldst) { widen_indices(c.colidx, nnz, tmp); d2h(coldst, tmp, (size_t)nnz * 8); }
        stream_sync();
        dfree(tmp); dfree(rows);
    }
    if (X) {
        size_t xs = kTypeSize[xcode];
        if (c.iso) {
            void *one = dmalloc(xs);
            cast_array(c.vals, code, one, xcode, 1);
            std::vector<char> h(xs);
            d2h(h.data(), one, xs);
            stream_sync();
            dfree(one);
            for (int64_t k = 0; k < nnz; k++) memcpy((char *)X + k * xs, h.data(), xs);
        } else if (xcode == code) {
            d2h(X, c.vals, (size_t)nnz * xs);
            stream_sync();
        } else {
            void *cx = dmalloc((size_t)nnz * xs);
            cast_array(c.vals, code, cx, xcode, nnz);
            d2h(X, cx, (size_t)nnz * xs);
            stream_sync();
            dfree(cx);
        }
    }
}

template <class T>
static GrB_Info matrix_extract_tuples(GrB_Index *I, GrB_Index *J, T *X, GrB_Index *nvals, const GrB_Matrix A_, int xcode) {
    API_BEGIN
    GB_Object *A = obj_check(A_, "A");
    if (!nvals) grb_throw(GrB_NULL_POINTER, "nvals is NULL");
    int64_t nnz = obj_nvals(A);
    if ((int64_t)*nvals < nnz) grb_throw(GrB_INSUFFICIENT_SPACE, "output arrays are too small");
    if (A->kind == K_MATRIX) csr_extract_host(A->csr, A->type->code, I, J, X, xcode, false);
    else { Csr c = vec_as_csr_column(A); csr_extract_host(c, A->type->code, I, J, X, xcode, false); csr_free(c); }
    *nvals = (GrB_Index)nnz;
    API_END(A_)
}
template <class T>
static GrB_Info vector_extract_tuples(GrB_Index *I, T *X, GrB_Index *nvals, const GrB_Vector v_, int xcode) {
    return matrix_extract_tuples<T>(I, nullptr, X, nvals, (GrB_Matrix)v_, xcode);
}

// ---- single elements: slow paths (tests / small updates), each a rebuild or a tiny kernel ----
__global__ void k_find_entry(const int64_t *rowptr, const uint32_t *col, int64_t i, uint32_t j, int64_t *out) {
    int64_t lo = rowptr[i], hi = rowptr[i + 1];
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (col[mid] < j) lo = mid + 1; else hi = mid;
    }
    out[0] = (lo < rowptr[i + 1] && col[lo] == j) ? lo : -1;
}
static int64_t csr_find(const Csr &c, int64_t i, int64_t j) {
    GRB_LAUNCH("find_entry", k_find_entry, 1, 1, 0, c.rowptr, c.colidx, i, (uint32_t)j, g.d_scratch + 8);
    return read_i64(g.d_scratch + 8);
}
__global__ void k_append_key(uint64_t *keys, int64_t n, uint64_t key) { keys[n] = key; }
__global__ void k_keys_from_csr(const uint32_t *rows, const uint32_t *cols, int64_t nnz, uint64_t ncols, uint64_t *keys) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < nnz) keys[k] = (uint64_t)rows[k] * ncols + cols[k];
}
template <class T>
__global__ void k_expand_iso(const T *one, int64_t n, T *out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < n) out[k] = one[0];
}
static void csr_expand_iso(Csr &c, int code) {
    if (!c.iso) return;
    size_t ts = kTypeSize[code];
    void *nv = dmalloc((size_t)(c.nnz > 0 ? c.nnz : 1) * ts);
    GRB_TYPE_SWITCH(code, T, if (c.nnz) GRB_LAUNCH("expand_iso", k_expand_iso<T>, nblk(c.nnz), 256, 0, (const T *)c.vals, c.nnz, (T *)nv));
    dfree(c.vals);
    c.vals = nv; c.iso = false;
}

template <class T>
static void matrix_set_element_impl(GB_Object *C, T x, int xcode, uint64_t i, uint64_t j) {
    if (i >= (uint64_t)C->nrows || j >= (uint64_t)obj_ncols(C)) grb_throw(GrB_INVALID_INDEX, "index out of bounds");
    int code = C->type->code;
    size_t ts = C->type->size;
    // value in C's type, staged on the device
    void *dx = dmalloc(8), *dxc = dmalloc(8);
    h2d(dx, &x, sizeof(T));
    stream_sync();
    cast_array(dx, xcode, dxc, code, 1);
    if (C->kind != K_MATRIX) {
        d2d((char *)C->vals + i * ts, dxc, ts);
        uint8_t one = 1;
        h2d(C->present + i, &one, 1);
        stream_sync();
        count_present(C->present, C->nrows, C->d_nvals);
        vec_set_unknown_nvals(C);
        dfree(dx); dfree(dxc);
        return;
    }
    csr_expand_iso(C->csr, code);
    int64_t pos = csr_find(C->csr, (int64_t)i, (int64_t)j);
    if (pos >= 0) {
        d2d((char *)C->csr.vals + pos * ts, dxc, ts);
        // an in-place edit: every cache that holds a copy of the values (transpose, typed copy of the values,
        // column panels, stream form) is stale now
        obj_invalidate_caches(C);
    } else {
        int64_t nnz = C->csr.nnz;
        uint64_t *keys = dalloc<uint64_t>(nnz + 1);
        void *vals = dmalloc((size_t)(nnz + 1) * ts);
        uint32_t *rows = csr_row_ids(C->csr);
        if (nnz) GRB_LAUNCH("keys_from_csr", k_keys_from_csr, nblk(nnz), 256, 0, rows, C->csr.colidx, nnz,
                            (uint64_t)C->ncols, keys);
        GRB_LAUNCH("append_key", k_append_key, 1, 1, 0, keys, nnz, i * (uint64_t)C->ncols + j);
        d2d(vals, C->csr.vals, (size_t)nnz * ts);
        d2d((char *)vals + nnz * ts, dxc, ts);
        Csr c = coo_to_csr(C->nrows, C->ncols, keys, vals, false, code, nnz + 1, OP_SECOND, false);
        dfree(keys); dfree(vals); dfree(rows);
        mat_set_csr(C, c);
    }
    dfree(dx); dfree(dxc);
}
template <class T>
static GrB_Info matrix_set_element(GrB_Matrix C_, T x, GrB_Index i, GrB_Index j, int xcode) {
    API_BEGIN
    matrix_set_element_impl<T>(obj_check(C_, "C"), x, xcode, i, j);
    API_END(C_)
}

template <class T>
static GrB_Info matrix_extract_element(T *x, const GrB_Matrix A_, GrB_Index i, GrB_Index j, int xcode) {
    bool found = false;
    try {
        GB_Object *A = obj_check(A_, "A");
        if (!x) grb_throw(GrB_NULL_POINTER, "x is NULL");
        if (i >= (uint64_t)A->nrows || j >= (uint64_t)obj_ncols(A)) grb_throw(GrB_INVALID_INDEX, "index out of bounds");
        int code = A->type->code;
        size_t ts = A->type->size;
        const void *src = nullptr;
        if (A->kind == K_MATRIX) {
            int64_t pos = csr_find(A->csr, (int64_t)i, (int64_t)j);
            if (pos >= 0) { found = true; src = (char *)A->csr.vals + (A->csr.iso ? 0 : pos) * ts; }
        } else {
            uint8_t p = 0;
            d2h(&p, A->present + i, 1);
            stream_sync();
            if (p) { found = true; src = (char *)A->vals + i * ts; }
        }
        if (found) {
            void *dxc = dmalloc(8);
            cast_array(src, code, dxc, xcode, 1);
            d2h(x, dxc, sizeof(T));
            stream_sync();
            dfree(dxc);
        }
    } catch (const GrbError &e) {
        GB_Object *o = (GB_Object *)A_;
        if (o && o->magic == GRB_MAGIC) o->err = e.msg;
        return e.info;
    }
    return found ? GrB_SUCCESS : GrB_NO_VALUE;
}

// ---------------------------------------------------------------- import / export (copy semantics)
// Host layout per reference graphblas/core/matrix.py:1057-1133: indptr/indices uint64, values typed.
static Csr csr_from_host(int64_t nrows, int64_t ncols, const uint64_t *Ap, const uint64_t *Ai, const void *Ax,
                         int64_t ap_len, int64_t nnz, int64_t ax_len, int xcode, int out_code, bool iso,
                         bool maybe_jumbled) {
    if (ap_len < nrows + 1) grb_throw(GrB_INVALID_VALUE, "Ap is shorter than nrows+1");
    Csr c;
    c.nrows = nrows; c.ncols = ncols; c.nnz = nnz; c.iso = iso;
    c.rowptr = dalloc<int64_t>(nrows + 1);
    c.colidx = dalloc<uint32_t>(nnz);
    int64_t nv = iso ? 1 : nnz;
    if (ax_len < nv) grb_throw(GrB_INVALID_VALUE, "Ax is shorter than the number of entries");
    c.vals = dmalloc((size_t)(nv > 0 ? nv : 1) * kTypeSize[out_code]);
    h2d(c.rowptr, Ap, (size_t)(nrows + 1) * 8);   // uint64 -> int64: same bits below 2^63
    uint64_t *tmp = dalloc<uint64_t>(nnz);
    h2d(tmp, Ai, (size_t)nnz * 8);
    try { narrow_indices(tmp, nnz, c.colidx, (uint64_t)ncols, "import"); }
    catch (...) { dfree(tmp); csr_free(c); throw; }
    dfree(tmp);
    // The values cross PCIe on a second stream, behind the indices, while this stream validates the pattern and -- for a matrix
    // large enough to be pulled from repeatedly -- builds the stream form of the pull kernels (hub ranking, row flags): on the
    // PageRank graph 1.7 ms of checking + 1.0 ms of building + the slow first pull used to FOLLOW the 15.4 ms of copies
    // (tools/debug/e2e_breakdown.py); now they sit under the 4.7 ms the values take.
    const bool overlap = (xcode == out_code) && nv >= (1 << 20) && g.copy_stream != nullptr;
    if (overlap) {
        CUDA_TRY(cudaEventRecord(g.copy_ev_a, g.stream));            // c.vals exists and the index copy is queued
        CUDA_TRY(cudaStreamWaitEvent(g.copy_stream, g.copy_ev_a, 0));
        CUDA_TRY(cudaMemcpyAsync(c.vals, Ax, (size_t)nv * kTypeSize[out_code], cudaMemcpyHostToDevice, g.copy_stream));
        CUDA_TRY(cudaEventRecord(g.copy_ev_b, g.copy_stream));
    } else if (xcode == out_code) h2d(c.vals, Ax, (size_t)nv * kTypeSize[out_code]);
    else {
        void *sx = dmalloc((size_t)(nv > 0 ? nv : 1) * kTypeSize[xcode]);
        h2d(sx, Ax, (size_t)nv * kTypeSize[xcode]);
        cast_array(sx, xcode, c.vals, out_code, nv);
        dfree(sx);
    }
    // join: the main stream waits for the values, and so does the host -- the caller may reuse or free Ax when this returns
    auto join = [&]() { if (overlap) { cudaStreamWaitEvent(g.stream, g.copy_ev_b, 0); cudaEventSynchronize(g.copy_ev_b); } };
    int64_t last = read_i64(c.rowptr + nrows);
    if (last != nnz) {
        if (last > nnz || last < 0) { join(); stream_sync(); csr_free(c); grb_throw(GrB_INVALID_VALUE, "Ap[nrows] exceeds the index array"); }
        c.nnz = last;
    }
    try {
        bool sorted = csr_is_sorted(c);
        if (!sorted) { (void)maybe_jumbled; join(); csr_sort_rows(c, out_code); }      // sorting permutes the values: they must have arrived
        else if (overlap && g.hub_mode && c.nnz >= (1 << 20)) csr_ensure_hubs(c, hub_capacity_for(kTypeSize[out_code]));
    } catch (...) { join(); stream_sync(); csr_free(c); throw; }
    join();
    return c;
}

static void coo_from_host_into(GB_Object *A, const uint64_t *I, const uint64_t *J, const void *X, int64_t n, int xcode) {
    Csr c = build_from_host(A->nrows, A->ncols, I, J, X, xcode, A->type->code, n, -1);
    mat_set_csr(A, c);
}

template <class T>
static GrB_Info matrix_import(GrB_Matrix *A_, GrB_Type type, GrB_Index nrows, GrB_Index ncols, const GrB_Index *Ap,
                              const GrB_Index *Ai, const T *Ax, GrB_Index Ap_len, GrB_Index Ai_len, GrB_Index Ax_len,
                              GrB_Format format, int xcode) {
    if (!A_) return GrB_NULL_POINTER;
    *A_ = nullptr;
    GB_Object *A = nullptr;
    API_BEGIN
    if (!Ap || (Ai_len && !Ai) || (Ax_len && !Ax)) grb_throw(GrB_NULL_POINTER, "Ap, Ai or Ax is NULL");
    A = obj_new(K_MATRIX, type, nrows, ncols);
    try {
        int code = type->code;
        if (format == GrB_CSR_FORMAT) {
            Csr c = csr_from_host(A->nrows, A->ncols, Ap, Ai, Ax, (int64_t)Ap_len, (int64_t)Ai_len, (int64_t)Ax_len,
                                  xcode, code, false, true);
            mat_set_csr(A, c);
        } else if (format == GrB_CSC_FORMAT) {
            Csr t = csr_from_host(A->ncols, A->nrows, Ap, Ai, Ax, (int64_t)Ap_len, (int64_t)Ai_len, (int64_t)Ax_len,
                                  xcode, code, false, true);
            // keep the columns only; rows are built on first use (A.T.mxv / u.vxm(A) never need them)
            A->csc = t; A->csc_valid = true; A->csr_stale = true;
        } else if (format == GrB_COO_FORMAT) {
            // C API 2.0 COO import: Ap = row indices, Ai = column indices
            coo_from_host_into(A, Ap, Ai, Ax, (int64_t)Ax_len, xcode);
        } else grb_throw(GrB_INVALID_VALUE, "unknown format");
    } catch (...) { obj_free(A); A = nullptr; throw; }
    *A_ = (GrB_Matrix)A;
    API_END(nullptr)
}

static void csr_to_host(const Csr &c, int code, uint64_t *Ap, uint64_t *Ai, void *Ax, int xcode) {
    d2h(Ap, c.rowptr, (size_t)(c.nrows + 1) * 8);
    stream_sync();
    csr_extract_host(c, code, nullptr, Ai, Ax, xcode, false);
}

template <class T>
static GrB_Info matrix_export(GrB_Index *Ap, GrB_Index *Ai, T *Ax, GrB_Index *Ap_len, GrB_Index *Ai_len,
                              GrB_Index *Ax_len, GrB_Format format, const GrB_Matrix A_, int xcode) {
    API_BEGIN
    GB_Object *A = obj_check(A_, "A");
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "export needs a matrix");
    if (!Ap_len || !Ai_len || !Ax_len) grb_throw(GrB_NULL_POINTER, "length pointers are NULL");
    int64_t nnz = A->csr.nnz;
    int code = A->type->code;
    if (format == GrB_CSR_FORMAT || format == GrB_CSC_FORMAT) {
        if (format == GrB_CSC_FORMAT) mat_ensure_csc(A);
        const Csr &c = (format == GrB_CSR_FORMAT) ? A->csr : A->csc;
        if ((int64_t)*Ap_len < c.nrows + 1 || (int64_t)*Ai_len < nnz || (int64_t)*Ax_len < nnz)
            grb_throw(GrB_INSUFFICIENT_SPACE, "output arrays are too small");
        if (!Ap || (nnz && (!Ai || !Ax))) grb_throw(GrB_NULL_POINTER, "output arrays are NULL");
        csr_to_host(c, code, Ap, Ai, Ax, xcode);
        *Ap_len = (GrB_Index)(c.nrows + 1);
    } else if (format == GrB_COO_FORMAT) {
        if ((int64_t)*Ap_len < nnz || (int64_t)*Ai_len < nnz || (int64_t)*Ax_len < nnz)
            grb_throw(GrB_INSUFFICIENT_SPACE, "output arrays are too small");
        csr_extract_host(A->csr, code, Ap, Ai, Ax, xcode, false);
        *Ap_len = (GrB_Index)nnz;
    } else grb_throw(GrB_INVALID_VALUE, "unknown format");
    *Ai_len = (GrB_Index)nnz; *Ax_len = (GrB_Index)nnz;
    API_END(A_)
}

template <class T>
static GrB_Info scalar_set(GrB_Scalar s_, T x, int xcode) {
    return matrix_set_element<T>((GrB_Matrix)s_, x, 0, 0, xcode);
}
template <class T>
static GrB_Info scalar_get(T *x, const GrB_Scalar s_, int xcode) {
    return matrix_extract_element<T>(x, (GrB_Matrix)s_, 0, 0, xcode);
}

extern "C" {

GrB_Info GrB_Matrix_exportSize(GrB_Index *Ap_len, GrB_Index *Ai_len, GrB_Index *Ax_len, GrB_Format format,
                               const GrB_Matrix A_) {
    API_BEGIN
    GB_Object *A = obj_check(A_, "A");
    if (!Ap_len || !Ai_len || !Ax_len) grb_throw(GrB_NULL_POINTER, "length pointers are NULL");
    int64_t nnz = obj_nvals(A);
    *Ai_len = *Ax_len = (GrB_Index)nnz;
    if (format == GrB_CSR_FORMAT) *Ap_len = (GrB_Index)(A->nrows + 1);
    else if (format == GrB_CSC_FORMAT) *Ap_len = (GrB_Index)(A->ncols + 1);
    else if (format == GrB_COO_FORMAT) *Ap_len = (GrB_Index)nnz;
    else grb_throw(GrB_INVALID_VALUE, "unknown format");
    API_END(A_)
}
GrB_Info GrB_Matrix_exportHint(GrB_Format *format, const GrB_Matrix A_) {
    if (!format) return GrB_NULL_POINTER;
    API_BEGIN
    obj_check(A_, "A");
    *format = GrB_CSR_FORMAT;
    API_END(A_)
}

GrB_Info GrB_Matrix_removeElement(GrB_Matrix C_, GrB_Index i, GrB_Index j) {
    API_BEGIN
    GB_Object *C = obj_check(C_, "C");
    if (i >= (uint64_t)C->nrows || j >= (uint64_t)obj_ncols(C)) grb_throw(GrB_INVALID_INDEX, "index out of bounds");
    if (C->kind != K_MATRIX) {
        uint8_t zero = 0;
        h2d(C->present + i, &zero, 1);
        stream_sync();
        count_present(C->present, C->nrows, C->d_nvals);
        vec_set_unknown_nvals(C);
    } else {
        int64_t pos = csr_find(C->csr, (int64_t)i, (int64_t)j);
        if (pos >= 0) {
            // rebuild without the entry: shift the tail down on the host side of the ABI is not possible,
            // so rebuild from keys with the victim dropped (slow path, used by tests only)
            int code = C->type->code;
            size_t ts = C->type->size;
            csr_expand_iso(C->csr, code);
            int64_t nnz = C->csr.nnz;
            uint64_t *keys = dalloc<uint64_t>(nnz);
            uint32_t *rows = csr_row_ids(C->csr);
            GRB_LAUNCH("keys_from_csr", k_keys_from_csr, nblk(nnz), 256, 0, rows, C->csr.colidx, nnz, (uint64_t)C->ncols, keys);
            void *vals = dmalloc((size_t)nnz * ts);
            d2d(vals, C->csr.vals, (size_t)pos * ts);
            d2d((char *)vals + pos * ts, (char *)C->csr.vals + (pos + 1) * ts, (size_t)(nnz - pos - 1) * ts);
            uint64_t *keys2 = dalloc<uint64_t>(nnz);
            d2d(keys2, keys, (size_t)pos * 8);
            d2d(keys2 + pos, keys + pos + 1, (size_t)(nnz - pos - 1) * 8);
            Csr c = coo_to_csr(C->nrows, C->ncols, keys2, vals, false, code, nnz - 1, OP_SECOND, false);
            dfree(keys); dfree(keys2); dfree(rows); dfree(vals);
            mat_set_csr(C, c);
        }
    }
    API_END(C_)
}
GrB_Info GrB_Vector_removeElement(GrB_Vector w, GrB_Index i) { return GrB_Matrix_removeElement((GrB_Matrix)w, i, 0); }

// ---------------------------------------------------------------- dense vector hand-off (copy semantics)
GrB_Info GxB_b200_Vector_import_dense(GrB_Vector v_, const void *values, const int8_t *present, GrB_Index n) {
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (v->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "dense import needs a vector");
    if ((int64_t)n != v->nrows) grb_throw(GrB_DIMENSION_MISMATCH, "length differs from the vector size");
    if (!values && n) grb_throw(GrB_NULL_POINTER, "values is NULL");
    h2d(v->vals, values, (size_t)n * v->type->size);
    if (present) {
        h2d(v->present, present, (size_t)n);
        count_present(v->present, v->nrows, v->d_nvals);
        vec_set_unknown_nvals(v);
    } else {
        CUDA_TRY(cudaMemsetAsync(v->present, 1, (size_t)n, g.stream));
        v->nvals = (int64_t)n; v->nvals_known = true; v->list_valid = false;
    }
    stream_sync();     // the caller may reuse or free its (possibly pinned) buffers as soon as this returns
    API_END(v_)
}
GrB_Info GxB_b200_Vector_export_dense(void *values, int8_t *present, GrB_Index n, const GrB_Vector v_) {
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (v->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "dense export needs a vector");
    if ((int64_t)n != v->nrows) grb_throw(GrB_DIMENSION_MISMATCH, "length differs from the vector size");
    if (values) d2h(values, v->vals, (size_t)n * v->type->size);
    if (present) d2h(present, v->present, (size_t)n);
    stream_sync();
    API_END(v_)
}
GrB_Info GxB_b200_Vector_device_ptrs(void **values, void **present, const GrB_Vector v_) {
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (v->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "needs a vector");
    if (values) *values = v->vals;
    if (present) *present = v->present;
    API_END(v_)
}
// after writing the device buffers directly (multi-GPU exchange): tell the object how many entries it holds
GrB_Info GxB_b200_Vector_set_dense_nvals(GrB_Vector v_, int64_t nvals) {
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (nvals < 0) { count_present(v->present, v->nrows, v->d_nvals); vec_set_unknown_nvals(v); }
    else { v->nvals = nvals; v->nvals_known = true; v->list_valid = false; }
    API_END(v_)
}
GrB_Info GxB_b200_Matrix_device_ptrs(void **rowptr, void **colidx, void **values, const GrB_Matrix A_) {
    API_BEGIN
    GB_Object *A = obj_check(A_, "A");
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "needs a matrix");
    if (rowptr) *rowptr = A->csr.rowptr;
    if (colidx) *colidx = A->csr.colidx;
    if (values) *values = A->csr.vals;
    API_END(A_)
}
GrB_Info GxB_b200_Matrix_prepare(GrB_Matrix A_, int want_transpose) {
    API_BEGIN
    GB_Object *A = obj_check(A_, "A");
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "needs a matrix");
    if (want_transpose) mat_ensure_csc(A);
    API_END(A_)
}

}  // extern "C"

template <class T> __global__ void k_fill(T *p, int64_t n, const T *one) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < n) p[k] = one[0];
}

// ---------------------------------------------------------------- SuiteSparse-style pack / unpack
// Ownership transfer as in reference graphblas/core/ss/matrix.py:1282-1352 (import_csr -> pack) and :744-1162
// (export -> unpack): host buffers come from / go to the allocator given to GxB_init (numpy's, so that
// claim_buffer/unclaim_buffer keep working).  HBM is the only place matrix data lives afterwards.
void pack_csx(GB_Object *A, bool by_col, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index Ap_size,
                     GrB_Index Ai_size, GrB_Index Ax_size, bool iso) {
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "pack needs a matrix");
    if (!Ap || !Ai || !Ax || !*Ap || !*Ax) grb_throw(GrB_NULL_POINTER, "Ap, Ai or Ax is NULL");
    int64_t nr = by_col ? A->ncols : A->nrows, nc = by_col ? A->nrows : A->ncols;
    if ((int64_t)(Ap_size / 8) < nr + 1) grb_throw(GrB_INVALID_VALUE, "Ap is too small");
    int64_t nnz = (int64_t)(*Ap)[nr];
    if ((int64_t)(Ai_size / 8) < nnz) grb_throw(GrB_INVALID_VALUE, "Ai is too small");
    int code = A->type->code;
    int64_t ax_len = (int64_t)(Ax_size / A->type->size);
    static const uint64_t zero_index = 0;
    Csr t = csr_from_host(nr, nc, *Ap, nnz ? *Ai : &zero_index, *Ax, nr + 1, nnz, ax_len, code, code, iso, true);
    if (!by_col) mat_set_csr(A, t);
    else {
        Csr c;
        csr_transpose(t, c, A->type->size);
        mat_set_csr(A, c);
        A->csc = t; A->csc_valid = true;
    }
    stream_sync();
    g.h_free(*Ap); *Ap = nullptr;
    if (*Ai) { g.h_free(*Ai); *Ai = nullptr; }
    g.h_free(*Ax); *Ax = nullptr;
}
void unpack_csx(GB_Object *A, bool by_col, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index *Ap_size,
                       GrB_Index *Ai_size, GrB_Index *Ax_size, bool *iso, bool *jumbled) {
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "unpack needs a matrix");
    if (!Ap || !Ai || !Ax || !Ap_size || !Ai_size || !Ax_size) grb_throw(GrB_NULL_POINTER, "output pointers are NULL");
    if (by_col) mat_ensure_csc(A);
    const Csr &c = by_col ? A->csc : A->csr;
    size_t ts = A->type->size;
    bool give_iso = iso && c.iso;
    int64_t nv = give_iso ? 1 : c.nnz;
    size_t pb = (size_t)(c.nrows + 1) * 8, ib = (size_t)(c.nnz > 0 ? c.nnz : 1) * 8, xb = (size_t)(nv > 0 ? nv : 1) * ts;
    uint64_t *hp = (uint64_t *)g.h_malloc(pb), *hi = (uint64_t *)g.h_malloc(ib);
    void *hx = g.h_malloc(xb);
    if (!hp || !hi || !hx) grb_throw(GrB_OUT_OF_MEMORY, "host allocation failed in unpack");
    d2h(hp, c.rowptr, pb);
    stream_sync();
    if (give_iso) { d2h(hx, c.vals, ts); stream_sync(); csr_extract_host(c, A->type->code, nullptr, hi, nullptr, A->type->code, false); }
    else csr_extract_host(c, A->type->code, nullptr, hi, hx, A->type->code, false);
    *Ap = hp; *Ai = hi; *Ax = hx; *Ap_size = pb; *Ai_size = ib; *Ax_size = xb;
    if (iso) *iso = give_iso;
    if (jumbled) *jumbled = false;
    obj_clear(A);
}

extern "C" {
GrB_Info GxB_Matrix_pack_CSR(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Aj, void **Ax, GrB_Index Ap_size,
                             GrB_Index Aj_size, GrB_Index Ax_size, bool iso, bool jumbled, const GrB_Descriptor desc) {
    (void)jumbled; (void)desc;
    API_BEGIN
    pack_csx(obj_check(A, "A"), false, Ap, Aj, Ax, Ap_size, Aj_size, Ax_size, iso);
    API_END(A)
}
GrB_Info GxB_Matrix_pack_CSC(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index Ap_size,
                             GrB_Index Ai_size, GrB_Index Ax_size, bool iso, bool jumbled, const GrB_Descriptor desc) {
    (void)jumbled; (void)desc;
    API_BEGIN
    pack_csx(obj_check(A, "A"), true, Ap, Ai, Ax, Ap_size, Ai_size, Ax_size, iso);
    API_END(A)
}
GrB_Info GxB_Matrix_unpack_CSR(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Aj, void **Ax, GrB_Index *Ap_size,
                               GrB_Index *Aj_size, GrB_Index *Ax_size, bool *iso, bool *jumbled,
                               const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    unpack_csx(obj_check(A, "A"), false, Ap, Aj, Ax, Ap_size, Aj_size, Ax_size, iso, jumbled);
    API_END(A)
}
GrB_Info GxB_Matrix_unpack_CSC(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index *Ap_size,
                               GrB_Index *Ai_size, GrB_Index *Ax_size, bool *iso, bool *jumbled,
                               const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    unpack_csx(obj_check(A, "A"), true, Ap, Ai, Ax, Ap_size, Ai_size, Ax_size, iso, jumbled);
    API_END(A)
}

GrB_Info GxB_Vector_pack_Full(GrB_Vector v_, void **vx, GrB_Index vx_size, bool iso, const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (!vx || !*vx) grb_throw(GrB_NULL_POINTER, "vx is NULL");
    size_t ts = v->type->size;
    int64_t n = v->nrows;
    if ((int64_t)(vx_size / ts) < (iso ? 1 : n)) grb_throw(GrB_INVALID_VALUE, "vx is too small");
    if (iso) {
        void *one = dmalloc(ts);
        h2d(one, *vx, ts);
        GRB_TYPE_SWITCH(v->type->code, T, GRB_LAUNCH("fill", k_fill<T>, nblk(n), 256, 0, (T *)v->vals, n, (const T *)one));
        dfree(one);
    } else h2d(v->vals, *vx, (size_t)n * ts);
    CUDA_TRY(cudaMemsetAsync(v->present, 1, (size_t)n, g.stream));
    v->nvals = n; v->nvals_known = true; v->list_valid = false;
    stream_sync();
    g.h_free(*vx); *vx = nullptr;
    API_END(v_)
}
GrB_Info GxB_Vector_unpack_Full(GrB_Vector v_, void **vx, GrB_Index *vx_size, bool *iso, const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (!vx || !vx_size) grb_throw(GrB_NULL_POINTER, "vx is NULL");
    if (vec_nvals(v) != v->nrows) grb_throw(GrB_INVALID_VALUE, "vector is not full");
    size_t b = (size_t)(v->nrows > 0 ? v->nrows : 1) * v->type->size;
    void *h = g.h_malloc(b);
    if (!h) grb_throw(GrB_OUT_OF_MEMORY, "host allocation failed");
    d2h(h, v->vals, (size_t)v->nrows * v->type->size);
    stream_sync();
    *vx = h; *vx_size = b;
    if (iso) *iso = false;
    obj_clear(v);
    API_END(v_)
}
GrB_Info GxB_Vector_pack_Bitmap(GrB_Vector v_, int8_t **vb, void **vx, GrB_Index vb_size, GrB_Index vx_size,
                                bool iso, GrB_Index nvals, const GrB_Descriptor desc) {
    (void)desc; (void)nvals;
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (!vb || !vx || !*vb || !*vx) grb_throw(GrB_NULL_POINTER, "vb or vx is NULL");
    size_t ts = v->type->size;
    int64_t n = v->nrows;
    if ((int64_t)vb_size < n || (int64_t)(vx_size / ts) < (iso ? 1 : n)) grb_throw(GrB_INVALID_VALUE, "buffers too small");
    if (iso) {
        void *one = dmalloc(ts);
        h2d(one, *vx, ts);
        GRB_TYPE_SWITCH(v->type->code, T, GRB_LAUNCH("fill", k_fill<T>, nblk(n), 256, 0, (T *)v->vals, n, (const T *)one));
        dfree(one);
    } else h2d(v->vals, *vx, (size_t)n * ts);
    h2d(v->present, *vb, (size_t)n);
    count_present(v->present, n, v->d_nvals);
    vec_set_unknown_nvals(v);
    stream_sync();
    g.h_free(*vb); *vb = nullptr;
    g.h_free(*vx); *vx = nullptr;
    API_END(v_)
}
GrB_Info GxB_Vector_unpack_Bitmap(GrB_Vector v_, int8_t **vb, void **vx, GrB_Index *vb_size, GrB_Index *vx_size,
                                  bool *iso, GrB_Index *nvals, const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    GB_Object *v = obj_check(v_, "v");
    if (!vb || !vx || !vb_size || !vx_size) grb_throw(GrB_NULL_POINTER, "output pointers are NULL");
    int64_t n = v->nrows;
    size_t bb = (size_t)(n > 0 ? n : 1), xb = bb * v->type->size;
    int8_t *hb = (int8_t *)g.h_malloc(bb);
    void *hx = g.h_malloc(xb);
    if (!hb || !hx) grb_throw(GrB_OUT_OF_MEMORY, "host allocation failed");
    d2h(hb, v->present, (size_t)n);
    d2h(hx, v->vals, (size_t)n * v->type->size);
    stream_sync();
    if (nvals) *nvals = (GrB_Index)vec_nvals(v);
    *vb = hb; *vx = hx; *vb_size = bb; *vx_size = xb;
    if (iso) *iso = false;
    obj_clear(v);
    API_END(v_)
}

// ---------------------------------------------------------------- per-type entry points
#define DEF_TYPED(NAME, CT, CODE)                                                                                   \
    GrB_Info GrB_Matrix_build_##NAME(GrB_Matrix C, const GrB_Index *I, const GrB_Index *J, const CT *X,             \
                                     GrB_Index nvals, const GrB_BinaryOp dup) {                                     \
        return matrix_build<CT>(C, I, J, X, nvals, dup, CODE);                                                      \
    }                                                                                                               \
    GrB_Info GrB_Matrix_extractTuples_##NAME(GrB_Index *I, GrB_Index *J, CT *X, GrB_Index *nvals, const GrB_Matrix A) { \
        return matrix_extract_tuples<CT>(I, J, X, nvals, A, CODE);                                                  \
    }                                                                                                               \
    GrB_Info GrB_Matrix_setElement_##NAME(GrB_Matrix C, CT x, GrB_Index i, GrB_Index j) {                           \
        return matrix_set_element<CT>(C, x, i, j, CODE);                                                            \
    }                                                                                                               \
    GrB_Info GrB_Matrix_extractElement_##NAME(CT *x, const GrB_Matrix A, GrB_Index i, GrB_Index j) {                \
        return matrix_extract_element<CT>(x, A, i, j, CODE);                                                        \
    }                                                                                                               \
    GrB_Info GrB_Matrix_import_##NAME(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols,               \
                                      const GrB_Index *Ap, const GrB_Index *Ai, const CT *Ax, GrB_Index Ap_len,     \
                                      GrB_Index Ai_len, GrB_Index Ax_len, GrB_Format format) {                      \
        return matrix_import<CT>(A, type, nrows, ncols, Ap, Ai, Ax, Ap_len, Ai_len, Ax_len, format, CODE);          \
    }                                                                                                               \
    GrB_Info GrB_Matrix_export_##NAME(GrB_Index *Ap, GrB_Index *Ai, CT *Ax, GrB_Index *Ap_len, GrB_Index *Ai_len,   \
                                      GrB_Index *Ax_len, GrB_Format format, const GrB_Matrix A) {                   \
        return matrix_export<CT>(Ap, Ai, Ax, Ap_len, Ai_len, Ax_len, format, A, CODE);                              \
    }                                                                                                               \
    GrB_Info GrB_Vector_build_##NAME(GrB_Vector w, const GrB_Index *I, const CT *X, GrB_Index nvals,                \
                                     const GrB_BinaryOp dup) {                                                      \
        return vector_build<CT>(w, I, X, nvals, dup, CODE);                                                         \
    }                                                                                                               \
    GrB_Info GrB_Vector_extractTuples_##NAME(GrB_Index *I, CT *X, GrB_Index *nvals, const GrB_Vector v) {           \
        return vector_extract_tuples<CT>(I, X, nvals, v, CODE);                                                     \
    }                                                                                                               \
    GrB_Info GrB_Vector_setElement_##NAME(GrB_Vector w, CT x, GrB_Index i) {                                        \
        return matrix_set_element<CT>((GrB_Matrix)w, x, i, 0, CODE);                                                \
    }                                                                                                               \
    GrB_Info GrB_Vector_extractElement_##NAME(CT *x, const GrB_Vector v, GrB_Index i) {                             \
        return matrix_extract_element<CT>(x, (GrB_Matrix)v, i, 0, CODE);                                            \
    }                                                                                                               \
    GrB_Info GrB_Scalar_setElement_##NAME(GrB_Scalar s, CT x) { return scalar_set<CT>(s, x, CODE); }               \
    GrB_Info GrB_Scalar_extractElement_##NAME(CT *x, const GrB_Scalar s) { return scalar_get<CT>(x, s, CODE); }

#include "grb_abi_lists.inc"
GRB_TYPE_LIST(DEF_TYPED)

}  // extern "C"
