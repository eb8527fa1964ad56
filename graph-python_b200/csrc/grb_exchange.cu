// grb_exchange.cu — the SuiteSparse-style storage entry points python-graphblas's `.ss` namespace calls:
//   Matrix.ss.import_csr/csc  -> GxB_Matrix_import_CSR/CSC       (reference graphblas/core/ss/matrix.py:1282-1352, 1472-1543)
//   Matrix.ss.export/unpack   -> GxB_Matrix_{export,unpack}_{CSR,CSC,BitmapR,BitmapC,FullR,FullC}   (:744-1162)
//   Matrix.ss.import_bitmapr/fullr ... -> GxB_Matrix_{import,pack}_{BitmapR,BitmapC,FullR,FullC}    (:2161-2680)
//   Vector.ss.import_sparse/bitmap/full -> GxB_Vector_{import,pack}_{CSC,Bitmap,Full}  (core/ss/vector.py:946-1016, 1128-1202, 1303-1360)
//   Vector.ss.export/unpack   -> GxB_Vector_{export,unpack}_{CSC,Bitmap,Full}           (core/ss/vector.py:494-636)
// import = GrB_*_new + pack; export = unpack + GrB_*_free.  Host buffers come from / go to the allocator given to GxB_init
// (numpy's), ownership moves exactly as SuiteSparse's does: pack frees the caller's buffers and NULLs the pointers, unpack
// returns freshly allocated ones.  Inside the library data lives in HBM only (CSR / dense+presence), so bitmap and full
// matrices are converted on the device; they are meant for small dense operands (nrows * ncols <= 2^31).
#include "grb_internal.cuh"
#include "grb_matops.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

// ---------------------------------------------------------------- dense (bitmap / full) matrix <-> CSR on the device
// pos(i, j) in the dense arrays: row-major (BitmapR/FullR) i * ncols + j, column-major (BitmapC/FullC) j * nrows + i
__global__ void k_dense_rowcount(const int8_t *bitmap, int64_t nrows, int64_t ncols, int rowmajor, int64_t *counts) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nrows) return;
    int64_t c = 0;
    if (!bitmap) c = ncols;
    else for (int64_t j = 0; j < ncols; j++) c += bitmap[rowmajor ? i * ncols + j : j * nrows + i] != 0;
    counts[i] = c;
}
template <class T>
__global__ void k_dense_to_csr(const int8_t *bitmap, const T *vals, int iso, int64_t nrows, int64_t ncols, int rowmajor,
                               const int64_t *rowptr, uint32_t *colidx, T *out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nrows) return;
    int64_t p = rowptr[i];
    for (int64_t j = 0; j < ncols; j++) {
        const int64_t d = rowmajor ? i * ncols + j : j * nrows + i;
        if (bitmap && !bitmap[d]) continue;
        colidx[p] = (uint32_t)j;
        if (!iso) out[p] = vals[d];
        p++;
    }
    if (iso && i == 0) out[0] = vals[0];
}
template <class T>
__global__ void k_csr_to_dense(const int64_t *rowptr, const uint32_t *colidx, const T *vals, int iso, int64_t nrows,
                               int64_t ncols, int rowmajor, int8_t *bitmap, T *out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nrows) return;
    for (int64_t p = rowptr[i]; p < rowptr[i + 1]; p++) {
        const int64_t d = rowmajor ? i * ncols + colidx[p] : (int64_t)colidx[p] * nrows + i;
        if (bitmap) bitmap[d] = 1;
        out[d] = vals[iso ? 0 : p];
    }
}

static int64_t dense_cells(GB_Object *A) {
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "bitmap / full matrix formats need a matrix");
    const int64_t cells = A->nrows * A->ncols;
    if (A->nrows && cells / A->nrows != A->ncols) grb_throw(GrB_OUT_OF_MEMORY, "nrows * ncols overflows");
    if (cells > (1ll << 31)) grb_throw(GrB_OUT_OF_MEMORY, "bitmap / full matrix formats are limited to 2^31 cells");
    return cells;
}

// host dense arrays -> A's CSR.  Ab may be NULL (full).  Frees the host buffers (ownership transfer) and NULLs the pointers.
static void pack_dense(GB_Object *A, bool rowmajor, int8_t **Ab, void **Ax, GrB_Index Ab_size, GrB_Index Ax_size, bool iso) {
    const int64_t cells = dense_cells(A);
    if (!Ax || !*Ax || (Ab && !*Ab)) grb_throw(GrB_NULL_POINTER, "Ab or Ax is NULL");
    const size_t ts = A->type->size;
    const int64_t nv = iso ? 1 : cells;
    if ((Ab && (int64_t)Ab_size < cells) || (int64_t)(Ax_size / ts) < nv) grb_throw(GrB_INVALID_VALUE, "bitmap / value buffers are too small");
    int8_t *db = nullptr;
    if (Ab) { db = dalloc<int8_t>(cells); h2d(db, *Ab, (size_t)cells); }
    void *dx = dmalloc((size_t)(nv > 0 ? nv : 1) * ts);
    h2d(dx, *Ax, (size_t)nv * ts);
    Csr c;
    c.nrows = A->nrows; c.ncols = A->ncols; c.iso = iso;
    int64_t *counts = dalloc<int64_t>(c.nrows + 1);
    c.rowptr = dalloc<int64_t>(c.nrows + 1);
    GRB_LAUNCH("dense_rowcount", k_dense_rowcount, nblk(c.nrows), 256, 0, db, c.nrows, c.ncols, (int)rowmajor, counts);
    c.nnz = scan_counts(counts, c.rowptr, c.nrows);
    c.colidx = dalloc<uint32_t>(c.nnz);
    c.vals = dmalloc((size_t)((iso ? 1 : c.nnz) > 0 ? (iso ? 1 : c.nnz) : 1) * ts);
    GRB_TYPE_SWITCH(A->type->code, T, {
        auto kern = k_dense_to_csr<T>;
        GRB_LAUNCH("dense_to_csr", kern, nblk(c.nrows), 256, 0, db, (const T *)dx, (int)iso, c.nrows, c.ncols, (int)rowmajor,
                   c.rowptr, c.colidx, (T *)c.vals);
    });
    dfree(counts); dfree(db); dfree(dx);
    mat_set_csr(A, c);
    stream_sync();
    if (Ab) { g.h_free(*Ab); *Ab = nullptr; }
    g.h_free(*Ax); *Ax = nullptr;
}

// A's CSR -> freshly allocated host dense arrays; A becomes empty.  want_bitmap = false: A must be full.
static void unpack_dense(GB_Object *A, bool rowmajor, int8_t **Ab, void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso,
                         GrB_Index *nvals, bool want_bitmap) {
    const int64_t cells = dense_cells(A);
    if (!Ax || !Ax_size || (want_bitmap && (!Ab || !Ab_size))) grb_throw(GrB_NULL_POINTER, "output pointers are NULL");
    const Csr &c = A->csr;
    if (!want_bitmap && c.nnz != cells) grb_throw(GrB_INVALID_VALUE, "matrix is not full");
    const size_t ts = A->type->size;
    const bool give_iso = iso && c.iso && (!want_bitmap || c.nnz > 0);
    const size_t bb = (size_t)(cells > 0 ? cells : 1), xb = (give_iso ? 1 : bb) * ts;
    int8_t *hb = want_bitmap ? (int8_t *)g.h_malloc(bb) : nullptr;
    void *hx = g.h_malloc(xb);
    if ((want_bitmap && !hb) || !hx) { if (hb) g.h_free(hb); if (hx) g.h_free(hx); grb_throw(GrB_OUT_OF_MEMORY, "host allocation failed in unpack"); }
    int8_t *db = nullptr;
    if (want_bitmap) { db = dalloc<int8_t>(cells); dzero(db, (size_t)cells); }
    if (give_iso) {
        // presence only; the single value travels alone
        if (want_bitmap) {
            void *scratch = dmalloc(bb * ts);
            GRB_TYPE_SWITCH(A->type->code, T, {
                auto kern = k_csr_to_dense<T>;
                GRB_LAUNCH("csr_to_dense", kern, nblk(c.nrows), 256, 0, c.rowptr, c.colidx, (const T *)c.vals, 1, c.nrows, c.ncols,
                           (int)rowmajor, db, (T *)scratch);
            });
            dfree(scratch);
        }
        d2h(hx, c.vals, ts);
    } else {
        void *dx = dmalloc(bb * ts);
        dzero(dx, bb * ts);
        GRB_TYPE_SWITCH(A->type->code, T, {
            auto kern = k_csr_to_dense<T>;
            GRB_LAUNCH("csr_to_dense", kern, nblk(c.nrows), 256, 0, c.rowptr, c.colidx, (const T *)c.vals, (int)c.iso, c.nrows,
                       c.ncols, (int)rowmajor, db, (T *)dx);
        });
        d2h(hx, dx, (size_t)cells * ts);
        dfree(dx);
    }
    if (want_bitmap) d2h(hb, db, (size_t)cells);
    stream_sync();
    dfree(db);
    if (want_bitmap) { *Ab = hb; *Ab_size = bb; }
    *Ax = hx; *Ax_size = xb;
    if (iso) *iso = give_iso;
    if (nvals) *nvals = (GrB_Index)c.nnz;
    obj_clear(A);
}

// ---------------------------------------------------------------- sparse vectors (the "CSC" format of an n x 1 matrix)
template <class T>
__global__ void k_vec_scatter(const uint64_t *idx, const T *vals, int iso, int64_t nvals, int64_t n, T *out, uint8_t *present,
                              int *bad) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= nvals) return;
    const uint64_t i = idx[k];
    if (i >= (uint64_t)n) { *bad = 1; return; }
    out[i] = vals[iso ? 0 : k];
    present[i] = 1;
}
template <class T>
__global__ void k_vec_gather_list(const uint32_t *list, const T *vals, int64_t nvals, uint64_t *idx, T *out) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= nvals) return;
    const uint32_t i = list[k];
    idx[k] = i;
    out[k] = vals[i];
}

static void need_vector(GB_Object *v) {
    if (v->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "a vector is required");
}

static void vec_pack_sparse(GB_Object *v, GrB_Index **vi, void **vx, GrB_Index vi_size, GrB_Index vx_size, bool iso, GrB_Index nvals) {
    need_vector(v);
    if (!vi || !vx || !*vx || (nvals && !*vi)) grb_throw(GrB_NULL_POINTER, "vi or vx is NULL");
    const size_t ts = v->type->size;
    const int64_t n = v->nrows, nv = (int64_t)nvals, nx = iso ? 1 : nv;
    if ((int64_t)(vi_size / 8) < nv || (int64_t)(vx_size / ts) < nx) grb_throw(GrB_INVALID_VALUE, "vi or vx is too small for nvals entries");
    dzero(v->present, (size_t)n);
    if (nv) {
        uint64_t *di = dalloc<uint64_t>(nv);
        void *dx = dmalloc((size_t)(nx > 0 ? nx : 1) * ts);
        int *bad = (int *)(g.d_scratch + 8);
        dzero(bad, sizeof(int));
        h2d(di, *vi, (size_t)nv * 8);
        h2d(dx, *vx, (size_t)nx * ts);
        GRB_TYPE_SWITCH(v->type->code, T, {
            auto kern = k_vec_scatter<T>;
            GRB_LAUNCH("vec_scatter", kern, nblk(nv), 256, 0, di, (const T *)dx, (int)iso, nv, n, (T *)v->vals, v->present, bad);
        });
        int hbad = 0;
        d2h(&hbad, bad, sizeof(int));
        stream_sync();
        dfree(di); dfree(dx);
        if (hbad) { obj_clear(v); grb_throw(GrB_INDEX_OUT_OF_BOUNDS, "vector index out of range in pack/import"); }
    }
    count_present(v->present, n, v->d_nvals);      // duplicates collapse: the count comes from the presence bytes
    vec_set_unknown_nvals(v);
    stream_sync();
    if (*vi) { g.h_free(*vi); *vi = nullptr; }
    g.h_free(*vx); *vx = nullptr;
}

static void vec_unpack_sparse(GB_Object *v, GrB_Index **vi, void **vx, GrB_Index *vi_size, GrB_Index *vx_size, bool *iso,
                              GrB_Index *nvals, bool *jumbled) {
    need_vector(v);
    if (!vi || !vx || !vi_size || !vx_size) grb_throw(GrB_NULL_POINTER, "output pointers are NULL");
    vec_build_list(v);
    const int64_t nv = v->nvals;
    const size_t ts = v->type->size;
    const size_t ib = (size_t)(nv > 0 ? nv : 1) * 8, xb = (size_t)(nv > 0 ? nv : 1) * ts;
    uint64_t *hi = (uint64_t *)g.h_malloc(ib);
    void *hx = g.h_malloc(xb);
    if (!hi || !hx) { if (hi) g.h_free(hi); if (hx) g.h_free(hx); grb_throw(GrB_OUT_OF_MEMORY, "host allocation failed in unpack"); }
    if (nv) {
        uint64_t *di = dalloc<uint64_t>(nv);
        void *dx = dmalloc((size_t)nv * ts);
        GRB_TYPE_SWITCH(v->type->code, T, {
            auto kern = k_vec_gather_list<T>;
            GRB_LAUNCH("vec_gather_list", kern, nblk(nv), 256, 0, v->list, (const T *)v->vals, nv, di, (T *)dx);
        });
        d2h(hi, di, (size_t)nv * 8);
        d2h(hx, dx, (size_t)nv * ts);
        stream_sync();
        dfree(di); dfree(dx);
    }
    *vi = hi; *vx = hx; *vi_size = ib; *vx_size = xb;
    if (iso) *iso = false;
    if (nvals) *nvals = (GrB_Index)nv;
    if (jumbled) *jumbled = false;
    obj_clear(v);
}

// ---------------------------------------------------------------- C-ABI
extern "C" {

// ---- bitmap / full matrices
#define DEF_DENSE_MATRIX(SUFFIX, ROWMAJOR)                                                                                     \
    GrB_Info GxB_Matrix_pack_Bitmap##SUFFIX(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index Ab_size, GrB_Index Ax_size,       \
                                            bool iso, GrB_Index nvals, const GrB_Descriptor desc) {                           \
        (void)nvals; (void)desc;                                                                                               \
        API_BEGIN                                                                                                              \
        if (!Ab) grb_throw(GrB_NULL_POINTER, "Ab is NULL");                                                                    \
        pack_dense(obj_check(A, "A"), ROWMAJOR, Ab, Ax, Ab_size, Ax_size, iso);                                                \
        API_END(A)                                                                                                             \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_unpack_Bitmap##SUFFIX(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size,   \
                                              bool *iso, GrB_Index *nvals, const GrB_Descriptor desc) {                       \
        (void)desc;                                                                                                            \
        API_BEGIN                                                                                                              \
        unpack_dense(obj_check(A, "A"), ROWMAJOR, Ab, Ax, Ab_size, Ax_size, iso, nvals, true);                                 \
        API_END(A)                                                                                                             \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_pack_Full##SUFFIX(GrB_Matrix A, void **Ax, GrB_Index Ax_size, bool iso, const GrB_Descriptor desc) {  \
        (void)desc;                                                                                                            \
        API_BEGIN                                                                                                              \
        pack_dense(obj_check(A, "A"), ROWMAJOR, nullptr, Ax, 0, Ax_size, iso);                                                 \
        API_END(A)                                                                                                             \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_unpack_Full##SUFFIX(GrB_Matrix A, void **Ax, GrB_Index *Ax_size, bool *iso,                           \
                                            const GrB_Descriptor desc) {                                                      \
        (void)desc;                                                                                                            \
        API_BEGIN                                                                                                              \
        unpack_dense(obj_check(A, "A"), ROWMAJOR, nullptr, Ax, nullptr, Ax_size, iso, nullptr, false);                         \
        API_END(A)                                                                                                             \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_import_Bitmap##SUFFIX(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, int8_t **Ab,    \
                                              void **Ax, GrB_Index Ab_size, GrB_Index Ax_size, bool iso, GrB_Index nvals,     \
                                              const GrB_Descriptor desc) {                                                    \
        if (!A) return GrB_NULL_POINTER;                                                                                       \
        GrB_Info info = GrB_Matrix_new(A, type, nrows, ncols);                                                                 \
        if (info != GrB_SUCCESS) return info;                                                                                  \
        info = GxB_Matrix_pack_Bitmap##SUFFIX(*A, Ab, Ax, Ab_size, Ax_size, iso, nvals, desc);                                 \
        if (info != GrB_SUCCESS) GrB_Matrix_free(A);                                                                           \
        return info;                                                                                                           \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_import_Full##SUFFIX(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, void **Ax,        \
                                            GrB_Index Ax_size, bool iso, const GrB_Descriptor desc) {                         \
        if (!A) return GrB_NULL_POINTER;                                                                                       \
        GrB_Info info = GrB_Matrix_new(A, type, nrows, ncols);                                                                 \
        if (info != GrB_SUCCESS) return info;                                                                                  \
        info = GxB_Matrix_pack_Full##SUFFIX(*A, Ax, Ax_size, iso, desc);                                                       \
        if (info != GrB_SUCCESS) GrB_Matrix_free(A);                                                                           \
        return info;                                                                                                           \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_export_Bitmap##SUFFIX(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, int8_t **Ab, \
                                              void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso,                   \
                                              GrB_Index *nvals, const GrB_Descriptor desc) {                                  \
        if (!A || !*A) return GrB_NULL_POINTER;                                                                                \
        GrB_Info info = export_header((GB_Object *)*A, type, nrows, ncols);                                                    \
        if (info == GrB_SUCCESS) info = GxB_Matrix_unpack_Bitmap##SUFFIX(*A, Ab, Ax, Ab_size, Ax_size, iso, nvals, desc);      \
        if (info == GrB_SUCCESS) info = GrB_Matrix_free(A);                                                                    \
        return info;                                                                                                           \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_export_Full##SUFFIX(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, void **Ax,     \
                                            GrB_Index *Ax_size, bool *iso, const GrB_Descriptor desc) {                       \
        if (!A || !*A) return GrB_NULL_POINTER;                                                                                \
        GrB_Info info = export_header((GB_Object *)*A, type, nrows, ncols);                                                    \
        if (info == GrB_SUCCESS) info = GxB_Matrix_unpack_Full##SUFFIX(*A, Ax, Ax_size, iso, desc);                            \
        if (info == GrB_SUCCESS) info = GrB_Matrix_free(A);                                                                    \
        return info;                                                                                                           \
    }

// type / shape outputs of the export forms (the object is freed by the caller of this helper once the unpack succeeded)
static GrB_Info export_header(GB_Object *o, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols) {
    if (o->magic != GRB_MAGIC) return GrB_INVALID_OBJECT;
    if (type) *type = o->type;
    if (nrows) *nrows = (GrB_Index)o->nrows;
    if (ncols) *ncols = (GrB_Index)o->ncols;
    return GrB_SUCCESS;
}

DEF_DENSE_MATRIX(R, true)
DEF_DENSE_MATRIX(C, false)

// ---- CSR / CSC import + export (pack / unpack themselves live in grb_containers.cu)
#define DEF_CSX_MATRIX(FMT)                                                                                                    \
    GrB_Info GxB_Matrix_import_##FMT(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, GrB_Index **Ap,          \
                                     GrB_Index **Ai, void **Ax, GrB_Index Ap_size, GrB_Index Ai_size, GrB_Index Ax_size,      \
                                     bool iso, bool jumbled, const GrB_Descriptor desc) {                                     \
        if (!A) return GrB_NULL_POINTER;                                                                                       \
        GrB_Info info = GrB_Matrix_new(A, type, nrows, ncols);                                                                 \
        if (info != GrB_SUCCESS) return info;                                                                                  \
        info = GxB_Matrix_pack_##FMT(*A, Ap, Ai, Ax, Ap_size, Ai_size, Ax_size, iso, jumbled, desc);                           \
        if (info != GrB_SUCCESS) GrB_Matrix_free(A);                                                                           \
        return info;                                                                                                           \
    }                                                                                                                          \
    GrB_Info GxB_Matrix_export_##FMT(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, GrB_Index **Ap,       \
                                     GrB_Index **Ai, void **Ax, GrB_Index *Ap_size, GrB_Index *Ai_size,                       \
                                     GrB_Index *Ax_size, bool *iso, bool *jumbled, const GrB_Descriptor desc) {               \
        if (!A || !*A) return GrB_NULL_POINTER;                                                                                \
        GrB_Info info = export_header((GB_Object *)*A, type, nrows, ncols);                                                    \
        if (info == GrB_SUCCESS) info = GxB_Matrix_unpack_##FMT(*A, Ap, Ai, Ax, Ap_size, Ai_size, Ax_size, iso, jumbled, desc); \
        if (info == GrB_SUCCESS) info = GrB_Matrix_free(A);                                                                    \
        return info;                                                                                                           \
    }
DEF_CSX_MATRIX(CSR)
DEF_CSX_MATRIX(CSC)

// ---- vectors: sparse ("CSC"), and import / export forms of bitmap and full
GrB_Info GxB_Vector_pack_CSC(GrB_Vector v, GrB_Index **vi, void **vx, GrB_Index vi_size, GrB_Index vx_size, bool iso,
                             GrB_Index nvals, bool jumbled, const GrB_Descriptor desc) {
    (void)jumbled; (void)desc;
    API_BEGIN
    vec_pack_sparse(obj_check(v, "v"), vi, vx, vi_size, vx_size, iso, nvals);
    API_END(v)
}
GrB_Info GxB_Vector_unpack_CSC(GrB_Vector v, GrB_Index **vi, void **vx, GrB_Index *vi_size, GrB_Index *vx_size, bool *iso,
                               GrB_Index *nvals, bool *jumbled, const GrB_Descriptor desc) {
    (void)desc;
    API_BEGIN
    vec_unpack_sparse(obj_check(v, "v"), vi, vx, vi_size, vx_size, iso, nvals, jumbled);
    API_END(v)
}
GrB_Info GxB_Vector_import_CSC(GrB_Vector *v, GrB_Type type, GrB_Index n, GrB_Index **vi, void **vx, GrB_Index vi_size,
                               GrB_Index vx_size, bool iso, GrB_Index nvals, bool jumbled, const GrB_Descriptor desc) {
    if (!v) return GrB_NULL_POINTER;
    GrB_Info info = GrB_Vector_new(v, type, n);
    if (info != GrB_SUCCESS) return info;
    info = GxB_Vector_pack_CSC(*v, vi, vx, vi_size, vx_size, iso, nvals, jumbled, desc);
    if (info != GrB_SUCCESS) GrB_Vector_free(v);
    return info;
}
GrB_Info GxB_Vector_export_CSC(GrB_Vector *v, GrB_Type *type, GrB_Index *n, GrB_Index **vi, void **vx, GrB_Index *vi_size,
                               GrB_Index *vx_size, bool *iso, GrB_Index *nvals, bool *jumbled, const GrB_Descriptor desc) {
    if (!v || !*v) return GrB_NULL_POINTER;
    GrB_Info info = export_header((GB_Object *)*v, type, n, nullptr);
    if (info == GrB_SUCCESS) info = GxB_Vector_unpack_CSC(*v, vi, vx, vi_size, vx_size, iso, nvals, jumbled, desc);
    if (info == GrB_SUCCESS) info = GrB_Vector_free(v);
    return info;
}
GrB_Info GxB_Vector_import_Bitmap(GrB_Vector *v, GrB_Type type, GrB_Index n, int8_t **vb, void **vx, GrB_Index vb_size,
                                  GrB_Index vx_size, bool iso, GrB_Index nvals, const GrB_Descriptor desc) {
    if (!v) return GrB_NULL_POINTER;
    GrB_Info info = GrB_Vector_new(v, type, n);
    if (info != GrB_SUCCESS) return info;
    info = GxB_Vector_pack_Bitmap(*v, vb, vx, vb_size, vx_size, iso, nvals, desc);
    if (info != GrB_SUCCESS) GrB_Vector_free(v);
    return info;
}
GrB_Info GxB_Vector_export_Bitmap(GrB_Vector *v, GrB_Type *type, GrB_Index *n, int8_t **vb, void **vx, GrB_Index *vb_size,
                                  GrB_Index *vx_size, bool *iso, GrB_Index *nvals, const GrB_Descriptor desc) {
    if (!v || !*v) return GrB_NULL_POINTER;
    GrB_Info info = export_header((GB_Object *)*v, type, n, nullptr);
    if (info == GrB_SUCCESS) info = GxB_Vector_unpack_Bitmap(*v, vb, vx, vb_size, vx_size, iso, nvals, desc);
    if (info == GrB_SUCCESS) info = GrB_Vector_free(v);
    return info;
}
GrB_Info GxB_Vector_import_Full(GrB_Vector *v, GrB_Type type, GrB_Index n, void **vx, GrB_Index vx_size, bool iso,
                                const GrB_Descriptor desc) {
    if (!v) return GrB_NULL_POINTER;
    GrB_Info info = GrB_Vector_new(v, type, n);
    if (info != GrB_SUCCESS) return info;
    info = GxB_Vector_pack_Full(*v, vx, vx_size, iso, desc);
    if (info != GrB_SUCCESS) GrB_Vector_free(v);
    return info;
}
GrB_Info GxB_Vector_export_Full(GrB_Vector *v, GrB_Type *type, GrB_Index *n, void **vx, GrB_Index *vx_size, bool *iso,
                                const GrB_Descriptor desc) {
    if (!v || !*v) return GrB_NULL_POINTER;
    GrB_Info info = export_header((GB_Object *)*v, type, n, nullptr);
    if (info == GrB_SUCCESS) info = GxB_Vector_unpack_Full(*v, vx, vx_size, iso, desc);
    if (info == GrB_SUCCESS) info = GrB_Vector_free(v);
    return info;
}

}  // extern "C"
