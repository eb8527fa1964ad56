// grb_matops.cu — the matrix-side plumbing around GrB_mxm: the masked/accumulated matrix write-back
// C<M,replace> = accum(C, T), plus GrB_transpose, matrix eWiseMult/eWiseAdd (needed by the reference's
// Matrix.isequal, graphblas/core/matrix.py:357-398), select(tril/triu/diag/offdiag) for building the
// triangle-count operand L (core/matrix.py:2452), and the row-block extraction used for 1-D sharding.
// All are row-wise sorted merges: one pass sizes every output row, a scan builds rowptr, a second pass fills.
#define HAVE_MATOPS 1
#include <cub/cub.cuh>

#include "grb_internal.cuh"
#include "grb_abi_lists.inc"
#include "grb_matops.cuh"
#include "grb_ops.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

int accum_code_of(const GrB_BinaryOp accum) {
    if (!accum) return -1;
    op_need_kernel(accum->op, accum->name);
    if (op_is_cmp(accum->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot be accumulators");
    return accum->op;
}

int64_t scan_counts(int64_t *counts, int64_t *out, int64_t n) {
    dzero(counts + n, sizeof(int64_t));
    size_t tb = 0;
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tb, counts, out, n + 1, g.stream));
    void *tmp = dmalloc(tb);
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, counts, out, n + 1, g.stream));
    g.launches += 2;
    dfree(tmp);
    return read_i64(out + n);
}

__device__ __forceinline__ bool mval_true(const void *mv, int vsize, int fp, int64_t i) {
    switch (vsize) {
    case 0: return true;
    case 1: return ((const uint8_t *)mv)[i] != 0;
    case 2: return ((const uint16_t *)mv)[i] != 0;
    case 4: { uint32_t b = ((const uint32_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    default: { uint64_t b = ((const uint64_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    }
}

// ---------------------------------------------------------------- write-back
template <class T> struct WbArgs {
    const int64_t *Cp; const uint32_t *Cj; const T *Cx; int c_iso;
    const int64_t *Tp; const uint32_t *Tj; const T *Tx; int t_iso;
    const int64_t *Mp; const uint32_t *Mj; const void *Mx; int m_vsize, m_fp, m_iso;   // Mp == nullptr: no mask
    int comp, replace, accum;
    int64_t nrows;
    int64_t *rownnz;            // pass 1 out
    const int64_t *Op; uint32_t *Oj; T *Ox;   // pass 2
};

// one thread per row: three-way sorted merge of C(i,:), T(i,:) and M(i,:)
template <class T, bool FILL>
__global__ void __launch_bounds__(128) k_mat_writeback(WbArgs<T> a) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= a.nrows) return;
    int64_t pc = a.Cp[i], ce = a.Cp[i + 1], pt = a.Tp[i], te = a.Tp[i + 1];
    int64_t pm = a.Mp ? a.Mp[i] : 0, me = a.Mp ? a.Mp[i + 1] : 0;
    int64_t o = FILL ? a.Op[i] : 0, cnt = 0;
    while (pc < ce || pt < te) {
        const uint32_t jc = pc < ce ? a.Cj[pc] : 0xFFFFFFFFu, jt = pt < te ? a.Tj[pt] : 0xFFFFFFFFu;
        const uint32_t j = jc < jt ? jc : jt;
        const bool has_c = (pc < ce && jc == j), has_t = (pt < te && jt == j);
        bool m = true;
        if (a.Mp) {
            while (pm < me && a.Mj[pm] < j) pm++;
            m = (pm < me && a.Mj[pm] == j) && mval_true(a.Mx, a.m_vsize, a.m_fp, a.m_iso ? 0 : pm);
        }
        const bool allowed = (m != (a.comp != 0));
        bool emit = false; T v = (T)0;
        if (allowed) {
            if (has_t) {
                if (FILL) {
                    v = a.Tx[a.t_iso ? 0 : pt];
                    if (a.accum >= 0 && has_c) v = binop_rt<T>(a.accum, a.Cx[a.c_iso ? 0 : pc], v);
                }
                emit = true;
            } else if (a.accum >= 0) { emit = true; if (FILL) v = a.Cx[a.c_iso ? 0 : pc]; }   // Z = C u T keeps C's entry
        } else if (has_c && !a.replace) { emit = true; if (FILL) v = a.Cx[a.c_iso ? 0 : pc]; }
        if (emit) { if (FILL) { a.Oj[o] = j; a.Ox[o] = v; o++; } else cnt++; }
        if (has_c) pc++;
        if (has_t) pt++;
    }
    if (!FILL) a.rownnz[i] = cnt;
}

static void set_result(GB_Object *C, Csr &R, bool c_is_vec) {
    if (c_is_vec) { csr_to_vec(R, C, true); csr_free(R); }
    else mat_set_csr(C, R);
}

void mat_writeback(GB_Object *C, GB_Object *mask, const DescFlags &f, int accum, Csr &T, int tcode, bool t_in_mask) {
    const int ccode = C->type->code;
    const bool c_is_vec = (C->kind != K_MATRIX);
    csr_cast_values(T, tcode, ccode);
    Csr Cv; bool cv_on = false;
    if (c_is_vec) { Cv = vec_as_csr_column(C); cv_on = true; }
    const Csr &Cc = c_is_vec ? Cv : C->csr;
    // fast paths: nothing of the old C survives, and T needs no filtering
    const bool t_clean = (!mask && !f.comp) || (mask && t_in_mask);
    const bool c_gone = (Cc.nnz == 0) || (accum < 0 && (!mask ? !f.comp : f.replace));
    if (t_clean && c_gone) {
        if (cv_on) csr_free(Cv);
        set_result(C, T, c_is_vec);
        return;
    }
    Csr Mv; bool mv_on = false;
    if (mask && mask->kind != K_MATRIX) { Mv = vec_as_csr_column(mask); mv_on = true; }
    const Csr *Mc = mask ? (mv_on ? &Mv : &mask->csr) : nullptr;
    const int64_t n = Cc.nrows;
    Csr R;
    R.nrows = Cc.nrows; R.ncols = Cc.ncols;
    int64_t *counts = dalloc<int64_t>(n + 1);
    R.rowptr = dalloc<int64_t>(n + 1);
    GRB_TYPE_SWITCH(ccode, W, {
        WbArgs<W> a;
        a.Cp = Cc.rowptr; a.Cj = Cc.colidx; a.Cx = (const W *)Cc.vals; a.c_iso = Cc.iso;
        a.Tp = T.rowptr; a.Tj = T.colidx; a.Tx = (const W *)T.vals; a.t_iso = T.iso;
        a.Mp = Mc ? Mc->rowptr : nullptr; a.Mj = Mc ? Mc->colidx : nullptr; a.Mx = Mc ? Mc->vals : nullptr;
        a.m_vsize = (mask && !f.structure) ? (int)mask->type->size : 0;
        a.m_fp = mask ? (mask->type->code == TC_FP32 || mask->type->code == TC_FP64) : 0;
        a.m_iso = Mc ? Mc->iso : 0;
        a.comp = f.comp; a.replace = f.replace; a.accum = accum;
        a.nrows = n; a.rownnz = counts; a.Op = nullptr; a.Oj = nullptr; a.Ox = nullptr;
        auto k1 = k_mat_writeback<W, false>;
        GRB_LAUNCH("mat_writeback_count", k1, nblk(n, 128), 128, 0, a);
        R.nnz = scan_counts(counts, R.rowptr, n);
        R.colidx = dalloc<uint32_t>(R.nnz);
        R.vals = dmalloc((size_t)(R.nnz > 0 ? R.nnz : 1) * sizeof(W));
        a.Op = R.rowptr; a.Oj = R.colidx; a.Ox = (W *)R.vals;
        auto k2 = k_mat_writeback<W, true>;
        GRB_LAUNCH("mat_writeback_fill", k2, nblk(n, 128), 128, 0, a);
    });
    dfree(counts);
    csr_free(T);
    if (cv_on) csr_free(Cv);
    if (mv_on) csr_free(Mv);
    set_result(C, R, c_is_vec);
}

// ---------------------------------------------------------------- operand views
// CSR of op(A) (A or A') as a borrowed pointer; vectors cast to n x 1 matrices are materialised into `tmp`.
struct MatView {
    const Csr *csr = nullptr;
    Csr tmp; bool tmp_on = false;
    int64_t nrows = 0, ncols = 0;
    ~MatView() { if (tmp_on) csr_free(tmp); }
};
static void mat_view(MatView &v, GB_Object *A, bool transpose) {
    if (A->kind != K_MATRIX) {
        v.tmp = transpose ? vec_as_csr_row(A) : vec_as_csr_column(A);
        v.tmp_on = true; v.csr = &v.tmp;
    } else if (transpose) { mat_ensure_csc(A); v.csr = &A->csc; }
    else v.csr = &A->csr;
    v.nrows = v.csr->nrows; v.ncols = v.csr->ncols;
}
static inline int64_t obj_rows(GB_Object *o) { return o->nrows; }
static inline int64_t obj_cols(GB_Object *o) { return o->kind == K_MATRIX ? o->ncols : 1; }

// ---------------------------------------------------------------- eWise
template <class X> struct EwArgs {
    const int64_t *Ap; const uint32_t *Aj; const X *Ax; int a_iso;
    const int64_t *Bp; const uint32_t *Bj; const X *Bx; int b_iso;
    int64_t nrows; int op;
    int64_t *rownnz; const int64_t *Op; uint32_t *Oj; void *Ox;
};
template <class X, bool UNION, bool CMP, bool FILL>
__global__ void __launch_bounds__(128) k_mat_ewise(EwArgs<X> a) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= a.nrows) return;
    int64_t pa = a.Ap[i], ae = a.Ap[i + 1], pb = a.Bp[i], be = a.Bp[i + 1];
    int64_t o = FILL ? a.Op[i] : 0, cnt = 0;
    while (UNION ? (pa < ae || pb < be) : (pa < ae && pb < be)) {
        const uint32_t ja = pa < ae ? a.Aj[pa] : 0xFFFFFFFFu, jb = pb < be ? a.Bj[pb] : 0xFFFFFFFFu;
        const uint32_t j = ja < jb ? ja : jb;
        const bool ha = (pa < ae && ja == j), hb = (pb < be && jb == j);
        if (UNION || (ha && hb)) {
            if (FILL) {
                const X x = ha ? a.Ax[a.a_iso ? 0 : pa] : (X)0, y = hb ? a.Bx[a.b_iso ? 0 : pb] : (X)0;
                a.Oj[o] = j;
                if (CMP) ((bool *)a.Ox)[o] = (ha && hb) ? cmp_rt<X>(a.op, x, y) : (ha ? (x != (X)0) : (y != (X)0));
                else ((X *)a.Ox)[o] = (ha && hb) ? binop_rt<X>(a.op, x, y) : (ha ? x : y);
                o++;
            } else cnt++;
        }
        if (ha) pa++;
        if (hb) pb++;
    }
    if (!FILL) a.rownnz[i] = cnt;
}

static void matrix_ewise(GrB_Matrix C_, const GrB_Matrix M_, const GrB_BinaryOp accum, const GrB_BinaryOp op,
                         const GrB_Matrix A_, const GrB_Matrix B_, const GrB_Descriptor desc, bool is_union) {
    require_init();
    GB_Object *C = obj_check(C_, "C"), *A = obj_check(A_, "A"), *B = obj_check(B_, "B");
    GB_Object *mask = M_ ? obj_check(M_, "mask") : nullptr;
    if (!op) grb_throw(GrB_NULL_POINTER, "op is NULL");
    DescFlags f = desc_flags(desc);
    const int acc = accum_code_of(accum);
    MatView va, vb;
    mat_view(va, A, f.tran0); mat_view(vb, B, f.tran1);
    if (va.nrows != vb.nrows || va.ncols != vb.ncols || obj_rows(C) != va.nrows || obj_cols(C) != va.ncols ||
        (mask && (obj_rows(mask) != va.nrows || obj_cols(mask) != va.ncols)))
        grb_throw(GrB_DIMENSION_MISMATCH, "eWise: matrix shapes differ");
    const int xcode = op->xcode, zcode = op->zcode;
    op_need_kernel(op->op, op->name);
    const bool cmp = op_is_cmp(op->op);
    Csr ac, bc; bool ac_on = false, bc_on = false;
    const Csr *pa = va.csr, *pb = vb.csr;
    if (A->type->code != xcode) { ac = csr_dup(*pa, A->type->size); csr_cast_values(ac, A->type->code, xcode); ac_on = true; pa = &ac; }
    if (B->type->code != xcode) { bc = csr_dup(*pb, B->type->size); csr_cast_values(bc, B->type->code, xcode); bc_on = true; pb = &bc; }
    const int64_t n = va.nrows;
    Csr T;
    T.nrows = va.nrows; T.ncols = va.ncols;
    int64_t *counts = dalloc<int64_t>(n + 1);
    T.rowptr = dalloc<int64_t>(n + 1);
    GRB_TYPE_SWITCH(xcode, X, {
        EwArgs<X> a;
        a.Ap = pa->rowptr; a.Aj = pa->colidx; a.Ax = (const X *)pa->vals; a.a_iso = pa->iso;
        a.Bp = pb->rowptr; a.Bj = pb->colidx; a.Bx = (const X *)pb->vals; a.b_iso = pb->iso;
        a.nrows = n; a.op = op->op; a.rownnz = counts; a.Op = nullptr; a.Oj = nullptr; a.Ox = nullptr;
        auto k1 = is_union ? k_mat_ewise<X, true, false, false> : k_mat_ewise<X, false, false, false>;
        GRB_LAUNCH("mat_ewise_count", k1, nblk(n, 128), 128, 0, a);
        T.nnz = scan_counts(counts, T.rowptr, n);
        T.colidx = dalloc<uint32_t>(T.nnz);
        T.vals = dmalloc((size_t)(T.nnz > 0 ? T.nnz : 1) * kTypeSize[zcode]);
        a.Op = T.rowptr; a.Oj = T.colidx; a.Ox = T.vals;
        auto k2 = cmp ? (is_union ? k_mat_ewise<X, true, true, true> : k_mat_ewise<X, false, true, true>)
                      : (is_union ? k_mat_ewise<X, true, false, true> : k_mat_ewise<X, false, false, true>);
        GRB_LAUNCH(is_union ? "mat_ewise_add" : "mat_ewise_mult", k2, nblk(n, 128), 128, 0, a);
    });
    dfree(counts);
    if (ac_on) csr_free(ac);
    if (bc_on) csr_free(bc);
    mat_writeback(C, mask, f, acc, T, zcode, false);
    if (g.blocking) stream_sync();
}

// ---------------------------------------------------------------- select (tril / triu / diag / offdiag)
// Rows are sorted by column, so every positional predicate keeps one contiguous run of a row (tril: a prefix, triu: a
// suffix, diag: at most one entry) or everything but one entry (offdiag): two binary searches per row instead of a
// walk over its entries (a thread per row walking a 27 K-entry hub row made select 7 of the 42 ms of triangle counting).
//   first[i] = position of the first kept entry, skip[i] = position of the dropped entry for offdiag (else -1)
__global__ void __launch_bounds__(256) k_select_ranges(const int64_t *Ap, const uint32_t *Aj, int64_t nrows, int op, int64_t y,
                                                       int64_t *rownnz, int64_t *first, int64_t *skip) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nrows) return;
    const int64_t s = Ap[i], e = Ap[i + 1];
    const int64_t t = i + y;                          // target column: j - i (<=, >=, ==, !=) y
    // lb = first position with j >= t, ub = first position with j > t
    int64_t lo = s, hi = e;
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if ((int64_t)Aj[mid] < t) lo = mid + 1; else hi = mid; }
    const int64_t lb = lo;
    const int64_t ub = (lb < e && (int64_t)Aj[lb] == t) ? lb + 1 : lb;
    int64_t f = s, cnt = 0, sk = -1;
    switch (op) {
    case IDX_TRIL: f = s; cnt = ub - s; break;
    case IDX_TRIU: f = lb; cnt = e - lb; break;
    case IDX_DIAG: f = lb; cnt = ub - lb; break;
    default: f = s; cnt = (e - s) - (ub - lb); sk = (ub > lb) ? lb : -1; break;
    }
    rownnz[i] = cnt; first[i] = f; skip[i] = sk;
}
__global__ void __launch_bounds__(256) k_select_src(const uint32_t *rowid, const int64_t *Op, const int64_t *first, const int64_t *skip,
                                                    int64_t n, int64_t *src) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t i = rowid[k];
    int64_t p = first[i] + (k - Op[i]);
    const int64_t sk = skip[i];
    if (sk >= 0 && p >= sk) p++;
    src[k] = p;
}
template <class T>
__global__ void k_gather_entries(const int64_t *src, int64_t n, const uint32_t *Aj, const T *Ax, int iso, uint32_t *Oj, T *Ox) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int64_t p = src[k];
    Oj[k] = Aj[p];
    if (!iso) Ox[k] = Ax[p];
}

static void matrix_select(GB_Object *C, GB_Object *mask, int acc, int idxop, GB_Object *A, int64_t y, const DescFlags &f) {
    MatView va;
    mat_view(va, A, f.tran0);
    if (obj_rows(C) != va.nrows || obj_cols(C) != va.ncols ||
        (mask && (obj_rows(mask) != va.nrows || obj_cols(mask) != va.ncols)))
        grb_throw(GrB_DIMENSION_MISMATCH, "select: shapes differ");
    const Csr &S = *va.csr;
    const int64_t n = S.nrows;
    Csr T;
    T.nrows = S.nrows; T.ncols = S.ncols; T.iso = S.iso;
    int64_t *counts = dalloc<int64_t>(n + 1);
    T.rowptr = dalloc<int64_t>(n + 1);
    int64_t *first = dalloc<int64_t>(n + 1), *skip = dalloc<int64_t>(n + 1);
    GRB_LAUNCH("mat_select_ranges", k_select_ranges, nblk(n, 256), 256, 0, S.rowptr, S.colidx, n, idxop, y, counts, first, skip);
    T.nnz = scan_counts(counts, T.rowptr, n);
    int64_t *src = dalloc<int64_t>(T.nnz);
    if (T.nnz) {
        uint32_t *rowid = csr_row_ids(T);
        GRB_LAUNCH("mat_select_src", k_select_src, nblk(T.nnz, 256), 256, 0, rowid, (const int64_t *)T.rowptr, first, skip, T.nnz, src);
        dfree(rowid);
    }
    dfree(first); dfree(skip);
    T.colidx = dalloc<uint32_t>(T.nnz);
    const size_t ts = A->type->size;
    T.vals = dmalloc(T.iso ? ts : (size_t)(T.nnz > 0 ? T.nnz : 1) * ts);
    if (T.iso) d2d(T.vals, S.vals, ts);
    GRB_TYPE_SWITCH(A->type->code, X,
        if (T.nnz) GRB_LAUNCH("gather_entries", k_gather_entries<X>, nblk(T.nnz), 256, 0, src, T.nnz, S.colidx,
                              (const X *)S.vals, (int)S.iso, T.colidx, (X *)T.vals));
    dfree(counts); dfree(src);
    mat_writeback(C, mask, f, acc, T, A->type->code, false);
}

// rows [r0, r1) of A as a new (r1-r0) x ncols matrix (1-D row-block sharding)
__global__ void k_shift_rowptr(const int64_t *in, int64_t n1, int64_t base, int64_t *out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n1) out[i] = in[i] - base;
}

extern "C" {

GrB_Info GrB_Matrix_eWiseMult_BinaryOp(GrB_Matrix C, const GrB_Matrix M, const GrB_BinaryOp accum, const GrB_BinaryOp op,
                                       const GrB_Matrix A, const GrB_Matrix B, const GrB_Descriptor desc) {
    API_BEGIN
    matrix_ewise(C, M, accum, op, A, B, desc, false);
    API_END(C)
}
GrB_Info GrB_Matrix_eWiseAdd_BinaryOp(GrB_Matrix C, const GrB_Matrix M, const GrB_BinaryOp accum, const GrB_BinaryOp op,
                                      const GrB_Matrix A, const GrB_Matrix B, const GrB_Descriptor desc) {
    API_BEGIN
    matrix_ewise(C, M, accum, op, A, B, desc, true);
    API_END(C)
}

GrB_Info GrB_transpose(GrB_Matrix C_, const GrB_Matrix M_, const GrB_BinaryOp accum, const GrB_Matrix A_,
                       const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *C = obj_check(C_, "C"), *A = obj_check(A_, "A");
    GB_Object *mask = M_ ? obj_check(M_, "mask") : nullptr;
    DescFlags f = desc_flags(desc);
    const int acc = accum_code_of(accum);
    MatView va;
    mat_view(va, A, !f.tran0);    // GrB_transpose with INP0 = TRAN is the identity copy
    if (obj_rows(C) != va.nrows || obj_cols(C) != va.ncols ||
        (mask && (obj_rows(mask) != va.nrows || obj_cols(mask) != va.ncols)))
        grb_throw(GrB_DIMENSION_MISMATCH, "transpose: shapes differ");
    Csr T = csr_dup(*va.csr, A->type->size);
    mat_writeback(C, mask, f, acc, T, A->type->code, false);
    if (g.blocking) stream_sync();
    API_END(C_)
}

GrB_Info GrB_Matrix_select_INT64(GrB_Matrix C_, const GrB_Matrix M_, const GrB_BinaryOp accum, const GrB_IndexUnaryOp op,
                                 const GrB_Matrix A_, int64_t y, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *C = obj_check(C_, "C"), *A = obj_check(A_, "A");
    GB_Object *mask = M_ ? obj_check(M_, "mask") : nullptr;
    if (!op) grb_throw(GrB_NULL_POINTER, "op is NULL");
    op_need_kernel(op->op, op->name);
    matrix_select(C, mask, accum_code_of(accum), op->op, A, y, desc_flags(desc));
    if (g.blocking) stream_sync();
    API_END(C_)
}

GrB_Info GxB_b200_Matrix_rowblock(GrB_Matrix *C_, const GrB_Matrix A_, GrB_Index row_begin, GrB_Index row_end) {
    if (!C_) return GrB_NULL_POINTER;
    *C_ = nullptr;
    GB_Object *o = nullptr;
    API_BEGIN
    require_init();
    GB_Object *A = obj_check(A_, "A");
    if (A->kind != K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "rowblock needs a matrix");
    if (row_begin > row_end || (int64_t)row_end > A->nrows) grb_throw(GrB_INVALID_INDEX, "rowblock: bad row range");
    const Csr &S = A->csr;
    const int64_t r0 = (int64_t)row_begin, r1 = (int64_t)row_end, n = r1 - r0;
    int64_t ends[2];
    d2h(&ends[0], S.rowptr + r0, sizeof(int64_t));
    d2h(&ends[1], S.rowptr + r1, sizeof(int64_t));
    stream_sync();
    Csr c;
    c.nrows = n; c.ncols = S.ncols; c.nnz = ends[1] - ends[0]; c.iso = S.iso;
    c.rowptr = dalloc<int64_t>(n + 1);
    GRB_LAUNCH("shift_rowptr", k_shift_rowptr, nblk(n + 1), 256, 0, S.rowptr + r0, n + 1, ends[0], c.rowptr);
    c.colidx = dalloc<uint32_t>(c.nnz);
    d2d(c.colidx, S.colidx + ends[0], (size_t)c.nnz * sizeof(uint32_t));
    const size_t ts = A->type->size;
    c.vals = dmalloc(c.iso ? ts : (size_t)(c.nnz > 0 ? c.nnz : 1) * ts);
    d2d(c.vals, c.iso ? S.vals : (const char *)S.vals + (size_t)ends[0] * ts, c.iso ? ts : (size_t)c.nnz * ts);
    o = new GB_Matrix_opaque();
    o->kind = K_MATRIX; o->type = A->type; o->nrows = n; o->ncols = S.ncols;
    mat_set_csr(o, c);
    *C_ = (GrB_Matrix)o;
    API_END(A_)
}

}  // extern "C"
