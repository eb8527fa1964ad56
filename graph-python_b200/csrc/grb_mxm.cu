// grb_mxm.cu — GrB_mxm: argument checking, method choice (masked dot product vs two-phase hash Gustavson),
// row binning, the launch sequence, and the hand-over to the matrix write-back.
// Replaces what python-graphblas reaches through `call("GrB_mxm", [C, Mask, accum, op, A, B, desc])`
// (reference graphblas/core/base.py:476-483; Matrix.mxm core/matrix.py:2206-2251; Vector.outer core/vector.py:1683-1692).
#define HAVE_MXM 1
#include <cub/cub.cuh>

#include "grb_internal.cuh"
#include "grb_gemm_launch.cuh"
#include "grb_matops.cuh"
#include "grb_ops.cuh"
#include "grb_spgemm.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

bool gemm_launch_numeric(int mon, int mul, int tcode, const GemmLaunch &L) {
    switch (mon) {
    case OP_PLUS: return gemm_numeric_PLUS(mul, tcode, L);
    case OP_MIN: return gemm_numeric_MIN(mul, tcode, L);
    case OP_MAX: return gemm_numeric_MAX(mul, tcode, L);
    case OP_ANY: return gemm_numeric_ANY(mul, tcode, L);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return gemm_numeric_BOOLMON(mul | (mon << 8), tcode, L);
    }
    return false;
}
bool gemm_launch_dot(int mon, int mul, int tcode, const void *da, int64_t mnnz) {
    switch (mon) {
    case OP_PLUS: return gemm_dot_PLUS(mul, tcode, da, mnnz);
    case OP_MIN: return gemm_dot_MIN(mul, tcode, da, mnnz);
    case OP_MAX: return gemm_dot_MAX(mul, tcode, da, mnnz);
    case OP_ANY: return gemm_dot_ANY(mul, tcode, da, mnnz);
    case OP_LOR: case OP_LAND: case OP_LXOR: case OP_LXNOR: return gemm_dot_BOOLMON(mul | (mon << 8), tcode, da, mnnz);
    }
    return false;
}

// ---------------------------------------------------------------- operand views (rows of op(A))
struct View {
    const Csr *csr = nullptr;
    Csr tmp; bool tmp_on = false;
    ~View() { if (tmp_on) csr_free(tmp); }
    void adopt(Csr &c) { if (tmp_on) csr_free(tmp); tmp = c; c = Csr(); tmp_on = true; csr = &tmp; }
};
static void view_of(View &v, GB_Object *A, bool transpose) {
    if (A->kind != K_MATRIX) { Csr c = transpose ? vec_as_csr_row(A) : vec_as_csr_column(A); v.adopt(c); }
    else if (transpose) { mat_ensure_csc(A); v.csr = &A->csc; }
    else { mat_ensure_csr(A); v.csr = &A->csr; }
}
static void view_cast(View &v, int from_code, int to_code, size_t from_size, int via_code = -1) {
    if (via_code == from_code) via_code = -1;
    if (from_code == to_code && via_code < 0) return;
    Csr c = csr_dup(*v.csr, from_size);
    if (via_code >= 0) { csr_cast_values(c, from_code, via_code); csr_cast_values(c, via_code, to_code); }
    else csr_cast_values(c, from_code, to_code);
    v.adopt(c);
}

// ---------------------------------------------------------------- row work estimate + binning
// one warp per row of A: ub[i] = min(sum_k nnz(B(k,:)), ncols)
__global__ void __launch_bounds__(256) k_row_flops(const int64_t *Ap, const uint32_t *Aj, const int64_t *Bp, int64_t nrows,
                                                   int64_t ncols, int64_t *ub, unsigned long long *total) {
    const int64_t row = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= nrows) return;
    int64_t s = 0;
    for (int64_t p = Ap[row] + lane, e = Ap[row + 1]; p < e; p += 32) { const uint32_t k = Aj[p]; s += Bp[k + 1] - Bp[k]; }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) { ub[row] = s < ncols ? s : ncols; if (s) atomicAdd(total, (unsigned long long)s); }
}

constexpr int NBINS = 7;   // 0: empty, 1: warp, 2..5: CTA hash 512/2048/8192/16384 slots, 6: dense
__host__ __device__ static inline int bin_of(int64_t u) {
    return u == 0 ? 0 : u <= 32 ? 1 : u <= 256 ? 2 : u <= 1024 ? 3 : u <= 4096 ? 4 : u <= 8192 ? 5 : 6;
}
static const int kBinTable[NBINS] = {0, HW_TABLE, 512, 2048, 8192, 16384, 0};
static const int kBinThreads[NBINS] = {0, HW_ROWS * 32, 64, 128, 256, 512, 1024};

__global__ void __launch_bounds__(256) k_bin_count(const int64_t *ub, int64_t nrows, unsigned long long *counts) {
    __shared__ int h[NBINS];
    if (threadIdx.x < NBINS) h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < nrows) atomicAdd(&h[bin_of(ub[i])], 1);
    __syncthreads();
    if (threadIdx.x < NBINS && h[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)h[threadIdx.x]);
}
// cursors[b] starts at the bin's offset; order inside a bin is arbitrary (results do not depend on it)
__global__ void __launch_bounds__(256) k_bin_scatter(const int64_t *ub, int64_t nrows, unsigned long long *cursors, int32_t *rows) {
    __shared__ int h[NBINS];
    __shared__ unsigned long long base[NBINS];
    if (threadIdx.x < NBINS) h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int b = -1, slot = 0;
    if (i < nrows) { b = bin_of(ub[i]); slot = atomicAdd(&h[b], 1); }
    __syncthreads();
    if (threadIdx.x < NBINS && h[threadIdx.x]) base[threadIdx.x] = atomicAdd(&cursors[threadIdx.x], (unsigned long long)h[threadIdx.x]);
    __syncthreads();
    if (b >= 0) rows[base[b] + slot] = (int32_t)i;
}

struct Bins {
    int32_t *rows = nullptr;       // row ids grouped by bin
    int64_t off[NBINS + 1];
    ~Bins() { dfree(rows); }
};
static void make_bins(const int64_t *work, int64_t nrows, Bins &b) {
    unsigned long long *d = (unsigned long long *)(g.d_scratch + 24);   // 2*NBINS counters
    dzero(d, sizeof(unsigned long long) * 2 * NBINS);
    GRB_LAUNCH("bin_count", k_bin_count, nblk(nrows), 256, 0, work, nrows, d);
    unsigned long long h[NBINS];
    d2h(h, d, sizeof h);
    stream_sync();
    b.off[0] = 0;
    for (int k = 0; k < NBINS; k++) b.off[k + 1] = b.off[k] + (int64_t)h[k];
    unsigned long long cur[NBINS];
    for (int k = 0; k < NBINS; k++) cur[k] = (unsigned long long)b.off[k];
    h2d(d + NBINS, cur, sizeof cur);
    b.rows = dalloc<int32_t>(nrows);
    GRB_LAUNCH("bin_scatter", k_bin_scatter, nblk(nrows), 256, 0, work, nrows, d + NBINS, b.rows);
}

template <class T> __global__ void k_fill_value(T *p, int64_t n, T v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// ---------------------------------------------------------------- Gustavson: T = A (+).(x) B
using SymSR = Semiring<OP_PLUS, OP_ONEB, int32_t>;   // the symbolic pass only touches patterns

template <class T>
static void gustavson_typed(const Csr &A, const Csr &B, int mon, int mul, int tcode, Csr &Tout) {
    const int64_t m = A.nrows, n = B.ncols;
    Tout = Csr();
    Tout.nrows = m; Tout.ncols = n;
    int64_t *ub = dalloc<int64_t>(m + 1);
    unsigned long long *d_total = (unsigned long long *)(g.d_scratch + 40);
    dzero(d_total, sizeof(unsigned long long));
    GRB_LAUNCH("row_flops", k_row_flops, nblk(m * 32), 256, 0, A.rowptr, A.colidx, B.rowptr, m, n, ub, d_total);
    int64_t *rownnz = dalloc<int64_t>(m + 1);
    dzero(rownnz, (size_t)(m + 1) * sizeof(int64_t));
    const int64_t words = cdiv(n, 32);
    const size_t bm_bytes = (size_t)words * 4;
    const bool bm_in_smem = bm_bytes <= 200 * 1024;
    // heavy rows: shared-memory column panels (k_panel_symbolic / k_panel_row) when the row bitmap fits shared memory
    const int np = (int)cdiv(n > 0 ? n : 1, PANEL_W);
    const bool panels = g.spgemm_panels && bm_in_smem && np <= PANEL_MAX_NP;
    uint16_t *rowS = panels ? dalloc<uint16_t>(m) : nullptr;
    unsigned long long *cursor = (unsigned long long *)(g.d_scratch + 48);
    uint32_t *ptab = nullptr;
    // ---- symbolic ----
    {
        Bins bins;
        make_bins(ub, m, bins);
        GemmArgs<int32_t> a;
        a.Ap = A.rowptr; a.Aj = A.colidx; a.Ax = nullptr; a.a_iso = 1;
        a.Bp = B.rowptr; a.Bj = B.colidx; a.Bx = nullptr; a.b_iso = 1;
        a.nrows = m; a.ncols = n; a.rownnz = rownnz; a.Cp = nullptr; a.Cj = nullptr; a.Cx = nullptr;
        a.gbitmap = nullptr; a.gacc = nullptr; a.words = words;
        a.ptab = nullptr; a.np = np; a.rowS = rowS; a.cursor = cursor;
        for (int b = 1; b < NBINS; b++) {
            const int64_t nb = bins.off[b + 1] - bins.off[b];
            if (!nb) continue;
            a.rows = bins.rows + bins.off[b]; a.nbin = nb;
            if (b == NBINS - 1 && panels) {
                auto kern = k_panel_symbolic;
                if (bm_bytes > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bm_bytes));
                dzero(cursor, sizeof(unsigned long long));
                GRB_LAUNCH("spgemm_symbolic_panel", kern, (int)(nb < g.sm_count ? nb : g.sm_count), PSYM_THREADS, bm_bytes, a);
            } else if (b == 1) {
                auto kern = k_hash_warp<SymSR, false>;
                GRB_LAUNCH("spgemm_symbolic_warp", kern, (int)cdiv(nb, HW_ROWS), HW_ROWS * 32, 0, a);
            } else if (b < NBINS - 1) {
                auto kern = k_hash_block<SymSR, false>;
                const size_t smem = (size_t)kBinTable[b] * 4;
                if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                GRB_LAUNCH("spgemm_symbolic_block", kern, (int)nb, kBinThreads[b], smem, a, kBinTable[b]);
            } else {
                auto kern = k_dense_block<SymSR, false>;
                const int grid = (int)(nb < 2 * g.sm_count ? nb : 2 * g.sm_count);
                uint32_t *gb = nullptr;
                size_t smem = bm_bytes;
                if (!bm_in_smem) { gb = dalloc<uint32_t>((int64_t)grid * words); smem = 0; }
                a.gbitmap = gb;
                if (smem > 48 * 1024) CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                GRB_LAUNCH("spgemm_symbolic_dense", kern, grid, 1024, smem, a);
                dfree(gb);
                a.gbitmap = nullptr;
            }
        }
    }
    Tout.rowptr = dalloc<int64_t>(m + 1);
    // keep the counts: the numeric pass re-bins rows by their exact size (tighter tables than the flop bound)
    int64_t *counts = dalloc<int64_t>(m + 1);
    d2d(counts, rownnz, (size_t)(m + 1) * sizeof(int64_t));
    Tout.nnz = scan_counts(rownnz, Tout.rowptr, m);
    dfree(rownnz);
    Tout.colidx = dalloc<uint32_t>(Tout.nnz);
    Tout.vals = dmalloc((size_t)(Tout.nnz > 0 ? Tout.nnz : 1) * sizeof(T));
    // ---- numeric ----
    if (Tout.nnz > 0) {
        Bins bins;
        make_bins(counts, m, bins);
        GemmArgs<T> a;
        a.Ap = A.rowptr; a.Aj = A.colidx; a.Ax = (const T *)A.vals; a.a_iso = A.iso;
        a.Bp = B.rowptr; a.Bj = B.colidx; a.Bx = (const T *)B.vals; a.b_iso = B.iso;
        a.nrows = m; a.ncols = n; a.rownnz = nullptr; a.Cp = Tout.rowptr; a.Cj = Tout.colidx; a.Cx = (T *)Tout.vals;
        a.gbitmap = nullptr; a.gacc = nullptr; a.words = words;
        a.ptab = nullptr; a.np = np; a.rowS = rowS; a.cursor = cursor;
        for (int b = 1; b < NBINS; b++) {
            const int64_t nb = bins.off[b + 1] - bins.off[b];
            if (!nb) continue;
            a.rows = bins.rows + bins.off[b]; a.nbin = nb;
            GemmLaunch L;
            L.args = &a; L.table = kBinTable[b]; L.threads = kBinThreads[b]; L.smem = 0;
            uint32_t *gb = nullptr; T *gacc = nullptr;
            if (b == NBINS - 1 && panels) {
                if (np > 1) {
                    ptab = dalloc<uint32_t>(B.nrows * (int64_t)(np + 1));
                    GRB_LAUNCH("spgemm_panel_table", k_panel_table, nblk(B.nrows * 32), 256, 0, B.rowptr, B.colidx, B.nrows, np, ptab);
                    a.ptab = ptab;
                }
                dzero(cursor, sizeof(unsigned long long));
                L.kernel = GK_PANEL; L.grid = (int)(nb < PANEL_CTAS_PER_SM * g.sm_count ? nb : PANEL_CTAS_PER_SM * g.sm_count); L.threads = PANEL_THREADS;
                L.smem = panel_row_smem<T>();
                if (g.spgemm_panels == 2) { gacc = dalloc<T>((int64_t)L.grid * PANEL_TABLE); a.gacc = gacc; }
            } else if (b == 1) { L.kernel = GK_WARP; L.grid = (int)cdiv(nb, HW_ROWS); }
            else if (b < NBINS - 1) { L.kernel = GK_BLOCK; L.grid = (int)nb; L.smem = hash_block_smem(kBinTable[b], sizeof(T)); }
            else {
                L.kernel = GK_DENSE;
                int64_t grid = nb < 2 * g.sm_count ? nb : 2 * g.sm_count;
                const int64_t max_grid = (int64_t)((8ull << 30) / ((size_t)n * sizeof(T)));   // cap the accumulators at 8 GiB
                if (grid > max_grid) grid = max_grid > 0 ? max_grid : 1;
                L.grid = (int)grid; L.threads = 1024; L.smem = bm_in_smem ? bm_bytes : 0;
                if (!bm_in_smem) gb = dalloc<uint32_t>(grid * words);
                gacc = dalloc<T>(grid * n);
                auto fill = k_fill_value<T>;
                GRB_LAUNCH("fill_identity", fill, 148 * 8, 256, 0, gacc, grid * n, monoid_identity_rt<T>(mon));
                a.gbitmap = gb; a.gacc = gacc;
            }
            const bool ok = gemm_launch_numeric(mon, mul, tcode, L);
            dfree(gb); dfree(gacc); dfree(ptab); ptab = nullptr;
            a.gbitmap = nullptr; a.gacc = nullptr; a.ptab = nullptr;
            if (!ok) { dfree(ub); dfree(counts); dfree(rowS); csr_free(Tout); grb_throw(GrB_NOT_IMPLEMENTED, "semiring has no sm_100a kernel"); }
        }
    }
    dfree(ub); dfree(counts); dfree(rowS);
}

// ---------------------------------------------------------------- masked dot: T<M> = A (+).(x) B with B given as rows of B'
__global__ void k_dot_rowptr(const int64_t *Mp, const int64_t *pos, int64_t nrows, int64_t *Tp) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i <= nrows) Tp[i] = pos[Mp[i]];
}
__global__ void k_has_to_i64(const uint8_t *has, int64_t n, int64_t *flag) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i <= n) flag[i] = (i < n && has[i]) ? 1 : 0;
}
template <class T>
__global__ void k_dot_compact(const uint8_t *has, const int64_t *pos, const uint32_t *Mj, const T *val, int64_t mnnz,
                              uint32_t *Tj, T *Tx) {
    const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= mnnz || !has[e]) return;
    const int64_t o = pos[e];
    Tj[o] = Mj[e]; Tx[o] = val[e];
}

// hash sets of the rows of R longer than DOT_HASH_MIN (grb_spgemm.cuh); caller frees off and tab
static void build_row_hashes(const Csr &R, int64_t *&off, uint32_t *&tab) {
    int64_t *sizes = dalloc<int64_t>(R.nrows + 1);
    off = dalloc<int64_t>(R.nrows + 1);
    GRB_LAUNCH("rowhash_sizes", k_rowhash_sizes, nblk(R.nrows), 256, 0, R.rowptr, R.nrows, sizes);
    const int64_t total = scan_counts(sizes, off, R.nrows);
    dfree(sizes);
    tab = dalloc<uint32_t>(total);
    if (total > 0) {
        CUDA_TRY(cudaMemsetAsync(tab, 0xFF, (size_t)total * sizeof(uint32_t), g.stream));
        uint32_t *rowid = csr_row_ids(R);
        GRB_LAUNCH("rowhash_build", k_rowhash_build, nblk(R.nnz), 256, 0, rowid, R.colidx, R.nnz, off, tab);
        dfree(rowid);
    }
}

template <class T>
static void dot_typed(const Csr &M, int m_vsize, int m_fp, const Csr &A, const Csr &Bt, int mon, int mul, int tcode, Csr &Tout) {
    const int64_t mnnz = M.nnz;
    Tout = Csr();
    Tout.nrows = M.nrows; Tout.ncols = M.ncols;
    uint32_t *mrow = csr_row_ids(M);
    T *val = dalloc<T>(mnnz);
    uint8_t *has = dalloc<uint8_t>(mnnz);
    DotArgs<T> d;
    d.Mrow = mrow; d.Mj = M.colidx; d.Mv = M.vals; d.m_vsize = m_vsize; d.m_fp = m_fp; d.mnnz = mnnz;
    d.Ap = A.rowptr; d.Aj = A.colidx; d.Ax = (const T *)A.vals; d.a_iso = A.iso;
    d.Bp = Bt.rowptr; d.Bj = Bt.colidx; d.Bx = (const T *)Bt.vals; d.b_iso = Bt.iso;
    d.val = val; d.has = has;
    // multiplies that read no values (pair): membership through per-row hash sets instead of binary searches
    int64_t *ah_off = nullptr, *bh_off = nullptr; uint32_t *ah_tab = nullptr, *bh_tab = nullptr;
    d.Ah_off = d.Bh_off = nullptr; d.Ah_tab = d.Bh_tab = nullptr;
    if (mul == OP_ONEB && g.dot_hash && mnnz > 0) {
        build_row_hashes(A, ah_off, ah_tab);
        d.Ah_off = ah_off; d.Ah_tab = ah_tab;
        if (Bt.rowptr == A.rowptr && Bt.colidx == A.colidx) { d.Bh_off = ah_off; d.Bh_tab = ah_tab; }   // L.mxm(L.T): one set
        else { build_row_hashes(Bt, bh_off, bh_tab); d.Bh_off = bh_off; d.Bh_tab = bh_tab; }
    }
    const bool dot_ok = gemm_launch_dot(mon, mul, tcode, &d, mnnz);
    dfree(ah_off); dfree(ah_tab); dfree(bh_off); dfree(bh_tab);
    if (!dot_ok) {
        dfree(mrow); dfree(val); dfree(has);
        grb_throw(GrB_NOT_IMPLEMENTED, "semiring has no sm_100a kernel");
    }
    int64_t *flag = dalloc<int64_t>(mnnz + 1), *pos = dalloc<int64_t>(mnnz + 1);
    GRB_LAUNCH("has_flags", k_has_to_i64, nblk(mnnz + 1), 256, 0, has, mnnz, flag);
    Tout.nnz = scan_counts(flag, pos, mnnz);
    Tout.rowptr = dalloc<int64_t>(M.nrows + 1);
    GRB_LAUNCH("dot_rowptr", k_dot_rowptr, nblk(M.nrows + 1), 256, 0, M.rowptr, pos, M.nrows, Tout.rowptr);
    Tout.colidx = dalloc<uint32_t>(Tout.nnz);
    Tout.vals = dmalloc((size_t)(Tout.nnz > 0 ? Tout.nnz : 1) * sizeof(T));
    if (mnnz) {
        auto kern = k_dot_compact<T>;
        GRB_LAUNCH("dot_compact", kern, nblk(mnnz), 256, 0, has, pos, M.colidx, val, mnnz, Tout.colidx, (T *)Tout.vals);
    }
    dfree(mrow); dfree(val); dfree(has); dfree(flag); dfree(pos);
}

// ---------------------------------------------------------------- GrB_mxm
static void matrix_product(GB_Object *C, GB_Object *mask, const GrB_BinaryOp accum, const GrB_Semiring sr, GB_Object *A,
                           GB_Object *B, const DescFlags &f) {
    require_init();
    if (!sr) grb_throw(GrB_NULL_POINTER, "semiring is NULL");
    const int acc = accum_code_of(accum);
    op_need_kernel(sr->mon, sr->name);
    int tcode = sr->tcode, mul = sr->mul;
    const int mon = sr->mon;
    // INT8/INT16/UINT8/UINT16 semirings run in the 32-bit type of the same signedness (grb_internal.cuh)
    int narrow = -1;
    if (tc_is_narrow(tcode)) { narrow = tcode; mul = narrow_mul(mon, mul, narrow); tcode = tc_wide_of(narrow); }
    const bool uses_a = !(mul == OP_SECOND || mul == OP_ONEB), uses_b = !(mul == OP_FIRST || mul == OP_ONEB);
    View va, vb;
    view_of(va, A, f.tran0);
    view_of(vb, B, f.tran1);
    const int64_t m = va.csr->nrows, k = va.csr->ncols, n = vb.csr->ncols;
    const int64_t crows = C->nrows, ccols = C->kind == K_MATRIX ? C->ncols : 1;
    if (k != vb.csr->nrows || crows != m || ccols != n ||
        (mask && (mask->nrows != m || (mask->kind == K_MATRIX ? mask->ncols : 1) != n))) {
        char b[256];
        snprintf(b, sizeof b, "dimension mismatch: mxm of %lldx%lld by %lldx%lld into %lldx%lld%s", (long long)m,
                 (long long)k, (long long)vb.csr->nrows, (long long)n, (long long)crows, (long long)ccols, mask ? " (masked)" : "");
        grb_throw(GrB_DIMENSION_MISMATCH, b);
    }
    if (uses_a) view_cast(va, A->type->code, tcode, A->type->size, narrow);
    // ---- method: masked dot product when a non-complemented mask bounds the output ----
    bool use_dot = false;
    View vm;
    if (mask && !f.comp && f.axb != GxB_AxB_GUSTAVSON && f.axb != GxB_AxB_HASH && f.axb != GxB_AxB_SAXPY) {
        view_of(vm, mask, false);
        const bool value_iso = !f.structure && vm.csr->iso;
        if (!value_iso) {
            // dot visits nnz(M) pairs; Gustavson expands every product: prefer dot unless the mask dwarfs the inputs
            const int64_t in_nnz = va.csr->nnz + vb.csr->nnz;
            use_dot = (f.axb == GxB_AxB_DOT) || vm.csr->nnz <= 8 * in_nnz + 1024;
        }
    }
    Csr T;
    if (use_dot) {
        View vbt;
        view_of(vbt, B, !f.tran1);   // rows of B' = columns of B
        if (uses_b) view_cast(vbt, B->type->code, tcode, B->type->size, narrow);
        const int m_vsize = f.structure ? 0 : (int)mask->type->size;
        const int m_fp = (mask->type->code == TC_FP32 || mask->type->code == TC_FP64);
        GRB_TYPE_SWITCH(tcode, TT, dot_typed<TT>(*vm.csr, m_vsize, m_fp, *va.csr, *vbt.csr, mon, mul, tcode, T));
    } else {
        if (uses_b) view_cast(vb, B->type->code, tcode, B->type->size, narrow);
        GRB_TYPE_SWITCH(tcode, TT, gustavson_typed<TT>(*va.csr, *vb.csr, mon, mul, tcode, T));
    }
    if (narrow >= 0) { csr_cast_values(T, tcode, narrow); tcode = narrow; }      // wrap to the semiring's own type
    mat_writeback(C, mask, f, acc, T, tcode, use_dot);
    if (g.blocking) stream_sync();
}

extern "C" {

GrB_Info GrB_mxm(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum, const GrB_Semiring op, const GrB_Matrix A,
                 const GrB_Matrix B, const GrB_Descriptor desc) {
    API_BEGIN
    GB_Object *Cobj = obj_check(C, "C"), *Aobj = obj_check_lazy(A, "A"), *Bobj = obj_check_lazy(B, "B");
    GB_Object *M = Mask ? obj_check(Mask, "Mask") : nullptr;
    matrix_product(Cobj, M, accum, op, Aobj, Bobj, desc_flags(desc));
    API_END(C)
}

}  // extern "C"
