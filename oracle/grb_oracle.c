/* oracle/grb_oracle.c — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by or called from the
 * product path (graph-python_b200/); only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.
 *
 * A plain C (OpenMP) restatement of the arithmetic behind GrB_mxm / GrB_mxv / GrB_vxm as
 * python-graphblas calls them (reference graphblas/core/matrix.py:2163-2251, core/vector.py:1259-1306,
 * argument order base.py:476-483).  The reference's own arithmetic lives in SuiteSparse:GraphBLAS 7.4.x
 * (PyPI suitesparse-graphblas >=7.4.0.0,<7.5; reference pyproject.toml:65), which is absent from
 * /root/reference and from this image, so this follows the published GraphBLAS C API 2.0 semantics and
 * is pinned to the literal golden vectors in the reference's tests (tests/golden/reference_kats.json)
 * and cross-checked against oracle/pyref.py, scipy.sparse and scipy.sparse.csgraph.
 * At R-MAT scale the reference pins nothing: large-scale parity is transitive through this oracle.
 *
 * Layout on the host: CSR with int64 row pointers and int64 column indices, values typed by a type
 * code; dense vectors as values[n] + present[n] (uint8).  Indices are sorted within rows.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* type codes == GraphBLAS_b200.h / tools/gen_abi.py */
enum { T_BOOL = 0, T_INT8, T_INT16, T_INT32, T_INT64, T_UINT8, T_UINT16, T_UINT32, T_UINT64, T_FP32, T_FP64 };
/* binary-op codes shared with the CUDA side (graph-python_b200/csrc/grb_ops.cuh) */
enum { OP_FIRST = 0, OP_SECOND, OP_ONEB, OP_ANY, OP_PLUS, OP_MINUS, OP_TIMES, OP_DIV, OP_MIN, OP_MAX,
       OP_LOR, OP_LAND, OP_LXOR, OP_LXNOR, OP_EQ, OP_NE, OP_GT, OP_LT, OP_GE, OP_LE };

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}
void orc_free(void *p) { free(p); }

/* ---------- typed binary op: integer arithmetic wraps (computed unsigned), fp follows IEEE ---------- */
#define DEF_INT_OPS(NAME, T, U)                                                       \
    static inline T binop_##NAME(int op, T x, T y) {                                  \
        switch (op) {                                                                 \
        case OP_FIRST: return x;                                                      \
        case OP_SECOND: case OP_ANY: return y;                                        \
        case OP_ONEB: return (T)1;                                                    \
        case OP_PLUS: return (T)((U)x + (U)y);                                        \
        case OP_MINUS: return (T)((U)x - (U)y);                                       \
        case OP_TIMES: return (T)((U)x * (U)y);                                       \
        case OP_DIV: { /* SuiteSparse integer division (GB_IDIV_SIGNED / GB_IDIV_UNSIGNED): never traps; */ \
            /* x/0 = 0, type max or type min by the sign of x; x/-1 = -x computed unsigned */                    \
            const int is_signed = ((T)-1 < (T)0);                                     \
            const T tmax = (T)(((U)~(U)0) >> (is_signed ? 1 : 0));                    \
            const T tmin = is_signed ? (T)(~tmax) : (T)0;                             \
            if (y == 0) return x == 0 ? (T)0 : (x < (T)0 ? tmin : tmax);              \
            if (is_signed && y == (T)-1) return (T)((U)0 - (U)x);                     \
            return (T)(x / y); }                                                      \
        case OP_MIN: return x < y ? x : y;                                            \
        case OP_MAX: return x > y ? x : y;                                            \
        case OP_LOR: return (T)((x != 0) || (y != 0));                                \
        case OP_LAND: return (T)((x != 0) && (y != 0));                               \
        case OP_LXOR: return (T)((x != 0) != (y != 0));                               \
        case OP_LXNOR: return (T)((x != 0) == (y != 0));                              \
        default: return y;                                                            \
        }                                                                             \
    }
#define DEF_FP_OPS(NAME, T, FMIN, FMAX)                                               \
    static inline T binop_##NAME(int op, T x, T y) {                                  \
        switch (op) {                                                                 \
        case OP_FIRST: return x;                                                      \
        case OP_SECOND: case OP_ANY: return y;                                        \
        case OP_ONEB: return (T)1;                                                    \
        case OP_PLUS: return x + y;                                                   \
        case OP_MINUS: return x - y;                                                  \
        case OP_TIMES: return x * y;                                                  \
        case OP_DIV: return x / y;                                                    \
        case OP_MIN: return FMIN(x, y);                                               \
        case OP_MAX: return FMAX(x, y);                                               \
        default: return y;                                                            \
        }                                                                             \
    }
DEF_INT_OPS(BOOL, uint8_t, uint8_t)
DEF_INT_OPS(INT8, int8_t, uint8_t)
DEF_INT_OPS(INT16, int16_t, uint16_t)
DEF_INT_OPS(INT32, int32_t, uint32_t)
DEF_INT_OPS(INT64, int64_t, uint64_t)
DEF_INT_OPS(UINT8, uint8_t, uint8_t)
DEF_INT_OPS(UINT16, uint16_t, uint16_t)
DEF_INT_OPS(UINT32, uint32_t, uint32_t)
DEF_INT_OPS(UINT64, uint64_t, uint64_t)
DEF_FP_OPS(FP32, float, fminf, fmaxf)
DEF_FP_OPS(FP64, double, fmin, fmax)

#define FOR_ALL_TYPES(X) \
    X(BOOL, uint8_t) X(INT8, int8_t) X(INT16, int16_t) X(INT32, int32_t) X(INT64, int64_t) \
    X(UINT8, uint8_t) X(UINT16, uint16_t) X(UINT32, uint32_t) X(UINT64, uint64_t) X(FP32, float) X(FP64, double)

/* ---------- t = A (+).(x) u  (pull over CSR rows; u dense with presence bytes, NULL = full) ----------
 * flip = 0: z = mult(A(i,j), u(j))   (GrB_mxv, reference core/matrix.py:2163)
 * flip = 1: z = mult(u(j), A(i,j))   (GrB_vxm evaluated on the transposed CSR, core/vector.py:1259)
 * row_mask (optional): rows with row_mask[i]==0 are skipped (t absent) — lets the baseline use the
 * same mask early-out a CPU library would.  any_stop: stop a row at its first product (ANY monoid). */
#define DEF_SPMV(NAME, T)                                                                            \
    static void spmv_##NAME(int mon, int mul, int flip, int64_t nrows, const int64_t *Ap,            \
                            const int64_t *Aj, const T *Ax, int a_iso, const T *u,                   \
                            const uint8_t *up, const uint8_t *row_mask, T *t, uint8_t *tp) {         \
        int any_stop = (mon == OP_ANY);                                                              \
        if (mon == OP_PLUS && mul == OP_TIMES && !up && !a_iso && !row_mask) {                       \
            _Pragma("omp parallel for schedule(dynamic, 1024)")                                      \
            for (int64_t i = 0; i < nrows; i++) {                                                    \
                T acc = 0;                                                                           \
                for (int64_t p = Ap[i]; p < Ap[i + 1]; p++) acc = binop_##NAME(OP_PLUS, acc,         \
                        binop_##NAME(OP_TIMES, Ax[p], u[Aj[p]]));                                    \
                t[i] = acc; tp[i] = Ap[i + 1] > Ap[i];                                               \
            }                                                                                        \
            return;                                                                                  \
        }                                                                                            \
        _Pragma("omp parallel for schedule(dynamic, 1024)")                                          \
        for (int64_t i = 0; i < nrows; i++) {                                                        \
            T acc = 0; int has = 0;                                                                  \
            if (!row_mask || row_mask[i]) {                                                          \
                for (int64_t p = Ap[i]; p < Ap[i + 1]; p++) {                                        \
                    int64_t j = Aj[p];                                                               \
                    if (up && !up[j]) continue;                                                      \
                    T a = Ax[a_iso ? 0 : p];                                                         \
                    T z = flip ? binop_##NAME(mul, u[j], a) : binop_##NAME(mul, a, u[j]);            \
                    acc = has ? binop_##NAME(mon, acc, z) : z;                                       \
                    has = 1;                                                                         \
                    if (any_stop) break;                                                             \
                }                                                                                    \
            }                                                                                        \
            t[i] = acc; tp[i] = (uint8_t)has;                                                        \
        }                                                                                            \
    }
FOR_ALL_TYPES(DEF_SPMV)

int orc_spmv(int tcode, int mon, int mul, int flip, int64_t nrows, const int64_t *Ap, const int64_t *Aj,
             const void *Ax, int a_iso, const void *u, const uint8_t *up, const uint8_t *row_mask,
             void *t, uint8_t *tp) {
    switch (tcode) {
#define CASE(NAME, T) case T_##NAME: spmv_##NAME(mon, mul, flip, nrows, Ap, Aj, (const T *)Ax, a_iso, \
                                                 (const T *)u, up, row_mask, (T *)t, tp); return 0;
        FOR_ALL_TYPES(CASE)
#undef CASE
    }
    return -1;
}

/* ---------- w<m,replace> = accum(w, t) on dense vectors of one type ----------
 * m[i] is the already-evaluated mask truth (structure/value/complement applied by the caller),
 * NULL = no mask.  accum < 0 = no accumulator.  C API 2.0 "standard method for masks and accumulators". */
#define DEF_VWB(NAME, T)                                                                             \
    static void vwb_##NAME(int64_t n, T *w, uint8_t *wp, const T *t, const uint8_t *tp,              \
                           const uint8_t *m, int accum, int replace) {                               \
        _Pragma("omp parallel for schedule(static)")                                                 \
        for (int64_t i = 0; i < n; i++) {                                                            \
            if (!m || m[i]) {                                                                        \
                if (tp[i]) {                                                                         \
                    w[i] = (accum >= 0 && wp[i]) ? binop_##NAME(accum, w[i], t[i]) : t[i];           \
                    wp[i] = 1;                                                                       \
                } else if (accum < 0) wp[i] = 0;                                                     \
            } else if (replace) wp[i] = 0;                                                           \
        }                                                                                            \
    }
FOR_ALL_TYPES(DEF_VWB)

int orc_vec_writeback(int tcode, int64_t n, void *w, uint8_t *wp, const void *t, const uint8_t *tp,
                      const uint8_t *m, int accum, int replace) {
    switch (tcode) {
#define CASE(NAME, T) case T_##NAME: vwb_##NAME(n, (T *)w, wp, (const T *)t, tp, m, accum, replace); return 0;
        FOR_ALL_TYPES(CASE)
#undef CASE
    }
    return -1;
}

/* ---------- T = A (+).(x) B, Gustavson with a dense accumulator per thread; sorted CSR out ----------
 * If Mp/Mj are given, only positions allowed by the structural mask (xor comp) are produced
 * (value masks are resolved by the caller into a structural pattern first).
 * Output arrays are malloc'd here; the caller frees them with orc_free. */
static int cmp_i64(const void *a, const void *b) {
    int64_t x = *(const int64_t *)a, y = *(const int64_t *)b;
    return (x > y) - (x < y);
}

#define DEF_SPGEMM(NAME, T)                                                                          \
    static int spgemm_##NAME(int mon, int mul, int64_t nrows, int64_t ncols, const int64_t *Ap,      \
                             const int64_t *Aj, const T *Ax, int a_iso, const int64_t *Bp,           \
                             const int64_t *Bj, const T *Bx, int b_iso, const int64_t *Mp,           \
                             const int64_t *Mj, int comp, int64_t **Cp_out, int64_t **Cj_out,        \
                             void **Cx_out) {                                                        \
        int64_t *Cp = (int64_t *)calloc((size_t)nrows + 1, sizeof(int64_t));                         \
        int64_t **rowj = (int64_t **)calloc((size_t)nrows, sizeof(int64_t *));                       \
        T **rowx = (T **)calloc((size_t)nrows, sizeof(T *));                                         \
        if (!Cp || !rowj || !rowx) return -1;                                                        \
        _Pragma("omp parallel")                                                                      \
        {                                                                                            \
            int64_t *mark = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ncols ? ncols : 1));        \
            T *acc = (T *)malloc(sizeof(T) * (size_t)(ncols ? ncols : 1));                           \
            int64_t *list = (int64_t *)malloc(sizeof(int64_t) * (size_t)(ncols ? ncols : 1));        \
            for (int64_t j = 0; j < ncols; j++) mark[j] = -1;                                        \
            _Pragma("omp for schedule(dynamic, 64)")                                                 \
            for (int64_t i = 0; i < nrows; i++) {                                                    \
                int64_t cnt = 0;                                                                     \
                /* mark[j] states for row i: 2i = allowed & untouched (mask), 2i+1 = touched */      \
                if (Mp && !comp) for (int64_t p = Mp[i]; p < Mp[i + 1]; p++) mark[Mj[p]] = 2 * i;    \
                if (Mp && comp) for (int64_t p = Mp[i]; p < Mp[i + 1]; p++) mark[Mj[p]] = -2 - i;    \
                for (int64_t p = Ap[i]; p < Ap[i + 1]; p++) {                                        \
                    int64_t k = Aj[p]; T a = Ax[a_iso ? 0 : p];                                      \
                    for (int64_t q = Bp[k]; q < Bp[k + 1]; q++) {                                    \
                        int64_t j = Bj[q];                                                           \
                        if (Mp && !comp && mark[j] != 2 * i && mark[j] != 2 * i + 1) continue;       \
                        if (Mp && comp && mark[j] == -2 - i) continue;                               \
                        T z = binop_##NAME(mul, a, Bx[b_iso ? 0 : q]);                               \
                        if (mark[j] == 2 * i + 1) { if (mon != OP_ANY) acc[j] = binop_##NAME(mon, acc[j], z); } \
                        else { mark[j] = 2 * i + 1; acc[j] = z; list[cnt++] = j; }                   \
                    }                                                                                \
                }                                                                                    \
                if (cnt) {                                                                           \
                    qsort(list, (size_t)cnt, sizeof(int64_t), cmp_i64);                              \
                    rowj[i] = (int64_t *)malloc(sizeof(int64_t) * (size_t)cnt);                      \
                    rowx[i] = (T *)malloc(sizeof(T) * (size_t)cnt);                                  \
                    for (int64_t c = 0; c < cnt; c++) { rowj[i][c] = list[c]; rowx[i][c] = acc[list[c]]; } \
                }                                                                                    \
                Cp[i + 1] = cnt;                                                                     \
            }                                                                                        \
            free(mark); free(acc); free(list);                                                       \
        }                                                                                            \
        for (int64_t i = 0; i < nrows; i++) Cp[i + 1] += Cp[i];                                      \
        int64_t nnz = Cp[nrows];                                                                     \
        int64_t *Cj = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nnz ? nnz : 1));                  \
        T *Cx = (T *)malloc(sizeof(T) * (size_t)(nnz ? nnz : 1));                                    \
        _Pragma("omp parallel for schedule(dynamic, 256)")                                           \
        for (int64_t i = 0; i < nrows; i++) {                                                        \
            int64_t cnt = Cp[i + 1] - Cp[i];                                                         \
            if (cnt) {                                                                               \
                memcpy(Cj + Cp[i], rowj[i], sizeof(int64_t) * (size_t)cnt);                          \
                memcpy(Cx + Cp[i], rowx[i], sizeof(T) * (size_t)cnt);                                \
                free(rowj[i]); free(rowx[i]);                                                        \
            }                                                                                        \
        }                                                                                            \
        free(rowj); free(rowx);                                                                      \
        *Cp_out = Cp; *Cj_out = Cj; *Cx_out = Cx;                                                    \
        return 0;                                                                                    \
    }
FOR_ALL_TYPES(DEF_SPGEMM)

int orc_spgemm(int tcode, int mon, int mul, int64_t nrows, int64_t ncols, const int64_t *Ap,
               const int64_t *Aj, const void *Ax, int a_iso, const int64_t *Bp, const int64_t *Bj,
               const void *Bx, int b_iso, const int64_t *Mp, const int64_t *Mj, int comp,
               int64_t **Cp, int64_t **Cj, void **Cx) {
    switch (tcode) {
#define CASE(NAME, T) case T_##NAME: return spgemm_##NAME(mon, mul, nrows, ncols, Ap, Aj, (const T *)Ax, \
                          a_iso, Bp, Bj, (const T *)Bx, b_iso, Mp, Mj, comp, Cp, Cj, Cx);
        FOR_ALL_TYPES(CASE)
#undef CASE
    }
    return -1;
}

/* ---------- T<M> = A (+).(x) B' by sorted-merge dot products under a structural mask ----------
 * For every mask entry (i,j): intersect row i of A with row j of B (B given by rows, i.e. B' is the
 * right operand — the triangle-count form L.mxm(L.T), reference notebooks / SURVEY 3.3).
 * Entries with an empty intersection are absent.  has[e], val[e] are aligned with the mask entries. */
#define DEF_DOT(NAME, T)                                                                             \
    static void dot_##NAME(int mon, int mul, int64_t nrows, const int64_t *Mp, const int64_t *Mj,    \
                           const int64_t *Ap, const int64_t *Aj, const T *Ax, int a_iso,             \
                           const int64_t *Bp, const int64_t *Bj, const T *Bx, int b_iso, T *val,     \
                           uint8_t *has) {                                                           \
        _Pragma("omp parallel for schedule(dynamic, 256)")                                           \
        for (int64_t i = 0; i < nrows; i++) {                                                        \
            for (int64_t e = Mp[i]; e < Mp[i + 1]; e++) {                                            \
                int64_t j = Mj[e], p = Ap[i], pe = Ap[i + 1], q = Bp[j], qe = Bp[j + 1];             \
                T acc = 0; int h = 0;                                                                \
                while (p < pe && q < qe) {                                                           \
                    int64_t ka = Aj[p], kb = Bj[q];                                                  \
                    if (ka < kb) p++; else if (kb < ka) q++;                                         \
                    else {                                                                           \
                        T z = binop_##NAME(mul, Ax[a_iso ? 0 : p], Bx[b_iso ? 0 : q]);               \
                        acc = h ? binop_##NAME(mon, acc, z) : z; h = 1; p++; q++;                    \
                        if (mon == OP_ANY) break;                                                    \
                    }                                                                                \
                }                                                                                    \
                val[e] = acc; has[e] = (uint8_t)h;                                                   \
            }                                                                                        \
        }                                                                                            \
    }
FOR_ALL_TYPES(DEF_DOT)

int orc_masked_dot(int tcode, int mon, int mul, int64_t nrows, const int64_t *Mp, const int64_t *Mj,
                   const int64_t *Ap, const int64_t *Aj, const void *Ax, int a_iso, const int64_t *Bp,
                   const int64_t *Bj, const void *Bx, int b_iso, void *val, uint8_t *has) {
    switch (tcode) {
#define CASE(NAME, T) case T_##NAME: dot_##NAME(mon, mul, nrows, Mp, Mj, Ap, Aj, (const T *)Ax, a_iso, \
                                                Bp, Bj, (const T *)Bx, b_iso, (T *)val, has); return 0;
        FOR_ALL_TYPES(CASE)
#undef CASE
    }
    return -1;
}

/* ---------- Graph500-style R-MAT edge generator: counter-based, integer-only ----------
 * The CUDA generator in graph-python_b200/csrc/grb_rmat.cu restates exactly this arithmetic, so the
 * host and device graphs are bit-identical.  (a,b,c,d) = (0.57,0.19,0.19,0.05) as 32-bit thresholds. */
static inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
/* bijection on [0, 2^scale): label shuffle so vertex id carries no degree information */
static inline uint64_t rmat_permute(uint64_t x, int scale, uint64_t seed) {
    uint64_t mask = (scale >= 64) ? ~0ull : ((1ull << scale) - 1);
    int h = scale / 2 > 0 ? scale / 2 : 1;
    uint64_t k1 = splitmix64(seed ^ 0xA5A5A5A5ull) | 1ull, k2 = splitmix64(seed ^ 0x5A5A5A5Aull) | 1ull;
    x = (x * k1) & mask; x ^= x >> h;
    x = (x * k2) & mask; x ^= x >> h;
    x = (x + (k1 >> 7)) & mask;
    return x;
}
void orc_rmat_edges(int scale, uint64_t nedges, uint64_t seed, uint64_t *src, uint64_t *dst) {
    const uint64_t TA = 2448131358ull, TAB = 3264175144ull, TABC = 4080218931ull;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < (int64_t)nedges; e++) {
        uint64_t st = splitmix64(seed ^ ((uint64_t)e * 0xD1342543DE82EF95ull));
        uint64_t r = 0, c = 0;
        for (int l = 0; l < scale; l++) {
            uint64_t u = splitmix64(st + (uint64_t)l) >> 32;
            uint64_t rb = (u >= TAB), cb = (u >= TA && u < TAB) || (u >= TABC);
            r = (r << 1) | rb; c = (c << 1) | cb;
        }
        src[e] = rmat_permute(r, scale, seed);
        dst[e] = rmat_permute(c, scale, seed);
    }
}

/* ---------- the same graph straight into CSR, in parallel (bench.py's CPU arm at scale 23-26, where the numpy path
 * np.unique(keys) of oracle.py:rmat_pattern takes minutes) ----------
 * Result == oracle.py:rmat_pattern(scale, edgefactor, seed, symmetrize) laid out by rows (transpose = 0) or by columns
 * (transpose = 1: rows of the result are the edges' destinations): self-loops dropped, duplicates collapsed, columns
 * sorted.  tests/test_oracle_golden.py checks it against the numpy path.  Buffers are malloc'd: release with orc_free. */
static int cmp_u32(const void *a, const void *b) {
    uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b;
    return x < y ? -1 : (x > y);
}
int orc_rmat_csr(int scale, uint64_t nedges, uint64_t seed, int transpose, int symmetrize, int64_t **Ap_out,
                 int64_t **Aj_out, int64_t *nnz_out) {
    if (scale < 1 || scale > 31) return -1;
    const int64_t n = (int64_t)1 << scale, ne = (int64_t)nedges;
    uint32_t *R = (uint32_t *)malloc((size_t)ne * 4), *C = (uint32_t *)malloc((size_t)ne * 4);
    int64_t *start = (int64_t *)calloc((size_t)n + 1, 8), *cursor = (int64_t *)malloc((size_t)(n + 1) * 8);
    if (!R || !C || !start || !cursor) { free(R); free(C); free(start); free(cursor); return -2; }
    const uint64_t TA = 2448131358ull, TAB = 3264175144ull, TABC = 4080218931ull;
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < ne; e++) {
        uint64_t st = splitmix64(seed ^ ((uint64_t)e * 0xD1342543DE82EF95ull));
        uint64_t r = 0, c = 0;
        for (int l = 0; l < scale; l++) {
            uint64_t u = splitmix64(st + (uint64_t)l) >> 32;
            uint64_t rb = (u >= TAB), cb = (u >= TA && u < TAB) || (u >= TABC);
            r = (r << 1) | rb; c = (c << 1) | cb;
        }
        const uint32_t s = (uint32_t)rmat_permute(r, scale, seed), d = (uint32_t)rmat_permute(c, scale, seed);
        R[e] = transpose ? d : s;
        C[e] = transpose ? s : d;
        if (s != d) {
#pragma omp atomic
            start[R[e] + 1]++;
            if (symmetrize) {
#pragma omp atomic
                start[C[e] + 1]++;
            }
        }
    }
    for (int64_t i = 0; i < n; i++) start[i + 1] += start[i];
    memcpy(cursor, start, (size_t)(n + 1) * 8);
    const int64_t total = start[n];
    uint32_t *col = (uint32_t *)malloc((size_t)(total > 0 ? total : 1) * 4);
    if (!col) { free(R); free(C); free(start); free(cursor); return -2; }
#pragma omp parallel for schedule(static)
    for (int64_t e = 0; e < ne; e++) {
        if (R[e] == C[e]) continue;
        int64_t p;
#pragma omp atomic capture
        p = cursor[R[e]]++;
        col[p] = C[e];
        if (symmetrize) {
#pragma omp atomic capture
            p = cursor[C[e]]++;
            col[p] = R[e];
        }
    }
    free(R); free(C);
    /* sort + dedup every row in place; cursor[i] becomes the row's final length */
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < n; i++) {
        uint32_t *a = col + start[i];
        const int64_t len = start[i + 1] - start[i];
        if (len > 24) qsort(a, (size_t)len, 4, cmp_u32);
        else for (int64_t x = 1; x < len; x++) { uint32_t v = a[x]; int64_t y = x; while (y > 0 && a[y - 1] > v) { a[y] = a[y - 1]; y--; } a[y] = v; }
        int64_t m = 0;
        for (int64_t x = 0; x < len; x++) if (x == 0 || a[x] != a[x - 1]) a[m++] = a[x];
        cursor[i] = m;
    }
    int64_t *Ap = (int64_t *)malloc((size_t)(n + 1) * 8);
    if (!Ap) { free(col); free(start); free(cursor); return -2; }
    Ap[0] = 0;
    for (int64_t i = 0; i < n; i++) Ap[i + 1] = Ap[i] + cursor[i];
    const int64_t nnz = Ap[n];
    int64_t *Aj = (int64_t *)malloc((size_t)(nnz > 0 ? nnz : 1) * 8);
    if (!Aj) { free(col); free(start); free(cursor); free(Ap); return -2; }
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < n; i++) {
        const uint32_t *a = col + start[i];
        int64_t *o = Aj + Ap[i];
        for (int64_t x = 0; x < cursor[i]; x++) o[x] = a[x];
    }
    free(col); free(start); free(cursor);
    *Ap_out = Ap; *Aj_out = Aj; *nnz_out = nnz;
    return 0;
}
/* per-entry values of oracle.py:edge_value for a CSR: the value belongs to the EDGE (src, dst); by_col = 1 when the CSR
 * rows are destinations (the transposed layout above).  mode 1: integer weight in [1,255]; mode 2: float in (0,1]. */
int orc_csr_edge_values(int tcode, int mode, int by_col, int64_t nrows, const int64_t *Ap, const int64_t *Aj, void *out) {
    if (tcode != T_FP32 && tcode != T_FP64 && tcode != T_INT64 && tcode != T_INT32 && tcode != T_UINT32) return -1;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < nrows; i++)
        for (int64_t p = Ap[i]; p < Ap[i + 1]; p++) {
            const uint64_t src = by_col ? (uint64_t)Aj[p] : (uint64_t)i, dst = by_col ? (uint64_t)i : (uint64_t)Aj[p];
            uint64_t h = (src * 0x9E3779B97F4A7C15ull) ^ (dst * 0xC2B2AE3D27D4EB4Full);
            h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
            if (mode == 1) {
                const uint64_t w = 1 + h % 255;
                if (tcode == T_INT64) ((int64_t *)out)[p] = (int64_t)w;
                else if (tcode == T_INT32) ((int32_t *)out)[p] = (int32_t)w;
                else if (tcode == T_UINT32) ((uint32_t *)out)[p] = (uint32_t)w;
                else if (tcode == T_FP32) ((float *)out)[p] = (float)w;
                else ((double *)out)[p] = (double)w;
            } else {
                const double x = (double)(h >> 40) + 1.0;
                if (tcode == T_FP32) ((float *)out)[p] = (float)x / 16777216.0f;
                else if (tcode == T_FP64) ((double *)out)[p] = x / 16777216.0;
                else if (tcode == T_INT64) ((int64_t *)out)[p] = 0;
                else if (tcode == T_INT32) ((int32_t *)out)[p] = 0;
                else ((uint32_t *)out)[p] = 0;
            }
        }
    return 0;
}
