// grb_ewise.cu — the loop-body operations that sit between the products in BFS / PageRank / SSSP / FastSV
// (SURVEY 8f rank 1): masked scalar / vector assign, eWiseMult / eWiseAdd, apply(bind-2nd), reduce.
// All of them are single-pass, bandwidth-trivial kernels over the dense vector layout and share the
// product kernels' write-back (grb_spmv.cuh: epi_store), so vectors never leave HBM between products.
// Reference call sites: graphblas/core/vector.py:1012-1140 (ewise), :1308 (apply), :1561 (reduce), :1772-1950 (assign).
#define HAVE_EWISE 1
#include "grb_internal.cuh"
#include "grb_abi_lists.inc"
#include "grb_ops.cuh"
#include "grb_spmv.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

template <class T>
static EpiArgs<T> epi_for(GB_Object *w, GB_Object *mask, const DescFlags &f, int accum) {
    EpiArgs<T> e;
    e.w_in = (const T *)w->vals; e.wp_in = w->present; e.w_out = (T *)w->vals; e.wp_out = w->present;
    e.mask_p = mask ? mask->present : nullptr;
    e.mask_v = mask ? mask->vals : nullptr;
    e.mask_vsize = (mask && !f.structure) ? (int)mask->type->size : 0;
    e.mask_fp = mask ? (mask->type->code == TC_FP32 || mask->type->code == TC_FP64) : 0;
    e.comp = f.comp; e.replace = f.replace; e.accum = accum;
    e.count = (unsigned long long *)w->d_nvals;
    return e;
}

static int accum_code(const GrB_BinaryOp accum) {
    if (!accum) return -1;
    op_need_kernel(accum->op, accum->name);
    if (op_is_cmp(accum->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot be accumulators");
    return accum->op;
}

static void check_vec(GB_Object *o, const char *what) {
    if (o->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, std::string(what) + " must be a vector");
}

// w<m,replace> = accum(w, t) where t (values of type tcode + presence) is cast to w's type if needed
void vec_writeback_any(GB_Object *w, GB_Object *mask, const DescFlags &f, int accum, const void *t, int tcode,
                       const uint8_t *tp) {
    int64_t n = w->nrows;
    int wcode = w->type->code;
    void *tc = nullptr;
    if (tcode != wcode) {
        tc = dmalloc((size_t)(n > 0 ? n : 1) * w->type->size);
        cast_array(t, tcode, tc, wcode, n);
        t = tc;
    }
    dzero(w->d_nvals, sizeof(int64_t));
    GRB_TYPE_SWITCH(wcode, W, {
        EpiArgs<W> e = epi_for<W>(w, mask, f, accum);
        launch_vec_writeback<W>(e, (const W *)t, tp, n);
    });
    dfree(tc);
    vec_set_unknown_nvals(w);
}

// C typecast of one scalar on the host, matching cast_array's device rules (float -> integer saturates, NaN -> 0,
// anything -> bool is "!= 0")
template <class W, class T> static inline W host_cast(T x) {
    if constexpr (std::is_same<W, bool>::value) return x != (T)0;
    else if constexpr (std::is_floating_point<T>::value && std::is_integral<W>::value) {
        if (x != x) return (W)0;
        const double lo = (double)std::numeric_limits<W>::lowest(), hi = (double)std::numeric_limits<W>::max();
        const double v = (double)x;
        if (v <= lo) return std::numeric_limits<W>::lowest();
        if (v >= hi) return std::numeric_limits<W>::max();
        return (W)v;
    } else return (W)x;
}

// ---------------------------------------------------------------- assign
template <class T>
__global__ void __launch_bounds__(256) k_assign_scalar(EpiArgs<T> e, T x, int64_t n) {
    int c = 0;
    const bool need_old = epi_needs_old(e);
    const int64_t base = blockIdx.x * (int64_t)(256 * EW_VEC) + threadIdx.x;
    // plain fill (no mask, no accumulator): nothing to read at all
    if (!e.mask_p && !e.comp && e.accum < 0) {
#pragma unroll
        for (int k = 0; k < EW_VEC; k++) { const int64_t i = base + k * 256; if (i < n) { e.w_out[i] = x; e.wp_out[i] = 1; c++; } }
        block_count_add(e.count, c);
        return;
    }
    bool al[EW_VEC], wh[EW_VEC]; T wo[EW_VEC];
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) {
        const int64_t i = base + k * 256;
        al[k] = wh[k] = false; wo[k] = (T)0;
        if (i < n) {
            al[k] = epi_allowed(e, i);
            if (epi_reads_old(e, al[k])) { wh[k] = e.wp_in[i] != 0; if (need_old) wo[k] = e.w_in[i]; }
        }
    }
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) { const int64_t i = base + k * 256; if (i < n) c += epi_apply(e, i, al[k], true, x, wh[k], wo[k]); }
    block_count_add(e.count, c);
}
// w(I)<m> = accum(w(I), x) for an explicit index list: positions outside I are left alone
template <class T>
__global__ void k_assign_scalar_list(EpiArgs<T> e, T x, const uint64_t *I, int64_t ni, int64_t n, int *bad) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= ni) return;
    uint64_t i = I[k];
    if (i >= (uint64_t)n) { *bad = 1; return; }
    epi_store(e, (int64_t)i, epi_allowed(e, (int64_t)i), true, x);
}

template <class T>
static GrB_Info vector_assign_scalar(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, T x, int xcode,
                                     const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc) {
    API_BEGIN_NOFLUSH
    require_init();
    // a plain fill right behind a deferred element-wise product joins it; anything else runs the queue first
    const bool may_join = g.pending_n == 1 && g.pending[0].kind == 1 && lazy_on() && !mask_ && !accum && I == GrB_ALL;
    if (g.pending_n && !may_join) pending_flush();
    GB_Object *w = obj_check(w_, "w");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    check_vec(w, "w");
    if (mask) { check_vec(mask, "mask"); if (mask->nrows != w->nrows) grb_throw(GrB_DIMENSION_MISMATCH, "mask size differs from w"); }
    DescFlags f = desc_flags(desc);
    int acc = accum_code(accum);
    int64_t n = w->nrows;
    if (may_join) {
        const PendingOp &a = g.pending[0];
        if (!f.comp && !f.replace && w->type->code == a.xcode && n == a.n && w->vals && w->kind != K_MATRIX) {
            PendingOp &p = g.pending[1];
            p.kind = 2; p.w = w; p.u = p.v = nullptr; p.op = 0; p.xcode = a.xcode; p.n = n; p.xbits = 0; p.is_union = false;
            GRB_TYPE_SWITCH(w->type->code, W, { W xv = host_cast<W, T>(x); memcpy(&p.xbits, &xv, sizeof(W)); });
            p.set_present = !(w->nvals_known && w->nvals == n);
            g.pending_n = 2;
            w->nvals = n; w->nvals_known = true; w->list_valid = false;
            return GrB_SUCCESS;
        }
        pending_flush();
    }
    // value in w's type (C typecast rules, on the host: no device round trip, no synchronisation)
    char hx[8];
    (void)xcode;
    GRB_TYPE_SWITCH(w->type->code, W, { W xv = host_cast<W, T>(x); memcpy(hx, &xv, sizeof(W)); });
    const bool all = (I == GrB_ALL);
    GRB_TYPE_SWITCH(w->type->code, W, {
        W xv; memcpy(&xv, hx, sizeof(W));
        EpiArgs<W> e = epi_for<W>(w, mask, f, acc);
        if (all && !mask && !f.comp) {
            // every position receives an entry: nvals is known without counting (and without a later read-back)
            e.count = nullptr;
            GRB_LAUNCH("assign_scalar", k_assign_scalar<W>, ew_grid4(n), 256, 0, e, xv, n);
            w->nvals = n; w->nvals_known = true; w->list_valid = false;
        } else if (all) {
            dzero(w->d_nvals, sizeof(int64_t));
            if (writeback16_ok(e, nullptr, nullptr, n)) {
                auto wide = k_writeback16<W, true>;
                GRB_LAUNCH("assign_scalar16", wide, writeback16_grid(n), 256, 0, e, (const W *)nullptr, (const uint8_t *)nullptr, xv, n);
            } else
                GRB_LAUNCH("assign_scalar", k_assign_scalar<W>, ew_grid4(n), 256, 0, e, xv, n);
            vec_set_unknown_nvals(w);
        } else {
            if (f.replace) grb_throw(GrB_NOT_IMPLEMENTED, "assign with an index list and replace");
            if (!I && ni) grb_throw(GrB_NULL_POINTER, "I is NULL");
            uint64_t *dI = dalloc<uint64_t>((int64_t)ni);
            h2d(dI, I, (size_t)ni * 8);
            int *dflag = (int *)g.d_scratch;
            dzero(dflag, sizeof(int));
            if (ni) GRB_LAUNCH("assign_scalar_list", k_assign_scalar_list<W>, nblk((int64_t)ni), 256, 0, e, xv, dI, (int64_t)ni, n, dflag);
            int bad = 0;
            d2h(&bad, dflag, sizeof(int));
            stream_sync();
            dfree(dI);
            if (bad) grb_throw(GrB_INDEX_OUT_OF_BOUNDS, "assign index out of bounds");
            count_present(w->present, n, w->d_nvals);
            vec_set_unknown_nvals(w);
        }
    });
    API_END(w_)
}

// ---------------------------------------------------------------- eWise
template <class T, bool UNION, bool CMP>
__global__ void __launch_bounds__(256) k_ewise(const T *u, const uint8_t *up, const T *v, const uint8_t *vp, int op,
                                               void *t_, uint8_t *tp, int64_t n) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool a = up[i] != 0, b = vp[i] != 0;
    bool has = UNION ? (a || b) : (a && b);
    tp[i] = has;
    if (!has) return;
    if (CMP) {
        bool *t = (bool *)t_;
        t[i] = (a && b) ? cmp_rt<T>(op, u[i], v[i]) : (a ? (u[i] != (T)0) : (v[i] != (T)0));
    } else {
        T *t = (T *)t_;
        t[i] = (a && b) ? binop_rt<T>(op, u[i], v[i]) : (a ? u[i] : v[i]);
    }
}

// fused form: t = u op v computed per element and handed straight to the write-back (w's type == the op's z type)
template <class X, class Z, bool UNION, bool CMP>
__global__ void __launch_bounds__(256) k_ewise_fused(const X *u, const uint8_t *up, const X *v, const uint8_t *vp, int op,
                                                     EpiArgs<Z> e, int64_t n) {
    int c = 0;
    const bool need_old = epi_needs_old(e);
    const int64_t base = blockIdx.x * (int64_t)(256 * EW_VEC) + threadIdx.x;
    bool pa[EW_VEC], pb[EW_VEC], al[EW_VEC], wh[EW_VEC]; X xv[EW_VEC], yv[EW_VEC]; Z wo[EW_VEC];
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) {      // all loads of the batch first: 4 x the memory-level parallelism
        const int64_t i = base + k * 256;
        pa[k] = pb[k] = al[k] = wh[k] = false; xv[k] = yv[k] = (X)0; wo[k] = (Z)0;
        if (i < n) {
            pa[k] = up[i] != 0; pb[k] = vp[i] != 0; xv[k] = u[i]; yv[k] = v[i];
            al[k] = epi_allowed(e, i);
            if (epi_reads_old(e, al[k])) { wh[k] = e.wp_in[i] != 0; if (need_old) wo[k] = e.w_in[i]; }
        }
    }
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) {
        const int64_t i = base + k * 256;
        if (i >= n) continue;
        const bool a = pa[k], b = pb[k];
        const bool has = UNION ? (a || b) : (a && b);
        Z t = (Z)0;
        if (has) {
            const X x = xv[k], y = yv[k];
            if constexpr (CMP) t = (Z)((a && b) ? cmp_rt<X>(op, x, y) : (a ? (x != (X)0) : (y != (X)0)));
            else t = (Z)((a && b) ? binop_rt<X>(op, x, y) : (a ? x : y));
        }
        c += epi_apply(e, i, al[k], has, t, wh[k], wo[k]);
    }
    block_count_add(e.count, c);
}

// every operand full, no mask, no accumulator: w = u op v with 128-bit accesses and no presence traffic
// (PageRank's `w << r.ewise_mult(d, truediv)` between two products)
template <class X>
__global__ void __launch_bounds__(256) k_ewise_full(const X *u, const X *v, int op, X *w, int64_t n) {
    constexpr int VEC = 16 / sizeof(X);
    const int64_t nvec = n / VEC;
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
        const uint4 a = reinterpret_cast<const uint4 *>(u)[q], b = reinterpret_cast<const uint4 *>(v)[q];
        X x[VEC], y[VEC], z[VEC];
        memcpy(x, &a, 16); memcpy(y, &b, 16);
#pragma unroll
        for (int k = 0; k < VEC; k++) z[k] = binop_rt<X>(op, x[k], y[k]);
        uint4 o;
        memcpy(&o, z, 16);
        reinterpret_cast<uint4 *>(w)[q] = o;
    }
    if (blockIdx.x == 0)
        for (int64_t i = nvec * VEC + threadIdx.x; i < n; i += blockDim.x) w[i] = binop_rt<X>(op, u[i], v[i]);
}

// the deferred pair in one pass: w = u op v, then r = x (r may be u, v or w: a thread reads its elements before it writes them)
template <class X>
__global__ void __launch_bounds__(256) k_lazy_pair(const X *u, const X *v, int op, X *w, X *r, uint8_t *rp, X x, int64_t n) {
    constexpr int VEC = 16 / sizeof(X);
    const int64_t nvec = n / VEC;
    X xs[VEC];
#pragma unroll
    for (int k = 0; k < VEC; k++) xs[k] = x;
    uint4 fill;
    memcpy(&fill, xs, 16);
    for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; q < nvec; q += (int64_t)gridDim.x * blockDim.x) {
        const uint4 a = reinterpret_cast<const uint4 *>(u)[q], b = reinterpret_cast<const uint4 *>(v)[q];
        X xa[VEC], ya[VEC], z[VEC];
        memcpy(xa, &a, 16); memcpy(ya, &b, 16);
#pragma unroll
        for (int k = 0; k < VEC; k++) z[k] = binop_rt<X>(op, xa[k], ya[k]);
        uint4 o;
        memcpy(&o, z, 16);
        reinterpret_cast<uint4 *>(w)[q] = o;
        reinterpret_cast<uint4 *>(r)[q] = fill;
        if (rp) {
#pragma unroll
            for (int k = 0; k < VEC; k++) rp[q * VEC + k] = 1;
        }
    }
    if (blockIdx.x == 0)
        for (int64_t i = nvec * VEC + threadIdx.x; i < n; i += blockDim.x) {
            w[i] = binop_rt<X>(op, u[i], v[i]);
            r[i] = x;
            if (rp) rp[i] = 1;
        }
}

static void pending_run_ewise(const PendingOp &p) {
    GRB_TYPE_SWITCH(p.xcode, X, {
        GRB_LAUNCH(p.is_union ? "ewise_add" : "ewise_mult", k_ewise_full<X>, ew_grid(cdiv(p.n, 16 / (int)sizeof(X))), 256, 0,
                   (const X *)p.u->vals, (const X *)p.v->vals, p.op, (X *)p.w->vals, p.n);
    });
}
static void pending_run_fill(const PendingOp &p) {
    GRB_TYPE_SWITCH(p.xcode, X, {
        X xv; memcpy(&xv, &p.xbits, sizeof(X));
        EpiArgs<X> e;
        e.w_out = (X *)p.w->vals; e.wp_out = p.w->present; e.w_in = e.w_out; e.wp_in = e.wp_out;
        e.mask_p = nullptr; e.mask_v = nullptr; e.mask_vsize = 0; e.mask_fp = 0; e.comp = 0; e.replace = 0; e.accum = -1;
        e.count = nullptr;
        GRB_LAUNCH("assign_scalar", k_assign_scalar<X>, ew_grid4(p.n), 256, 0, e, xv, p.n);
    });
}
void pending_flush() {
    const int np = g.pending_n;
    if (!np) return;
    g.pending_n = 0;               // (the launches below go through no entry point: nothing can re-enter)
    const PendingOp &a = g.pending[0];
    if (np == 2) {
        const PendingOp &b = g.pending[1];
        GRB_TYPE_SWITCH(a.xcode, X, {
            X xv; memcpy(&xv, &b.xbits, sizeof(X));
            GRB_LAUNCH("ewise_fill_pair", k_lazy_pair<X>, ew_grid(cdiv(a.n, 16 / (int)sizeof(X))), 256, 0,
                       (const X *)a.u->vals, (const X *)a.v->vals, a.op, (X *)a.w->vals, (X *)b.w->vals,
                       b.set_present ? b.w->present : (uint8_t *)nullptr, xv, a.n);
        });
        return;
    }
    if (a.kind == 1) pending_run_ewise(a); else pending_run_fill(a);
}

static void vector_ewise(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_BinaryOp op,
                         const GrB_Vector u_, const GrB_Vector v_, const GrB_Descriptor desc, bool is_union) {
    require_init();
    if (g.pending_n) pending_flush();          // an element-wise product only ever STARTS a deferred pair
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u"), *v = obj_check(v_, "v");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    if (!op) grb_throw(GrB_NULL_POINTER, "op is NULL");
    check_vec(w, "w"); check_vec(u, "u"); check_vec(v, "v");
    int64_t n = w->nrows;
    if (u->nrows != n || v->nrows != n || (mask && mask->nrows != n)) grb_throw(GrB_DIMENSION_MISMATCH, "eWise: sizes differ");
    DescFlags f = desc_flags(desc);
    int acc = accum_code(accum);
    int xcode = op->xcode, zcode = op->zcode;
    op_need_kernel(op->op, op->name);
    const bool cmp = op_is_cmp(op->op);
    void *uc = nullptr, *vc = nullptr;
    const void *uv = u->vals, *vv = v->vals;
    if (u->type->code != xcode) { uc = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[xcode]); cast_array(u->vals, u->type->code, uc, xcode, n); uv = uc; }
    if (v->type->code != xcode) { vc = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[xcode]); cast_array(v->vals, v->type->code, vc, xcode, n); vv = vc; }
    if (w->type->code == zcode) {
        // both operands full and no mask: the result is full, nvals needs no counting
        const bool known_full = !mask && !f.comp && u->nvals_known && v->nvals_known && u->nvals == n && v->nvals == n;
        if (known_full && acc < 0 && !cmp && n > 0) {
            const bool w_was_full = w->nvals_known && w->nvals == n;
            if (lazy_on() && w_was_full && !uc && !vc) {      // deferred: runs with the next entry point (grb_internal.cuh)
                PendingOp &p = g.pending[0];
                p.kind = 1; p.w = w; p.u = u; p.v = v; p.op = op->op; p.xcode = xcode; p.n = n; p.xbits = 0; p.set_present = false;
                p.is_union = is_union;
                g.pending_n = 1;
                w->list_valid = false;
                return;
            }
            GRB_TYPE_SWITCH(xcode, X, {
                GRB_LAUNCH(is_union ? "ewise_add" : "ewise_mult", k_ewise_full<X>, ew_grid(cdiv(n, 16 / (int)sizeof(X))), 256, 0,
                           (const X *)uv, (const X *)vv, op->op, (X *)w->vals, n);
            });
            if (!w_was_full) CUDA_TRY(cudaMemsetAsync(w->present, 1, (size_t)n, g.stream));
            w->nvals = n; w->nvals_known = true; w->list_valid = false;
            dfree(uc); dfree(vc);
            return;
        }
        if (!known_full) dzero(w->d_nvals, sizeof(int64_t));
        GRB_TYPE_SWITCH(xcode, X, {
            if (cmp) {
                EpiArgs<bool> e = epi_for<bool>(w, mask, f, acc);
                if (known_full) e.count = nullptr;
                auto kern = is_union ? k_ewise_fused<X, bool, true, true> : k_ewise_fused<X, bool, false, true>;
                GRB_LAUNCH(is_union ? "ewise_add" : "ewise_mult", kern, ew_grid4(n), 256, 0, (const X *)uv, u->present,
                           (const X *)vv, v->present, op->op, e, n);
            } else {
                EpiArgs<X> e = epi_for<X>(w, mask, f, acc);
                if (known_full) e.count = nullptr;
                auto kern = is_union ? k_ewise_fused<X, X, true, false> : k_ewise_fused<X, X, false, false>;
                GRB_LAUNCH(is_union ? "ewise_add" : "ewise_mult", kern, ew_grid4(n), 256, 0, (const X *)uv, u->present,
                           (const X *)vv, v->present, op->op, e, n);
            }
        });
        if (known_full) { w->nvals = n; w->nvals_known = true; w->list_valid = false; }
        else vec_set_unknown_nvals(w);
        dfree(uc); dfree(vc);
        return;
    }
    void *t = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[zcode]);
    uint8_t *tp = dalloc<uint8_t>(n);
    GRB_TYPE_SWITCH(xcode, X, {
        auto kern = cmp ? (is_union ? k_ewise<X, true, true> : k_ewise<X, false, true>)
                        : (is_union ? k_ewise<X, true, false> : k_ewise<X, false, false>);
        GRB_LAUNCH(is_union ? "ewise_add" : "ewise_mult", kern, nblk(n), 256, 0, (const X *)uv, u->present,
                   (const X *)vv, v->present, op->op, t, tp, n);
    });
    vec_writeback_any(w, mask, f, acc, t, zcode, tp);
    dfree(t); dfree(tp); dfree(uc); dfree(vc);
}

// ---------------------------------------------------------------- apply bind-2nd
template <class T, bool CMP>
__global__ void __launch_bounds__(256) k_apply2nd(const T *u, const uint8_t *up, T y, int op, void *t_, uint8_t *tp, int64_t n) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const bool a = up[i] != 0;
    tp[i] = a;
    if (!a) return;
    if (CMP) ((bool *)t_)[i] = cmp_rt<T>(op, u[i], y);
    else ((T *)t_)[i] = binop_rt<T>(op, u[i], y);
}

template <class T>
static GrB_Info vector_apply2nd(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_BinaryOp op,
                                const GrB_Vector u_, T y, int ycode, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    if (!op) grb_throw(GrB_NULL_POINTER, "op is NULL");
    check_vec(w, "w"); check_vec(u, "u");
    int64_t n = w->nrows;
    if (u->nrows != n || (mask && mask->nrows != n)) grb_throw(GrB_DIMENSION_MISMATCH, "apply: sizes differ");
    DescFlags f = desc_flags(desc);
    int acc = accum_code(accum);
    int xcode = op->xcode, zcode = op->zcode;
    op_need_kernel(op->op, op->name);
    const bool cmp = op_is_cmp(op->op);
    void *uc = nullptr;
    const void *uv = u->vals;
    if (u->type->code != xcode) { uc = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[xcode]); cast_array(u->vals, u->type->code, uc, xcode, n); uv = uc; }
    void *t = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[zcode]);
    uint8_t *tp = dalloc<uint8_t>(n);
    GRB_TYPE_SWITCH(xcode, X, {
        X yv = (X)y;
        (void)ycode;
        auto kern = cmp ? k_apply2nd<X, true> : k_apply2nd<X, false>;
        GRB_LAUNCH("apply2nd", kern, nblk(n), 256, 0, (const X *)uv, u->present, yv, op->op, t, tp, n);
    });
    vec_writeback_any(w, mask, f, acc, t, zcode, tp);
    dfree(t); dfree(tp); dfree(uc);
    API_END(w_)
}

// ---------------------------------------------------------------- reduce to a C scalar
template <class T>
__global__ void __launch_bounds__(256) k_reduce(const T *vals, const uint8_t *present, int iso, int64_t n, int op,
                                                T *partial, uint8_t *partial_has) {
    __shared__ T s_v[8];
    __shared__ uint8_t s_h[8];
    T acc = (T)0; bool has = false;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (present && !present[i]) continue;
        T x = vals[iso ? 0 : i];
        acc = has ? binop_rt<T>(op, acc, x) : x; has = true;
    }
    for (int o = 16; o; o >>= 1) {
        T oa = __shfl_xor_sync(0xffffffffu, acc, o);
        bool oh = __shfl_xor_sync(0xffffffffu, (int)has, o) != 0;
        if (oh) { acc = has ? binop_rt<T>(op, acc, oa) : oa; has = true; }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_v[warp] = acc; s_h[warp] = has; }
    __syncthreads();
    if (threadIdx.x == 0) {
        T a = (T)0; bool h = false;
        for (int k = 0; k < (int)(blockDim.x >> 5); k++)
            if (s_h[k]) { a = h ? binop_rt<T>(op, a, s_v[k]) : s_v[k]; h = true; }
        partial[blockIdx.x] = a; partial_has[blockIdx.x] = h;
    }
}

// reduce (vals[n], optional presence) with a monoid; returns has + value (in T) on the host
template <class T>
static bool reduce_device(const T *vals, const uint8_t *present, bool iso, int64_t n, int op, T *out) {
    if (n == 0) return false;
    int blocks = (int)(cdiv(n, 256) < 1184 ? cdiv(n, 256) : 1184);
    T *p1 = dalloc<T>(blocks + 1);
    uint8_t *h1 = dalloc<uint8_t>(blocks + 1);
    GRB_LAUNCH("reduce", k_reduce<T>, blocks, 256, 0, vals, present, (int)iso, n, op, p1, h1);
    GRB_LAUNCH("reduce_final", k_reduce<T>, 1, 256, 0, (const T *)p1, (const uint8_t *)h1, 0, (int64_t)blocks, op, p1 + blocks, h1 + blocks);
    T hv; uint8_t hh = 0;
    d2h(&hv, p1 + blocks, sizeof(T));
    d2h(&hh, h1 + blocks, 1);
    stream_sync();
    dfree(p1); dfree(h1);
    *out = hv;
    return hh != 0;
}

// times/plus over an iso matrix would need nnz repetitions: expand the count analytically only for idempotent ops
template <class T>
static GrB_Info reduce_to_scalar(T *c, const GrB_BinaryOp accum, const GrB_Monoid monoid, GB_Object *A, int ccode) {
    if (!c) grb_throw(GrB_NULL_POINTER, "c is NULL");
    if (!monoid) grb_throw(GrB_NULL_POINTER, "monoid is NULL");
    op_need_kernel(monoid->op, monoid->name);
    int mcode = monoid->tcode;
    const void *vals; const uint8_t *present; int64_t n; bool iso = false; int vcode = A->type->code;
    if (A->kind == K_MATRIX) { vals = A->csr.vals; present = nullptr; n = A->csr.nnz; iso = A->csr.iso; }
    else { vals = A->vals; present = A->present; n = A->nrows; }
    void *vc = nullptr;
    if (vcode != mcode) {
        int64_t m = iso ? 1 : n;
        vc = dmalloc((size_t)(m > 0 ? m : 1) * kTypeSize[mcode]);
        cast_array(vals, vcode, vc, mcode, m);
        vals = vc;
    }
    bool has = false;
    double as_double = 0; int64_t as_i64 = 0; uint64_t as_u64 = 0;
    GRB_TYPE_SWITCH(mcode, M, {
        M r = (M)0;
        has = reduce_device<M>((const M *)vals, present, iso, n, monoid->op, &r);
        as_double = (double)r; as_i64 = (int64_t)r; as_u64 = (uint64_t)r;
    });
    dfree(vc);
    if (has) {
        T r;
        if (mcode == TC_FP32 || mcode == TC_FP64) r = (T)as_double;
        else if (mcode == TC_UINT64) r = (T)as_u64;
        else r = (T)as_i64;
        if (accum) {
            op_need_kernel(accum->op, accum->name);
    if (op_is_cmp(accum->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot be accumulators");
            *c = binop_rt<T>(accum->op, *c, r);
        } else *c = r;
    } else if (!accum) {
        // empty input: the monoid identity (C API: reduce of nothing yields the identity)
        GRB_TYPE_SWITCH(ccode, C, {
            C idv = (C)0;
            switch (monoid->op) {
            case OP_MIN: if constexpr (std::is_same<C, bool>::value) idv = true;
                         else if constexpr (std::is_floating_point<C>::value) idv = std::numeric_limits<C>::infinity();
                         else idv = std::numeric_limits<C>::max(); break;
            case OP_MAX: if constexpr (std::is_same<C, bool>::value) idv = false;
                         else if constexpr (std::is_floating_point<C>::value) idv = -std::numeric_limits<C>::infinity();
                         else idv = std::numeric_limits<C>::lowest(); break;
            case OP_TIMES: case OP_LAND: case OP_LXNOR: idv = (C)1; break;
            default: idv = (C)0;
            }
            *c = (T)idv;
        });
    }
    return GrB_SUCCESS;
}

extern "C" {

GrB_Info GrB_Vector_eWiseMult_BinaryOp(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                                       const GrB_BinaryOp op, const GrB_Vector u, const GrB_Vector v,
                                       const GrB_Descriptor desc) {
    API_BEGIN_NOFLUSH
    vector_ewise(w, mask, accum, op, u, v, desc, false);
    API_END(w)
}
GrB_Info GrB_Vector_eWiseAdd_BinaryOp(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                                      const GrB_BinaryOp op, const GrB_Vector u, const GrB_Vector v,
                                      const GrB_Descriptor desc) {
    API_BEGIN_NOFLUSH
    vector_ewise(w, mask, accum, op, u, v, desc, true);
    API_END(w)
}
GrB_Info GrB_Vector_assign(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_Vector u_,
                           const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc) {
    (void)ni;
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    check_vec(w, "w"); check_vec(u, "u");
    if (I != GrB_ALL) {
        if (I == nullptr) grb_throw(GrB_NULL_POINTER, "index list is NULL");
        vector_assign_indexed(w, mask, accum_code(accum), u, I, (int64_t)ni, desc_flags(desc));
        return GrB_SUCCESS;
    }
    if (u->nrows != w->nrows || (mask && mask->nrows != w->nrows)) grb_throw(GrB_DIMENSION_MISMATCH, "assign: sizes differ");
    vec_writeback_any(w, mask, desc_flags(desc), accum_code(accum), u->vals, u->type->code, u->present);
    API_END(w_)
}

#define DEF_EWISE_TYPED(NAME, CT, CODE)                                                                              \
    GrB_Info GrB_Vector_assign_##NAME(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, CT x,          \
                                      const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc) {               \
        return vector_assign_scalar<CT>(w, mask, accum, x, CODE, I, ni, desc);                                      \
    }                                                                                                               \
    GrB_Info GrB_Vector_apply_BinaryOp2nd_##NAME(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,     \
                                                 const GrB_BinaryOp op, const GrB_Vector u, CT y,                   \
                                                 const GrB_Descriptor desc) {                                       \
        return vector_apply2nd<CT>(w, mask, accum, op, u, y, CODE, desc);                                           \
    }                                                                                                               \
    GrB_Info GrB_Vector_reduce_##NAME(CT *c, const GrB_BinaryOp accum, const GrB_Monoid monoid, const GrB_Vector u, \
                                      const GrB_Descriptor desc) {                                                  \
        (void)desc;                                                                                                 \
        API_BEGIN                                                                                                   \
        require_init();                                                                                             \
        return reduce_to_scalar<CT>(c, accum, monoid, obj_check(u, "u"), CODE);                                     \
        API_END(u)                                                                                                  \
    }                                                                                                               \
    GrB_Info GrB_Matrix_reduce_##NAME(CT *c, const GrB_BinaryOp accum, const GrB_Monoid monoid, const GrB_Matrix A, \
                                      const GrB_Descriptor desc) {                                                  \
        (void)desc;                                                                                                 \
        API_BEGIN                                                                                                   \
        require_init();                                                                                             \
        return reduce_to_scalar<CT>(c, accum, monoid, obj_check(A, "A"), CODE);                                     \
        API_END(A)                                                                                                  \
    }
GRB_TYPE_LIST(DEF_EWISE_TYPED)

}  // extern "C"
