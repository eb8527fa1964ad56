// grb_vecops.cu — loop-body vector operations that sit between the products in the reference's algorithms
// (SURVEY 8f rank 1 and 3): unary apply, extract w = u(I), indexed assign w(I) = u, row-wise reduce of a matrix
// (through the product kernels), and the option getters the reference queries at import / per object
// (graphblas/__init__.py:158, graphblas/core/ss/matrix.py:216-220).
#include "grb_spmv.cuh"
#include "grb_storage.cuh"

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

static int accum_code_of(const GrB_BinaryOp accum) {
    if (!accum) return -1;
    op_need_kernel(accum->op, accum->name);
    if (op_is_cmp(accum->op)) grb_throw(GrB_DOMAIN_MISMATCH, "comparison operators cannot be accumulators");
    return accum->op;
}
static void need_vec(GB_Object *o, const char *what) {
    if (o->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, std::string(what) + " must be a vector");
}

// ---------------------------------------------------------------- unary apply
// GrB_MINV on integers follows the library's integer division (1 / x through binop_rt: x == 0 saturates)
template <class T> __device__ __forceinline__ T unop_rt(int op, T x) {
    switch (op) {
        case UOP_IDENTITY: return x;
        case UOP_AINV:
            if constexpr (std::is_same<T, bool>::value) return x;
            else if constexpr (std::is_floating_point<T>::value) return -x;
            else { using U = typename std::make_unsigned<T>::type; return (T)((U)0 - (U)x); }
        case UOP_MINV:
            if constexpr (std::is_same<T, bool>::value) return true;
            else return binop_rt<T>(OP_DIV, (T)1, x);
        case UOP_ABS:
            if constexpr (std::is_same<T, bool>::value || std::is_unsigned<T>::value) return x;
            else if constexpr (std::is_floating_point<T>::value) return x < 0 ? -x : (x == 0 ? (T)0 : x);
            else { using U = typename std::make_unsigned<T>::type; return x < 0 ? (T)((U)0 - (U)x) : x; }
        case UOP_ONE: return (T)1;
        case UOP_BNOT:
            if constexpr (std::is_integral<T>::value && !std::is_same<T, bool>::value) return (T)~x;
            else return x;
        case UOP_LNOT:
            if constexpr (std::is_same<T, bool>::value) return !x;
            else return (T)(x == (T)0);
    }
    return x;
}
template <class T>
__global__ void __launch_bounds__(256) k_apply_unary(const T *u, const uint8_t *up, int op, T *t, uint8_t *tp, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool a = up[i] != 0;
        tp[i] = a;
        if (a) t[i] = unop_rt<T>(op, u[i]);
    }
}

// ---------------------------------------------------------------- extract / indexed assign
template <class T>
__global__ void __launch_bounds__(256) k_extract(const T *u, const uint8_t *up, const uint64_t *I, int64_t ni, int64_t n,
                                                 T *t, uint8_t *tp, int *bad) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < ni; k += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t i = I[k];
        if (i >= (uint64_t)n) { *bad = 1; tp[k] = 0; continue; }
        const bool a = up[i] != 0;
        tp[k] = a;
        if (a) t[k] = u[i];
    }
}
// z = w; z(I[k]) = accum(z(I[k]), u(k)); without accum an absent u(k) deletes z(I[k]).  Duplicate indices are
// undefined by the C API; the last writer wins here.
template <class T>
__global__ void __launch_bounds__(256) k_assign_scatter(const T *u, const uint8_t *up, const uint64_t *I, int64_t ni, int64_t n,
                                                        int accum, T *z, uint8_t *zp, int *bad) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < ni; k += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t i = I[k];
        if (i >= (uint64_t)n) { *bad = 1; continue; }
        if (up[k]) {
            z[i] = (accum >= 0 && zp[i]) ? binop_rt<T>(accum, z[i], u[k]) : u[k];
            zp[i] = 1;
        } else if (accum < 0) zp[i] = 0;
    }
}

// Duplicate indices with a monoid accumulator (FastSV's `f(min)[I] << mngp`, the reference notebook's stand-in for
// LAGraph's Reduce_assign32): every duplicate is folded in with the monoid's atomic, so the result is the
// accumulation over all of them and does not depend on thread order.  Absent targets are seeded with the identity
// by a first pass.
template <class T>
__global__ void __launch_bounds__(256) k_assign_seed(const uint8_t *up, const uint64_t *I, int64_t ni, int64_t n, T ident,
                                                     T *z, const uint8_t *zp, int *bad) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < ni; k += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t i = I[k];
        if (i >= (uint64_t)n) { *bad = 1; continue; }
        if (up[k] && !zp[i]) z[i] = ident;
    }
}
template <class Mon, class T>
__global__ void __launch_bounds__(256) k_assign_atomic(const T *u, const uint8_t *up, const uint64_t *I, int64_t ni, int64_t n,
                                                       T *z, uint8_t *zp) {
    for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < ni; k += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t i = I[k];
        if (i >= (uint64_t)n || !up[k]) continue;
        Mon::atomic(z + i, u[k]);
        zp[i] = 1;
    }
}

static uint64_t *upload_indices(const uint64_t *I, int64_t ni) {
    uint64_t *d = dalloc<uint64_t>(ni);
    h2d(d, I, (size_t)ni * sizeof(uint64_t));
    return d;
}
static void check_bad(int *d_bad, const char *what) {
    int bad = 0;
    d2h(&bad, d_bad, sizeof(int));
    stream_sync();
    dfree(d_bad);
    if (bad) grb_throw(GrB_INDEX_OUT_OF_BOUNDS, std::string(what) + ": index out of bounds");
}

static void vector_assign_indexed_dev(GB_Object *w, GB_Object *mask, int accum, GB_Object *u, uint64_t *dI, int64_t ni,
                                      const DescFlags &f);
void vector_assign_indexed(GB_Object *w, GB_Object *mask, int accum, GB_Object *u, const uint64_t *I, int64_t ni,
                           const DescFlags &f) {
    need_vec(w, "w"); need_vec(u, "u");
    if (u->nrows != ni) grb_throw(GrB_DIMENSION_MISMATCH, "assign: u must have one entry slot per index");
    vector_assign_indexed_dev(w, mask, accum, u, upload_indices(I, ni), ni, f);
}
// dI: device index array, consumed (freed) here
static void vector_assign_indexed_dev(GB_Object *w, GB_Object *mask, int accum, GB_Object *u, uint64_t *dI, int64_t ni,
                                      const DescFlags &f) {
    need_vec(w, "w"); need_vec(u, "u");
    const int64_t n = w->nrows;
    if (u->nrows != ni) grb_throw(GrB_DIMENSION_MISMATCH, "assign: u must have one entry slot per index");
    if (mask && mask->nrows != n) grb_throw(GrB_DIMENSION_MISMATCH, "assign: mask and w sizes differ");
    const int wcode = w->type->code;
    void *uc = nullptr;
    const void *uv = u->vals;
    if (u->type->code != wcode) {
        uc = dmalloc((size_t)(ni > 0 ? ni : 1) * w->type->size);
        cast_array(u->vals, u->type->code, uc, wcode, ni);
        uv = uc;
    }
    void *z = dmalloc((size_t)(n > 0 ? n : 1) * w->type->size);
    uint8_t *zp = dalloc<uint8_t>(n);
    d2d(z, w->vals, (size_t)n * w->type->size);
    d2d(zp, w->present, (size_t)n);
    int *bad = dalloc<int>(1);
    dzero(bad, sizeof(int));
    const bool atomic_ok = (accum == OP_MIN || accum == OP_MAX || accum == OP_PLUS) &&
                           (wcode == TC_INT32 || wcode == TC_INT64 || wcode == TC_UINT32 || wcode == TC_UINT64 ||
                            wcode == TC_FP32 || wcode == TC_FP64);
    GRB_TYPE_SWITCH(wcode, T, {
        if constexpr (sizeof(T) >= 4) {
            if (atomic_ok) {
                GRB_LAUNCH("assign_seed", k_assign_seed<T>, ew_grid(ni), 256, 0, u->present, dI, ni, n, monoid_identity_rt<T>(accum),
                           (T *)z, zp, bad);
                if (accum == OP_MIN)
                    GRB_LAUNCH("assign_atomic", (k_assign_atomic<Monoid<OP_MIN, T>, T>), ew_grid(ni), 256, 0, (const T *)uv, u->present, dI, ni, n, (T *)z, zp);
                else if (accum == OP_MAX)
                    GRB_LAUNCH("assign_atomic", (k_assign_atomic<Monoid<OP_MAX, T>, T>), ew_grid(ni), 256, 0, (const T *)uv, u->present, dI, ni, n, (T *)z, zp);
                else
                    GRB_LAUNCH("assign_atomic", (k_assign_atomic<Monoid<OP_PLUS, T>, T>), ew_grid(ni), 256, 0, (const T *)uv, u->present, dI, ni, n, (T *)z, zp);
            }
        }
        if (!atomic_ok || sizeof(T) < 4)
            GRB_LAUNCH("assign_scatter", k_assign_scatter<T>, ew_grid(ni), 256, 0, (const T *)uv, u->present, dI, ni, n, accum,
                       (T *)z, zp, bad);
    });
    try { check_bad(bad, "assign"); } catch (...) { dfree(z); dfree(zp); dfree(dI); dfree(uc); throw; }
    vec_writeback_any(w, mask, f, -1, z, wcode, zp);       // w<m,replace> = z
    dfree(z); dfree(zp); dfree(dI); dfree(uc);
}

extern "C" {

GrB_Info GrB_Vector_apply(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_UnaryOp op,
                          const GrB_Vector u_, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    if (!op) grb_throw(GrB_NULL_POINTER, "op is NULL");
    need_vec(w, "w"); need_vec(u, "u");
    const int64_t n = w->nrows;
    if (u->nrows != n || (mask && mask->nrows != n)) grb_throw(GrB_DIMENSION_MISMATCH, "apply: sizes differ");
    op_need_kernel(op->op, op->name);
    const int tcode = op->tcode;
    void *uc = nullptr;
    const void *uv = u->vals;
    if (u->type->code != tcode) {
        uc = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[tcode]);
        cast_array(u->vals, u->type->code, uc, tcode, n);
        uv = uc;
    }
    void *t = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[tcode]);
    uint8_t *tp = dalloc<uint8_t>(n);
    GRB_TYPE_SWITCH(tcode, T, {
        GRB_LAUNCH("apply_unary", k_apply_unary<T>, ew_grid(n), 256, 0, (const T *)uv, u->present, op->op, (T *)t, tp, n);
    });
    vec_writeback_any(w, mask, desc_flags(desc), accum_code_of(accum), t, tcode, tp);
    dfree(t); dfree(tp); dfree(uc);
    API_END(w_)
}

GrB_Info GrB_Vector_extract(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_Vector u_,
                            const GrB_Index *I, GrB_Index ni_, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    need_vec(w, "w"); need_vec(u, "u");
    const int64_t n = w->nrows;
    if (mask && mask->nrows != n) grb_throw(GrB_DIMENSION_MISMATCH, "extract: mask and w sizes differ");
    if (I == GrB_ALL) {
        if (u->nrows != n) grb_throw(GrB_DIMENSION_MISMATCH, "extract: u and w sizes differ");
        vec_writeback_any(w, mask, desc_flags(desc), accum_code_of(accum), u->vals, u->type->code, u->present);
        return GrB_SUCCESS;
    }
    if (!I) grb_throw(GrB_NULL_POINTER, "index list is NULL");
    const int64_t ni = (int64_t)ni_;
    if (ni != n) grb_throw(GrB_DIMENSION_MISMATCH, "extract: w must have one entry slot per index");
    uint64_t *dI = upload_indices(I, ni);
    void *t = dmalloc((size_t)(ni > 0 ? ni : 1) * u->type->size);
    uint8_t *tp = dalloc<uint8_t>(ni);
    int *bad = dalloc<int>(1);
    dzero(bad, sizeof(int));
    GRB_TYPE_SWITCH(u->type->code, T, {
        GRB_LAUNCH("extract", k_extract<T>, ew_grid(ni), 256, 0, (const T *)u->vals, u->present, dI, ni, u->nrows, (T *)t, tp, bad);
    });
    try { check_bad(bad, "extract"); } catch (...) { dfree(t); dfree(tp); dfree(dI); throw; }
    vec_writeback_any(w, mask, desc_flags(desc), accum_code_of(accum), t, u->type->code, tp);
    dfree(t); dfree(tp); dfree(dI);
    API_END(w_)
}

// ---- index lists that already live in HBM (the values of a full integer vector), as SuiteSparse's GxB_Vector_assign_Vector /
// GxB_Vector_extract_Vector take them.  FastSV's `f(min)[I] << mngp` and `gp << f[I]` use I = the parent vector f itself: with
// the host-array forms the notebook pulls f to the host and pushes it back every iteration (1.3 GB over PCIe per iteration at
// scale 26); these keep the loop on the device.
static uint64_t *indices_of_vector(GB_Object *Iv, int64_t ni) {
    need_vec(Iv, "I");
    if (Iv->nrows != ni) grb_throw(GrB_DIMENSION_MISMATCH, "index vector and operand sizes differ");
    if (vec_nvals(Iv) != ni) grb_throw(GrB_INVALID_VALUE, "the index vector must be full");
    const int c = Iv->type->code;
    if (c == TC_FP32 || c == TC_FP64 || c == TC_BOOL) grb_throw(GrB_DOMAIN_MISMATCH, "the index vector must have an integer type");
    uint64_t *d = dalloc<uint64_t>(ni);
    cast_array(Iv->vals, c, d, TC_UINT64, ni);
    return d;
}
extern "C" GrB_Info GxB_b200_Vector_assign_Vector(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_Vector u_,
                                       const GrB_Vector I_, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u"), *Iv = obj_check(I_, "I");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    need_vec(u, "u");
    vector_assign_indexed_dev(w, mask, accum_code_of(accum), u, indices_of_vector(Iv, u->nrows), u->nrows, desc_flags(desc));
    if (g.blocking) stream_sync();
    API_END(w_)
}
extern "C" GrB_Info GxB_b200_Vector_extract_Vector(GrB_Vector w_, const GrB_Vector mask_, const GrB_BinaryOp accum, const GrB_Vector u_,
                                        const GrB_Vector I_, const GrB_Descriptor desc) {
    API_BEGIN
    require_init();
    GB_Object *w = obj_check(w_, "w"), *u = obj_check(u_, "u"), *Iv = obj_check(I_, "I");
    GB_Object *mask = mask_ ? obj_check(mask_, "mask") : nullptr;
    need_vec(w, "w"); need_vec(u, "u");
    const int64_t ni = w->nrows;
    if (mask && mask->nrows != ni) grb_throw(GrB_DIMENSION_MISMATCH, "extract: mask and w sizes differ");
    uint64_t *dI = indices_of_vector(Iv, ni);
    void *t = dmalloc((size_t)(ni > 0 ? ni : 1) * u->type->size);
    uint8_t *tp = dalloc<uint8_t>(ni);
    int *bad = dalloc<int>(1);
    dzero(bad, sizeof(int));
    GRB_TYPE_SWITCH(u->type->code, T, {
        GRB_LAUNCH("extract", k_extract<T>, ew_grid(ni), 256, 0, (const T *)u->vals, u->present, dI, ni, u->nrows, (T *)t, tp, bad);
    });
    try { check_bad(bad, "extract"); } catch (...) { dfree(t); dfree(tp); dfree(dI); throw; }
    vec_writeback_any(w, mask, desc_flags(desc), accum_code_of(accum), t, u->type->code, tp);
    dfree(t); dfree(tp); dfree(dI);
    API_END(w_)
}

// ---- option getters: this library stores matrices by row (with a cached by-column copy), vectors as bitmaps ----
GrB_Info GxB_Global_Option_get_INT32(GxB_Option_Field field, int32_t *value) {
    if (!value) return GrB_NULL_POINTER;
    switch (field) {
        case GxB_MODE: *value = g.blocking ? GrB_BLOCKING : GrB_NONBLOCKING; return g.initialized ? GrB_SUCCESS : GrB_PANIC;
        case GxB_FORMAT: *value = GxB_BY_ROW; return GrB_SUCCESS;
        default: return GrB_INVALID_VALUE;
    }
}
GrB_Info GxB_Matrix_Option_get_INT32(GrB_Matrix A, GxB_Option_Field field, int32_t *value) {
    if (!value) return GrB_NULL_POINTER;
    API_BEGIN
    GB_Object *o = obj_check_lazy(A, "A");
    const bool is_vec = o->kind != K_MATRIX;
    switch (field) {
        case GxB_FORMAT: *value = (is_vec || o->csr_stale) ? GxB_BY_COL : GxB_BY_ROW; break;
        case GxB_SPARSITY_STATUS: *value = is_vec ? GxB_BITMAP : GxB_SPARSE; break;
        case GxB_SPARSITY_CONTROL: *value = is_vec ? GxB_BITMAP : GxB_SPARSE; break;
        default: return GrB_INVALID_VALUE;
    }
    API_END(A)
}
GrB_Info GxB_Vector_Option_get_INT32(GrB_Vector v, GxB_Option_Field field, int32_t *value) {
    return GxB_Matrix_Option_get_INT32((GrB_Matrix)v, field, value);
}

}  // extern "C"

// ---------------------------------------------------------------- fused isequal
// What python-graphblas's Vector.isequal computes with eWiseMult(eq) + nvals + reduce(land) (reference
// graphblas/core/vector.py isequal -> core/base.py _isequal), i.e. four host synchronisations per call, in one pass and
// one read-back: same pattern, and `eq` true on every common entry.  The fixed-point loops (Bellman-Ford, FastSV) call
// it every iteration.
template <class T>
__global__ void __launch_bounds__(256) k_vec_isequal(const T *uv, const uint8_t *up, const T *vv, const uint8_t *vp, int64_t n, int *differ) {
    bool d = false;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const bool pu = up[i] != 0, pv = vp[i] != 0;
        d |= (pu != pv) || (pu && !(uv[i] == vv[i]));
    }
    if (__any_sync(0xffffffffu, d) && (threadIdx.x & 31) == 0) *differ = 1;
}

extern "C" GrB_Info GxB_b200_Vector_isequal(bool *result, const GrB_Vector u_, const GrB_Vector v_) {
    if (!result) return GrB_NULL_POINTER;
    API_BEGIN
    require_init();
    GB_Object *u = obj_check(u_, "u"), *v = obj_check(v_, "v");
    if (u->kind == K_MATRIX || v->kind == K_MATRIX) grb_throw(GrB_INVALID_OBJECT, "isequal: needs vectors");
    if (u->type->code != v->type->code) grb_throw(GrB_DOMAIN_MISMATCH, "isequal: types differ (compare through eWiseMult(eq))");
    if (u->nrows != v->nrows) { *result = false; return GrB_SUCCESS; }
    const int64_t n = u->nrows;
    if (n == 0 || u == v) { *result = true; return GrB_SUCCESS; }
    int *differ = (int *)(g.d_scratch + 56);
    dzero(differ, sizeof(int));
    const int blocks = (int)(cdiv(n, 256) < 148 * 8 ? cdiv(n, 256) : 148 * 8);
    GRB_TYPE_SWITCH(u->type->code, T,
        GRB_LAUNCH("vec_isequal", k_vec_isequal<T>, blocks, 256, 0, (const T *)u->vals, u->present, (const T *)v->vals, v->present, n, differ));
    int h = 0;
    d2h(&h, differ, sizeof(int));
    stream_sync();
    *result = (h == 0);
    API_END(u_)
}
