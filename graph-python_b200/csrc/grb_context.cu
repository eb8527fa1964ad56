// grb_context.cu — library state, built-in handle tables, descriptors, GxB_b200_* plumbing.
// Reference seams: graphblas/__init__.py:118-197 (_init -> initialize(blocking, memory_manager="numpy")),
// graphblas/core/descriptor.py:51-156, graphblas/exceptions.py:93-155.
#include "grb_internal.cuh"
#include "grb_abi_lists.inc"

GrbGlobal g;

// ---------------- device memory ----------------
void *dmalloc(size_t bytes) {
    void *p = nullptr;
    if (bytes == 0) bytes = 16;
    CUDA_TRY(cudaMallocAsync(&p, bytes, g.stream));
    return p;
}
void dfree(void *p) {
    if (p) cudaFreeAsync(p, g.stream);
}
void h2d(void *dst, const void *src, size_t bytes) {
    if (bytes) CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, g.stream));
}
void d2h(void *dst, const void *src, size_t bytes) {
    if (bytes) CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, g.stream));
}
void d2d(void *dst, const void *src, size_t bytes) {
    if (bytes) CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, g.stream));
}
void dzero(void *dst, size_t bytes) {
    if (bytes) CUDA_TRY(cudaMemsetAsync(dst, 0, bytes, g.stream));
}
void stream_sync() { CUDA_TRY(cudaStreamSynchronize(g.stream)); }
int64_t read_i64(const int64_t *dptr) {
    CUDA_TRY(cudaMemcpyAsync(g.h_pinned, dptr, sizeof(int64_t), cudaMemcpyDeviceToHost, g.stream));
    stream_sync();
    return g.h_pinned[0];
}

// ---------------- built-in handle tables ----------------
static GB_Type_opaque type_tab[TC_COUNT] = {
#define X(NAME, CT, CODE) {CODE, sizeof(CT), "GrB_" #NAME},
    GRB_TYPE_LIST(X)
#undef X
};
#define TCODE_BOOL TC_BOOL
#define TCODE_INT8 TC_INT8
#define TCODE_INT16 TC_INT16
#define TCODE_INT32 TC_INT32
#define TCODE_INT64 TC_INT64
#define TCODE_UINT8 TC_UINT8
#define TCODE_UINT16 TC_UINT16
#define TCODE_UINT32 TC_UINT32
#define TCODE_UINT64 TC_UINT64
#define TCODE_FP32 TC_FP32
#define TCODE_FP64 TC_FP64

// ---------------------------------------------------------------- per-kernel device timing
// A ring of CUDA event pairs recorded on g.stream around each launch while g.timing != 0; folded into
// per-name totals when the ring fills or when a total is read.  bench.py uses it for roofline.achieved.
struct KtName { const char *name; double ms; uint64_t count; };
struct KtPending { cudaEvent_t a, b; int name_idx; cudaStream_t stream; };
static std::vector<KtName> kt_names;
static std::vector<KtPending> kt_ring;
static int kt_used = 0;
static const int KT_RING = 2048;
static void ktime_fold() {
    if (kt_used == 0) return;
    for (int i = 0; i < kt_used; i++) cudaEventSynchronize(kt_ring[i].b);      // launches may sit on different streams
    for (int i = 0; i < kt_used; i++) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, kt_ring[i].a, kt_ring[i].b) == cudaSuccess) {
            kt_names[kt_ring[i].name_idx].ms += ms;
            kt_names[kt_ring[i].name_idx].count++;
        }
    }
    kt_used = 0;
}
int ktime_begin(const char *name, cudaStream_t stream) {
    if ((int)kt_ring.size() < KT_RING) {
        kt_ring.resize(KT_RING);
        for (auto &p : kt_ring) { cudaEventCreate(&p.a); cudaEventCreate(&p.b); }
    }
    if (kt_used == KT_RING) ktime_fold();
    int idx = -1;
    for (size_t i = 0; i < kt_names.size(); i++) if (kt_names[i].name == name || !strcmp(kt_names[i].name, name)) { idx = (int)i; break; }
    if (idx < 0) { kt_names.push_back({name, 0.0, 0}); idx = (int)kt_names.size() - 1; }
    kt_ring[kt_used].name_idx = idx;
    kt_ring[kt_used].stream = stream;
    cudaEventRecord(kt_ring[kt_used].a, stream);
    return kt_used++;
}
int ktime_begin(const char *name) { return ktime_begin(name, g.stream); }
void ktime_end(int slot) { cudaEventRecord(kt_ring[slot].b, kt_ring[slot].stream); }

extern "C" {
#define X(NAME, CT, CODE) GrB_Type GrB_##NAME = &type_tab[CODE];
GRB_TYPE_LIST(X)
#undef X

#define X(SYM, OP, XT, ZT) static GB_BinaryOp_opaque obj_##SYM = {OP_##OP, TCODE_##XT, TCODE_##ZT, #SYM}; \
                           GrB_BinaryOp SYM = &obj_##SYM;
GRB_BINOP_LIST(X)
#undef X

#define X(SYM, OP, T) static GB_UnaryOp_opaque obj_##SYM = {UOP_##OP, TCODE_##T, #SYM}; GrB_UnaryOp SYM = &obj_##SYM;
GRB_UNOP_LIST(X)
#undef X

#define X(SYM, MON, T) static GB_Monoid_opaque obj_##SYM = {OP_##MON, TCODE_##T, #SYM}; GrB_Monoid SYM = &obj_##SYM;
GRB_MONOID_LIST(X)
#undef X

#define OP_PAIR OP_ONEB
#define X(SYM, MON, MUL, T) static GB_Semiring_opaque obj_##SYM = {OP_##MON, OP_##MUL, TCODE_##T, #SYM}; \
                            GrB_Semiring SYM = &obj_##SYM;
GRB_SEMIRING_LIST(X)
#undef X

// registry-only handles: distinct objects (the reference keys dictionaries by handle), all marked OP_UNSUPPORTED
#define X(SYM) static GB_UnaryOp_opaque obj_##SYM = {OP_UNSUPPORTED, -1, #SYM}; GrB_UnaryOp SYM = &obj_##SYM;
GRB_UNOP_STUB_LIST(X)
#undef X
#define X(SYM) static GB_BinaryOp_opaque obj_##SYM = {OP_UNSUPPORTED, -1, -1, #SYM}; GrB_BinaryOp SYM = &obj_##SYM;
GRB_BINOP_STUB_LIST(X)
#undef X
#define X(SYM) static GB_IndexUnaryOp_opaque obj_##SYM = {OP_UNSUPPORTED, #SYM}; GrB_IndexUnaryOp SYM = &obj_##SYM;
GRB_IDXOP_STUB_LIST(X)
#undef X
#define X(SYM) static GB_Monoid_opaque obj_##SYM = {OP_UNSUPPORTED, -1, #SYM}; GrB_Monoid SYM = &obj_##SYM;
GRB_MONOID_STUB_LIST(X)
#undef X
#define X(SYM) static GB_Semiring_opaque obj_##SYM = {OP_UNSUPPORTED, OP_UNSUPPORTED, -1, #SYM}; GrB_Semiring SYM = &obj_##SYM;
GRB_SEMIRING_STUB_LIST(X)
#undef X

static GB_IndexUnaryOp_opaque obj_TRIL = {IDX_TRIL, "GrB_TRIL"}, obj_TRIU = {IDX_TRIU, "GrB_TRIU"},
                              obj_DIAG = {IDX_DIAG, "GrB_DIAG"}, obj_OFFDIAG = {IDX_OFFDIAG, "GrB_OFFDIAG"};
GrB_IndexUnaryOp GrB_TRIL = &obj_TRIL, GrB_TRIU = &obj_TRIU, GrB_DIAG = &obj_DIAG, GrB_OFFDIAG = &obj_OFFDIAG;

extern "C" { double GxB_HYPER_DEFAULT = 0.0625, GxB_ALWAYS_HYPER = 1.0, GxB_NEVER_HYPER = -1.0; }   // no hypersparse format here

static GB_Descriptor_opaque *make_desc(bool r, bool c, bool s, bool t0, bool t1) {
    GB_Descriptor_opaque *d = new GB_Descriptor_opaque();
    d->replace = r; d->comp = c; d->structure = s; d->tran0 = t0; d->tran1 = t1; d->builtin = true;
    return d;
}
#define X(NAME, R, C, S, T0, T1) GrB_Descriptor GrB_DESC_##NAME = make_desc(R, C, S, T0, T1);
GRB_DESC_LIST(X)
#undef X

static const uint64_t grb_all_sentinel = 0;
const uint64_t *GrB_ALL = &grb_all_sentinel;
}  // extern "C"

// ---------------- context ----------------
static GrB_Info do_init(GrB_Mode mode) {
    if (g.initialized) return GrB_INVALID_VALUE;  // C API: GrB_init may be called once
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        fprintf(stderr, "[grb_b200] GrB_init: no CUDA device (%s); libgrb_b200 has no CPU fallback\n",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
        return GrB_PANIC;
    }
    try {
        CUDA_TRY(cudaGetDevice(&g.device));
        cudaDeviceProp prop;
        CUDA_TRY(cudaGetDeviceProperties(&prop, g.device));
        g.sm_count = prop.multiProcessorCount;
        cudaMemPool_t pool;
        CUDA_TRY(cudaDeviceGetDefaultMemPool(&pool, g.device));
        uint64_t thresh = UINT64_MAX;  // keep freed blocks in the pool: iterative algorithms reuse them
        CUDA_TRY(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh));
        CUDA_TRY(cudaMalloc((void **)&g.d_scratch, 64 * sizeof(int64_t)));
        CUDA_TRY(cudaMallocHost((void **)&g.h_pinned, 64 * sizeof(int64_t)));
        CUDA_TRY(cudaStreamCreateWithFlags(&g.copy_stream, cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&g.copy_ev_a, cudaEventDisableTiming));
        CUDA_TRY(cudaEventCreateWithFlags(&g.copy_ev_b, cudaEventDisableTiming));
        const char *b = getenv("GRB_B200_BURBLE");
        g.burble = b ? atoi(b) : 0;
    } catch (const GrbError &err) {
        fprintf(stderr, "[grb_b200] GrB_init failed: %s\n", err.msg.c_str());
        return err.info;
    }
    g.blocking = (mode == GrB_BLOCKING);
    g.initialized = true;
    return GrB_SUCCESS;
}

extern "C" {

GrB_Info GrB_init(GrB_Mode mode) { return do_init(mode); }

GrB_Info GxB_init(GrB_Mode mode, void *(*user_malloc)(size_t), void *(*user_calloc)(size_t, size_t),
                  void *(*user_realloc)(void *, size_t), void (*user_free)(void *)) {
    (void)user_calloc; (void)user_realloc;
    if (!user_malloc || !user_free) return GrB_NULL_POINTER;
    GrB_Info info = do_init(mode);
    if (info == GrB_SUCCESS) { g.h_malloc = user_malloc; g.h_free = user_free; }
    return info;
}

GrB_Info GrB_finalize(void) {
    if (g.initialized) {
        if (g.pending_n) { try { pending_flush(); } catch (const GrbError &) {} }
        cudaStreamSynchronize(g.stream);
        g.initialized = false;
    }
    return GrB_SUCCESS;
}

GrB_Info GrB_getVersion(unsigned int *version, unsigned int *subversion) {
    if (!version || !subversion) return GrB_NULL_POINTER;
    *version = 2; *subversion = 0;
    return GrB_SUCCESS;
}

// ---------------- descriptors ----------------
GrB_Info GrB_Descriptor_new(GrB_Descriptor *desc) {
    if (!desc) return GrB_NULL_POINTER;
    *desc = new GB_Descriptor_opaque();
    return GrB_SUCCESS;
}
GrB_Info GrB_Descriptor_set(GrB_Descriptor d, GrB_Desc_Field field, GrB_Desc_Value val) {
    if (!d) return GrB_NULL_POINTER;
    if (d->magic != GRB_MAGIC) return GrB_INVALID_OBJECT;
    if (d->builtin) { d->err = "predefined descriptors are read-only"; return GrB_INVALID_VALUE; }
    switch (field) {
    case GrB_OUTP:
        if (val == GrB_REPLACE) d->replace = true; else if (val == GxB_DEFAULT) d->replace = false;
        else { d->err = "GrB_OUTP takes GrB_REPLACE or default"; return GrB_INVALID_VALUE; }
        return GrB_SUCCESS;
    case GrB_MASK:
        if (val == GrB_COMP) d->comp = true;
        else if (val == GrB_STRUCTURE) d->structure = true;
        else if (val == GxB_DEFAULT) d->comp = d->structure = false;
        else { d->err = "GrB_MASK takes GrB_COMP, GrB_STRUCTURE or default"; return GrB_INVALID_VALUE; }
        return GrB_SUCCESS;
    case GrB_INP0:
        if (val == GrB_TRAN) d->tran0 = true; else if (val == GxB_DEFAULT) d->tran0 = false;
        else { d->err = "GrB_INP0 takes GrB_TRAN or default"; return GrB_INVALID_VALUE; }
        return GrB_SUCCESS;
    case GrB_INP1:
        if (val == GrB_TRAN) d->tran1 = true; else if (val == GxB_DEFAULT) d->tran1 = false;
        else { d->err = "GrB_INP1 takes GrB_TRAN or default"; return GrB_INVALID_VALUE; }
        return GrB_SUCCESS;
    case GxB_AxB_METHOD:
        if (val == GxB_DEFAULT || val == GxB_AxB_GUSTAVSON || val == GxB_AxB_DOT || val == GxB_AxB_HASH ||
            val == GxB_AxB_SAXPY) { d->axb_method = (int)val; return GrB_SUCCESS; }
        d->err = "unknown GxB_AxB_METHOD"; return GrB_INVALID_VALUE;
    case GxB_DESCRIPTOR_NTHREADS: case GxB_DESCRIPTOR_CHUNK: case GxB_DESCRIPTOR_GPU_ID: case GxB_SORT:
    case GxB_COMPRESSION: case GxB_IMPORT:
        return GrB_SUCCESS;  // CPU-threading / lazy-sort hints: accepted, no effect on results
    }
    d->err = "unknown descriptor field";
    return GrB_INVALID_VALUE;
}
GrB_Info GxB_Desc_set_INT32(GrB_Descriptor d, GrB_Desc_Field field, int32_t val) {
    return GrB_Descriptor_set(d, field, (GrB_Desc_Value)val);
}
GrB_Info GrB_Descriptor_free(GrB_Descriptor *desc) {
    if (!desc || !*desc) return GrB_SUCCESS;
    if ((*desc)->magic == GRB_MAGIC && !(*desc)->builtin) { (*desc)->magic = GRB_FREED; delete *desc; }
    *desc = nullptr;
    return GrB_SUCCESS;
}
GrB_Info GrB_Descriptor_error(const char **error, const GrB_Descriptor d) {
    if (!error) return GrB_NULL_POINTER;
    *error = (d && d->magic == GRB_MAGIC) ? d->err.c_str() : "";
    return GrB_SUCCESS;
}

// GrB_Semiring_new composes BUILT-IN monoids and multiplies (the reference's registry does this at import for plus_pow,
// *_truediv, ... graphblas/core/operator/semiring.py:205-243, 370-390).  The handle always comes into being; whether it can
// be used is decided at the call: a (monoid, multiply, type) triple with an instantiated kernel runs, anything else --
// registry-only operators, comparison multiplies, and every user-defined function (there is no JIT and no CPU fallback,
// north_star) -- returns GrB_NOT_IMPLEMENTED from GrB_mxm / GrB_mxv / GrB_vxm.
GrB_Info GrB_Semiring_new(GrB_Semiring *semiring, GrB_Monoid add, GrB_BinaryOp multiply) {
    if (!semiring) return GrB_NULL_POINTER;
    *semiring = nullptr;
    if (!add || !multiply) return GrB_NULL_POINTER;
    GB_Semiring_opaque *s = new GB_Semiring_opaque{add->op, multiply->op, add->tcode, "user semiring", true};
    const bool usable = add->op != OP_UNSUPPORTED && multiply->op != OP_UNSUPPORTED && !op_is_cmp(multiply->op) &&
                        multiply->xcode == add->tcode && multiply->zcode == add->tcode;
    if (!usable) { s->mon = OP_UNSUPPORTED; s->mul = OP_UNSUPPORTED; }
    *semiring = s;
    return GrB_SUCCESS;
}
GrB_Info GrB_Semiring_free(GrB_Semiring *semiring) {
    if (semiring && *semiring && (*semiring)->heap) delete *semiring;     // built-ins are static
    if (semiring) *semiring = nullptr;
    return GrB_SUCCESS;
}
// user-defined operators need a JIT or a CPU: refused (the symbols exist because the reference looks them up by name)
GrB_Info GrB_UnaryOp_new(GrB_UnaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype) {
    (void)function; (void)ztype; (void)xtype;
    if (op) *op = nullptr;
    return GrB_NOT_IMPLEMENTED;
}
GrB_Info GrB_BinaryOp_new(GrB_BinaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype, GrB_Type ytype) {
    (void)function; (void)ztype; (void)xtype; (void)ytype;
    if (op) *op = nullptr;
    return GrB_NOT_IMPLEMENTED;
}
GrB_Info GrB_IndexUnaryOp_new(GrB_IndexUnaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype, GrB_Type ytype) {
    (void)function; (void)ztype; (void)xtype; (void)ytype;
    if (op) *op = nullptr;
    return GrB_NOT_IMPLEMENTED;
}
GrB_Info GrB_UnaryOp_free(GrB_UnaryOp *op) { if (op) *op = nullptr; return GrB_SUCCESS; }
GrB_Info GrB_BinaryOp_free(GrB_BinaryOp *op) { if (op) *op = nullptr; return GrB_SUCCESS; }
GrB_Info GrB_IndexUnaryOp_free(GrB_IndexUnaryOp *op) { if (op) *op = nullptr; return GrB_SUCCESS; }
GrB_Info GrB_Monoid_free(GrB_Monoid *op) { if (op) *op = nullptr; return GrB_SUCCESS; }
// operator handles carry no error state: the string is always empty (graphblas/exceptions.py:143-154 reads it)
GrB_Info GrB_Semiring_error(const char **error, const GrB_Semiring op) { (void)op; if (error) *error = ""; return GrB_SUCCESS; }
GrB_Info GrB_Monoid_error(const char **error, const GrB_Monoid op) { (void)op; if (error) *error = ""; return GrB_SUCCESS; }
GrB_Info GrB_BinaryOp_error(const char **error, const GrB_BinaryOp op) { (void)op; if (error) *error = ""; return GrB_SUCCESS; }
GrB_Info GrB_UnaryOp_error(const char **error, const GrB_UnaryOp op) { (void)op; if (error) *error = ""; return GrB_SUCCESS; }
GrB_Info GrB_IndexUnaryOp_error(const char **error, const GrB_IndexUnaryOp op) { (void)op; if (error) *error = ""; return GrB_SUCCESS; }
GrB_Info GrB_Type_error(const char **error, const GrB_Type t) { (void)t; if (error) *error = ""; return GrB_SUCCESS; }

// ---------------- B200 plumbing ----------------
GrB_Info GxB_b200_set_stream(void *cuda_stream) {
    if (g.initialized && g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return e.info; } }
    if (g.initialized) cudaStreamSynchronize(g.stream);
    g.stream = (cudaStream_t)cuda_stream;
    return GrB_SUCCESS;
}
GrB_Info GxB_b200_sync(void) {
    if (!g.initialized) return GrB_PANIC;
    if (g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return e.info; } }
    return cudaStreamSynchronize(g.stream) == cudaSuccess ? GrB_SUCCESS : GrB_PANIC;
}
GrB_Info GxB_b200_device_info(int *device, int *sm_count, size_t *free_bytes, size_t *total_bytes) {
    if (!g.initialized) return GrB_PANIC;
    if (device) *device = g.device;
    if (sm_count) *sm_count = g.sm_count;
    size_t f = 0, t = 0;
    cudaMemGetInfo(&f, &t);
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return GrB_SUCCESS;
}
uint64_t GxB_b200_launch_count(void) { return g.launches; }
const char *GxB_b200_last_kernel(void) { return g.last_kernel; }
GrB_Info GxB_b200_kernel_time(const char *name, double *total_ms, uint64_t *launches) {
    if (!name || !total_ms || !launches) return GrB_NULL_POINTER;
    ktime_fold();
    *total_ms = 0; *launches = 0;
    for (auto &k : kt_names)
        if (!strcmp(k.name, name)) { *total_ms += k.ms; *launches += k.count; }
    return GrB_SUCCESS;
}
GrB_Info GxB_b200_kernel_time_dump(void) {
    ktime_fold();
    for (auto &k : kt_names)
        fprintf(stderr, "[grb_b200 timing] %-28s launches=%8llu total_ms=%10.3f avg_us=%9.2f\n", k.name,
                (unsigned long long)k.count, k.ms, k.count ? 1e3 * k.ms / k.count : 0.0);
    return GrB_SUCCESS;
}
GrB_Info GxB_b200_kernel_time_reset(void) {
    ktime_fold();
    kt_names.clear();
    return GrB_SUCCESS;
}
GrB_Info GxB_b200_set_option(const char *name, int64_t value) {
    if (!name) return GrB_NULL_POINTER;
    if (!strcmp(name, "push_ratio")) g.push_ratio = value;
    else if (!strcmp(name, "force_direction")) g.force_direction = value;
    else if (!strcmp(name, "burble")) g.burble = value;
    else if (!strcmp(name, "hub_mode")) g.hub_mode = value;
    else if (!strcmp(name, "red_flags")) g.red_flags = value;
    else if (!strcmp(name, "spgemm_panels")) g.spgemm_panels = value;
    else if (!strcmp(name, "dot_hash")) g.dot_hash = value;
    else if (!strcmp(name, "panel_kb")) { if (value < 0) return GrB_INVALID_VALUE; g.panel_kb = value; }
    else if (!strcmp(name, "push_edge_ratio")) { if (value < 0) return GrB_INVALID_VALUE; g.push_edge_ratio = value; }
    else if (!strcmp(name, "smem_kb")) { if (value < 64 || value > 226) return GrB_INVALID_VALUE; g.smem_kb = value; }
    else if (!strcmp(name, "timing")) { if (!value) ktime_fold(); g.timing = value; }
    else if (!strcmp(name, "peer_carveout")) g.peer_carveout = value;
    else if (!strcmp(name, "lazy")) {       // 0: never defer element-wise work (grb_internal.cuh PendingOp)
        if (g.initialized && g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return e.info; } }
        g.lazy = value;
    }
    else return GrB_INVALID_VALUE;
    return GrB_SUCCESS;
}

}  // extern "C"
