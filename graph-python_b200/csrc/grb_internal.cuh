// grb_internal.cuh — object model, device memory and error plumbing of libgrb_b200.so.
//
// Design (B200-first, not a SuiteSparse port):
//   * every GrB_Matrix lives in HBM as CSR with 64-bit row pointers, 32-bit column indices and typed
//     (or iso) values; a transposed copy (CSR of A') is built lazily on the device and cached, so each
//     product can run in its bandwidth-friendly direction (pull = streaming, push = sparse frontier);
//   * every GrB_Vector lives in HBM as a dense value array + presence bytes (the GxB bitmap format),
//     double-buffered so `w(accum) << w.vxm(A)` ping-pongs without allocation, with an optional sorted
//     index list (the sparse frontier) and a device-side nvals counter;
//   * Scalar / Vector / Matrix handles share one struct so the reference's pointer casts
//     (Vector._as_matrix, reference graphblas/core/vector.py:194-202) stay legal;
//   * all work is enqueued on one CUDA stream (default: the legacy default stream, which is torch's
//     current stream unless changed; GxB_b200_set_stream overrides); the host only synchronises when a
//     value must cross the ABI (nvals, extract*, export*, wait).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/GraphBLAS_b200.h"

// ---- codes (type codes match tools/gen_abi.py; op codes match oracle/grb_oracle.c) ----
enum TypeCode { TC_BOOL = 0, TC_INT8, TC_INT16, TC_INT32, TC_INT64, TC_UINT8, TC_UINT16, TC_UINT32,
                TC_UINT64, TC_FP32, TC_FP64, TC_COUNT };
enum OpCode { OP_FIRST = 0, OP_SECOND, OP_ONEB, OP_ANY, OP_PLUS, OP_MINUS, OP_TIMES, OP_DIV, OP_MIN, OP_MAX,
              OP_LOR, OP_LAND, OP_LXOR, OP_LXNOR, OP_EQ, OP_NE, OP_GT, OP_LT, OP_GE, OP_LE,
              OP_RDIV,   // internal: div with the operands swapped (vxm hands the kernels u as the SECOND operand)
              OP_COUNT };
// ---- semirings on INT8 / INT16 / UINT8 / UINT16 (reference graphblas/core/operator/semiring.py:170-204 registers them).
// There are no 8/16-bit atomics, so they are computed in the 32-bit type of the same signedness and wrapped to the narrow
// type afterwards.  That is exact when the monoid is PLUS (mod 2^32 then mod 2^8/2^16) or when the multiply cannot leave the
// narrow range (first, second, pair, min, max); where it can and the monoid orders values (min_plus, min_times, max_plus,
// max_times) or saturates (div), the multiply itself wraps: internal multiply code = op | narrow id << 8.
enum NarrowId { NARROW_INT8 = 1, NARROW_INT16 = 2, NARROW_UINT8 = 3, NARROW_UINT16 = 4 };
#define GRB_NARROW_OPS(OPNAME)                                                                    \
    OP_##OPNAME##_N1 = OP_##OPNAME | (1 << 8), OP_##OPNAME##_N2 = OP_##OPNAME | (2 << 8),         \
    OP_##OPNAME##_N3 = OP_##OPNAME | (3 << 8), OP_##OPNAME##_N4 = OP_##OPNAME | (4 << 8)
enum NarrowOp { GRB_NARROW_OPS(PLUS), GRB_NARROW_OPS(TIMES), GRB_NARROW_OPS(DIV), GRB_NARROW_OPS(RDIV) };
inline bool tc_is_narrow(int c) { return c == TC_INT8 || c == TC_INT16 || c == TC_UINT8 || c == TC_UINT16; }
inline int tc_wide_of(int c) { return (c == TC_INT8 || c == TC_INT16) ? TC_INT32 : TC_UINT32; }
inline int narrow_id_of(int c) { return c == TC_INT8 ? NARROW_INT8 : c == TC_INT16 ? NARROW_INT16 : c == TC_UINT8 ? NARROW_UINT8 : NARROW_UINT16; }
inline int narrow_mul(int mon, int mul, int narrow_code) {
    const bool ordered = (mon == OP_MIN || mon == OP_MAX) && (mul == OP_PLUS || mul == OP_TIMES);
    return (ordered || mul == OP_DIV || mul == OP_RDIV) ? (mul | (narrow_id_of(narrow_code) << 8)) : mul;
}

// handles that exist for the reference's operator registry only (tools/gen_abi.py registry_only_*): no kernel
#define OP_UNSUPPORTED 1000
inline bool op_is_cmp(int op) { return op >= OP_EQ && op <= OP_LE; }
enum IdxOp { IDX_TRIL = 0, IDX_TRIU, IDX_DIAG, IDX_OFFDIAG };
enum ObjKind { K_SCALAR = 0, K_VECTOR = 1, K_MATRIX = 2 };

static const size_t kTypeSize[TC_COUNT] = {1, 1, 2, 4, 8, 1, 2, 4, 8, 4, 8};

struct GB_Type_opaque { int code; size_t size; const char *name; };
enum UnaryCode { UOP_IDENTITY = 0, UOP_AINV, UOP_MINV, UOP_ABS, UOP_ONE, UOP_BNOT, UOP_LNOT };
struct GB_UnaryOp_opaque { int op; int tcode; const char *name; };
struct GB_IndexUnaryOp_opaque { int op; const char *name; };
struct GB_BinaryOp_opaque { int op; int xcode; int zcode; const char *name; };
struct GB_Monoid_opaque { int op; int tcode; const char *name; };
struct GB_Semiring_opaque { int mon; int mul; int tcode; const char *name; bool heap; };   // heap: made by GrB_Semiring_new

#define GRB_MAGIC 0x42323030u   /* "B200" */
#define GRB_FREED 0x46524545u

struct GB_Descriptor_opaque {
    uint32_t magic = GRB_MAGIC;
    bool replace = false, comp = false, structure = false, tran0 = false, tran1 = false;
    int axb_method = 0;  // GxB_DEFAULT / GUSTAVSON / DOT / HASH / SAXPY (a hint; results identical)
    bool builtin = false;
    std::string err;
};

// CSR block in HBM.  rowptr int64[nrows+1]; colidx uint32[nnz] sorted within rows; vals type[nnz] or [1] if iso.
struct Csr {
    int64_t nrows = 0, ncols = 0, nnz = 0;
    int64_t *rowptr = nullptr;
    uint32_t *colidx = nullptr;
    void *vals = nullptr;
    bool iso = false;
    // tile metadata for the streaming SpMV (first owned row of every nnz tile), built lazily
    int32_t *tile_row = nullptr;
    int64_t ntiles = 0;
    int tile_nnz = 0;
    // hub cache for the pull kernel (csr_ensure_hubs): the hub_n most referenced columns, and a copy of colidx in
    // which those columns are replaced by 0x80000000 | rank.  Built after the matrix has been pulled from twice.
    uint32_t *hub_ids = nullptr;
    uint32_t *hub_col = nullptr;   // bit 31: hub (low bits = rank); bit 30: first entry of its row; else the column
    int hub_n = 0;
    int pull_uses = 0;
    uint32_t *nzrow = nullptr;     // ids of the rows that have entries, ascending
    int64_t n_nzrow = 0;
    int64_t *tile_first = nullptr; // [stream tiles + 1] rows of nzrow that start before the tile
    uint32_t *chunk_first = nullptr;   // [ceil(nnz / 8)] index in nzrow of the row that owns entry 8 * i
    int64_t n_stiles = 0;
    void *cast_vals = nullptr;     // the values in another type (semiring type != matrix type), kept across calls
    int cast_code = -1;
    // column panels (csr_ensure_panels): the same rows restricted to column ranges of `panel_width`, so that the
    // slice of the vector one pull gathers from stays L2-resident when the whole vector does not (multi-GPU row
    // blocks see all n columns).  Each panel is a full Csr with its own values (in type panel_code) and hub cache.
    void *hub_vbuf = nullptr; uint8_t *hub_pbuf = nullptr; int64_t hub_buf_bytes = 0;   // this call's values at the hub columns (reused)
    // per-tile row-block range + tiles per block of the stream form, for progress reporting (csr_ensure_tile_blocks)
    uint32_t *tile_blocks = nullptr;
    unsigned long long *tile_expected = nullptr;
    int64_t tile_blocks_rpb = 0, tile_blocks_tstep = 0;
    int tile_blocks_n = 0;
    Csr *panels = nullptr;
    int n_panels = 0;
    int panel_code = -1;
    int64_t panel_width = 0;
};

struct GB_Object {
    uint32_t magic = GRB_MAGIC;
    int kind = K_MATRIX;
    GrB_Type type = nullptr;
    int64_t nrows = 0, ncols = 0;
    // ---- matrix storage ----
    Csr csr;             // A by rows
    Csr csc;             // A' by rows (== A by columns), valid iff csc_valid
    bool csc_valid = false;
    bool csr_stale = false;   // imported by columns: only csc is populated until somebody needs rows (mat_ensure_csr)
    // ---- dense vector / scalar storage (n = nrows) ----
    void *vals = nullptr;
    uint8_t *present = nullptr;
    void *vals_alt = nullptr;        // ping-pong spares (lazily allocated)
    uint8_t *present_alt = nullptr;
    int64_t nvals = 0;               // host copy, valid iff nvals_known
    bool nvals_known = true;
    int64_t *d_nvals = nullptr;      // device counter written by the producing kernel
    uint32_t *list = nullptr;        // sorted indices of present entries (sparse frontier), iff list_valid
    int64_t list_cap = 0;
    bool list_valid = false;
    bool external_vals = false;      // vals belongs to the caller (GxB_b200_Vector_wrap_device): never freed, never swapped
    std::string err;
};
struct GB_Matrix_opaque : GB_Object {};
struct GB_Vector_opaque : GB_Object {};
struct GB_Scalar_opaque : GB_Object {};

struct GrbError {
    GrB_Info info;
    std::string msg;
};
[[noreturn]] inline void grb_throw(GrB_Info info, const std::string &msg) { throw GrbError{info, msg}; }

#define CUDA_TRY(expr)                                                                             \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            GrB_Info _i = (_e == cudaErrorMemoryAllocation) ? GrB_OUT_OF_MEMORY : GrB_PANIC;       \
            char _b[512];                                                                          \
            snprintf(_b, sizeof _b, "CUDA error %s at %s:%d (%s)", cudaGetErrorName(_e), __FILE__, \
                     __LINE__, cudaGetErrorString(_e));                                            \
            cudaGetLastError();                                                                    \
            grb_throw(_i, _b);                                                                     \
        }                                                                                          \
    } while (0)

// ---- deferred element-wise work (GrB_NONBLOCKING only) ----
// Two statements of an iterative loop are cheap, full-vector passes that always precede a product:
//     w << u.ewise_mult(v, op)        (every operand full, no mask, no accumulator)
//     r[:] << x                       (plain fill)
// In non-blocking mode the C API lets them wait: the first is queued, the second joins it, and the next entry point of the
// library (whatever it is) runs both as ONE kernel before doing anything else.  PageRank's iteration goes from four launches
// to three.  Nothing else is ever deferred; every API entry flushes the queue first (API_BEGIN), GrB_wait / GxB_b200_sync
// included, so an object is materialised whenever the caller can see it.  `GxB_b200_set_option("lazy", 0)` turns it off.
struct GB_Object;
struct PendingOp {
    int kind;                      // 1 = w = u op v (full), 2 = w[:] = x
    GB_Object *w, *u, *v;
    int op, xcode;
    int64_t n;
    unsigned long long xbits;      // the scalar of a fill, in w's type
    bool set_present;              // fill: w was not known to be full, presence bytes are written too
    bool is_union;
};

struct GrbGlobal {
    bool initialized = false;
    bool blocking = false;
    int64_t lazy = 1;
    int pending_n = 0;
    PendingOp pending[2];
    int device = -1;
    int sm_count = 148;
    cudaStream_t stream = 0;
    cudaStream_t copy_stream = nullptr;   // second stream: the values of an imported matrix cross PCIe while the indices are checked
    cudaEvent_t copy_ev_a = nullptr, copy_ev_b = nullptr;
    void *(*h_malloc)(size_t) = malloc;
    void (*h_free)(void *) = free;
    uint64_t launches = 0;
    const char *last_kernel = "";
    // tunables (GxB_b200_set_option)
    int64_t push_ratio = 24;       // push when nvals(u) * push_ratio < n
    int64_t force_direction = 0;   // 0 auto, 1 pull, 2 push
    int64_t hub_mode = 1;          // 0 never, 1 after the second pull of a matrix with >= 2^20 entries, 2 always,
                                   // 3 always + scan-based k_spmv_stream for masked / foreign-accumulator write-backs
    int64_t burble = 0;
    int64_t panel_kb = 24576;      // pull: split a matrix into column panels when the gathered vector exceeds this many KB
                                   // (0 = never); 24 MB keeps one panel's slice of the vector well inside the 126 MB L2
    int64_t push_edge_ratio = 4;   // push only while (edges leaving the frontier) * ratio <= nnz; beyond that the pull reads less
    int64_t smem_kb = 194;         // shared-memory budget of the hub-table kernels: <= 195 KB keeps the 196 KB carve-out, i.e. 60 KB of L1
                                   // for gather misses in flight (228 KB carve-out: 28 KB L1, gathers 1.5x slower, tools/microbench/gather_bench5)
    int64_t red_flags = 0;         // k_spmv_red: bit 0 = direct loads instead of TMA bulk copies (A/B measurements)
    int64_t dot_hash = 1;          // masked dot product with a pair multiply: per-row hash sets instead of binary searches
    int64_t spgemm_panels = 1;     // SpGEMM heavy rows: shared-memory column panels (0 = per-CTA dense accumulator rows in HBM)
    int64_t timing = 0;            // record CUDA events around every launch (GxB_b200_kernel_time reads them)
    int64_t peer_carveout = 86;    // preferred shared-memory carve-out (percent) of the fused peer exchange kernel: must match k_spmv_red's
    // one-shot request set by GxB_b200_next_product_progress: the next k_spmv_red launch reports per-row-block progress
    unsigned long long *progress_dev = nullptr, *progress_expected_dev = nullptr;
    int progress_blocks = 0;
    int64_t progress_rpb = 0;
    // scratch
    int64_t *d_scratch = nullptr;  // 64 int64 of device scratch (counters)
    int64_t *h_pinned = nullptr;   // 64 int64 of pinned host memory for small read-backs
};
extern GrbGlobal g;
// per-kernel-name device timing (grb_context.cu): events on g.stream around each launch while g.timing is set
int ktime_begin(const char *name);
int ktime_begin(const char *name, cudaStream_t stream);   // a launch on another stream (grb_peer.cu)
void ktime_end(int slot);

// every kernel launch goes through this macro: counts launches (bench.py's gpu_launches) and names the path
#define GRB_LAUNCH(name, kernel, grid, block, smem, ...)                             \
    do {                                                                             \
        const int _slot = g.timing ? ktime_begin(name) : -1;                         \
        kernel<<<(grid), (block), (smem), g.stream>>>(__VA_ARGS__);                  \
        if (_slot >= 0) ktime_end(_slot);                                            \
        g.launches++;                                                                \
        g.last_kernel = name;                                                        \
        if (g.burble) fprintf(stderr, "[grb_b200] %s grid=%lld block=%d\n", name, (long long)(grid), (int)(block)); \
        CUDA_TRY(cudaGetLastError());                                                \
    } while (0)

// ---- device memory (stream-ordered pool) ----
void *dmalloc(size_t bytes);
void dfree(void *p);
template <class T> inline T *dalloc(int64_t n) { return (T *)dmalloc((size_t)(n > 0 ? n : 1) * sizeof(T)); }
void h2d(void *dst, const void *src, size_t bytes);
void d2h(void *dst, const void *src, size_t bytes);
void d2d(void *dst, const void *src, size_t bytes);
void dzero(void *dst, size_t bytes);
void stream_sync();
int64_t read_i64(const int64_t *dptr);   // synchronising read-back of one device int64

// ---- object helpers (grb_containers.cu) ----
GB_Object *obj_check(const void *h, const char *what);   // NULL -> GrB_NULL_POINTER, bad magic -> INVALID_OBJECT; materialises a stale CSR
GB_Object *obj_check_lazy(const void *h, const char *what);   // same, but leaves a column-imported matrix as it is (hot paths pick a side)
void mat_ensure_csr(GB_Object *o);                        // rows of A from the cached columns, if A was imported by columns
void csr_free(Csr &c);
GB_Object *obj_new(int kind, GrB_Type type, uint64_t nrows, uint64_t ncols);
void obj_free(GB_Object *o);
void obj_clear(GB_Object *o);
// SuiteSparse-style hand-off with ownership transfer (grb_containers.cu; the import/export forms are in grb_exchange.cu)
void pack_csx(GB_Object *A, bool by_col, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index Ap_size, GrB_Index Ai_size,
              GrB_Index Ax_size, bool iso);
void unpack_csx(GB_Object *A, bool by_col, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index *Ap_size, GrB_Index *Ai_size,
                GrB_Index *Ax_size, bool *iso, bool *jumbled);
void obj_release_storage(GB_Object *o);
void obj_invalidate_caches(GB_Object *o);                 // after the CSR changed: drop csc + tiles
void vec_alloc(GB_Object *o);                             // allocate dense storage (zeroed presence)
void vec_ensure_alt(GB_Object *o);
int64_t vec_nvals(GB_Object *o);                          // syncs if the count is still on the device
struct DescFlags;
void vec_writeback_any(GB_Object *w, GB_Object *mask, const DescFlags &f, int accum, const void *t, int tcode,
                       const uint8_t *tp);                 // w<m,replace> = accum(w, t), t cast to w's type (grb_ewise.cu)
void vector_assign_indexed(GB_Object *w, GB_Object *mask, int accum, GB_Object *u, const uint64_t *I, int64_t ni,
                           const DescFlags &f);            // w<m>(I) = accum(w(I), u)  (grb_vecops.cu)
void vec_set_unknown_nvals(GB_Object *o);                 // producer kernels wrote d_nvals
void vec_build_list(GB_Object *o);                        // dense -> sorted index list (device compaction)
void mat_ensure_csc(GB_Object *o);                        // build + cache the transpose on the device
void csr_transpose(const Csr &a, Csr &t, size_t tsize);
void csr_ensure_tiles(Csr &c, int tile_nnz);
void csr_ensure_hubs(Csr &c, int capacity);
int hub_capacity_for(size_t tsize);                       // hub-table entries k_spmv_red holds for values of tsize bytes (grb_mxv.cu)
void csr_ensure_tile_blocks(Csr &c, int64_t rows_per_block, int nblocks, int64_t tstep);   // needs the stream form (csr_ensure_hubs)               // build / rebuild the hub cache for at most `capacity` hubs
void csr_ensure_panels(Csr &c, int npanels, const void *vals, int tcode);   // split by column ranges (values in tcode)
void csr_drop_caches(Csr &c);                             // tiles + hubs (after the pattern changed)
const void *csr_vals_as(Csr &c, int from_code, int to_code, int via_code = -1);   // c.vals, or a cached typed copy of them
void mat_set_csr(GB_Object *o, Csr &c);                   // take ownership of c as o's CSR
Csr csr_dup(const Csr &a, size_t tsize);
Csr vec_as_csr_column(GB_Object *v);                      // n x 1 matrix view of a vector (copy)
Csr vec_as_csr_row(GB_Object *v);                         // 1 x n
void csr_to_vec(const Csr &c, GB_Object *v, bool as_column);
void csr_cast_values(Csr &c, int from_code, int to_code);

struct DescFlags { bool replace, comp, structure, tran0, tran1; int axb; };
DescFlags desc_flags(const GrB_Descriptor d);

// API boundary: catch GrbError, stash the message on the first object for GrB_*_error
void pending_flush();             // grb_ewise.cu: run the deferred element-wise work, if any
inline bool lazy_on() { return g.lazy && !g.blocking; }
#define API_BEGIN try { if (g.pending_n) pending_flush();
#define API_BEGIN_NOFLUSH try {   // the two entry points that may defer flush by themselves
#define API_END(errobj)                                                          \
    }                                                                            \
    catch (const GrbError &e) {                                                  \
        GB_Object *_o = (GB_Object *)(errobj);                                   \
        if (_o && _o->magic == GRB_MAGIC) _o->err = e.msg;                       \
        if (g.burble) fprintf(stderr, "[grb_b200] error %d: %s\n", (int)e.info, e.msg.c_str()); \
        return e.info;                                                           \
    }                                                                            \
    catch (const std::bad_alloc &) { return GrB_OUT_OF_MEMORY; }                 \
    return GrB_SUCCESS;

inline void require_init() {
    if (!g.initialized) grb_throw(GrB_PANIC, "GrB_init has not been called (or no CUDA device: libgrb_b200 has no CPU fallback)");
}

// ---- type dispatch: binds T to the C++ type of a type code ----
#define GRB_TYPE_SWITCH(code, T, ...)                                              \
    switch (code) {                                                                \
    case TC_BOOL: { using T = bool; __VA_ARGS__; } break;                          \
    case TC_INT8: { using T = int8_t; __VA_ARGS__; } break;                        \
    case TC_INT16: { using T = int16_t; __VA_ARGS__; } break;                      \
    case TC_INT32: { using T = int32_t; __VA_ARGS__; } break;                      \
    case TC_INT64: { using T = int64_t; __VA_ARGS__; } break;                      \
    case TC_UINT8: { using T = uint8_t; __VA_ARGS__; } break;                      \
    case TC_UINT16: { using T = uint16_t; __VA_ARGS__; } break;                    \
    case TC_UINT32: { using T = uint32_t; __VA_ARGS__; } break;                    \
    case TC_UINT64: { using T = uint64_t; __VA_ARGS__; } break;                    \
    case TC_FP32: { using T = float; __VA_ARGS__; } break;                         \
    case TC_FP64: { using T = double; __VA_ARGS__; } break;                        \
    default: grb_throw(GrB_INVALID_VALUE, "unknown type code");                    \
    }

inline void op_need_kernel(int op, const char *name) {
    if (op == OP_UNSUPPORTED)
        grb_throw(GrB_NOT_IMPLEMENTED, std::string(name ? name : "operator") +
                  " has no sm_100a kernel in libgrb_b200 (registry-only handle; there is no CPU fallback)");
}
inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
