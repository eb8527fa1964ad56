// grb_noalloc.cu — GxB_b200_mxv_plus_times_FP32_noalloc: w = A (+).(x) u in FP32 with NO device or host allocation.
//
// No counterpart in the reference; added at the request of the run's operators for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Every other entry point of the library
// allocates from the stream-ordered pool (temporaries, per-tile carries, cached matrix forms); this one does not.
//
// It is a variant of the tiled pull kernel k_spmv_pull + k_spmv_fixup (grb_spmv.cuh), the path GrB_mxv takes on the first
// product with a matrix.  That kernel needs scratch memory: the first row of every 256-entry tile and, per tile, the
// partial sums of the rows that cross its borders.  GrB_mxv takes them from the pool; here they live in ONE file-scope
// __device__ array of fixed size, which bounds the matrix: at most GXB_B200_NOALLOC_MAX_NVALS = 2^26 (67 108 864) entries
// (262 144 tiles, 3.5 MB of scratch, held by every GPU that loads the library until the process exits).
//
// The array is shared by every call and keeps the last call's contents.  Nothing in it needs clearing except the output
// counter: each tile's carries are written before the fix-up pass reads them, and the tile table is rebuilt per call.
// Two calls that overlap on the GPU (different streams, or different host threads on different streams) RACE on the
// scratch: callers must serialise them (calls on the library's one stream are ordered and safe).
#include "grb_internal.cuh"
#include "grb_ops.cuh"
#include "grb_spmv.cuh"

namespace {

constexpr int64_t NOALLOC_MAX_NVALS = (int64_t)1 << 26;
constexpr int64_t NOALLOC_MAX_TILES = NOALLOC_MAX_NVALS / SPMV_TILE;      // 262 144
static_assert(NOALLOC_MAX_NVALS == GXB_B200_NOALLOC_MAX_NVALS, "header and kernel disagree on the limit");

struct NoallocScratch {
    int32_t tile_row[NOALLOC_MAX_TILES + 64];       // first row owned by each tile (+ the end marker)
    float carry_head[NOALLOC_MAX_TILES];            // partial of the row entering a tile from the left
    float carry_own[NOALLOC_MAX_TILES];             // partial of a tile's last row when it leaves the tile
    uint8_t carry_head_has[NOALLOC_MAX_TILES];
    uint8_t carry_own_has[NOALLOC_MAX_TILES];
    unsigned long long count;                       // entries of w (cleared by every call)
};

// In host code this name is a host-side stand-in, not a device address: the address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char noalloc_scratch[sizeof(NoallocScratch)];

constexpr int MAX_DEVICES = 64;
static NoallocScratch *scratch_of_device[MAX_DEVICES];       // resolved once per device, with that device current

// same search as k_tile_rows (grb_storage.cu), writing into the static scratch
__global__ void k_noalloc_tile_rows(const int64_t *rowptr, int64_t nrows, int64_t ntiles, int32_t *tile_row) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    if (t == ntiles) { tile_row[t] = (int32_t)nrows; return; }
    const int64_t target = t * (int64_t)SPMV_TILE;
    int64_t lo = 0, hi = nrows;      // first r in [0, nrows] with rowptr[r] >= target
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (rowptr[mid] >= target) hi = mid; else lo = mid + 1;
    }
    tile_row[t] = (int32_t)lo;
}

using SRf = Semiring<OP_PLUS, OP_TIMES, float>;

GrB_Info fail(const char **error, GrB_Info info, const char *msg) {
    if (error) *error = msg;
    return info;
}

}  // namespace

extern "C" GrB_Info GxB_b200_mxv_plus_times_FP32_noalloc(float *w, uint8_t *w_present, GrB_Index *w_nvals,
                                                         const int64_t *rowptr, const uint32_t *colidx, const float *values,
                                                         const float *u, GrB_Index nrows, GrB_Index ncols, GrB_Index nvals,
                                                         const char **error) {
    if (error) *error = "";
    if (g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return fail(error, e.info, "GxB_b200_mxv_plus_times_FP32_noalloc: deferred work failed"); } }
    if (!w || !w_present || !rowptr || (nvals && (!colidx || !values || !u)))
        return fail(error, GrB_NULL_POINTER, "GxB_b200_mxv_plus_times_FP32_noalloc: w, w_present, rowptr, colidx, values or u is NULL");
    if (nvals > (GrB_Index)NOALLOC_MAX_NVALS)
        return fail(error, GrB_INSUFFICIENT_SPACE,
                    "GxB_b200_mxv_plus_times_FP32_noalloc: the matrix has more than GXB_B200_NOALLOC_MAX_NVALS = 67108864 entries; "
                    "the static scratch holds the carries of 262144 tiles only (use GrB_mxv, which allocates)");
    if (nrows >= ((GrB_Index)1 << 31) || ncols >= ((GrB_Index)1 << 32))
        return fail(error, GrB_INVALID_VALUE, "GxB_b200_mxv_plus_times_FP32_noalloc: nrows must be < 2^31 and ncols < 2^32");
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES)
        return fail(error, GrB_PANIC, "GxB_b200_mxv_plus_times_FP32_noalloc: no current CUDA device");
    if (!scratch_of_device[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(error, GrB_PANIC, "GxB_b200_mxv_plus_times_FP32_noalloc: cudaGetSymbolAddress failed for the static scratch");
        }
        scratch_of_device[dev] = (NoallocScratch *)p;
    }
    NoallocScratch *s = scratch_of_device[dev];
    const int64_t nnz = (int64_t)nvals, n = (int64_t)nrows;
    const int64_t ntiles = nnz > 0 ? cdiv(nnz, SPMV_TILE) : 1;

    PullArgs<float> a;
    a.rowptr = rowptr; a.col = colidx; a.av = values; a.a_iso = 0;
    a.u = u; a.up = nullptr;
    a.hub_v = nullptr; a.hub_p = nullptr; a.hub_n = 0;
    a.nrows = n; a.nnz = nnz; a.ntiles = ntiles;
    a.tile_row = s->tile_row;
    a.carry_head = s->carry_head; a.carry_head_has = s->carry_head_has;
    a.carry_own = s->carry_own; a.carry_own_has = s->carry_own_has;
    a.epi.w_out = w; a.epi.wp_out = w_present; a.epi.w_in = w; a.epi.wp_in = w_present;
    a.epi.mask_p = nullptr; a.epi.mask_v = nullptr; a.epi.mask_vsize = 0; a.epi.mask_fp = 0;
    a.epi.comp = 0; a.epi.replace = 0; a.epi.accum = -1;
    a.epi.count = &s->count;

    cudaStream_t st = g.stream;
    if (cudaMemsetAsync(&s->count, 0, sizeof(unsigned long long), st) != cudaSuccess) {
        cudaGetLastError();
        return fail(error, GrB_PANIC, "GxB_b200_mxv_plus_times_FP32_noalloc: cudaMemsetAsync failed");
    }
    k_noalloc_tile_rows<<<(int)cdiv(ntiles + 1, 256), 256, 0, st>>>(rowptr, n, ntiles, s->tile_row);
    const int64_t want = cdiv(ntiles, SPMV_WARPS);
    const int grid = (int)(want < g.sm_count ? (want > 0 ? want : 1) : g.sm_count);
    k_spmv_pull<SRf><<<grid, SPMV_THREADS, spmv_pull_smem<float>(0, true), st>>>(a);
    k_spmv_fixup<SRf><<<(int)cdiv(ntiles * 32, 256), 256, 0, st>>>(a);
    g.launches += 3;
    g.last_kernel = "spmv_pull_noalloc";
    if (cudaGetLastError() != cudaSuccess)
        return fail(error, GrB_PANIC, "GxB_b200_mxv_plus_times_FP32_noalloc: kernel launch failed");
    if (w_nvals) {
        // the count crosses the ABI through the caller's own word: one small copy into host memory the caller owns
        unsigned long long c = 0;
        if (cudaMemcpyAsync(&c, &s->count, sizeof c, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            cudaStreamSynchronize(st) != cudaSuccess) {
            cudaGetLastError();
            return fail(error, GrB_PANIC, "GxB_b200_mxv_plus_times_FP32_noalloc: reading the entry count failed");
        }
        *w_nvals = (GrB_Index)c;
    }
    return GrB_SUCCESS;
}
