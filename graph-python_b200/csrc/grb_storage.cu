// grb_storage.cu — device-side storage plumbing: COO -> CSR build (sort + dedup with dup op), transpose,
// vector <-> n x 1 / 1 x n CSR conversions, index narrowing, SpMV tile metadata.
// This is ingest / format plumbing, not the hot loop; the sort and
// scan primitives come from CUB (shipped with the CUDA toolkit), everything else is written here.
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>

#include "grb_internal.cuh"
#include "grb_ops.cuh"
#include "grb_storage.cuh"

// ---------------------------------------------------------------- small kernels
__global__ void k_iota_u32(uint32_t *p, int64_t n) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) p[i] = (uint32_t)i;
}
__global__ void k_make_keys(const uint64_t *I, const uint64_t *J, uint64_t nrows, uint64_t ncols, int64_t n,
                            uint64_t *keys, int *bad) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint64_t i = I[k], j = J ? J[k] : 0;
    if (i >= nrows || j >= ncols) { *bad = 1; keys[k] = 0; return; }
    keys[k] = i * ncols + j;
}
__global__ void k_head_flags(const uint64_t *keys, int64_t n, int64_t *flag) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < n) flag[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1 : 0;
}
// one thread per distinct key: fold its duplicates in input order with the dup op, emit col + value
template <class T>
__global__ void k_dedup(const uint64_t *keys, const uint32_t *perm, const int64_t *pos, int64_t n,
                        const T *vin, int iso_in, int dup_op, uint64_t ncols, uint32_t *col_out,
                        uint32_t *row_out, T *vout, int *dup_seen) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    if (k != 0 && keys[k] == keys[k - 1]) return;
    uint64_t key = keys[k];
    T acc = vin ? vin[iso_in ? 0 : perm[k]] : (T)0;
    int64_t q = k + 1;
    while (q < n && keys[q] == key) {
        if (dup_op < 0) *dup_seen = 1;
        else if (vin) acc = binop_rt<T>(dup_op, acc, vin[iso_in ? 0 : perm[q]]);
        q++;
    }
    int64_t o = pos[k];
    col_out[o] = (uint32_t)(key % ncols);
    row_out[o] = (uint32_t)(key / ncols);
    if (vout) vout[o] = acc;
}
// rowptr from sorted row ids: rowptr[r] = first entry whose row >= r
__global__ void k_rowptr_from_rows(const uint32_t *rows, int64_t nnz, int64_t nrows, int64_t *rowptr) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k > nnz) return;
    int64_t lo = (k == 0) ? 0 : (int64_t)rows[k - 1] + 1;
    int64_t hi = (k == nnz) ? nrows : (int64_t)rows[k];
    for (int64_t r = lo; r <= hi; r++) rowptr[r] = k;
}
__global__ void k_expand_rows(const int64_t *rowptr, int64_t nrows, uint32_t *rows) {
    // one warp per row
    int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (w >= nrows) return;
    for (int64_t p = rowptr[w] + lane; p < rowptr[w + 1]; p += 32) rows[p] = (uint32_t)w;
}
__global__ void k_transpose_keys(const uint32_t *rows, const uint32_t *cols, int64_t nnz, uint64_t nrows,
                                 uint64_t *keys) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < nnz) keys[k] = (uint64_t)cols[k] * nrows + rows[k];
}
__global__ void k_narrow_u64(const uint64_t *in, int64_t n, uint32_t *out, uint64_t bound, int *bad) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint64_t v = in[k];
    if (v >= bound) *bad = 1;
    out[k] = (uint32_t)v;
}
__global__ void k_widen_u32(const uint32_t *in, int64_t n, uint64_t *out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < n) out[k] = in[k];
}
__global__ void k_check_csr(const int64_t *rowptr, const uint32_t *col, int64_t nrows, int64_t nnz, int *flags) {
    // flags[0]: rowptr not monotone / out of range; flags[1]: some row unsorted or with duplicates
    int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (w >= nrows) return;
    int64_t a = rowptr[w], b = rowptr[w + 1];
    if (a > b || a < 0 || b > nnz) { flags[0] = 1; return; }
    for (int64_t p = a + 1 + lane; p < b; p += 32)
        if (col[p] <= col[p - 1]) flags[1] = 1;
}
template <class T>
__global__ void k_gather(const T *in, const uint32_t *perm, int64_t n, T *out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < n) out[k] = in[perm[k]];
}

static inline int nblk(int64_t n, int b = 256) { return (int)cdiv(n > 0 ? n : 1, b); }

// ---------------------------------------------------------------- COO -> CSR
// keys = row*ncols+col (device, unsorted, n of them); vals may be null (pattern only) or iso (1 value).
// dup_op >= 0: fold duplicates; dup_op == -1: duplicates are an error; result rows sorted, no duplicates.
Csr coo_to_csr(int64_t nrows, int64_t ncols, uint64_t *keys, const void *vals, bool iso_in, int tcode,
               int64_t n, int dup_op, bool keep_iso) {
    Csr c;
    c.nrows = nrows; c.ncols = ncols;
    size_t tsize = kTypeSize[tcode];
    c.rowptr = dalloc<int64_t>(nrows + 1);
    if (n == 0) {
        dzero(c.rowptr, (size_t)(nrows + 1) * sizeof(int64_t));
        c.colidx = dalloc<uint32_t>(1);
        c.vals = dmalloc(tsize);
        return c;
    }
    uint64_t *keys2 = dalloc<uint64_t>(n);
    uint32_t *perm = dalloc<uint32_t>(n), *perm2 = dalloc<uint32_t>(n);
    GRB_LAUNCH("iota", k_iota_u32, nblk(n), 256, 0, perm, n);
    int bits = 1;
    {
        unsigned __int128 span = (unsigned __int128)(uint64_t)nrows * (uint64_t)(ncols > 0 ? ncols : 1);
        while (bits < 64 && ((unsigned __int128)1 << bits) < span) bits++;
    }
    size_t tmp_bytes = 0;
    cub::DoubleBuffer<uint64_t> dk(keys, keys2);
    cub::DoubleBuffer<uint32_t> dv(perm, perm2);
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, n, 0, bits, g.stream));
    void *tmp = dmalloc(tmp_bytes);
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk, dv, n, 0, bits, g.stream));
    g.launches += 2 * ((bits + 7) / 8);
    dfree(tmp);
    uint64_t *sk = dk.Current();
    uint32_t *sp = dv.Current();
    // distinct keys -> output positions
    int64_t *flag = dalloc<int64_t>(n + 1), *pos = dalloc<int64_t>(n + 1);
    GRB_LAUNCH("head_flags", k_head_flags, nblk(n), 256, 0, sk, n, flag);
    dzero(flag + n, sizeof(int64_t));
    tmp_bytes = 0;
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag, pos, n + 1, g.stream));
    tmp = dmalloc(tmp_bytes);
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flag, pos, n + 1, g.stream));
    g.launches += 2;
    dfree(tmp);
    int64_t nnz = read_i64(pos + n);
    c.nnz = nnz;
    c.colidx = dalloc<uint32_t>(nnz);
    uint32_t *rows = dalloc<uint32_t>(nnz);
    bool iso_out = keep_iso && iso_in;
    c.iso = iso_out;
    c.vals = dmalloc(iso_out ? tsize : (size_t)(nnz > 0 ? nnz : 1) * tsize);
    int *dflag = (int *)g.d_scratch;
    dzero(dflag, sizeof(int));
    GRB_TYPE_SWITCH(tcode, T,
        GRB_LAUNCH("dedup", k_dedup<T>, nblk(n), 256, 0, sk, sp, pos, n, (const T *)vals, (int)iso_in,
                   dup_op, (uint64_t)(ncols > 0 ? ncols : 1), c.colidx, rows,
                   iso_out ? (T *)nullptr : (T *)c.vals, dflag));
    if (iso_out && vals) d2d(c.vals, vals, tsize);
    GRB_LAUNCH("rowptr", k_rowptr_from_rows, nblk(nnz + 1), 256, 0, rows, nnz, nrows, c.rowptr);
    int hflag = 0;
    d2h(&hflag, dflag, sizeof(int));
    stream_sync();
    dfree(rows); dfree(flag); dfree(pos); dfree(keys2); dfree(perm); dfree(perm2);
    if (hflag) {
        csr_free(c);
        grb_throw(GrB_INVALID_VALUE, "duplicate indices in build and no dup_op given");
    }
    return c;
}

// ---------------------------------------------------------------- helpers on CSR
void csr_drop_caches(Csr &c) {
    dfree(c.tile_row); c.tile_row = nullptr; c.ntiles = 0; c.tile_nnz = 0;
    dfree(c.hub_ids); dfree(c.hub_col); c.hub_ids = nullptr; c.hub_col = nullptr; c.hub_n = 0; c.pull_uses = 0;
    dfree(c.nzrow); dfree(c.tile_first); dfree(c.chunk_first); c.nzrow = nullptr; c.tile_first = nullptr; c.chunk_first = nullptr;
    c.n_nzrow = 0; c.n_stiles = 0;
    dfree(c.cast_vals); c.cast_vals = nullptr; c.cast_code = -1;
    dfree(c.hub_vbuf); dfree(c.hub_pbuf); c.hub_vbuf = nullptr; c.hub_pbuf = nullptr; c.hub_buf_bytes = 0;
    dfree(c.tile_blocks); dfree(c.tile_expected); c.tile_blocks = nullptr; c.tile_expected = nullptr; c.tile_blocks_rpb = 0; c.tile_blocks_n = 0;
    for (int p = 0; p < c.n_panels; p++) csr_free(c.panels[p]);
    delete[] c.panels; c.panels = nullptr; c.n_panels = 0; c.panel_code = -1; c.panel_width = 0;
}
// Matrices are replaced (mat_set_csr) or, for GrB_Matrix_setElement on an existing entry, edited in place followed by
// obj_invalidate_caches, so a typed copy of the values stays valid for the
// life of the Csr: BOOL adjacency x INT32 vector products cast once, not per call.
// via_code: cast through that type first (a narrow-integer semiring on a wider matrix: INT32 -> INT8 -> INT32)
const void *csr_vals_as(Csr &c, int from_code, int to_code, int via_code) {
    if (via_code == from_code) via_code = -1;
    if (from_code == to_code && via_code < 0) return c.vals;
    const int key = to_code | ((via_code + 1) << 8);
    if (c.cast_vals && c.cast_code == key) return c.cast_vals;
    dfree(c.cast_vals);
    const int64_t n = c.iso ? 1 : c.nnz;
    c.cast_vals = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[to_code]);
    if (via_code < 0) cast_array(c.vals, from_code, c.cast_vals, to_code, n);
    else {
        void *mid = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[via_code]);
        cast_array(c.vals, from_code, mid, via_code, n);
        cast_array(mid, via_code, c.cast_vals, to_code, n);
        dfree(mid);
    }
    c.cast_code = key;
    return c.cast_vals;
}
void csr_free(Csr &c) {
    csr_drop_caches(c);
    dfree(c.rowptr); dfree(c.colidx); dfree(c.vals);
    c = Csr();
}

// ---------------------------------------------------------------- progress table of the stream form
// tile t of the flagged stream covers entries [256 t, 256 t + 256): which row blocks (of rows_per_block rows) do its rows lie in?
__global__ void k_tile_blocks(const uint32_t *pcol, const uint32_t *nzrow, const uint32_t *chunk_first, int64_t nnz, int64_t ntiles,
                              int64_t rpb, uint32_t *tile_blocks, unsigned long long *expected) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    const int64_t e0 = t * 256, e1 = (e0 + 256 < nnz ? e0 + 256 : nnz) - 1;
    const uint32_t first_row = nzrow[chunk_first[e0 >> 3]];
    uint32_t j1 = chunk_first[e1 >> 3];
    for (int64_t e = (e1 & ~(int64_t)7) + 1; e <= e1; e++) j1 += (pcol[e] >> 30) & 1u;      // row starts after the chunk's first entry
    const uint32_t last_row = nzrow[j1];
    const uint32_t b0 = (uint32_t)(first_row / rpb), b1 = (uint32_t)(last_row / rpb);
    tile_blocks[t] = b0 | (b1 << 16);
}
// expected[b] = number of warps (out of `tstep` = grid x warps-per-CTA, tiles dealt round-robin) that touch row block b
__global__ void k_tile_expected(const uint32_t *tile_blocks, int64_t ntiles, int64_t tstep, unsigned long long *expected) {
    const int64_t w = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (w >= tstep) return;
    int64_t done_below = -1;            // blocks <= done_below are already counted for this warp
    for (int64_t t = w; t < ntiles; t += tstep) {
        const uint32_t tb = tile_blocks[t];
        int64_t b0 = tb & 0xffffu;
        const int64_t b1 = tb >> 16;
        if (b0 <= done_below) b0 = done_below + 1;
        for (int64_t b = b0; b <= b1; b++) atomicAdd(expected + b, 1ull);
        if (b1 > done_below) done_below = b1;
    }
}
void csr_ensure_tile_blocks(Csr &c, int64_t rpb, int nblocks, int64_t tstep) {
    if (c.tile_blocks && c.tile_blocks_rpb == rpb && c.tile_blocks_n == nblocks && c.tile_blocks_tstep == tstep) return;
    if (!c.hub_col || !c.nzrow || !c.chunk_first) grb_throw(GrB_INVALID_OBJECT, "progress table: the stream form of the matrix is missing");
    if (nblocks < 1 || nblocks > 65535 || rpb < 1 || cdiv(c.nrows, rpb) > nblocks) grb_throw(GrB_INVALID_VALUE, "progress table: bad row blocks");
    dfree(c.tile_blocks); dfree(c.tile_expected);
    c.tile_blocks = dalloc<uint32_t>(c.n_stiles);
    c.tile_expected = dalloc<unsigned long long>(nblocks);
    dzero(c.tile_expected, (size_t)nblocks * 8);
    GRB_LAUNCH("tile_blocks", k_tile_blocks, nblk(c.n_stiles), 256, 0, c.hub_col, c.nzrow, c.chunk_first, c.nnz, c.n_stiles, rpb,
               c.tile_blocks, c.tile_expected);
    GRB_LAUNCH("tile_expected", k_tile_expected, nblk(tstep), 256, 0, c.tile_blocks, c.n_stiles, tstep, c.tile_expected);
    c.tile_blocks_rpb = rpb; c.tile_blocks_n = nblocks; c.tile_blocks_tstep = tstep;
}

// ---------------------------------------------------------------- column panels
constexpr int MAX_PANELS = 8;
struct PanelPtrs { int64_t *rowptr[MAX_PANELS]; uint32_t *col[MAX_PANELS]; unsigned char *val[MAX_PANELS]; };
__device__ __forceinline__ int64_t panel_lower_bound(const uint32_t *col, int64_t lo, int64_t hi, uint32_t key) {
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (col[mid] < key) lo = mid + 1; else hi = mid; }
    return lo;
}
// one thread per row: where each panel starts inside the row (absolute entry position) and how many entries it holds
__global__ void k_panel_split(const int64_t *rowptr, const uint32_t *col, int64_t nrows, int P, int64_t width,
                              int64_t *start /*[P][nrows]*/, int64_t *count /*[P][nrows+1]*/) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const int64_t e = rowptr[r + 1];
    int64_t b = rowptr[r];
    for (int p = 0; p < P; p++) {
        const int64_t nb = (p + 1 == P) ? e : panel_lower_bound(col, b, e, (uint32_t)((p + 1) * width));
        start[p * nrows + r] = b;
        count[p * (nrows + 1) + r] = nb - b;
        b = nb;
    }
}
__global__ void k_panel_fill(const uint32_t *rows, const uint32_t *col, const unsigned char *vals, int tsize, int iso, int64_t nnz,
                             int64_t nrows, int64_t width, const int64_t *start, PanelPtrs out) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < nnz; e += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t r = rows[e], c = col[e];
        const int p = (int)(c / width);
        const int64_t d = out.rowptr[p][r] + (e - start[p * nrows + r]);
        out.col[p][d] = c;
        if (!iso) {
            if (tsize == 4) ((uint32_t *)out.val[p])[d] = ((const uint32_t *)vals)[e];
            else if (tsize == 8) ((unsigned long long *)out.val[p])[d] = ((const unsigned long long *)vals)[e];
            else if (tsize == 2) ((uint16_t *)out.val[p])[d] = ((const uint16_t *)vals)[e];
            else out.val[p][d] = vals[e];
        }
    }
}
void csr_ensure_panels(Csr &c, int npanels, const void *vals, int tcode) {
    // vals == nullptr: the multiply ignores the matrix values (second, pair): pattern-only panels, code 100 + tcode
    if (npanels > MAX_PANELS) npanels = MAX_PANELS;
    const int code = vals ? tcode : 100 + tcode;
    if (c.n_panels == npanels && c.panel_code == code) return;
    for (int p = 0; p < c.n_panels; p++) csr_free(c.panels[p]);
    delete[] c.panels; c.panels = nullptr; c.n_panels = 0;
    const int64_t nrows = c.nrows;
    const int tsize = (int)kTypeSize[tcode];
    const int64_t width = ((cdiv(c.ncols, npanels) + 1023) / 1024) * 1024;
    int64_t *start = dalloc<int64_t>((int64_t)npanels * nrows), *count = dalloc<int64_t>((int64_t)npanels * (nrows + 1));
    dzero(count, (size_t)npanels * (nrows + 1) * sizeof(int64_t));
    GRB_LAUNCH("panel_split", k_panel_split, nblk(nrows), 256, 0, c.rowptr, c.colidx, nrows, npanels, width, start, count);
    c.panels = new Csr[npanels];
    PanelPtrs pp{};
    size_t tb = 0;
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tb, count, count, nrows + 1, g.stream));
    void *tmp = dmalloc(tb);
    for (int p = 0; p < npanels; p++) {
        Csr &q = c.panels[p];
        q.nrows = nrows; q.ncols = c.ncols; q.iso = c.iso || vals == nullptr;
        q.rowptr = dalloc<int64_t>(nrows + 1);
        CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, count + (int64_t)p * (nrows + 1), q.rowptr, nrows + 1, g.stream));
        g.launches++;
        q.nnz = read_i64(q.rowptr + nrows);
        q.colidx = dalloc<uint32_t>(q.nnz);
        q.vals = dmalloc(q.iso ? (size_t)tsize : (size_t)(q.nnz > 0 ? q.nnz : 1) * tsize);
        if (q.iso) { if (vals) d2d(q.vals, vals, (size_t)tsize); else dzero(q.vals, (size_t)tsize); }
        pp.rowptr[p] = q.rowptr; pp.col[p] = q.colidx; pp.val[p] = (unsigned char *)q.vals;
    }
    dfree(tmp);
    uint32_t *rows = csr_row_ids(c);
    if (c.nnz)
        GRB_LAUNCH("panel_fill", k_panel_fill, 148 * 16, 256, 0, rows, c.colidx, (const unsigned char *)vals, tsize, (int)(c.iso || vals == nullptr), c.nnz,
                   nrows, width, start, pp);
    dfree(rows); dfree(start); dfree(count);
    c.n_panels = npanels; c.panel_code = code; c.panel_width = width;
}

// ---------------------------------------------------------------- hub cache of the pull kernel
// The pull SpMV on a power-law graph is bound by the divergent gather of u, and a small set of columns takes a
// large share of the references (R-MAT scale 22: the top 32 K of 4 M columns hold about half of the entries).
// csr_ensure_hubs ranks the columns by reference count, keeps the top `capacity`, and writes a second index
// array in which a hub column c reads 0x80000000 | rank(c); k_spmv_pull serves those from shared memory.
__global__ void k_col_histogram(const uint32_t *col, int64_t nnz, uint32_t *count) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&count[col[p]], 1u);
}
__global__ void k_hub_rank(const uint32_t *sorted_ids, const uint32_t *sorted_counts, int h, uint32_t *rank, uint32_t *hub_ids) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= h) return;
    hub_ids[i] = sorted_ids[i];
    (void)sorted_counts;
    rank[sorted_ids[i]] = 0x80000000u | (uint32_t)i;
}
__global__ void k_hub_remap(const uint32_t *col, int64_t nnz, const uint32_t *rank, uint32_t *out) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t c = col[p], r = rank[c];
        out[p] = r ? r : c;
    }
}
struct NonEmptyRow {
    const int64_t *rowptr;
    __host__ __device__ bool operator()(const uint32_t &r) const { return rowptr[r + 1] > rowptr[r]; }
};
__global__ void k_row_start_flags(const uint32_t *nzrow, int64_t n, const int64_t *rowptr, uint32_t *pcol) {
    const int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (j < n) pcol[rowptr[nzrow[j]]] |= 0x40000000u;
}
// tile_first[t] = number of non-empty rows that start before entry t*256 (t = ntiles: all of them)
__global__ void k_tile_first(const uint32_t *nzrow, int64_t n, const int64_t *rowptr, int64_t ntiles, int64_t *tile_first) {
    const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    if (t == ntiles) { tile_first[t] = n; return; }
    const int64_t target = t * 256;
    int64_t lo = 0, hi = n;   // first j with rowptr[nzrow[j]] >= target
    while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if (rowptr[nzrow[mid]] >= target) hi = mid; else lo = mid + 1; }
    tile_first[t] = lo;
}
// chunk_first[i] = largest j with rowptr[nzrow[j]] <= 8 * i
__global__ void k_chunk_first(const uint32_t *nzrow, int64_t n, const int64_t *rowptr, int64_t nchunks, uint32_t *chunk_first) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= nchunks) return;
    const int64_t target = i * 8;
    int64_t lo = 0, hi = n - 1;
    while (lo < hi) { const int64_t mid = (lo + hi + 1) >> 1; if (rowptr[nzrow[mid]] <= target) lo = mid; else hi = mid - 1; }
    chunk_first[i] = (uint32_t)lo;
}
void csr_ensure_hubs(Csr &c, int capacity) {
    int h = (int)(c.ncols < capacity ? c.ncols : capacity);
    if (h <= 0 || c.nnz == 0 || c.ncols > (int64_t)0x3fffffff || c.nrows > (int64_t)0xffffffff) return;
    if (c.hub_col && c.hub_n <= capacity && (c.hub_n == h || c.hub_n >= capacity / 2)) return;
    dfree(c.hub_ids); dfree(c.hub_col); c.hub_ids = nullptr; c.hub_col = nullptr; c.hub_n = 0;
    const int64_t n = c.ncols;
    uint32_t *count = dalloc<uint32_t>(n), *count2 = dalloc<uint32_t>(n), *ids = dalloc<uint32_t>(n), *ids2 = dalloc<uint32_t>(n);
    dzero(count, (size_t)n * sizeof(uint32_t));
    GRB_LAUNCH("col_histogram", k_col_histogram, 148 * 16, 256, 0, c.colidx, c.nnz, count);
    GRB_LAUNCH("iota", k_iota_u32, nblk(n), 256, 0, ids, n);
    cub::DoubleBuffer<uint32_t> dk(count, count2), dv(ids, ids2);
    size_t tb = 0;
    CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(nullptr, tb, dk, dv, n, 0, 32, g.stream));
    void *tmp = dmalloc(tb);
    CUDA_TRY(cub::DeviceRadixSort::SortPairsDescending(tmp, tb, dk, dv, n, 0, 32, g.stream));
    g.launches += 8;
    dfree(tmp);
    uint32_t *rank = dk.Alternate();   // reuse the spare key buffer as the rank table
    dzero(rank, (size_t)n * sizeof(uint32_t));
    c.hub_ids = dalloc<uint32_t>(h);
    GRB_LAUNCH("hub_rank", k_hub_rank, nblk(h), 256, 0, dv.Current(), dk.Current(), h, rank, c.hub_ids);
    c.hub_col = dalloc<uint32_t>(c.nnz);
    GRB_LAUNCH("hub_remap", k_hub_remap, 148 * 16, 256, 0, c.colidx, c.nnz, rank, c.hub_col);
    c.hub_n = h;
    dfree(count); dfree(count2); dfree(ids); dfree(ids2);
    // row structure of the flagged stream (k_spmv_stream): non-empty rows, row-start flags, rows before each tile
    c.nzrow = dalloc<uint32_t>(c.nrows);
    int64_t *d_n = g.d_scratch + 8;
    NonEmptyRow pred{c.rowptr};
    cub::CountingInputIterator<uint32_t> it(0);
    tb = 0;
    CUDA_TRY(cub::DeviceSelect::If(nullptr, tb, it, c.nzrow, d_n, c.nrows, pred, g.stream));
    tmp = dmalloc(tb);
    CUDA_TRY(cub::DeviceSelect::If(tmp, tb, it, c.nzrow, d_n, c.nrows, pred, g.stream));
    g.launches += 1;
    dfree(tmp);
    c.n_nzrow = read_i64(d_n);
    GRB_LAUNCH("row_start_flags", k_row_start_flags, nblk(c.n_nzrow), 256, 0, c.nzrow, c.n_nzrow, c.rowptr, c.hub_col);
    c.n_stiles = cdiv(c.nnz, 256);
    c.tile_first = dalloc<int64_t>(c.n_stiles + 1);
    const int64_t nchunks = cdiv(c.nnz, 8);
    c.chunk_first = dalloc<uint32_t>(nchunks);
    GRB_LAUNCH("chunk_first", k_chunk_first, nblk(nchunks), 256, 0, c.nzrow, c.n_nzrow, c.rowptr, nchunks, c.chunk_first);
    GRB_LAUNCH("tile_first", k_tile_first, nblk(c.n_stiles + 1), 256, 0, c.nzrow, c.n_nzrow, c.rowptr, c.n_stiles, c.tile_first);
}

uint32_t *csr_row_ids(const Csr &c) {
    uint32_t *rows = dalloc<uint32_t>(c.nnz);
    if (c.nnz) GRB_LAUNCH("expand_rows", k_expand_rows, nblk(c.nrows * 32), 256, 0, c.rowptr, c.nrows, rows);
    return rows;
}

void narrow_indices(const uint64_t *d_in, int64_t n, uint32_t *d_out, uint64_t bound, const char *what) {
    if (n == 0) return;
    int *dflag = (int *)g.d_scratch;
    dzero(dflag, sizeof(int));
    GRB_LAUNCH("narrow", k_narrow_u64, nblk(n), 256, 0, d_in, n, d_out, bound, dflag);
    int h = 0;
    d2h(&h, dflag, sizeof(int));
    stream_sync();
    if (h) grb_throw(GrB_INDEX_OUT_OF_BOUNDS, std::string("index out of bounds in ") + what);
}
void widen_indices(const uint32_t *d_in, int64_t n, uint64_t *d_out) {
    if (n) GRB_LAUNCH("widen", k_widen_u32, nblk(n), 256, 0, d_in, n, d_out);
}

bool csr_is_sorted(const Csr &c) {
    if (c.nnz == 0) return true;
    int *dflag = (int *)g.d_scratch;
    dzero(dflag, 2 * sizeof(int));
    GRB_LAUNCH("check_csr", k_check_csr, nblk(c.nrows * 32), 256, 0, c.rowptr, c.colidx, c.nrows, c.nnz, dflag);
    int h[2] = {0, 0};
    d2h(h, dflag, 2 * sizeof(int));
    stream_sync();
    if (h[0]) grb_throw(GrB_INVALID_VALUE, "row pointers are not monotone or exceed the index array");
    return h[1] == 0;
}

Csr csr_dup(const Csr &a, size_t tsize) {
    Csr c;
    c.nrows = a.nrows; c.ncols = a.ncols; c.nnz = a.nnz; c.iso = a.iso;
    c.rowptr = dalloc<int64_t>(a.nrows + 1);
    c.colidx = dalloc<uint32_t>(a.nnz);
    size_t vb = a.iso ? tsize : (size_t)(a.nnz > 0 ? a.nnz : 1) * tsize;
    c.vals = dmalloc(vb);
    d2d(c.rowptr, a.rowptr, (size_t)(a.nrows + 1) * sizeof(int64_t));
    d2d(c.colidx, a.colidx, (size_t)a.nnz * sizeof(uint32_t));
    d2d(c.vals, a.vals, a.iso ? tsize : (size_t)a.nnz * tsize);
    return c;
}

static int tcode_of_size(size_t tsize) {
    return tsize == 1 ? TC_UINT8 : tsize == 2 ? TC_UINT16 : tsize == 4 ? TC_UINT32 : TC_UINT64;
}

// t = a' : keys (col*nrows+row) radix-sorted; values moved as opaque words of their size
void csr_transpose(const Csr &a, Csr &t, size_t tsize) {
    uint32_t *rows = csr_row_ids(a);
    uint64_t *keys = dalloc<uint64_t>(a.nnz);
    if (a.nnz)
        GRB_LAUNCH("transpose_keys", k_transpose_keys, nblk(a.nnz), 256, 0, rows, a.colidx, a.nnz,
                   (uint64_t)(a.nrows > 0 ? a.nrows : 1), keys);
    dfree(rows);
    t = coo_to_csr(a.ncols, a.nrows, keys, a.vals, a.iso, tcode_of_size(tsize), a.nnz, OP_SECOND, true);
    dfree(keys);
}

void csr_sort_rows(Csr &c, int tcode) {
    uint32_t *rows = csr_row_ids(c);
    uint64_t *keys = dalloc<uint64_t>(c.nnz);
    // reuse transpose-key kernel with roles swapped: key = row*ncols + col
    if (c.nnz)
        GRB_LAUNCH("sort_keys", k_transpose_keys, nblk(c.nnz), 256, 0, c.colidx, rows, c.nnz,
                   (uint64_t)(c.ncols > 0 ? c.ncols : 1), keys);
    dfree(rows);
    Csr s = coo_to_csr(c.nrows, c.ncols, keys, c.vals, c.iso, tcode, c.nnz, OP_SECOND, true);
    dfree(keys);
    csr_free(c);
    c = s;
}

// ---------------------------------------------------------------- SpMV tile metadata
// tile t covers entries [t*TN, (t+1)*TN); tile_row[t] = first row r with rowptr[r] >= t*TN (row "owned"
// by the tile where it starts).  tile_row[ntiles] = nrows.
__global__ void k_tile_rows(const int64_t *rowptr, int64_t nrows, int64_t ntiles, int tile_nnz, int32_t *tile_row) {
    int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    if (t == ntiles) { tile_row[t] = (int32_t)nrows; return; }
    int64_t target = t * (int64_t)tile_nnz;
    int64_t lo = 0, hi = nrows;  // first r in [0,nrows] with rowptr[r] >= target
    while (lo < hi) {
        int64_t mid = (lo + hi) >> 1;
        if (rowptr[mid] >= target) hi = mid; else lo = mid + 1;
    }
    tile_row[t] = (int32_t)lo;
}
void csr_ensure_tiles(Csr &c, int tile_nnz) {
    if (c.tile_row && c.tile_nnz == tile_nnz) return;
    dfree(c.tile_row);
    c.ntiles = c.nnz > 0 ? cdiv(c.nnz, tile_nnz) : 1;
    c.tile_nnz = tile_nnz;
    c.tile_row = dalloc<int32_t>(c.ntiles + 1);
    GRB_LAUNCH("tile_rows", k_tile_rows, nblk(c.ntiles + 1), 256, 0, c.rowptr, c.nrows, c.ntiles, tile_nnz, c.tile_row);
}

// ---------------------------------------------------------------- dense vector <-> CSR
__global__ void k_count_present(const uint8_t *p, int64_t n, int64_t *count) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int c = 0;
    for (; i < n; i += (int64_t)gridDim.x * blockDim.x) c += p[i] ? 1 : 0;
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd((unsigned long long *)count, (unsigned long long)c);
}
int64_t count_present(const uint8_t *present, int64_t n, int64_t *d_count) {
    dzero(d_count, sizeof(int64_t));
    int blocks = (int)(cdiv(n, 256) < 148 * 8 ? cdiv(n > 0 ? n : 1, 256) : 148 * 8);
    GRB_LAUNCH("count_present", k_count_present, blocks, 256, 0, present, n, d_count);
    return 0;
}
template <class T>
__global__ void k_vec_compact(const uint8_t *present, const T *vals, const int64_t *pos, int64_t n,
                              uint32_t *idx_out, T *val_out) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n || !present[i]) return;
    int64_t o = pos[i];
    if (idx_out) idx_out[o] = (uint32_t)i;
    if (val_out) val_out[o] = vals[i];
}
// pos[i] = number of present entries before i; pos has n+1 entries, pos[n] = total
__global__ void k_present_to_i64(const uint8_t *present, int64_t n, int64_t *flag) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i <= n) flag[i] = (i < n && present[i]) ? 1 : 0;
}
static int64_t *present_scan(const uint8_t *present, int64_t n) {
    int64_t *flag = dalloc<int64_t>(n + 1), *pos = dalloc<int64_t>(n + 1);
    GRB_LAUNCH("present_flags", k_present_to_i64, nblk(n + 1), 256, 0, present, n, flag);
    size_t tb = 0;
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tb, flag, pos, n + 1, g.stream));
    void *tmp = dmalloc(tb);
    CUDA_TRY(cub::DeviceScan::ExclusiveSum(tmp, tb, flag, pos, n + 1, g.stream));
    g.launches += 2;
    dfree(tmp); dfree(flag);
    return pos;
}

void vec_build_list(GB_Object *o) {
    if (o->list_valid) return;
    int64_t n = o->nrows;
    if (n < ((int64_t)1 << 31) - 1) {
        // one selection pass over the presence bytes (n bytes read, 4 nvals written) instead of 8-byte flags, an 8-byte scan and
        // a compaction (33 n bytes of traffic): the frontier list of a BFS level went from 3 launches / 98 us to 1 / ~15 us
        // at 16.7 M vertices
        const int64_t cap = o->nvals_known ? o->nvals : n;
        if (o->list_cap < cap || !o->list) {
            dfree(o->list);
            o->list_cap = cap > 1024 ? cap : 1024;
            o->list = dalloc<uint32_t>(o->list_cap);
        }
        int64_t *d_num = dalloc<int64_t>(1);
        thrust::counting_iterator<uint32_t> ids(0);
        size_t tb = 0;
        CUDA_TRY(cub::DeviceSelect::Flagged(nullptr, tb, ids, o->present, o->list, d_num, (int)n, g.stream));
        void *tmp = dmalloc(tb);
        const int slot = g.timing ? ktime_begin("vec_select_list") : -1;
        CUDA_TRY(cub::DeviceSelect::Flagged(tmp, tb, ids, o->present, o->list, d_num, (int)n, g.stream));
        if (slot >= 0) ktime_end(slot);
        g.launches += 1;
        if (!o->nvals_known) { o->nvals = read_i64(d_num); o->nvals_known = true; }     // known count: no host round trip
        dfree(tmp); dfree(d_num);
        o->list_valid = true;
        return;
    }
    int64_t *pos = present_scan(o->present, n);
    int64_t nv = read_i64(pos + n);
    o->nvals = nv; o->nvals_known = true;
    if (o->list_cap < nv || !o->list) {
        dfree(o->list);
        o->list_cap = nv > 1024 ? nv : 1024;
        o->list = dalloc<uint32_t>(o->list_cap);
    }
    if (nv) GRB_LAUNCH("vec_compact", k_vec_compact<uint8_t>, nblk(n), 256, 0, o->present, (const uint8_t *)nullptr,
                       pos, n, o->list, (uint8_t *)nullptr);
    dfree(pos);
    o->list_valid = true;
}

Csr vec_as_csr_column(GB_Object *v) {
    int64_t n = v->nrows;
    size_t ts = v->type->size;
    Csr c;
    c.nrows = n; c.ncols = 1;
    int64_t *pos = present_scan(v->present, n);
    c.nnz = read_i64(pos + n);
    c.rowptr = pos;
    c.colidx = dalloc<uint32_t>(c.nnz);
    dzero(c.colidx, (size_t)c.nnz * sizeof(uint32_t));
    c.vals = dmalloc((size_t)(c.nnz > 0 ? c.nnz : 1) * ts);
    GRB_TYPE_SWITCH(v->type->code, T,
        GRB_LAUNCH("vec_compact", k_vec_compact<T>, nblk(n), 256, 0, v->present, (const T *)v->vals, pos, n,
                   (uint32_t *)nullptr, (T *)c.vals));
    return c;
}
__global__ void k_row_ptr2(int64_t *rowptr, int64_t nnz) { rowptr[0] = 0; rowptr[1] = nnz; }
Csr vec_as_csr_row(GB_Object *v) {
    int64_t n = v->nrows;
    size_t ts = v->type->size;
    Csr c;
    c.nrows = 1; c.ncols = n;
    int64_t *pos = present_scan(v->present, n);
    c.nnz = read_i64(pos + n);
    c.rowptr = dalloc<int64_t>(2);
    GRB_LAUNCH("row_ptr2", k_row_ptr2, 1, 1, 0, c.rowptr, c.nnz);
    c.colidx = dalloc<uint32_t>(c.nnz);
    c.vals = dmalloc((size_t)(c.nnz > 0 ? c.nnz : 1) * ts);
    GRB_TYPE_SWITCH(v->type->code, T,
        GRB_LAUNCH("vec_compact", k_vec_compact<T>, nblk(n), 256, 0, v->present, (const T *)v->vals, pos, n,
                   c.colidx, (T *)c.vals));
    dfree(pos);
    return c;
}
template <class T>
__global__ void k_csr_col_to_vec(const int64_t *rowptr, const T *cv, int iso, int64_t n, T *vals, uint8_t *present) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t a = rowptr[i], b = rowptr[i + 1];
    present[i] = b > a;
    if (b > a) vals[i] = cv[iso ? 0 : a];
}
template <class T>
__global__ void k_csr_row_to_vec(const uint32_t *col, const T *cv, int iso, int64_t nnz, T *vals, uint8_t *present) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    present[col[k]] = 1;
    vals[col[k]] = cv[iso ? 0 : k];
}
// c must already hold values of v's type
void csr_to_vec(const Csr &c, GB_Object *v, bool as_column) {
    int64_t n = v->nrows;
    GRB_TYPE_SWITCH(v->type->code, T, {
        if (as_column) {
            GRB_LAUNCH("csr_col_to_vec", k_csr_col_to_vec<T>, nblk(n), 256, 0, c.rowptr, (const T *)c.vals,
                       (int)c.iso, n, (T *)v->vals, v->present);
        } else {
            dzero(v->present, (size_t)n);
            if (c.nnz) GRB_LAUNCH("csr_row_to_vec", k_csr_row_to_vec<T>, nblk(c.nnz), 256, 0, c.colidx,
                                  (const T *)c.vals, (int)c.iso, c.nnz, (T *)v->vals, v->present);
        }
    });
    v->nvals = c.nnz; v->nvals_known = true; v->list_valid = false;
}

// ---------------------------------------------------------------- value casts (C typecast rules)
template <class S, class D>
__global__ void k_cast(const S *in, int64_t n, D *out) {
    int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k >= n) return;
    if constexpr (std::is_same<D, bool>::value) out[k] = in[k] != (S)0;
    else out[k] = (D)in[k];
}
void cast_array(const void *in, int from_code, void *out, int to_code, int64_t n) {
    if (n == 0) return;
    GRB_TYPE_SWITCH(from_code, S, {
        GRB_TYPE_SWITCH(to_code, D, {
            auto kern = k_cast<S, D>;
            GRB_LAUNCH("cast", kern, nblk(n), 256, 0, (const S *)in, n, (D *)out);
        });
    });
}
void csr_cast_values(Csr &c, int from_code, int to_code) {
    if (from_code == to_code) return;
    int64_t n = c.iso ? 1 : c.nnz;
    void *nv = dmalloc((size_t)(n > 0 ? n : 1) * kTypeSize[to_code]);
    cast_array(c.vals, from_code, nv, to_code, n);
    dfree(c.vals);
    c.vals = nv;
}
