// grb_spgemm.cuh — kernels behind GrB_mxm.
//
// Gustavson, two-phase (symbolic sizes every output row, numeric fills it), rows binned by an upper bound
// of their products ub = min(flops(i), ncols):
//   k_hash_warp    ub <= 32     one warp per row, 64-slot shared-memory hash table
//   k_hash_block   ub <= 8192   one CTA per row, 1K / 8K / 16K-slot shared-memory hash table (up to 192 KB)
//   k_panel_symbolic / k_panel_row   ub > 8192, the rows that hold > 95 % of the products of an R-MAT A*A:
//                               symbolic = whole-row presence bitmap in shared memory (ncols bits), which also
//                               yields the entries per 16 K-column base panel; numeric = the row's columns are cut
//                               into ranges of s base panels (s the largest power of two whose ranges each hold
//                               <= 10 240 entries), one 16 K-slot shared-memory hash table per range, or a dense
//                               16 K-entry accumulator + bitmap when a single base panel overflows the table.
//                               B rows are cut at panel boundaries through a per-row offset table (ptab), so a
//                               range only reads its own slice of B.  Accumulators never leave shared memory.
//   k_dense_block  fallback of the above when ncols bits do not fit shared memory: per-CTA bitmap + dense
//                               accumulator row in HBM; ordered extraction by popcount prefix sums
// Hash rows are sorted in shared memory (bitonic) before they are written, so every output row is sorted and
// output patterns are bit-exact against the oracle.  Algorithmic bytes:
//   nnz(A)*(4+V) + flops*(4+V) + nnz(C)*(4+V) + 2*n*8  (+ flops*4 + n*8 for the symbolic pass).
//
//   k_dot_masked   C<M> = A*B' under a structural/valued mask (triangle counting): per mask entry, the
//                  shorter row is searched in the longer one (galloping lower bound); long pairs are handed
//                  to whole warps.  bytes = nnz(M)*8 + sum min(da,db)*4*(1+log) + nnz(C)*(4+V).
#pragma once
#include "grb_internal.cuh"
#include "grb_ops.cuh"

constexpr uint32_t HASH_EMPTY = 0xFFFFFFFFu;

template <class T> struct GemmArgs {
    const int64_t *Ap; const uint32_t *Aj; const T *Ax; int a_iso;
    const int64_t *Bp; const uint32_t *Bj; const T *Bx; int b_iso;
    int64_t nrows, ncols;
    const int32_t *rows;     // row ids of this bin
    int64_t nbin;
    int64_t *rownnz;         // symbolic out: entries per row
    const int64_t *Cp;       // numeric in: row offsets
    uint32_t *Cj; T *Cx;
    // dense path
    uint32_t *gbitmap;       // per-CTA bitmaps in HBM when ncols bits do not fit shared memory (else nullptr)
    T *gacc;                 // per-CTA dense accumulator rows, pre-filled with the monoid identity
    int64_t words;           // bitmap words = ceil(ncols / 32)
    // panel path (k_panel_symbolic / k_panel_row)
    const uint32_t *ptab;    // [B rows][np + 1]: offset inside B row k of the first entry with column >= q * PANEL_W
    int np;                  // base panels = ceil(ncols / PANEL_W)
    uint16_t *rowS;          // per row: base panels per hash range (power of two), 0 = dense accumulator per base panel
    unsigned long long *cursor;   // dynamic row scheduler
};

template <class T> __device__ __forceinline__ T ld_val(const T *p, int iso, int64_t k) { return p[iso ? 0 : k]; }
template <class T> struct Product { uint32_t j; T z; };   // one product A(i,k) (x) B(k,j)
template <class SR, bool NUMERIC>
__device__ __forceinline__ Product<typename SR::type> load_product(const GemmArgs<typename SR::type> &a, int64_t p, int64_t q) {
    using T = typename SR::type;
    Product<T> r;
    r.j = a.Bj[q];
    const T av = (NUMERIC && SR::kUsesX) ? ld_val(a.Ax, a.a_iso, p) : (T)0;
    const T bv = (NUMERIC && SR::kUsesY) ? ld_val(a.Bx, a.b_iso, q) : (T)0;
    r.z = NUMERIC ? SR::mult(av, bv) : (T)0;
    return r;
}

template <class SR, bool NUMERIC>
__device__ __forceinline__ void hash_insert(uint32_t *keys, typename SR::type *vals, uint32_t mask, uint32_t j,
                                            typename SR::type z) {
    uint32_t h = (j * 2654435761u) & mask;
    while (true) {
        uint32_t old = keys[h];
        if (old == j) break;
        if (old == HASH_EMPTY) {
            old = atomicCAS(&keys[h], HASH_EMPTY, j);
            if (old == HASH_EMPTY || old == j) break;
        }
        h = (h + 1) & mask;
    }
    if (NUMERIC) SR::Mon::atomic(&vals[h], z);
}

// bitonic sort of (keys, vals) by key in shared memory; EMPTY keys sort to the end.  SYNC() separates stages.
template <class T, bool WARP>
__device__ __forceinline__ void bitonic_sort(uint32_t *keys, T *vals, int n, int tid, int nthreads) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < n; i += nthreads) {
                int ixj = i ^ j;
                if (ixj > i) {
                    bool up = ((i & k) == 0);
                    uint32_t a = keys[i], b = keys[ixj];
                    if ((a > b) == up) {
                        keys[i] = b; keys[ixj] = a;
                        T va = vals[i]; vals[i] = vals[ixj]; vals[ixj] = va;
                    }
                }
            }
            if (WARP) __syncwarp(); else __syncthreads();
        }
    }
}

__device__ __forceinline__ int block_exclusive_scan(int v, int *total, int *s_warp /* >= 33 ints */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    if (warp == 0) {
        int w = lane < nwarps ? s_warp[lane] : 0;
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        s_warp[lane] = w;          // inclusive warp totals
    }
    __syncthreads();
    const int base = warp ? s_warp[warp - 1] : 0;
    *total = s_warp[nwarps - 1];
    __syncthreads();
    return base + x - v;
}


// Emit a shared-memory hash table in key order without sorting the table: keys are dealt into nbk order-preserving
// buckets ((key - min) >> shift, about 8 entries each), a counting pass places slot numbers bucket by bucket, one
// thread per entry ranks it inside its bucket, and the row leaves in order.  O(table) work against the
// O(table log^2 table) of a bitonic sort (measured: 140 us -> a few us per 16 K-slot table); clustered keys cost
// O(bucket^2) compares spread over the bucket's threads.  Returns the entry count.
//   slot: >= (entries) uint16;  bpos: >= nbk uint32;  s_tmp: >= 40 ints;  nbk a power of two
template <class T> __device__ __forceinline__ T ld_acc(const T *p, bool global_vals) { return global_vals ? *(const volatile T *)p : *p; }
template <class T, bool GV = false>
__device__ __forceinline__ int hash_emit_sorted(uint32_t *keys, T *vals, int table, uint16_t *slot, uint32_t *bpos, int nbk,
                                                uint32_t *Cj, T *Cx, int64_t out, int *s_tmp) {
    const int tid = threadIdx.x, nthreads = blockDim.x, lane = tid & 31;
    uint32_t *s_min = (uint32_t *)&s_tmp[33], *s_max = (uint32_t *)&s_tmp[34];
    for (int b = tid; b < nbk; b += nthreads) bpos[b] = 0;
    if (tid == 0) { *s_min = 0xFFFFFFFFu; *s_max = 0; }
    __syncthreads();
    uint32_t lo = 0xFFFFFFFFu, hi = 0;
    for (int k = tid; k < table; k += nthreads) { const uint32_t key = keys[k]; if (key != HASH_EMPTY) { lo = min(lo, key); hi = max(hi, key); } }
    for (int o = 16; o; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
    if (lane == 0 && lo != 0xFFFFFFFFu) { atomicMin(s_min, lo); atomicMax(s_max, hi); }
    __syncthreads();
    lo = *s_min; hi = *s_max;
    if (lo == 0xFFFFFFFFu) return 0;
    int sh = 0;
    while (((hi - lo) >> sh) >= (uint32_t)nbk) sh++;
    for (int k = tid; k < table; k += nthreads) { const uint32_t key = keys[k]; if (key != HASH_EMPTY) atomicAdd(&bpos[(key - lo) >> sh], 1u); }
    __syncthreads();
    // exclusive scan of the bucket counts (each thread owns `per` consecutive buckets)
    const int per = (nbk + nthreads - 1) / nthreads;
    int c = 0;
    for (int b = tid * per; b < min(nbk, (tid + 1) * per); b++) c += (int)bpos[b];
    int total;
    int off = block_exclusive_scan(c, &total, s_tmp);
    for (int b = tid * per; b < min(nbk, (tid + 1) * per); b++) { const int n = (int)bpos[b]; bpos[b] = (uint32_t)off; off += n; }
    __syncthreads();
    for (int k = tid; k < table; k += nthreads) {
        const uint32_t key = keys[k];
        if (key != HASH_EMPTY) slot[atomicAdd(&bpos[(key - lo) >> sh], 1u)] = (uint16_t)k;
    }
    __syncthreads();
    // bpos[b] is now the end of bucket b.  One thread per ENTRY: its rank inside its bucket is the number of smaller
    // keys there (keys of a row are distinct), which is its output position -- ordering and emission in one pass, spread
    // over all threads whatever the bucket sizes are (one thread ordering each bucket by insertion was 11 % of the
    // instructions of k_panel_row and left a clustered bucket to a single thread).
    for (int p = tid; p < total; p += nthreads) {
        const int sl = slot[p];
        const uint32_t key = keys[sl];
        const uint32_t b = (key - lo) >> sh;
        const int beg = b ? (int)bpos[b - 1] : 0, end = (int)bpos[b];
        int rank = 0;
        for (int j = beg; j < end; j++) rank += keys[slot[j]] < key;
        Cj[out + beg + rank] = key;
        Cx[out + beg + rank] = ld_acc(&vals[sl], GV);
    }
    return total;
}

// Every product of one row of A with the slice [q0, q1) (base panels) of the B rows it selects: f(pA, qB).
// A CTA takes blockDim.x entries of the A row at a time, one per thread, so the dependent loads Aj -> Bp -> ptab of a
// whole batch are in flight together.  Tiny B segments are walked by their lane, mid-sized ones by the warp,
// longer ones go to a shared-memory queue that all warps of the CTA drain in 32-entry chunks dealt round-robin: a hub
// row of B met by one lane must not keep the other 31 warps waiting at the barrier that ends the range (ncu: 91 % of
// the stall samples of the first version sat on that barrier).
// (Flattening the 32 segments of a warp with a scan + shuffle search was measured 1.4-4x SLOWER on R-MAT: the search
// costs more than the idle lanes it saves.)
// queue capacity: 384 for the 1024-thread heavy-row kernels; 64 for the 64..512-thread hash-row kernel, whose occupancy the
// 6 KB of a large queue would cut (mid-size rows 35.7 -> 27.8 ms at scale 20 with a smaller queue)
template <int CAP> struct SegQueueT { static constexpr int kCap = CAP; long long s0[CAP]; int len[CAP]; int poff[CAP]; int n[2]; };
using SegQueue = SegQueueT<384>;
using SegQueueSmall = SegQueueT<64>;
// ld(pA, qB) loads one product (column + value), put(prod) accumulates it.  The two are split so that PANEL_UNROLL loads
// are in flight per thread before the first shared-memory atomic: the compiler cannot move loads across atomics itself.
// PANEL_UNROLL = products loaded ahead of their inserts: 4 pays in the symbolic pass (8.5 vs 10.5 ms), not in the numeric one.
// PANEL_SHORT: a lane walks a B segment of up to this many entries alone; above PANEL_LONG a segment goes to the CTA queue
// (1024-thread CTAs: 8 / 8, i.e. everything that is not tiny is shared -- 83 -> 64 ms; 64..512-thread CTAs: 16 / 64).
template <int PANEL_UNROLL, int PANEL_SHORT, int PANEL_LONG, class Q, class LD, class PUT>
__device__ __forceinline__ void panel_products(const int64_t *Ap, const uint32_t *Aj, const int64_t *Bp, int64_t row,
                                               const uint32_t *ptab, int np, int q0, int q1, Q *sq, LD &&ld, PUT &&put) {
    constexpr int SEGQ_CAP = Q::kCap;
    using Prod = decltype(ld((int64_t)0, (int64_t)0));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;   // nwarps: a power of two
    const int64_t pb = Ap[row], pe = Ap[row + 1];
    if (tid == 0) sq->n[0] = 0;
    __syncthreads();
    int par = 0;
    for (int64_t base = pb; base < pe; base += blockDim.x, par ^= 1) {
        const int64_t p = base + tid;
        int64_t s0 = 0; int len = 0;
        if (p < pe) {
            const uint32_t k = Aj[p];
            const int64_t bs = Bp[k];
            if (ptab) { const uint32_t *t = ptab + (int64_t)k * (np + 1); const uint32_t o0 = t[q0]; s0 = bs + o0; len = (int)(t[q1] - o0); }
            else { s0 = bs; len = (int)(Bp[k + 1] - bs); }
        }
        if (len > PANEL_LONG) {
            const int e = atomicAdd(&sq->n[par], 1);
            if (e < SEGQ_CAP) { sq->s0[e] = s0; sq->len[e] = len; sq->poff[e] = (int)(p - pb); len = 0; }   // else: the warp keeps it
        }
        if (len <= PANEL_SHORT) {
            for (int q = 0; q < len; q += PANEL_UNROLL) {
                Prod r[PANEL_UNROLL];
#pragma unroll
                for (int u = 0; u < PANEL_UNROLL; u++) if (q + u < len) r[u] = ld(p, s0 + q + u);
#pragma unroll
                for (int u = 0; u < PANEL_UNROLL; u++) if (q + u < len) put(r[u]);
            }
        }
        unsigned m = __ballot_sync(0xffffffffu, len > PANEL_SHORT);
        while (m) {
            const int l = __ffs(m) - 1; m &= m - 1;
            const int64_t ws = __shfl_sync(0xffffffffu, s0, l), wp = __shfl_sync(0xffffffffu, p, l);
            const int wl = __shfl_sync(0xffffffffu, len, l);
            for (int q = lane; q < wl; q += 32 * PANEL_UNROLL) {
                Prod r[PANEL_UNROLL];
#pragma unroll
                for (int u = 0; u < PANEL_UNROLL; u++) if (q + 32 * u < wl) r[u] = ld(wp, ws + q + 32 * u);
#pragma unroll
                for (int u = 0; u < PANEL_UNROLL; u++) if (q + 32 * u < wl) put(r[u]);
            }
        }
        __syncthreads();
        // drain the queue: 32-entry chunks dealt round-robin over the warps
        if (tid == 0) sq->n[par ^ 1] = 0;
        const int nq = sq->n[par] < SEGQ_CAP ? sq->n[par] : SEGQ_CAP;
        // 32 queue entries at a time, one per lane: a warp scan numbers their chunks, and each lane works out whether
        // its entry holds a chunk this warp owns (chunk number == warp mod nwarps) -- no walk over the whole queue
        int cum = 0;
        for (int e0 = 0; e0 < nq; e0 += 32) {
            const int e = e0 + lane;
            const int L = e < nq ? sq->len[e] : 0, nch = (L + 31) >> 5;
            int incl = nch;
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
            const int first = cum + incl - nch;                                   // number of this entry's first chunk
            const int c0 = (warp - first) & (nwarps - 1);                         // its first chunk owned by this warp
            unsigned m = __ballot_sync(0xffffffffu, c0 < nch);
            cum += __shfl_sync(0xffffffffu, incl, 31);
            while (m) {
                const int l = __ffs(m) - 1; m &= m - 1;
                const int wL = __shfl_sync(0xffffffffu, L, l), wc = __shfl_sync(0xffffffffu, c0, l);
                const int64_t es = sq->s0[e0 + l], ep = pb + sq->poff[e0 + l];
                for (int c = wc; c * 32 < wL; c += nwarps * PANEL_UNROLL) {
                    Prod r[PANEL_UNROLL];                  // PANEL_UNROLL of this warp's chunks in flight before the inserts
#pragma unroll
                    for (int u = 0; u < PANEL_UNROLL; u++) { const int q = (c + u * nwarps) * 32 + lane; if (q < wL) r[u] = ld(ep, es + q); }
#pragma unroll
                    for (int u = 0; u < PANEL_UNROLL; u++) { const int q = (c + u * nwarps) * 32 + lane; if (q < wL) put(r[u]); }
                }
            }
        }
        __syncthreads();
    }
}

// ---- one warp per row, TABLE = 64 ----
constexpr int HW_TABLE = 64;
constexpr int HW_ROWS = 4;   // rows (warps) per CTA

template <class SR, bool NUMERIC>
__global__ void __launch_bounds__(HW_ROWS * 32) k_hash_warp(GemmArgs<typename SR::type> a) {
    using T = typename SR::type;
    __shared__ uint32_t s_keys[HW_ROWS * HW_TABLE];
    __shared__ T s_vals[HW_ROWS * HW_TABLE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r = blockIdx.x * (int64_t)HW_ROWS + warp;
    if (r >= a.nbin) return;
    const int64_t row = a.rows[r];
    uint32_t *keys = s_keys + warp * HW_TABLE;
    T *vals = s_vals + warp * HW_TABLE;
    for (int k = lane; k < HW_TABLE; k += 32) { keys[k] = HASH_EMPTY; if (NUMERIC) vals[k] = SR::Mon::identity(); }
    __syncwarp();
    for (int64_t p = a.Ap[row], pe = a.Ap[row + 1]; p < pe; p++) {
        const uint32_t k = a.Aj[p];
        const T av = (NUMERIC && SR::kUsesX) ? ld_val(a.Ax, a.a_iso, p) : (T)0;
        for (int64_t q = a.Bp[k] + lane, qe = a.Bp[k + 1]; q < qe; q += 32) {
            const T bv = (NUMERIC && SR::kUsesY) ? ld_val(a.Bx, a.b_iso, q) : (T)0;
            hash_insert<SR, NUMERIC>(keys, vals, HW_TABLE - 1, a.Bj[q], SR::mult(av, bv));
        }
    }
    __syncwarp();
    if (!NUMERIC) {
        int c = 0;
        for (int k = lane; k < HW_TABLE; k += 32) c += keys[k] != HASH_EMPTY;
        for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0) a.rownnz[row] = c;
    } else {
        bitonic_sort<T, true>(keys, vals, HW_TABLE, lane, 32);
        const int64_t o = a.Cp[row];
        const int cnt = (int)(a.Cp[row + 1] - o);
        for (int k = lane; k < cnt; k += 32) { a.Cj[o + k] = keys[k]; a.Cx[o + k] = vals[k]; }
    }
}

// ---- one CTA per row, table size given at launch (power of two), dynamic shared memory ----
template <class SR, bool NUMERIC>
__global__ void k_hash_block(GemmArgs<typename SR::type> a, int table) {
    using T = typename SR::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *keys = (uint32_t *)smem_raw;
    T *vals = (T *)(smem_raw + (size_t)table * sizeof(uint32_t));
    __shared__ int s_cnt;
    __shared__ int s_tmp[40];
    __shared__ SegQueueSmall s_q;
    const int tid = threadIdx.x, nthreads = blockDim.x, lane = tid & 31;
    const int64_t row = a.rows[blockIdx.x];
    for (int k = tid; k < table; k += nthreads) { keys[k] = HASH_EMPTY; if (NUMERIC) vals[k] = SR::Mon::identity(); }
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    panel_products<(NUMERIC ? 1 : 4), 16, 64>(a.Ap, a.Aj, a.Bp, row, nullptr, 1, 0, 1, &s_q, [&](int64_t p, int64_t q) { return load_product<SR, NUMERIC>(a, p, q); },
                   [&](const Product<T> &r) { hash_insert<SR, NUMERIC>(keys, vals, (uint32_t)table - 1, r.j, r.z); });
    __syncthreads();
    if (!NUMERIC) {
        int c = 0;
        for (int k = tid; k < table; k += nthreads) c += keys[k] != HASH_EMPTY;
        for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0 && c) atomicAdd(&s_cnt, c);
        __syncthreads();
        if (tid == 0) a.rownnz[row] = s_cnt;
    } else {
        // numeric rows are binned by their exact size: at most table / 2 entries
        uint16_t *slot = (uint16_t *)(smem_raw + (size_t)table * (sizeof(uint32_t) + sizeof(T)));
        uint32_t *bpos = (uint32_t *)(smem_raw + (size_t)table * (sizeof(uint32_t) + sizeof(T) + 1));
        hash_emit_sorted<T>(keys, vals, table, slot, bpos, table / 8, a.Cj, a.Cx, a.Cp[row], s_tmp);
    }
}
// dynamic shared memory of k_hash_block<SR, true>: table, slot list (table / 2 x uint16), bucket cursors (table / 8 x uint32)
inline size_t hash_block_smem(int table, size_t tsize) { return (size_t)table * (4 + tsize + 1) + (size_t)(table / 8) * 4; }

// ---- hub rows (fallback): bitmap + dense accumulator in HBM, CTAs loop over the bin ----
template <class SR, bool NUMERIC>
__global__ void k_dense_block(GemmArgs<typename SR::type> a) {
    using T = typename SR::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_scan[33];
    uint32_t *bm = a.gbitmap ? a.gbitmap + blockIdx.x * a.words : (uint32_t *)smem_raw;
    T *acc = NUMERIC ? a.gacc + blockIdx.x * a.ncols : nullptr;
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
    for (int64_t w = tid; w < a.words; w += nthreads) bm[w] = 0;
    __syncthreads();
    const int64_t chunk = (a.words + nthreads - 1) / nthreads;
    for (int64_t r = blockIdx.x; r < a.nbin; r += gridDim.x) {
        const int64_t row = a.rows[r];
        for (int64_t p = a.Ap[row] + warp, pe = a.Ap[row + 1]; p < pe; p += nwarps) {
            const uint32_t k = a.Aj[p];
            const T av = (NUMERIC && SR::kUsesX) ? ld_val(a.Ax, a.a_iso, p) : (T)0;
            for (int64_t q = a.Bp[k] + lane, qe = a.Bp[k + 1]; q < qe; q += 32) {
                const uint32_t j = a.Bj[q];
                const uint32_t bit = 1u << (j & 31);
                if (!(bm[j >> 5] & bit)) atomicOr(&bm[j >> 5], bit);
                if (NUMERIC) {
                    const T bv = SR::kUsesY ? ld_val(a.Bx, a.b_iso, q) : (T)0;
                    SR::Mon::atomic(&acc[j], SR::mult(av, bv));
                }
            }
        }
        __threadfence_block();
        __syncthreads();
        // ordered extraction: each thread owns a contiguous chunk of bitmap words
        const int64_t w0 = tid * chunk, w1 = (w0 + chunk < a.words) ? w0 + chunk : a.words;
        int c = 0;
        for (int64_t w = w0; w < w1; w++) c += __popc(bm[w]);
        int total;
        int off = block_exclusive_scan(c, &total, s_scan);
        if (!NUMERIC) {
            if (tid == 0) a.rownnz[row] = total;
            for (int64_t w = w0; w < w1; w++) bm[w] = 0;
        } else {
            int64_t o = a.Cp[row] + off;
            for (int64_t w = w0; w < w1; w++) {
                uint32_t bits = bm[w];
                bm[w] = 0;
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const uint32_t j = (uint32_t)(w * 32 + b);
                    a.Cj[o] = j; a.Cx[o] = acc[j]; acc[j] = SR::Mon::identity();
                    o++;
                }
            }
        }
        __threadfence_block();
        __syncthreads();
    }
}

// ---- heavy rows: column panels, accumulators in shared memory ----
// One 1024-thread CTA per SM around a 16 K-slot table.  (Four 256-thread CTAs per SM with 4 K-slot tables were measured
// 25 % slower on R-MAT scale 18-20, two 512-thread CTAs with 8 K-slot tables 12 % slower at scale 20 (425 vs 378 ms): more
// ranges, each re-walking the A row, cost more than the overlap of the barrier-separated phases gains.)
constexpr int PANEL_W = 16384;                  // columns per base panel = entries of the dense accumulator
constexpr int PANEL_TABLE = PANEL_W;            // hash slots per range
constexpr int PANEL_CAP = PANEL_W / 2;          // entries a range may hold: load factor 0.5 (0.375: 402 ms, 0.75: 435 ms, 0.5: 378 ms at scale 20)
constexpr int PANEL_NBK = PANEL_TABLE / 8;      // buckets of the ordered emission
constexpr int PANEL_THREADS = PANEL_W / 16;     // the dense extraction gives every thread 16 columns
constexpr int PANEL_CTAS_PER_SM = 1;
constexpr int PSYM_THREADS = 1024;              // symbolic pass: one CTA per SM around the whole-row bitmap
constexpr int PANEL_MAX_NP = 256;     // base panels (ncols <= 4 Mi); beyond that the HBM dense path runs

// ptab[k][q] for one B row per warp
static __global__ void __launch_bounds__(256) k_panel_table(const int64_t *Bp, const uint32_t *Bj, int64_t brows, int np, uint32_t *ptab) {
    const int64_t k = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (k >= brows) return;
    const int64_t bs = Bp[k], be = Bp[k + 1];
    for (int q = lane; q <= np; q += 32) {
        const uint64_t key = (uint64_t)q * PANEL_W;
        int64_t lo = bs, hi = be;
        if (q == 0) hi = lo;
        while (lo < hi) { const int64_t mid = (lo + hi) >> 1; if ((uint64_t)Bj[mid] < key) lo = mid + 1; else hi = mid; }
        ptab[k * (np + 1) + q] = (uint32_t)(lo - bs);
    }
}

__device__ __forceinline__ int warp_max(int v) { for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }

// symbolic pass of the heavy rows: entries per row, and how the numeric pass will cut the row (rowS)
static __global__ void __launch_bounds__(PSYM_THREADS, 1) k_panel_symbolic(GemmArgs<int32_t> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *bm = (uint32_t *)smem_raw;
    __shared__ int s_pc[PANEL_MAX_NP];
    __shared__ SegQueue s_q;
    __shared__ long long s_r;
    const int tid = threadIdx.x, nthreads = blockDim.x, lane = tid & 31;
    for (int64_t w = tid; w < a.words; w += nthreads) bm[w] = 0;
    const int64_t chunk = (a.words + nthreads - 1) / nthreads;
    while (true) {
        if (tid == 0) s_r = (long long)atomicAdd(a.cursor, 1ull);
        if (tid < PANEL_MAX_NP) s_pc[tid] = 0;
        __syncthreads();
        const int64_t r = s_r;
        if (r >= a.nbin) break;
        const int64_t row = a.rows[r];
        const uint32_t *Bj = a.Bj;
        panel_products<4, 8, 8>(a.Ap, a.Aj, a.Bp, row, nullptr, 1, 0, 1, &s_q, [&](int64_t, int64_t q) { return Bj[q]; },
                       [&](uint32_t j) {
                           const uint32_t bit = 1u << (j & 31);
                           if (!(bm[j >> 5] & bit)) atomicOr(&bm[j >> 5], bit);
                       });
        __syncthreads();
        // entries per base panel (PANEL_W / 32 words each); the bitmap is cleared on the way
        const int64_t w0 = tid * chunk, w1 = (w0 + chunk < a.words) ? w0 + chunk : a.words;
        constexpr int PW = PANEL_W / 32;       // bitmap words per base panel
        int c = 0; int64_t cur = w0 / PW;
        for (int64_t w = w0; w < w1; w++) {
            if ((w / PW) != cur) { if (c) atomicAdd(&s_pc[cur], c); c = 0; cur = w / PW; }
            c += __popc(bm[w]);
            bm[w] = 0;
        }
        if (c) atomicAdd(&s_pc[cur], c);
        __syncthreads();
        if (tid < 32) {
            // lane owns base panels 8*lane .. 8*lane+7; m[g] = the fullest range when ranges are 2^g base panels wide
            int cnt[8];
            for (int i = 0; i < 8; i++) cnt[i] = s_pc[lane * 8 + i];
            int m[9];
            m[0] = 0; for (int i = 0; i < 8; i++) m[0] = max(m[0], cnt[i]);
            m[1] = 0; for (int i = 0; i < 8; i += 2) m[1] = max(m[1], cnt[i] + cnt[i + 1]);
            m[2] = max(cnt[0] + cnt[1] + cnt[2] + cnt[3], cnt[4] + cnt[5] + cnt[6] + cnt[7]);
            int f = cnt[0] + cnt[1] + cnt[2] + cnt[3] + cnt[4] + cnt[5] + cnt[6] + cnt[7];
            m[3] = f;
            for (int g = 4; g < 9; g++) { f += __shfl_xor_sync(0xffffffffu, f, 1 << (g - 4)); m[g] = f; }
            const int total = m[8];
            int best = -1;
            for (int g = 8; g >= 0; g--) { const int mx = warp_max(m[g]); if (best < 0 && mx <= PANEL_CAP) best = g; }
            // ranges of one base panel always take the dense accumulator (it holds any count and needs no hashing)
            if (lane == 0) { a.rownnz[row] = total; a.rowS[row] = best < 2 ? (uint16_t)1 : (uint16_t)(1u << best); }
        }
        __syncthreads();
    }
}

// numeric pass of the heavy rows
// dynamic shared memory: hash view = keys + values (PANEL_TABLE each) + slot list (PANEL_CAP x uint16) + bucket cursors;
//                        dense view = PANEL_W values + PANEL_W presence bits
template <class T> constexpr size_t panel_row_smem() {
    constexpr size_t hash_b = (size_t)PANEL_TABLE * (4 + sizeof(T)) + (size_t)PANEL_CAP * 2 + PANEL_NBK * 4;
    constexpr size_t dense_b = (size_t)PANEL_W * sizeof(T) + PANEL_W / 8;
    return hash_b > dense_b ? hash_b : dense_b;
}
// GV: the values live in a per-CTA PANEL_TABLE-entry buffer in global memory (a.gacc; stays in L2) and are combined
// with L2 atomics (RED), keys / presence bits stay in shared memory.  Shared-memory atomics on 8-byte floating-point
// values are compare-and-swap loops; the L2 path is fire-and-forget.
template <class SR, bool GV>
__global__ void __launch_bounds__(PANEL_THREADS, PANEL_CTAS_PER_SM) k_panel_row(GemmArgs<typename SR::type> a) {
    using T = typename SR::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    // hash view
    uint32_t *keys = (uint32_t *)smem_raw;
    T *vals = GV ? a.gacc + (size_t)blockIdx.x * PANEL_TABLE : (T *)(smem_raw + (size_t)PANEL_TABLE * sizeof(uint32_t));
    uint16_t *slot = (uint16_t *)(smem_raw + (size_t)PANEL_TABLE * (sizeof(uint32_t) + sizeof(T)));
    uint32_t *bpos = (uint32_t *)(smem_raw + (size_t)PANEL_TABLE * (sizeof(uint32_t) + sizeof(T)) + (size_t)PANEL_CAP * 2);
    // dense view
    T *acc = GV ? a.gacc + (size_t)blockIdx.x * PANEL_TABLE : (T *)smem_raw;
    uint32_t *bm = (uint32_t *)(smem_raw + (size_t)PANEL_W * sizeof(T));
    __shared__ int s_tmp[40];
    __shared__ long long s_r;
    __shared__ SegQueue s_q;
    const int tid = threadIdx.x;
    while (true) {
        if (tid == 0) s_r = (long long)atomicAdd(a.cursor, 1ull);
        __syncthreads();
        const int64_t r = s_r;
        if (r >= a.nbin) break;
        const int64_t row = a.rows[r];
        const int s = a.rowS[row];
        const bool dense = (s == 1);
        int64_t out = a.Cp[row];
        for (int q0 = 0; q0 < a.np; q0 += s) {
            const int q1 = (q0 + s < a.np) ? q0 + s : a.np;
            const uint32_t c0 = (uint32_t)q0 * PANEL_W;
            const uint32_t *ptab = (q0 == 0 && q1 == a.np) ? nullptr : a.ptab;   // a range over every panel reads whole B rows
            int total;
            if (dense) {
                for (int k = tid; k < PANEL_W; k += PANEL_THREADS) acc[k] = SR::Mon::identity();
                if (tid < PANEL_W / 32) bm[tid] = 0;
                __syncthreads();
                panel_products<1, 8, 8>(a.Ap, a.Aj, a.Bp, row, ptab, a.np, q0, q1, &s_q, [&](int64_t p, int64_t q) { return load_product<SR, true>(a, p, q); },
                               [&](const Product<T> &r) {
                                   const uint32_t j = r.j - c0;
                                   const uint32_t bit = 1u << (j & 31);
                                   if (!(bm[j >> 5] & bit)) atomicOr(&bm[j >> 5], bit);
                                   SR::Mon::atomic(&acc[j], r.z);
                               });
                if (GV) __threadfence();
                __syncthreads();
                // ordered extraction: thread t owns columns 16 t .. 16 t + 15 of the panel
                uint32_t bits = (bm[tid >> 1] >> ((tid & 1) * 16)) & 0xffffu;
                int64_t o = out + block_exclusive_scan(__popc(bits), &total, s_tmp);
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    a.Cj[o] = c0 + tid * 16 + b; a.Cx[o] = ld_acc(&acc[tid * 16 + b], GV);
                    o++;
                }
            } else {
                for (int k = tid; k < PANEL_TABLE; k += PANEL_THREADS) { keys[k] = HASH_EMPTY; vals[k] = SR::Mon::identity(); }
                __syncthreads();
                panel_products<1, 8, 8>(a.Ap, a.Aj, a.Bp, row, ptab, a.np, q0, q1, &s_q, [&](int64_t p, int64_t q) { return load_product<SR, true>(a, p, q); },
                               [&](const Product<T> &r) { hash_insert<SR, true>(keys, vals, PANEL_TABLE - 1, r.j, r.z); });
                if (GV) __threadfence();
                __syncthreads();
                total = hash_emit_sorted<T, GV>(keys, vals, PANEL_TABLE, slot, bpos, PANEL_NBK, a.Cj, a.Cx, out, s_tmp);
            }
            out += total;
            __syncthreads();
        }
    }
}

// ---- masked dot product ----
template <class T> struct DotArgs {
    const uint32_t *Mrow; const uint32_t *Mj; const void *Mv; int m_vsize, m_fp; int64_t mnnz;
    const int64_t *Ap; const uint32_t *Aj; const T *Ax; int a_iso;     // A by rows
    const int64_t *Bp; const uint32_t *Bj; const T *Bx; int b_iso;     // B' by rows (columns of B)
    T *val; uint8_t *has;
    // per-row hash sets of the column indices (rows longer than DOT_HASH_MIN), for multiplies that read no values
    // (pair): off[r] .. off[r+1] is row r's power-of-two table inside tab; nullptr = not built
    const int64_t *Ah_off; const uint32_t *Ah_tab;
    const int64_t *Bh_off; const uint32_t *Bh_tab;
};

__device__ __forceinline__ bool mask_true_at(const void *mv, int vsize, int fp, int64_t i) {
    switch (vsize) {
    case 0: return true;
    case 1: return ((const uint8_t *)mv)[i] != 0;
    case 2: return ((const uint16_t *)mv)[i] != 0;
    case 4: { uint32_t b = ((const uint32_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    default: { uint64_t b = ((const uint64_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    }
}
// first position in [lo, hi) with col[pos] >= key
__device__ __forceinline__ int64_t lower_bound_u32(const uint32_t *col, int64_t lo, int64_t hi, uint32_t key) {
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (col[mid] < key) lo = mid + 1; else hi = mid; }
    return lo;
}

// ---- per-row hash sets: membership in one probe instead of a 10-15 step binary search through L2 ----
constexpr int DOT_HASH_MIN = 32;     // rows longer than this get a table of next_pow2(2 * length) slots
static __global__ void __launch_bounds__(256) k_rowhash_sizes(const int64_t *Rp, int64_t nrows, int64_t *sizes) {
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (r >= nrows) return;
    const int64_t d = Rp[r + 1] - Rp[r];
    sizes[r] = d > DOT_HASH_MIN ? (int64_t)1 << (64 - __clzll((unsigned long long)(2 * d - 1))) : 0;
}
static __global__ void __launch_bounds__(256) k_rowhash_build(const uint32_t *rowid, const uint32_t *Rj, int64_t nnz,
                                                               const int64_t *off, uint32_t *tab) {
    const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (e >= nnz) return;
    const uint32_t r = rowid[e];
    const int64_t o = off[r], size = off[r + 1] - o;
    if (!size) return;
    const uint32_t mask = (uint32_t)size - 1, key = Rj[e];
    uint32_t h = (key * 2654435761u) & mask;
    while (atomicCAS(&tab[o + h], HASH_EMPTY, key) != HASH_EMPTY) h = (h + 1) & mask;   // keys of a row are distinct
}
__device__ __forceinline__ bool rowhash_has(const uint32_t *tab, int64_t o, uint32_t mask, uint32_t key) {
    uint32_t h = (key * 2654435761u) & mask;
    while (true) {
        const uint32_t v = tab[o + h];
        if (v == key) return true;
        if (v == HASH_EMPTY) return false;
        h = (h + 1) & mask;
    }
}

// hits among sc[sp], sc[sp + stride], ... (up to U entries below se): the U first probes are issued together, the
// rare collision chains are followed afterwards -- a probe is one random L2 / HBM sector, so loads in flight are the lever
template <int U>
__device__ __forceinline__ int rowhash_hits(const uint32_t *tab, int64_t o, uint32_t mask, const uint32_t *sc, int64_t sp,
                                            int64_t se, int64_t stride) {
    uint32_t key[U], h[U], v[U];
#pragma unroll
    for (int u = 0; u < U; u++) { key[u] = (sp + u * stride < se) ? sc[sp + u * stride] : 0u; h[u] = (key[u] * 2654435761u) & mask; }
#pragma unroll
    for (int u = 0; u < U; u++) v[u] = (sp + u * stride < se) ? tab[o + h[u]] : HASH_EMPTY;
    int hits = 0;
#pragma unroll
    for (int u = 0; u < U; u++) {
        if (sp + u * stride >= se) continue;
        uint32_t hh = h[u], vv = v[u];
        while (vv != key[u] && vv != HASH_EMPTY) { hh = (hh + 1) & mask; vv = tab[o + hh]; }
        hits += vv == key[u];
    }
    return hits;
}

constexpr int DOT_THREADS = 256;
constexpr int DOT_SERIAL_MAX = 32;   // a thread handles the pair alone if the shorter row has <= this many entries

template <class SR>
__global__ void __launch_bounds__(DOT_THREADS, 8) k_dot_masked(DotArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    __shared__ int q_e[DOT_THREADS];
    __shared__ int q_n;
    if (threadIdx.x == 0) q_n = 0;
    __syncthreads();
    const int64_t e0 = blockIdx.x * (int64_t)DOT_THREADS;
    const int64_t e = e0 + threadIdx.x;
    if (e < a.mnnz) {
        if (!mask_true_at(a.Mv, a.m_vsize, a.m_fp, e)) { a.has[e] = 0; }
        else {
            const uint32_t i = a.Mrow[e], j = a.Mj[e];
            const int64_t as = a.Ap[i], ae = a.Ap[i + 1], bs = a.Bp[j], be = a.Bp[j + 1];
            const int64_t la = ae - as, lb = be - bs;
            const int64_t ls = la < lb ? la : lb;
            if (ls > DOT_SERIAL_MAX) q_e[atomicAdd(&q_n, 1)] = threadIdx.x;
            else {
                T acc = Mon::identity(); bool has = false;
                const bool a_short = la <= lb;
                const uint32_t *sc = a_short ? a.Aj : a.Bj, *lc = a_short ? a.Bj : a.Aj;
                int64_t sp = a_short ? as : bs, se = a_short ? ae : be, lo = a_short ? bs : as, le = a_short ? be : ae;
                if constexpr (!SR::kUsesX && !SR::kUsesY) {
                    // the longer row has a hash set: one probe per entry of the shorter row
                    const int64_t *hoff = a_short ? a.Bh_off : a.Ah_off;
                    if (hoff) {
                        const uint32_t lr = a_short ? j : i;
                        const int64_t ho = hoff[lr], hs = hoff[lr + 1] - ho;
                        if (hs) {
                            const uint32_t *tab = a_short ? a.Bh_tab : a.Ah_tab;
                            const T one = SR::mult((T)0, (T)0);
                            for (; sp < se; sp += 4) {
                                const int hits = rowhash_hits<4>(tab, ho, (uint32_t)hs - 1, sc, sp, se, 1);
                                for (int c = 0; c < hits; c++) { acc = has ? Mon::apply(acc, one) : one; has = true; }
                                if (Mon::kAny && has) break;
                            }
                            sp = se;       // done: skip the search loop below
                        }
                    }
                }
                for (; sp < se && lo < le; sp++) {
                    const uint32_t key = sc[sp];
                    lo = lower_bound_u32(lc, lo, le, key);
                    if (lo < le && lc[lo] == key) {
                        const int64_t pa = a_short ? sp : lo, pb = a_short ? lo : sp;
                        const T z = SR::mult(SR::kUsesX ? ld_val(a.Ax, a.a_iso, pa) : (T)0,
                                             SR::kUsesY ? ld_val(a.Bx, a.b_iso, pb) : (T)0);
                        acc = has ? Mon::apply(acc, z) : z; has = true;
                        if (Mon::kAny) break;
                        lo++;
                    }
                }
                a.val[e] = acc; a.has[e] = has;
            }
        }
    }
    __syncthreads();
    // long pairs: one warp each, lanes stride the shorter row, each lane searches the longer row
    const int nq = q_n, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int q = warp; q < nq; q += DOT_THREADS / 32) {
        const int64_t ee = e0 + q_e[q];
        const uint32_t i = a.Mrow[ee], j = a.Mj[ee];
        const int64_t as = a.Ap[i], ae = a.Ap[i + 1], bs = a.Bp[j], be = a.Bp[j + 1];
        const bool a_short = (ae - as) <= (be - bs);
        const uint32_t *sc = a_short ? a.Aj : a.Bj, *lc = a_short ? a.Bj : a.Aj;
        const int64_t ss = a_short ? as : bs, se = a_short ? ae : be, ls = a_short ? bs : as, le = a_short ? be : ae;
        T acc = Mon::identity(); bool has = false;
        bool hashed = false;
        if constexpr (!SR::kUsesX && !SR::kUsesY) {
            const int64_t *hoff = a_short ? a.Bh_off : a.Ah_off;
            if (hoff) {
                const uint32_t lr = a_short ? j : i;
                const int64_t ho = hoff[lr], hs = hoff[lr + 1] - ho;
                if (hs) {
                    hashed = true;
                    const uint32_t *tab = a_short ? a.Bh_tab : a.Ah_tab;
                    const T one = SR::mult((T)0, (T)0);
                    for (int64_t sp = ss + lane; sp < se; sp += 128) {
                        const int hits = rowhash_hits<4>(tab, ho, (uint32_t)hs - 1, sc, sp, se, 32);
                        for (int c = 0; c < hits; c++) { acc = has ? Mon::apply(acc, one) : one; has = true; }
                    }
                }
            }
        }
        // (Running 2-4 of these searches per lane in lockstep was measured SLOWER, 49 vs 34 ms at scale 20: the kernel is
        // bound by L1 wavefronts of the diverging lower levels of the search, not by the latency of one chain.)
        for (int64_t sp = ss + lane; sp < se && !hashed; sp += 32) {
            const uint32_t key = sc[sp];
            const int64_t lo = lower_bound_u32(lc, ls, le, key);
            if (lo < le && lc[lo] == key) {
                const int64_t pa = a_short ? sp : lo, pb = a_short ? lo : sp;
                const T z = SR::mult(SR::kUsesX ? ld_val(a.Ax, a.a_iso, pa) : (T)0,
                                     SR::kUsesY ? ld_val(a.Bx, a.b_iso, pb) : (T)0);
                acc = has ? Mon::apply(acc, z) : z; has = true;
            }
        }
        for (int off = 16; off; off >>= 1) {
            T oacc = __shfl_xor_sync(0xffffffffu, acc, off);
            bool ohas = __shfl_xor_sync(0xffffffffu, (int)has, off) != 0;
            if (ohas) { acc = has ? Mon::apply(acc, oacc) : oacc; has = true; }
        }
        if (lane == 0) { a.val[ee] = acc; a.has[ee] = has; }
    }
}
