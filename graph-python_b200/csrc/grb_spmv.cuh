// grb_spmv.cuh — the vector-product kernels behind GrB_mxv / GrB_vxm.
//
//   k_spmv_pull      streaming pull: every CTA owns one contiguous tile of TILE_NNZ matrix entries, reads
//                    indices/values fully coalesced exactly once, gathers u (L2-resident), stages the
//                    products in shared memory, reduces the rows that start in the tile and runs the fused
//                    write-back w<m,replace> = accum(w, t) for them.  Rows that cross tile borders leave a
//                    partial in carry[]; k_spmv_fixup (one warp per tile) finishes those rows.
//                    HBM-bound: algorithmic bytes = nnz*(4 + V + X) + n*(8 + Y) (+ mask / accum reads).
//   k_pull_any       masked pull with early exit for the ANY monoid (BFS bottom-up): one thread per
//                    allowed row stops at its first parent; bytes = edges actually scanned.
//   k_push_scatter   sparse-frontier push: edge-balanced (load-balanced search over the frontier's degree
//                    prefix sums), monoid atomics into a dense temporary, then k_vec_writeback.
//
// Operand convention inside the kernels: z = mult(a, x) with a = matrix value, x = vector value.  GrB_vxm
// (z = mult(u, A)) is mapped onto it by swapping FIRST<->SECOND at dispatch (all other built-in multiplies
// commute), so no extra template parameter is needed.
#pragma once
#include "grb_internal.cuh"
#include "grb_ops.cuh"

constexpr int SPMV_THREADS = 256;
constexpr int SPMV_PER_THREAD = 8;
constexpr int SPMV_TILE = SPMV_THREADS * SPMV_PER_THREAD;   // 2048 entries per CTA

// read-only load: LDG.NC where the type has an intrinsic, plain load for bool
template <class T> __device__ __forceinline__ T ld_ro(const T *p) {
    if constexpr (std::is_same<T, bool>::value) return *p;
    else return __ldg(p);
}

// ---- fused write-back ----
template <class T> struct EpiArgs {
    T *w_out; uint8_t *wp_out;            // destination buffers (== w_in for in-place, else the ping-pong spare)
    const T *w_in; const uint8_t *wp_in;  // current content of w
    const uint8_t *mask_p;                // mask presence bytes, nullptr = no mask
    const void *mask_v;                   // mask values (value masks only)
    int mask_vsize;                       // 0 = structural; else element size of the mask values
    int mask_fp;                          // mask values are floating point (-0.0 is false)
    int comp, replace;
    int accum;                            // op code or -1
    unsigned long long *count;            // device nvals counter of w (zeroed before launch)
};

__device__ __forceinline__ bool mask_value_true(const void *mv, int vsize, int fp, int64_t i) {
    switch (vsize) {
    case 1: return ((const uint8_t *)mv)[i] != 0;
    case 2: return ((const uint16_t *)mv)[i] != 0;
    case 4: { uint32_t b = ((const uint32_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    default: { uint64_t b = ((const uint64_t *)mv)[i]; return fp ? (b << 1) != 0 : b != 0; }
    }
}
template <class T> __device__ __forceinline__ bool epi_allowed(const EpiArgs<T> &e, int64_t i) {
    bool m = true;
    if (e.mask_p) {
        m = e.mask_p[i] != 0;
        if (m && e.mask_vsize) m = mask_value_true(e.mask_v, e.mask_vsize, e.mask_fp, i);
    }
    return m != (e.comp != 0);
}
// C API 2.0 write-back for one position; returns the final presence (for the nvals counter)
template <class T>
__device__ __forceinline__ int epi_store(const EpiArgs<T> &e, int64_t i, bool allowed, bool t_has, T t) {
    bool w_has = e.wp_in[i] != 0;
    bool inplace = (e.w_out == e.w_in);
    if (allowed) {
        if (t_has) {
            T v = t;
            if (e.accum >= 0 && w_has) v = binop_rt<T>(e.accum, e.w_in[i], t);
            e.w_out[i] = v; e.wp_out[i] = 1;
            return 1;
        }
        if (e.accum < 0) { e.wp_out[i] = 0; return 0; }     // Z = T has no entry here: delete
    } else if (e.replace) { e.wp_out[i] = 0; return 0; }
    if (!inplace) { e.wp_out[i] = w_has; if (w_has) e.w_out[i] = e.w_in[i]; }
    return w_has ? 1 : 0;
}
__device__ __forceinline__ void block_count_add(unsigned long long *count, int c) {
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, (unsigned long long)c);
}

// ---- pull ----
template <class T> struct PullArgs {
    const int64_t *rowptr; const uint32_t *col; const T *av; int a_iso;
    const T *u; const uint8_t *up;        // up == nullptr: u is full
    int64_t nrows, nnz, ntiles;
    const int32_t *tile_row;
    T *carry_head; uint8_t *carry_head_has;   // [ntiles] partial of the row that entered the tile from the left
    T *carry_own; uint8_t *carry_own_has;     // [ntiles] partial of the tile's last owned row when it leaves the tile
    EpiArgs<T> epi;
};

template <class SR>
__global__ void __launch_bounds__(SPMV_THREADS) k_spmv_pull(PullArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    __shared__ T prod[SPMV_TILE];
    __shared__ uint8_t flag[SPMV_TILE];
    __shared__ int q_rows[64];
    __shared__ int q_n;
    const int tid = threadIdx.x;
    const int64_t t = blockIdx.x;
    const int64_t t0 = t * (int64_t)SPMV_TILE;
    const int64_t t1 = (t0 + SPMV_TILE < a.nnz) ? t0 + SPMV_TILE : a.nnz;
    const bool full = (a.up == nullptr);
    if (tid == 0) q_n = 0;

    // phase 1: coalesced stream of the tile, gather u, multiply
    uint32_t c[SPMV_PER_THREAD];
    T av[SPMV_PER_THREAD];
#pragma unroll
    for (int k = 0; k < SPMV_PER_THREAD; k++) {
        int64_t p = t0 + k * SPMV_THREADS + tid;
        c[k] = (p < t1) ? __ldg(a.col + p) : 0u;
        if (SR::kUsesX && !a.a_iso) av[k] = (p < t1) ? ld_ro(a.av + p) : (T)0;
    }
    if (SR::kUsesX && a.a_iso) {
        T v = a.av[0];
#pragma unroll
        for (int k = 0; k < SPMV_PER_THREAD; k++) av[k] = v;
    }
#pragma unroll
    for (int k = 0; k < SPMV_PER_THREAD; k++) {
        int64_t p = t0 + k * SPMV_THREADS + tid;
        if (p < t1) {
            T x = SR::kUsesY ? ld_ro(a.u + c[k]) : (T)0;
            prod[k * SPMV_THREADS + tid] = SR::mult(SR::kUsesX ? av[k] : (T)0, x);
            if (!full) flag[k * SPMV_THREADS + tid] = __ldg(a.up + c[k]);
        }
    }
    __syncthreads();

    // phase 2: rows.  item -1 = head partial (row r_first-1 continuing from the left), items 0..nr-1 = owned rows
    const int64_t r_first = a.tile_row[t], r_last = a.tile_row[t + 1];
    const int nr = (int)(r_last - r_first);
    const int64_t head_end = (r_first < a.nrows) ? a.rowptr[r_first] : a.nnz;  // entries [t0, head_end) are the head
    int my_count = 0;
    for (int it = tid - 1; it < nr; it += SPMV_THREADS) {
        int64_t s, e;
        if (it < 0) { s = t0; e = head_end < t1 ? head_end : t1; if (e <= s) continue; }
        else { s = a.rowptr[r_first + it]; e = a.rowptr[r_first + it + 1]; if (e > t1) e = t1; }
        int len = (int)(e - s);
        if (len > 64) {   // long segment: hand to the warps
            int q = atomicAdd(&q_n, 1);
            if (q < 64) { q_rows[q] = it; continue; }
        }
        T acc = Mon::identity(); bool has = false;
        int o = (int)(s - t0);
        for (int k = 0; k < len; k++) {
            bool f = full || flag[o + k];
            if (f) { acc = has ? Mon::apply(acc, prod[o + k]) : prod[o + k]; has = true; if (Mon::kAny) break; }
        }
        if (it < 0) { a.carry_head[t] = acc; a.carry_head_has[t] = has; }
        else {
            int64_t r = r_first + it;
            if (a.rowptr[r + 1] > t1) { a.carry_own[t] = acc; a.carry_own_has[t] = has; }
            else my_count += epi_store(a.epi, r, epi_allowed(a.epi, r), has, acc);
        }
    }
    __syncthreads();
    // long segments: one warp each, lanes stride the segment, shuffle-reduce
    const int nq = q_n < 64 ? q_n : 64;
    const int warp = tid >> 5, lane = tid & 31;
    for (int q = warp; q < nq; q += SPMV_THREADS / 32) {
        int it = q_rows[q];
        int64_t s, e;
        if (it < 0) { s = t0; e = head_end < t1 ? head_end : t1; }
        else { s = a.rowptr[r_first + it]; e = a.rowptr[r_first + it + 1]; if (e > t1) e = t1; }
        int len = (int)(e - s), o = (int)(s - t0);
        T acc = Mon::identity(); bool has = false;
        for (int k = lane; k < len; k += 32) {
            bool f = full || flag[o + k];
            if (f) { acc = has ? Mon::apply(acc, prod[o + k]) : prod[o + k]; has = true; }
        }
        for (int off = 16; off; off >>= 1) {
            T oacc = __shfl_xor_sync(0xffffffffu, acc, off);
            bool ohas = __shfl_xor_sync(0xffffffffu, (int)has, off) != 0;
            if (ohas) { acc = has ? Mon::apply(acc, oacc) : oacc; has = true; }
        }
        if (lane == 0) {
            if (it < 0) { a.carry_head[t] = acc; a.carry_head_has[t] = has; }
            else {
                int64_t r = r_first + it;
                if (a.rowptr[r + 1] > t1) { a.carry_own[t] = acc; a.carry_own_has[t] = has; }
                else my_count += epi_store(a.epi, r, epi_allowed(a.epi, r), has, acc);
            }
        }
    }
    block_count_add(a.epi.count, my_count);
}

// one warp per tile: if the tile's last owned row runs past the tile, fold the following tiles' head partials
template <class SR>
__global__ void k_spmv_fixup(PullArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    const int64_t t = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    int my_count = 0;
    if (t < a.ntiles) {
        const int64_t r_first = a.tile_row[t], r_last = a.tile_row[t + 1];
        const int64_t t1 = (t + 1) * (int64_t)SPMV_TILE;
        if (r_last > r_first) {
            const int64_t r = r_last - 1, rend = a.rowptr[r + 1];
            if (rend > t1) {
                const int64_t t_end = (rend - 1) / SPMV_TILE;   // last tile holding entries of row r
                T acc = Mon::identity(); bool has = false;
                for (int64_t k = t + 1 + lane; k <= t_end; k += 32)
                    if (a.carry_head_has[k]) { acc = has ? Mon::apply(acc, a.carry_head[k]) : a.carry_head[k]; has = true; }
                for (int off = 16; off; off >>= 1) {
                    T oacc = __shfl_xor_sync(0xffffffffu, acc, off);
                    bool ohas = __shfl_xor_sync(0xffffffffu, (int)has, off) != 0;
                    if (ohas) { acc = has ? Mon::apply(acc, oacc) : oacc; has = true; }
                }
                if (lane == 0) {
                    if (a.carry_own_has[t]) { acc = has ? Mon::apply(a.carry_own[t], acc) : a.carry_own[t]; has = true; }
                    my_count = epi_store(a.epi, r, epi_allowed(a.epi, r), has, acc);
                }
            }
        }
    }
    block_count_add(a.epi.count, my_count);
}

// ---- masked pull with early exit (ANY monoid): one thread per row, stop at the first hit ----
template <class SR>
__global__ void __launch_bounds__(256) k_pull_any(PullArgs<typename SR::type> a) {
    using T = typename SR::type;
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int my_count = 0;
    if (r < a.nrows) {
        bool allowed = epi_allowed(a.epi, r);
        bool has = false; T acc = (T)0;
        if (allowed) {
            const bool full = (a.up == nullptr);
            for (int64_t p = a.rowptr[r], e = a.rowptr[r + 1]; p < e; p++) {
                uint32_t c = __ldg(a.col + p);
                if (full || __ldg(a.up + c)) {
                    T av = SR::kUsesX ? a.av[a.a_iso ? 0 : p] : (T)0;
                    T x = SR::kUsesY ? a.u[c] : (T)0;
                    acc = SR::mult(av, x); has = true;
                    break;
                }
            }
        }
        my_count = epi_store(a.epi, r, allowed, has, acc);
    }
    block_count_add(a.epi.count, my_count);
}

// ---- dense write-back w<m,replace> = accum(w, t) (after push, or when types had to be cast) ----
template <class T>
__global__ void __launch_bounds__(256) k_vec_writeback(EpiArgs<T> e, const T *t, const uint8_t *tp, int64_t n) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    int my_count = 0;
    if (i < n) my_count = epi_store(e, i, epi_allowed(e, i), tp[i] != 0, t[i]);
    block_count_add(e.count, my_count);
}

// ---- push ----
template <class T>
__global__ void k_fill_identity(T *t, uint8_t *tp, int64_t n, T ident) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i < n) { t[i] = ident; tp[i] = 0; }
}
static __global__ void k_frontier_degrees(const uint32_t *list, int64_t nf, const int64_t *rowptr, int64_t *deg) {
    const int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (k < nf) { uint32_t v = list[k]; deg[k] = rowptr[v + 1] - rowptr[v]; }
    else if (k == nf) deg[k] = 0;
}

constexpr int PUSH_THREADS = 256;
constexpr int PUSH_PER_THREAD = 8;
constexpr int PUSH_CHUNK = PUSH_THREADS * PUSH_PER_THREAD;

template <class T> struct PushArgs {
    const int64_t *rowptr; const uint32_t *col; const T *av; int a_iso;   // rows indexed by u's index
    const T *u;                                                            // dense values of u
    const uint32_t *list; int64_t nf;                                      // frontier (sorted indices of u)
    const int64_t *offs;                                                   // exclusive degree prefix sums [nf+1]
    int64_t nedges;
    T *t; uint8_t *tp;                                                     // dense temporary (identity-filled)
    const uint8_t *mask_p; int comp;                                       // optional structural pre-filter
};

// edge-balanced scatter: chunk of PUSH_CHUNK frontier edges per CTA; owner of each edge by binary search in offs
template <class SR>
__global__ void __launch_bounds__(PUSH_THREADS) k_push_scatter(PushArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    __shared__ int64_t s_lo, s_hi;
    const int64_t e0 = blockIdx.x * (int64_t)PUSH_CHUNK;
    const int64_t e1 = (e0 + PUSH_CHUNK < a.nedges) ? e0 + PUSH_CHUNK : a.nedges;
    if (threadIdx.x == 0) {
        // frontier range whose edges intersect [e0, e1): last k with offs[k] <= e0 ... first k with offs[k] >= e1
        int64_t lo = 0, hi = a.nf;
        while (lo < hi) { int64_t mid = (lo + hi + 1) >> 1; if (a.offs[mid] <= e0) lo = mid; else hi = mid - 1; }
        s_lo = lo;
        int64_t lo2 = lo, hi2 = a.nf;
        while (lo2 < hi2) { int64_t mid = (lo2 + hi2) >> 1; if (a.offs[mid] >= e1) hi2 = mid; else lo2 = mid + 1; }
        s_hi = lo2;
    }
    __syncthreads();
    const int64_t klo = s_lo, khi = s_hi;
#pragma unroll 2
    for (int j = 0; j < PUSH_PER_THREAD; j++) {
        const int64_t e = e0 + j * PUSH_THREADS + threadIdx.x;
        if (e >= e1) break;
        int64_t lo = klo, hi = khi;   // last k in [klo, khi] with offs[k] <= e
        while (lo < hi) { int64_t mid = (lo + hi + 1) >> 1; if (__ldg(a.offs + mid) <= e) lo = mid; else hi = mid - 1; }
        const uint32_t src = __ldg(a.list + lo);
        const int64_t p = a.rowptr[src] + (e - __ldg(a.offs + lo));
        const uint32_t dst = __ldg(a.col + p);
        if (a.mask_p && ((a.mask_p[dst] != 0) == (a.comp != 0))) continue;   // structural mask says "not writable"
        T av = SR::kUsesX ? a.av[a.a_iso ? 0 : p] : (T)0;
        T x = SR::kUsesY ? a.u[src] : (T)0;
        Mon::atomic(a.t + dst, SR::mult(av, x));
        a.tp[dst] = 1;
    }
}
