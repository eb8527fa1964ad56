// grb_spgemm.cuh — kernels behind GrB_mxm.
//
// Gustavson, two-phase (symbolic sizes every output row, numeric fills it), rows binned by an upper bound
// of their products ub = min(flops(i), ncols):
//   k_hash_warp    ub <= 32     one warp per row, 64-slot shared-memory hash table
//   k_hash_block   ub <= 8192   one CTA per row, 1K / 8K / 16K-slot shared-memory hash table (up to 192 KB)
//   k_dense_block  ub  > 8192   hub rows: presence bitmap in shared memory (ncols bits) + a dense accumulator
//                               row per CTA in HBM (L2-resident); ordered extraction by popcount prefix sums
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
};

template <class T> __device__ __forceinline__ T ld_val(const T *p, int iso, int64_t k) { return p[iso ? 0 : k]; }

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
    const int tid = threadIdx.x, nthreads = blockDim.x;
    const int warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
    const int64_t row = a.rows[blockIdx.x];
    for (int k = tid; k < table; k += nthreads) { keys[k] = HASH_EMPTY; if (NUMERIC) vals[k] = SR::Mon::identity(); }
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    for (int64_t p = a.Ap[row] + warp, pe = a.Ap[row + 1]; p < pe; p += nwarps) {
        const uint32_t k = a.Aj[p];
        const T av = (NUMERIC && SR::kUsesX) ? ld_val(a.Ax, a.a_iso, p) : (T)0;
        for (int64_t q = a.Bp[k] + lane, qe = a.Bp[k + 1]; q < qe; q += 32) {
            const T bv = (NUMERIC && SR::kUsesY) ? ld_val(a.Bx, a.b_iso, q) : (T)0;
            hash_insert<SR, NUMERIC>(keys, vals, (uint32_t)table - 1, a.Bj[q], SR::mult(av, bv));
        }
    }
    __syncthreads();
    if (!NUMERIC) {
        int c = 0;
        for (int k = tid; k < table; k += nthreads) c += keys[k] != HASH_EMPTY;
        for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
        if (lane == 0 && c) atomicAdd(&s_cnt, c);
        __syncthreads();
        if (tid == 0) a.rownnz[row] = s_cnt;
    } else {
        bitonic_sort<T, false>(keys, vals, table, tid, nthreads);
        const int64_t o = a.Cp[row];
        const int cnt = (int)(a.Cp[row + 1] - o);
        for (int k = tid; k < cnt; k += nthreads) { a.Cj[o + k] = keys[k]; a.Cx[o + k] = vals[k]; }
    }
}

// ---- hub rows: bitmap + dense accumulator, CTAs loop over the bin ----
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

// ---- masked dot product ----
template <class T> struct DotArgs {
    const uint32_t *Mrow; const uint32_t *Mj; const void *Mv; int m_vsize, m_fp; int64_t mnnz;
    const int64_t *Ap; const uint32_t *Aj; const T *Ax; int a_iso;     // A by rows
    const int64_t *Bp; const uint32_t *Bj; const T *Bx; int b_iso;     // B' by rows (columns of B)
    T *val; uint8_t *has;
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

constexpr int DOT_THREADS = 256;
constexpr int DOT_SERIAL_MAX = 32;   // a thread handles the pair alone if the shorter row has <= this many entries

template <class SR>
__global__ void __launch_bounds__(DOT_THREADS) k_dot_masked(DotArgs<typename SR::type> a) {
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
        for (int64_t sp = ss + lane; sp < se; sp += 32) {
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
