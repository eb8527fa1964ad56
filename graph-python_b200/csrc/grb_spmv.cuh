// grb_spmv.cuh — the vector-product kernels behind GrB_mxv / GrB_vxm.
//
//   k_spmv_pull      persistent streaming pull: one CTA per SM, every warp walks tiles of 256 consecutive matrix
//                    entries, reads indices/values coalesced exactly once, gathers u (the most referenced "hub"
//                    entries from a shared-memory table, the rest from L2), stages the products in its private
//                    shared-memory strip, reduces the rows that start in the tile and runs the fused write-back
//                    w<m,replace> = accum(w, t) for them.  Rows that cross tile borders leave a partial in
//                    carry[]; k_spmv_fixup (one warp per tile) finishes those rows.
//                    Algorithmic bytes = nnz*(4 + V + X) + n*(8 + Y) (+ mask / accum reads).
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


// read-only load: LDG.NC where the type has an intrinsic, plain load for bool
template <class T> __device__ __forceinline__ T ld_ro(const T *p) {
    if constexpr (std::is_same<T, bool>::value) return *p;
    else return __ldg(p);
}

// shared-space load from a 32-bit shared address: keeps the compiler from re-deriving the shared window base
// (S2R SR_CgaCtaId + LEA per access under register pressure, seen in the sm_100a SASS of the pull kernels)
template <class T> __device__ __forceinline__ T lds_at(uint32_t addr) {
    if constexpr (sizeof(T) == 1) { uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr)); uint8_t b = (uint8_t)v; T o; memcpy(&o, &b, 1); return o; }
    else if constexpr (sizeof(T) == 2) { uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr)); T o; memcpy(&o, &v, 2); return o; }
    else if constexpr (sizeof(T) == 4) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr)); T o; memcpy(&o, &v, 4); return o; }
    else { unsigned long long v; asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr)); T o; memcpy(&o, &v, 8); return o; }
}
// pin a value in a register: the compiler may not rematerialise it at each use
__device__ __forceinline__ uint32_t pinned(uint32_t x) { uint32_t y; asm volatile("mov.u32 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint4 lds128_at(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// (value, presence) pairs of a sparse vector, 2 * sizeof(T) bytes each: one gather (one sector) fetches both
template <int BYTES> struct PairBits;
template <> struct PairBits<2> { using type = uint16_t; };
template <> struct PairBits<4> { using type = uint32_t; };
template <> struct PairBits<8> { using type = unsigned long long; };
template <> struct PairBits<16> { using type = ulonglong2; };
template <class T> using pair_bits_t = typename PairBits<2 * sizeof(T)>::type;
template <class T> __device__ __forceinline__ pair_bits_t<T> make_pair_bits(T v, bool present) {
    pair_bits_t<T> out;
    unsigned char raw[2 * sizeof(T)] = {};
    memcpy(raw, &v, sizeof(T));
    raw[sizeof(T)] = present ? 1 : 0;
    memcpy(&out, raw, sizeof(out));
    return out;
}
template <class T> __device__ __forceinline__ bool split_pair_bits(pair_bits_t<T> p, T &v) {
    unsigned char raw[2 * sizeof(T)];
    memcpy(raw, &p, sizeof(p));
    memcpy(&v, raw, sizeof(T));
    return raw[sizeof(T)] != 0;
}
template <class T> __device__ __forceinline__ pair_bits_t<T> ld_pair(const pair_bits_t<T> *p) {
    if constexpr (sizeof(T) == 8) {
        ulonglong2 v;
        asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
        return v;
    } else return __ldg(p);
}
template <class T>
__global__ void __launch_bounds__(256) k_pack_pairs(const T *u, const uint8_t *up, int64_t n, pair_bits_t<T> *out) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = make_pair_bits<T>(u[i], up[i] != 0);
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
    // w's old presence is read only on the paths that need it.  Reading it up front made an in-place masked write-back
    // (`t(~v.S, replace) << q.vxm(A)`, t != q) 2.7x slower than the same product written to a spare buffer: every warp loaded a
    // line of w's presence bytes that neighbouring warps were storing to (same instructions, same DRAM bytes, three times the
    // load latency in ncu, profiles/r02_bfs_pull_any_inplace.md).
    if (allowed) {
        if (t_has) {
            T v = t;
            if (e.accum >= 0 && e.wp_in[i] != 0) v = binop_rt<T>(e.accum, e.w_in[i], t);
            e.w_out[i] = v; e.wp_out[i] = 1;
            return 1;
        }
        if (e.accum < 0) { e.wp_out[i] = 0; return 0; }     // Z = T has no entry here: delete
    } else if (e.replace) { e.wp_out[i] = 0; return 0; }
    const bool w_has = e.wp_in[i] != 0;
    if (e.w_out != e.w_in) { e.wp_out[i] = w_has; if (w_has) e.w_out[i] = e.w_in[i]; }
    return w_has ? 1 : 0;
}
// one global atomic per CTA (every thread of the CTA must call it): same-address atomics from every warp of a
// large grid serialise in L2 and used to cost more than the kernels themselves
// same, with w's old state (presence, value) fetched early by the caller so the loads overlap other work
template <class T>
__device__ __forceinline__ int epi_store_pre(const EpiArgs<T> &e, int64_t i, bool allowed, T t, bool w_has, T w_old) {
    if (allowed) {
        T v = t;
        if (e.accum >= 0 && w_has) v = binop_rt<T>(e.accum, w_old, t);
        e.w_out[i] = v; e.wp_out[i] = 1;
        return 1;
    }
    if (e.replace) { e.wp_out[i] = 0; return 0; }
    if (e.w_out != e.w_in) { e.wp_out[i] = w_has; if (w_has) e.w_out[i] = w_old; }
    return w_has ? 1 : 0;
}
// general form with everything pre-loaded (element-wise kernels batch the loads of several elements first)
template <class T>
__device__ __forceinline__ int epi_apply(const EpiArgs<T> &e, int64_t i, bool allowed, bool t_has, T t, bool w_has, T w_old) {
    if (allowed) {
        if (t_has) {
            T v = t;
            if (e.accum >= 0 && w_has) v = binop_rt<T>(e.accum, w_old, t);
            e.w_out[i] = v; e.wp_out[i] = 1;
            return 1;
        }
        if (e.accum < 0) { e.wp_out[i] = 0; return 0; }
    } else if (e.replace) { e.wp_out[i] = 0; return 0; }
    if (e.w_out != e.w_in) { e.wp_out[i] = w_has; if (w_has) e.w_out[i] = w_old; }
    return w_has ? 1 : 0;
}
template <class T> __device__ __forceinline__ bool epi_needs_old(const EpiArgs<T> &e) { return e.accum >= 0 || e.w_out != e.w_in; }
// does the write-back of one position look at w's old presence / value at all?  An allowed position without accumulator is
// overwritten or deleted, a masked-out one under `replace` is deleted: neither reads w (see epi_store on why that matters)
template <class T> __device__ __forceinline__ bool epi_reads_old(const EpiArgs<T> &e, bool allowed) { return allowed ? e.accum >= 0 : !e.replace; }
constexpr int EW_VEC = 4;            // elements per thread per pass of the element-wise kernels (loads batched)
static inline int ew_grid4(int64_t n) { int64_t b = (n + 256 * EW_VEC - 1) / (256 * EW_VEC); return (int)(b < 1 ? 1 : b); }
__device__ __forceinline__ void block_count_add(unsigned long long *count, int c) {
    __shared__ int s_block_count;
    if (count == nullptr) return;           // the caller knows the result's nvals analytically (uniform for the grid)
    if (threadIdx.x == 0) s_block_count = 0;
    __syncthreads();
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_block_count, c);
    __syncthreads();
    if (threadIdx.x == 0 && s_block_count) atomicAdd(count, (unsigned long long)s_block_count);
}
// grid size of the element-wise kernels: grid-stride loops over at most 8 CTAs per SM
static inline int ew_grid(int64_t n) {
    int64_t b = (n + 255) / 256;
    if (b < 1) b = 1;
    return (int)(b < 148 * 8 ? b : 148 * 8);
}

// ---- pull ----
// Persistent, warp-autonomous streaming pull.  One CTA of SPMV_THREADS threads per SM; every warp walks tiles of
// SPMV_TILE consecutive matrix entries (static round-robin), so there is no block barrier in the main loop.
//   * the tile's indices / values are read once, fully coalesced;
//   * gathers of u are what bounds this kernel on a power-law graph (one 32 B sector through L1TEX per lane, not
//     HBM): the `hub_n` most frequently referenced entries of u (chosen once per matrix, grb_storage.cu:
//     csr_ensure_hubs) are copied to shared memory at kernel start, and the matrix carries a remapped index
//     array in which a hub column c is stored as HUB_FLAG | rank(c).  Hub lanes read shared memory and stay out
//     of the divergent global load;
//   * products are staged in the warp's private shared-memory strip (padded against bank conflicts); each lane
//     reduces the rows that start in the tile, long segments are reduced by the whole warp;
//   * rows completed inside the tile run the fused write-back w<m,replace> = accum(w, t); rows that cross tile
//     borders leave partials in carry[] and k_spmv_fixup (one warp per tile) finishes them.
constexpr int SPMV_THREADS = 1024;
constexpr int SPMV_WARPS = SPMV_THREADS / 32;
constexpr int SPMV_EPL = 8;                        // entries per lane
constexpr int SPMV_TILE = 32 * SPMV_EPL;           // 256 entries per warp tile
constexpr int SPMV_STRIP = SPMV_TILE + SPMV_TILE / 32;   // padded strip: index e lives at e + (e >> 5)
constexpr int SPMV_LONG = 64;                      // segments longer than this are reduced by the whole warp
constexpr uint32_t HUB_FLAG = 0x80000000u;

template <class T> struct PullArgs {
    const int64_t *rowptr; const uint32_t *col; const T *av; int a_iso;   // col: hub-remapped copy when hub_n > 0
    const T *u; const uint8_t *up;        // up == nullptr: u is full
    const T *hub_v; const uint8_t *hub_p; int hub_n;   // compact copy of u at the hub columns (this call's values)
    int64_t nrows, nnz, ntiles;
    const int32_t *tile_row;
    T *carry_head; uint8_t *carry_head_has;   // [ntiles] partial of the row that entered the tile from the left
    T *carry_own; uint8_t *carry_own_has;     // [ntiles] partial of the tile's last owned row when it leaves the tile
    EpiArgs<T> epi;
};
// dynamic shared memory of k_spmv_pull for a hub table of h entries
template <class T> static inline size_t spmv_pull_smem(int h, bool full) {
    size_t b = (size_t)h * sizeof(T);
    b = (b + 15) & ~(size_t)15;
    b += (size_t)SPMV_WARPS * SPMV_STRIP * sizeof(T);
    if (!full) b += (size_t)h + (size_t)SPMV_WARPS * SPMV_STRIP;
    return b + 16;
}
template <class T> static inline int spmv_hub_capacity() {   // hub entries that fit next to the strips (both layouts)
    const size_t budget = 216 * 1024;
    const size_t fixed = (size_t)SPMV_WARPS * SPMV_STRIP * (sizeof(T) + 1) + 64;
    return (int)(((budget - fixed) / (sizeof(T) + 1)) & ~(size_t)1023);
}

__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
template <class T> __device__ __forceinline__ T ld_stream(const T *p) {
    if constexpr (sizeof(T) == 4) {
        uint32_t v;
        asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
        return *reinterpret_cast<T *>(&v);
    } else if constexpr (sizeof(T) == 8) {
        unsigned long long v;
        asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(v) : "l"(p));
        return *reinterpret_cast<T *>(&v);
    } else return *p;
}

template <class SR>
__global__ void __launch_bounds__(SPMV_THREADS, 1) k_spmv_pull(PullArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const bool full = (a.up == nullptr);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    T *hub = (T *)smem_raw;
    size_t off = ((size_t)a.hub_n * sizeof(T) + 15) & ~(size_t)15;
    T *strip = (T *)(smem_raw + off) + warp * SPMV_STRIP;
    off += (size_t)SPMV_WARPS * SPMV_STRIP * sizeof(T);
    uint8_t *hubp = smem_raw + off;                                   // only when !full
    uint8_t *fstrip = hubp + a.hub_n + warp * SPMV_STRIP;             // only when !full
    if (SR::kUsesY) for (int i = tid; i < a.hub_n; i += SPMV_THREADS) hub[i] = a.hub_v[i];
    if (!full) for (int i = tid; i < a.hub_n; i += SPMV_THREADS) hubp[i] = a.hub_p[i];
    __syncthreads();

    int my_count = 0;
    for (int64_t t = blockIdx.x * (int64_t)SPMV_WARPS + warp; t < a.ntiles; t += (int64_t)gridDim.x * SPMV_WARPS) {
        const int64_t t0 = t * (int64_t)SPMV_TILE;
        const int64_t t1 = (t0 + SPMV_TILE < a.nnz) ? t0 + SPMV_TILE : a.nnz;
        // phase 1: coalesced stream of the tile, gather u (hubs from shared memory), multiply, stage
        uint32_t c[SPMV_EPL];
        T av[SPMV_EPL];
#pragma unroll
        for (int k = 0; k < SPMV_EPL; k++) {
            const int64_t p = t0 + k * 32 + lane;
            c[k] = (p < t1) ? ld_stream_u32(a.col + p) : 0u;
            if (SR::kUsesX && !a.a_iso) av[k] = (p < t1) ? ld_stream(a.av + p) : (T)0;
        }
        if (SR::kUsesX && a.a_iso) {
            const T v = a.av[0];
#pragma unroll
            for (int k = 0; k < SPMV_EPL; k++) av[k] = v;
        }
        T x[SPMV_EPL];
        if (SR::kUsesY) {
#pragma unroll
            for (int k = 0; k < SPMV_EPL; k++) {
                const int64_t p = t0 + k * 32 + lane;
                x[k] = (T)0;
                if (p < t1 && !(c[k] & HUB_FLAG)) x[k] = ld_ro(a.u + c[k]);
            }
#pragma unroll
            for (int k = 0; k < SPMV_EPL; k++)
                if (c[k] & HUB_FLAG) x[k] = hub[c[k] & ~HUB_FLAG];
        }
#pragma unroll
        for (int k = 0; k < SPMV_EPL; k++) {
            const int e = k * 32 + lane;
            const int64_t p = t0 + e;
            if (p < t1) {
                strip[e + k] = SR::mult(SR::kUsesX ? av[k] : (T)0, SR::kUsesY ? x[k] : (T)0);   // e >> 5 == k
                if (!full) fstrip[e + k] = (c[k] & HUB_FLAG) ? hubp[c[k] & ~HUB_FLAG] : __ldg(a.up + c[k]);
            }
        }
        __syncwarp();

        // phase 2: rows.  item -1 = head partial (row r_first-1 continuing from the left), items 0..nr-1 = owned rows
        const int64_t r_first = a.tile_row[t], r_last = a.tile_row[t + 1];
        const int nr = (int)(r_last - r_first);
        const int64_t head_end = (r_first < a.nrows) ? a.rowptr[r_first] : a.nnz;   // entries [t0, head_end) are the head
        for (int base = -1; base < nr; base += 32) {
            const int it = base + lane;
            int64_t s = 0, e = 0;
            bool valid = it < nr;
            if (valid) {
                if (it < 0) { s = t0; e = head_end < t1 ? head_end : t1; valid = e > s; }
                else { s = a.rowptr[r_first + it]; e = a.rowptr[r_first + it + 1]; if (e > t1) e = t1; }
            }
            const int len = (int)(e - s), o = (int)(s - t0);
            const bool is_long = valid && len > SPMV_LONG;
            T acc = Mon::identity(); bool has = false;
            if (valid && !is_long) {
                for (int k = 0; k < len; k++) {
                    const int q = o + k, qq = q + (q >> 5);
                    if (full || fstrip[qq]) { acc = has ? Mon::apply(acc, strip[qq]) : strip[qq]; has = true; if (Mon::kAny) break; }
                }
            }
            // long segments: the whole warp strides the segment and shuffle-reduces; the owner lane keeps the result
            unsigned lm = __ballot_sync(0xffffffffu, is_long);
            while (lm) {
                const int src = __ffs(lm) - 1;
                lm &= lm - 1;
                const int lo = __shfl_sync(0xffffffffu, o, src), ll = __shfl_sync(0xffffffffu, len, src);
                T pacc = Mon::identity(); bool phas = false;
                for (int k = lane; k < ll; k += 32) {
                    const int q = lo + k, qq = q + (q >> 5);
                    if (full || fstrip[qq]) { pacc = phas ? Mon::apply(pacc, strip[qq]) : strip[qq]; phas = true; }
                }
                for (int sh = 16; sh; sh >>= 1) {
                    const T oacc = __shfl_xor_sync(0xffffffffu, pacc, sh);
                    const bool ohas = __shfl_xor_sync(0xffffffffu, (int)phas, sh) != 0;
                    if (ohas) { pacc = phas ? Mon::apply(pacc, oacc) : oacc; phas = true; }
                }
                if (lane == src) { acc = pacc; has = phas; }
            }
            if (valid) {
                if (it < 0) { a.carry_head[t] = acc; a.carry_head_has[t] = has; }
                else {
                    const int64_t r = r_first + it;
                    if (a.rowptr[r + 1] > t1) { a.carry_own[t] = acc; a.carry_own_has[t] = has; }
                    else my_count += epi_store(a.epi, r, epi_allowed(a.epi, r), has, acc);
                }
            }
        }
        __syncwarp();
    }
    block_count_add(a.epi.count, my_count);
}

// one warp per tile: if the tile's last owned row runs past the tile, fold the following tiles' head partials
template <class SR>
__global__ void __launch_bounds__(256) k_spmv_fixup(PullArgs<typename SR::type> a) {
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

// ---- flagged-stream pull (u full): the fast path of PageRank / SSSP / FastSV iterations ----
// Works on the cached stream form of the matrix (grb_storage.cu: csr_ensure_hubs): pcol carries, per entry, the
// column (or 0x80000000 | hub rank) and the bit 0x40000000 on the first entry of every row; nzrow lists the rows
// that have entries; tile_first[t] counts the rows that start before tile t.  Every lane owns 8 CONSECUTIVE
// entries (one 256-bit load per array), reduces them in registers between row-start flags, and a 5-step
// segmented suffix scan over the lanes closes the rows that cross lanes.  Finished row sums are compacted into
// the warp's shared-memory strip in row order and written back, coalesced, through the fused epilogue.  A row
// left open at the tile end is finished by its owner if it ends within the next 32 entries (one extra coalesced
// load); longer rows leave carry_own / carry_head partials for k_stream_fixup.  About 1 warp instruction per
// entry, against 3.3 for the staged kernel below.
constexpr uint32_t ROW_FLAG = 0x40000000u;
constexpr uint32_t COL_MASK = 0x3fffffffu;
constexpr int ST_TILE = 256;
constexpr int ST_EPL = 8;
constexpr int RED_WARPS_8B = 24;      // 8-byte types: 24 warps x 85 registers (32 would spill, 16 leaves the LSU idle)
constexpr int ST_LOOK = 32;

template <class T> struct StreamArgs {
    const int64_t *rowptr; const uint32_t *pcol; const T *av; int a_iso;
    const T *u;
    const T *hub_v; int hub_n;
    const uint32_t *nzrow; const int64_t *tile_first;
    int64_t nrows, nnz, ntiles;
    T *carry_head; T *carry_own;          // [ntiles]
    int32_t *open_list; int *open_count;  // tiles whose last row stays open past the look-ahead window (k_stream_fixup)
    EpiArgs<T> epi;
};
template <class T> static inline int stream_warps() { return sizeof(T) > 4 ? 16 : 32; }
template <class T> static inline size_t stream_smem(int h) {
    return (((size_t)h * sizeof(T) + 15) & ~(size_t)15) + (size_t)stream_warps<T>() * ST_TILE * sizeof(T) + 16;
}
template <class T> static inline int stream_hub_capacity() {
    const size_t budget = 216 * 1024 - (size_t)stream_warps<T>() * ST_TILE * sizeof(T);
    return (int)((budget / sizeof(T)) & ~(size_t)1023);
}

__device__ __forceinline__ void ld256_u32(const uint32_t *p, uint32_t (&v)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(p));
}
template <class T> __device__ __forceinline__ void ld_tile_vals(const T *p, T (&v)[8]) {
    if constexpr (sizeof(T) == 4) {
        uint32_t r[8];
        ld256_u32((const uint32_t *)p, r);
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = *reinterpret_cast<T *>(&r[k]);
    } else if constexpr (sizeof(T) == 8) {
        unsigned long long r[8];
        asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r[0]), "=l"(r[1]), "=l"(r[2]), "=l"(r[3]) : "l"(p));
        asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r[4]), "=l"(r[5]), "=l"(r[6]), "=l"(r[7]) : "l"(p + 4));
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = *reinterpret_cast<T *>(&r[k]);
    } else {
#pragma unroll
        for (int k = 0; k < 8; k++) v[k] = p[k];
    }
}

template <class SR, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) k_spmv_stream(StreamArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    T *hub = (T *)smem_raw;
    T *strip = (T *)(smem_raw + (((size_t)a.hub_n * sizeof(T) + 15) & ~(size_t)15)) + warp * ST_TILE;
    if (SR::kUsesY) for (int i = tid; i < a.hub_n; i += WARPS * 32) hub[i] = a.hub_v[i];
    __syncthreads();
    const unsigned FULLM = 0xffffffffu;
    const int e0 = lane * ST_EPL;
    int my_count = 0;
    const int64_t tstep = (int64_t)gridDim.x * WARPS;
    int64_t t = blockIdx.x * (int64_t)WARPS + warp;
    // tile_first of the current tile is fetched one tile ahead (it heads the dependent chain of the write-back)
    int64_t tf_prev = 0, tf_cur = 0, tf_next = 0;
    if (t < a.ntiles) { tf_prev = t > 0 ? a.tile_first[t - 1] : 0; tf_cur = a.tile_first[t]; tf_next = a.tile_first[t + 1]; }
    for (; t < a.ntiles; t += tstep) {
        const int64_t t0 = t * (int64_t)ST_TILE;
        const int tl = (a.nnz - t0 < ST_TILE) ? (int)(a.nnz - t0) : ST_TILE;
        const int64_t t1 = t0 + tl;
        const int64_t first = tf_cur;
        const int nrows_t = (int)(tf_next - tf_cur);
        const bool prev_has_flag = tf_cur > tf_prev;
        // ---- loads: 8 consecutive entries per lane; the 32-entry look-ahead goes out with them.  Entries past
        //      the end of the matrix become one-entry dummy rows (flag + hub 0) whose sums are never emitted ----
        uint32_t c[ST_EPL];
        T av[ST_EPL];
        uint32_t la_c = 0;
        const bool la_valid = (tl == ST_TILE) && (t1 + lane < a.nnz);
        if (tl == ST_TILE) {
            ld256_u32(a.pcol + t0 + e0, c);
            if (SR::kUsesX && !a.a_iso) ld_tile_vals<T>(a.av + t0 + e0, av);
            if (la_valid) la_c = ld_stream_u32(a.pcol + t1 + lane);
        } else {
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) {
                const bool v = e0 + k < tl;
                c[k] = v ? a.pcol[t0 + e0 + k] : (HUB_FLAG | ROW_FLAG);
                if (SR::kUsesX && !a.a_iso) av[k] = v ? a.av[t0 + e0 + k] : (T)0;
            }
        }
        if (SR::kUsesX && a.a_iso) {
            const T v = a.av[0];
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) av[k] = v;
        }
        // the first 32 output rows of the tile: their ids, and (below) w's old state, are fetched while the
        // stream loads are in flight, so the write-back at the end of the tile does not start a new latency chain
        int64_t r_pre = 0;
        if (lane < nrows_t) r_pre = a.nzrow[first + lane];
        {   // next tile's row bookkeeping, off the critical path
            const int64_t tn = t + tstep;
            tf_prev = tf_cur;   // placeholder, fixed below
            if (tn < a.ntiles) { tf_prev = a.tile_first[tn - 1]; tf_cur = a.tile_first[tn]; tf_next = a.tile_first[tn + 1]; }
        }
        // ---- gather (hubs from shared memory) ----
        T x[ST_EPL];
        if (SR::kUsesY) {
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) {
                x[k] = (T)0;
                if ((int)c[k] >= 0) x[k] = ld_ro(a.u + (c[k] & COL_MASK));
            }
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) if ((int)c[k] < 0) x[k] = hub[c[k] & COL_MASK];
        }
        bool pre_allowed = false, pre_has = false; T pre_old = (T)0;
        if (lane < nrows_t) {
            pre_allowed = epi_allowed(a.epi, r_pre);
            pre_has = a.epi.wp_in[r_pre] != 0;
            if (a.epi.accum >= 0 || a.epi.w_out != a.epi.w_in) pre_old = a.epi.w_in[r_pre];
        }
        // ---- row-start flags -> slot of this lane's first row in the tile ----
        unsigned fbits = 0;
#pragma unroll
        for (int k = 0; k < ST_EPL; k++) fbits |= ((c[k] >> 30) & 1u) << k;
        const int cnt = __popc(fbits);
        int lane_rank = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULLM, lane_rank, d); if (lane >= d) lane_rank += o; }
        lane_rank -= cnt;
        const unsigned flagged_lanes = __ballot_sync(FULLM, fbits != 0);
        const bool seen = fbits != 0;
        // ---- lane-local segmented prefix: acc[k] = sum of the entries of k's row seen so far in this lane ----
        T acc[ST_EPL];
        acc[0] = SR::mult(SR::kUsesX ? av[0] : (T)0, SR::kUsesY ? x[0] : (T)0);
#pragma unroll
        for (int k = 1; k < ST_EPL; k++) {
            const T z = SR::mult(SR::kUsesX ? av[k] : (T)0, SR::kUsesY ? x[k] : (T)0);
            acc[k] = (fbits & (1u << k)) ? z : Mon::apply(acc[k - 1], z);
        }
        // rows that start AND end inside the lane go straight to their slot; the part before the first flag is `head`
        {
            int slot = lane_rank - 1;
#pragma unroll
            for (int k = 1; k < ST_EPL; k++) {
                slot += (fbits >> (k - 1)) & 1u;
                if ((fbits & (1u << k)) && slot >= lane_rank) strip[slot] = acc[k - 1];
            }
        }
        T head = acc[ST_EPL - 1];
#pragma unroll
        for (int k = ST_EPL - 1; k >= 1; k--) if (fbits & (1u << k)) head = acc[k - 1];
        const bool head_has = !(fbits & 1u);
        // ---- segmented suffix scan of the heads: S_i = head_i (+) (lane i has a row start ? nothing : S_{i+1}) ----
        T sv = head;
        int st = (head_has ? 1 : 0) | (seen ? 2 : 0);       // bit 0: has a value, bit 1: closed by a row start
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const T ov = __shfl_down_sync(FULLM, sv, d);
            const int os = __shfl_down_sync(FULLM, st, d);
            if (lane + d < 32 && !(st & 2)) {
                if (os & 1) { sv = (st & 1) ? Mon::apply(sv, ov) : ov; st |= 1; }
                st |= os & 2;
            }
        }
        const T nv = __shfl_down_sync(FULLM, sv, 1);
        const int ns = __shfl_down_sync(FULLM, st, 1);
        T total = acc[ST_EPL - 1];                          // the row opened by this lane's last flag
        if (lane < 31 && (ns & 1)) total = Mon::apply(total, nv);
        const int last_lane = 31 - __clz(flagged_lanes);     // -1 if the tile has no row start
        if (seen && lane != last_lane) strip[lane_rank + cnt - 1] = total;
        // ---- tile head: the part of a row that began in an earlier tile ----
        {
            int head_len = tl;
            if (flagged_lanes) {
                const int fl = __ffs(flagged_lanes) - 1;
                head_len = fl * ST_EPL + (__ffs(__shfl_sync(FULLM, fbits, fl)) - 1);
            }
            const bool owner_finishes = prev_has_flag && ((flagged_lanes && head_len < ST_LOOK) || (t0 + ST_LOOK >= a.nnz));
            if (lane == 0 && head_len > 0 && t > 0 && !owner_finishes) a.carry_head[t] = sv;
        }
        // ---- the row left open at the tile end: finish it from the look-ahead window, or leave a carry ----
        int emit = nrows_t;
        if (flagged_lanes) {
            bool closed = true; int f = 0;
            if (tl == ST_TILE && t1 < a.nnz) {
                const unsigned lm = __ballot_sync(FULLM, la_valid && (la_c & ROW_FLAG));
                if (lm) f = __ffs(lm) - 1;
                else if (t1 + ST_LOOK >= a.nnz) f = (int)(a.nnz - t1);
                else closed = false;
            }
            if (closed && f > 0) {   // warp-uniform
                T rem = Mon::identity(); int rhas = 0;
                if (lane < f) {
                    const T ra = SR::kUsesX ? a.av[a.a_iso ? 0 : t1 + lane] : (T)0;
                    T rx = (T)0;
                    if (SR::kUsesY) rx = ((int)la_c < 0) ? hub[la_c & COL_MASK] : ld_ro(a.u + (la_c & COL_MASK));
                    rem = SR::mult(ra, rx); rhas = 1;
                }
#pragma unroll
                for (int d = 16; d; d >>= 1) {
                    const T ov = __shfl_xor_sync(FULLM, rem, d);
                    const int oh = __shfl_xor_sync(FULLM, rhas, d);
                    if (oh) { rem = rhas ? Mon::apply(rem, ov) : ov; rhas = 1; }
                }
                total = Mon::apply(total, rem);
            }
            if (lane == last_lane) {
                if (closed) strip[lane_rank + cnt - 1] = total;
                else { a.carry_own[t] = total; a.open_list[atomicAdd(a.open_count, 1)] = (int32_t)t; }
            }
            if (!closed) emit--;
        }
        __syncwarp();
        // ---- fused write-back of the rows finished in this tile ----
        if (lane < emit) my_count += epi_store_pre(a.epi, r_pre, pre_allowed, strip[lane], pre_has, pre_old);
        for (int q = lane + 32; q < emit; q += 32) {
            const int64_t r = a.nzrow[first + q];
            my_count += epi_store(a.epi, r, epi_allowed(a.epi, r), true, strip[q]);
        }
        __syncwarp();
    }
    block_count_add(a.epi.count, my_count);
}

// one warp per open row (listed by the owners): carry_own of its tile (+) the heads of the tiles it runs through
template <class SR>
__global__ void __launch_bounds__(256) k_stream_fixup(StreamArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    const int lane = threadIdx.x & 31;
    const int nopen = *a.open_count;
    int my_count = 0;
    for (int w = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5); w < nopen; w += (int)((gridDim.x * (int64_t)blockDim.x) >> 5)) {
        const int64_t t = a.open_list[w];
        const int64_t r = a.nzrow[a.tile_first[t + 1] - 1], rend = a.rowptr[r + 1];
        const int64_t t_end = (rend - 1) / ST_TILE;
        T acc = Mon::identity(); bool has = false;
        int64_t k = t + 1 + lane;
        for (; k + 7 * 32 <= t_end; k += 8 * 32) {     // hub rows run through thousands of tiles: 8 loads in flight
            T v[8];
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = a.carry_head[k + j * 32];
#pragma unroll
            for (int j = 0; j < 8; j++) { acc = has ? Mon::apply(acc, v[j]) : v[j]; has = true; }
        }
        for (; k <= t_end; k += 32) { acc = has ? Mon::apply(acc, a.carry_head[k]) : a.carry_head[k]; has = true; }
        for (int off = 16; off; off >>= 1) {
            const T oacc = __shfl_xor_sync(0xffffffffu, acc, off);
            const bool ohas = __shfl_xor_sync(0xffffffffu, (int)has, off) != 0;
            if (ohas) { acc = has ? Mon::apply(acc, oacc) : oacc; has = true; }
        }
        if (lane == 0) {
            acc = has ? Mon::apply(a.carry_own[t], acc) : a.carry_own[t];
            my_count += epi_store(a.epi, r, epi_allowed(a.epi, r), true, acc);
        }
    }
    block_count_add(a.epi.count, my_count);
}

// rows without entries produce no t(i): the write-back still applies (delete without accum / under replace)
template <class T>
__global__ void __launch_bounds__(256) k_stream_empty_rows(const int64_t *rowptr, int64_t nrows, EpiArgs<T> e) {
    int c = 0;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x)
        if (rowptr[r + 1] == rowptr[r]) c += epi_store(e, r, epi_allowed(e, r), false, (T)0);
    block_count_add(e.count, c);
}

// ---- atomic-segment pull (u full, no mask, accumulator absent or equal to the monoid) ----
// The case every iteration of PageRank (`r(plus) << A.T.mxv(w, plus_times)`), Bellman-Ford
// (`w(min) << w.vxm(A, min_plus)`) and FastSV (`mngp << A.mxv(gp, min_second)`) hits.  Because the write-back is
// the monoid itself, partial row sums can be combined IN the output with the monoid's atomic: every lane reduces
// its 8 consecutive entries between row-start flags in registers and issues one reduction (RED) per row segment.
// No staging, no warp scans, no carry / fix-up pass; two dependent memory round trips per chunk
// (stream -> gather), the REDs are fire-and-forget.  Tiles that lie inside one hub row are pre-reduced by the warp.
// chunk_first[i] = index in nzrow of the row that owns entry 8*i.  The caller pre-fills absent positions of w with
// the monoid identity.  Floating-point PLUS results depend on the RED order in the last bits (north_star tolerance
// 1e-5); integer and MIN/MAX results are exact.
template <class T> struct RedArgs {
    const uint32_t *pcol; const T *av; int a_iso;
    const T *u; const uint8_t *up;              // up == nullptr: u is full (dense instantiation)
    const void *upairs;                         // sparse u: (value, presence) pairs, pair_bits_t<T>[n]
    const T *hub_v; const uint8_t *hub_p; int hub_n;
    const uint32_t *nzrow; const uint32_t *chunk_first;
    int64_t nnz, ntiles;
    T *w; uint8_t *wp;            // wp == nullptr: presence is already known to be all ones
    int flags;                    // bit 0: load tiles directly instead of by TMA (option red_flags = 1, for A/B measurements)
    unsigned long long *progress; // optional: per row block, the number of finished tiles that touch it (monotonic)
    const uint32_t *tile_blocks;  // [ntiles] first block | last block << 16 of each tile (csr_ensure_tile_blocks)
    unsigned long long *stamps;   // optional debugging aid: [0] = first CTA's start, [1] = last warp's end (globaltimer ns)
};
template <class T> static inline int red_warps() { return sizeof(T) > 4 ? RED_WARPS_8B : 32; }
// shared memory: hub table | per-warp stage (256 x (4 + sizeof(T)) bytes, filled by TMA bulk copies) | mbarriers
template <class T> static inline size_t red_stage_bytes() { return (size_t)ST_TILE * (4 + sizeof(T)) + 128; }   // columns, values, chunk_first
static inline size_t red_hub_bits_bytes(int h) { return ((size_t)((h + 31) / 32) * 4 + 127) & ~(size_t)127; }   // presence bitmap of the hubs
template <class T> static inline size_t red_smem(int h) {
    return (((size_t)h * sizeof(T) + 127) & ~(size_t)127) + red_hub_bits_bytes(h) + (size_t)red_warps<T>() * red_stage_bytes<T>() +
           (size_t)red_warps<T>() * 8 + 128;
}
template <class T> static inline int hub_capacity() {
    const size_t total = (size_t)g.smem_kb * 1024;
    const size_t budget = (total - (size_t)red_warps<T>() * (red_stage_bytes<T>() + 8) - 512) * 8 / (8 * sizeof(T) + 1) * sizeof(T);
    const size_t also_stream = total - 1024 - (size_t)stream_warps<T>() * ST_TILE * sizeof(T);     // k_spmv_stream shares the table
    const size_t b = budget < also_stream ? budget : also_stream;
    return (int)((b / sizeof(T)) & ~(size_t)1023);
}

// ---- TMA bulk copy + mbarrier (sm_90+ PTX; SASS: UBLKCP / SYNCS) ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}

// Shared-memory budget: the hub table competes with L1 for the 256 KB of each SM.  Gather misses in flight are
// bounded by the L1 that is left (measured, tools/microbench/gather_bench5 + bench.py --smem-kb): with the 228 KB
// carve-out (28 KB L1) the same kernel runs 1.3-1.5x slower than with <= 196 KB (60 KB L1), although the table
// is 20% smaller.  Issuing the next tile's gathers ahead of the reduction (look-ahead) made it slower at every
// budget for the same reason, so each warp keeps exactly one tile in flight.
// one warp reports "I am done with row blocks lo..hi" (see k_spmv_red)
__device__ __forceinline__ void red_report(unsigned long long *progress, uint32_t lo, uint32_t hi, int lane) {
    if (lo > hi) return;
    __syncwarp();
    if (lane == 0) {
        __threadfence();
        for (uint32_t b = lo; b <= hi; b++) atomicAdd(progress + b, 1ull);
    }
}

template <class SR, int WARPS, bool SPARSE, bool PROGRESS = false>
__global__ void __launch_bounds__(WARPS * 32, 1) k_spmv_red(RedArgs<typename SR::type> a) {
    using T = typename SR::type;
    using Mon = typename SR::Mon;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr bool kStageVals = SR::kUsesX;          // iso matrices are decided at run time (a.a_iso)
    constexpr uint32_t COL_BYTES = ST_TILE * 4, VAL_BYTES = ST_TILE * sizeof(T), CF_BYTES = 128;
    constexpr uint32_t WSTAGE = COL_BYTES + VAL_BYTES + CF_BYTES;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const size_t hub_bytes = ((size_t)a.hub_n * sizeof(T) + 127) & ~(size_t)127;
    const size_t bits_bytes = ((size_t)((a.hub_n + 31) / 32) * 4 + 127) & ~(size_t)127;
    T *hub = (T *)smem_raw;
    uint32_t *hub_bits = (uint32_t *)(smem_raw + hub_bytes);                       // SPARSE: presence of the hubs
    const uint32_t hub_a = pinned(smem_u32(smem_raw)), bits_a = hub_a + (uint32_t)hub_bytes;
    const uint32_t col_a = hub_a + (uint32_t)(hub_bytes + bits_bytes) + warp * WSTAGE, val_a = col_a + COL_BYTES, cf_a = val_a + VAL_BYTES;
    const uint32_t bar_a = hub_a + (uint32_t)(hub_bytes + bits_bytes) + WARPS * WSTAGE + warp * 8;
    if (SR::kUsesY) for (int i = tid; i < a.hub_n; i += WARPS * 32) hub[i] = a.hub_v[i];
    if (SPARSE) {
        const int padded = (a.hub_n + 31) & ~31;
        for (int i = tid; i < padded; i += WARPS * 32) {
            const unsigned word = __ballot_sync(0xffffffffu, i < a.hub_n && a.hub_p[i] != 0);
            if (lane == 0) hub_bits[i >> 5] = word;
        }
    }
    if (lane == 0) mbar_init(bar_a, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    // programmatic dependent launch: a kernel launched right behind this one with the programmatic-serialization attribute
    // (the fused peer exchange, grb_peer.cu) may start now, i.e. only once every CTA of this grid is resident -- the exchange
    // spins on this kernel's progress counters, so it must never occupy an SM before this kernel does.  No-op otherwise.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const unsigned FULLM = 0xffffffffu;
    const bool stage_vals = kStageVals && !a.a_iso;
    const int64_t nfull = (a.flags & 1) ? 0 : a.nnz / ST_TILE;   // tiles that arrive by bulk copy; a ragged last tile is loaded directly
    const int64_t tstep = (int64_t)gridDim.x * WARPS;
    int64_t t = blockIdx.x * (int64_t)WARPS + warp;
    auto issue = [&](int64_t tt) {                      // lane 0: arm the barrier, start the copies of tile tt
        mbar_expect_tx(bar_a, COL_BYTES + (stage_vals ? VAL_BYTES : 0) + CF_BYTES);
        tma_load_1d(col_a, a.pcol + tt * ST_TILE, COL_BYTES, bar_a);
        if (stage_vals) tma_load_1d(val_a, a.av + tt * ST_TILE, VAL_BYTES, bar_a);
        tma_load_1d(cf_a, a.chunk_first + tt * 32, CF_BYTES, bar_a);
    };
    if (lane == 0 && t < nfull) issue(t);
    uint32_t parity = 0;
    // progress reporting (multi-GPU exchange, grb_peer.cu): a warp's tiles ascend through the rows, so when it starts a tile
    // whose first row block is b0, it is done with every block below b0: it then counts itself once into each such block it
    // touched (one fence per block transition, i.e. a handful per warp -- a fence per tile made the kernel 3x slower).  A
    // block is complete when as many warps as touch it (csr_ensure_tile_blocks: `expected`) have reported.
    // PROGRESS is a compile-time variant (8 more registers, which the co-resident exchange kernel needs): the plain kernel is
    // untouched.  The block range of the NEXT tile is fetched one tile ahead: a warp has one tile in flight, so a dependent
    // load at the top of every tile costs its full latency (262 vs 199 us measured, profiles/r02i_*).
    uint32_t pend_lo = 0xffffffffu, pend_hi = 0, tb_next = 0;
    if (PROGRESS) tb_next = (t < a.ntiles) ? a.tile_blocks[t] : 0u;
    if (PROGRESS && a.stamps && blockIdx.x == 0 && tid == 0) { unsigned long long now; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now)); a.stamps[0] = now; a.stamps[1] = 0; }
    for (; t < a.ntiles; t += tstep) {
        if (PROGRESS) {
            const uint32_t tb = tb_next, b0 = tb & 0xffffu, b1 = tb >> 16;
            tb_next = (t + tstep < a.ntiles) ? a.tile_blocks[t + tstep] : 0u;
            if (pend_lo < b0) { red_report(a.progress, pend_lo, pend_hi < b0 - 1 ? pend_hi : b0 - 1, lane); pend_lo = b0; }
            pend_lo = pend_lo < b0 ? pend_lo : b0;
            pend_hi = pend_hi > b1 ? pend_hi : b1;
        }
      do {
        const int64_t chunk = t * 32 + lane;
        const int64_t p0 = chunk * ST_EPL;
        const bool tile_full = t < nfull;
        int nvalid = ST_EPL;
        uint32_t c[ST_EPL];
        T av[ST_EPL];
        uint32_t j0 = 0;
        if (tile_full) {
            mbar_wait(bar_a, parity);
            parity ^= 1;
            const uint4 c0 = lds128_at(col_a + lane * 32), c1 = lds128_at(col_a + lane * 32 + 16);
            c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
            if (stage_vals) {
                if constexpr (sizeof(T) == 4) {
                    const uint4 v0 = lds128_at(val_a + lane * 32), v1 = lds128_at(val_a + lane * 32 + 16);
                    const uint32_t raw[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                    for (int k = 0; k < ST_EPL; k++) memcpy(&av[k], &raw[k], 4);
                } else {
#pragma unroll
                    for (int k = 0; k < ST_EPL; k++) av[k] = lds_at<T>(val_a + (lane * ST_EPL + k) * (uint32_t)sizeof(T));
                }
            }
            j0 = lds_at<uint32_t>(cf_a + lane * 4);
            // the generic-proxy reads above must be ordered before the async-proxy (TMA) writes that re-fill the stage:
            // without this fence tiles were observed half-overwritten on B200 (tools/debug/dbg_red.py)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0 && t + tstep < nfull) issue(t + tstep);     // next tile streams in behind the gathers
        } else {
            nvalid = a.nnz - p0 < 0 ? 0 : (a.nnz - p0 > ST_EPL ? ST_EPL : (int)(a.nnz - p0));
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) {
                c[k] = (k < nvalid) ? a.pcol[p0 + k] : HUB_FLAG;      // past the end: hub 0, no flag, result unused
                if (stage_vals) av[k] = (k < nvalid) ? a.av[p0 + k] : (T)0;
            }
            if (nvalid) j0 = a.chunk_first[chunk];
        }
        if (kStageVals && a.a_iso) {
            const T v = a.av[0];
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) av[k] = v;
        }
        unsigned fbits = 0;
#pragma unroll
        for (int k = 0; k < ST_EPL; k++) fbits |= ((c[k] >> 30) & 1u) << k;
        const int nseg = __popc(fbits >> 1);                 // row starts after the lane's first entry
        // row ids of the lane's first three segments (more than three starts inside 8 entries: loaded on demand)
        uint32_t r0 = 0, r1 = 0, r2 = 0;
        if (nvalid) {
            r0 = a.nzrow[j0];
            if (nseg >= 1) r1 = a.nzrow[j0 + 1];
            if (nseg >= 2) r2 = a.nzrow[j0 + 2];
        }
        T x[ST_EPL];
        unsigned pbits = (nvalid >= ST_EPL) ? 0xffu : ((1u << nvalid) - 1u);     // entries that exist AND whose u(j) is present
        if (!SPARSE) {
            if (SR::kUsesY) {
#pragma unroll
                for (int k = 0; k < ST_EPL; k++) {
                    x[k] = (T)0;
                    if ((int)c[k] >= 0) x[k] = ld_ro(a.u + (c[k] & COL_MASK));
                }
            }
        } else {
            const pair_bits_t<T> *pairs = (const pair_bits_t<T> *)a.upairs;
            pair_bits_t<T> q[ST_EPL];
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) {
                q[k] = pair_bits_t<T>{};
                if ((int)c[k] >= 0) q[k] = ld_pair<T>(pairs + (c[k] & COL_MASK));
            }
            unsigned got = 0;
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) got |= (split_pair_bits<T>(q[k], x[k]) ? 1u : 0u) << k;
#pragma unroll
            for (int k = 0; k < ST_EPL; k++)
                if ((int)c[k] < 0) { const uint32_t h = c[k] & COL_MASK; got |= ((lds_at<uint32_t>(bits_a + (h >> 5) * 4) >> (h & 31)) & 1u) << k; }
            pbits &= got;
        }
        if (SR::kUsesY) {
#pragma unroll
            for (int k = 0; k < ST_EPL; k++) if ((int)c[k] < 0) x[k] = lds_at<T>(hub_a + (c[k] & COL_MASK) * (uint32_t)sizeof(T));
        }
        // a tile that lies inside a single row (hub rows): one RED for the warp instead of 32 on one address
        const uint32_t j_lane0 = __shfl_sync(FULLM, j0, 0);
        const bool one_row = tile_full && __all_sync(FULLM, (fbits >> 1) == 0 && j0 == j_lane0);
        // segment state: acc is meaningful iff has
        T acc = SR::mult(SR::kUsesX ? av[0] : (T)0, SR::kUsesY ? x[0] : (T)0);
        bool has = pbits & 1u;
        if (one_row) {
#pragma unroll
            for (int k = 1; k < ST_EPL; k++) {
                const T z = SR::mult(SR::kUsesX ? av[k] : (T)0, SR::kUsesY ? x[k] : (T)0);
                if (!SPARSE) acc = Mon::apply(acc, z);
                else if (pbits & (1u << k)) { acc = has ? Mon::apply(acc, z) : z; has = true; }
            }
#pragma unroll
            for (int d = 16; d; d >>= 1) {
                const T ov = __shfl_xor_sync(FULLM, acc, d);
                if (!SPARSE) acc = Mon::apply(acc, ov);
                else {
                    const bool oh = __shfl_xor_sync(FULLM, (int)has, d) != 0;
                    if (oh) { acc = has ? Mon::apply(acc, ov) : ov; has = true; }
                }
            }
            if (lane == 0 && (!SPARSE || has)) {
                Mon::atomic(a.w + r0, acc);
                if (a.wp && (SPARSE || (fbits & 1u))) a.wp[r0] = 1;
            }
            break;
        }
        if (nvalid == 0) break;
        if (!SPARSE && a.wp && (fbits & 1u)) a.wp[r0] = 1;
        int m = 0;
        uint32_t r = r0;
#pragma unroll
        for (int k = 1; k < ST_EPL; k++) {
            const T z = SR::mult(SR::kUsesX ? av[k] : (T)0, SR::kUsesY ? x[k] : (T)0);
            if (fbits & (1u << k)) {
                if (!SPARSE || has) {
                    Mon::atomic(a.w + r, acc);
                    if (SPARSE && a.wp) a.wp[r] = 1;
                }
                m++;
                r = (m == 1) ? r1 : (m == 2) ? r2 : a.nzrow[j0 + m];
                if (!SPARSE && a.wp) a.wp[r] = 1;
                acc = z; has = (pbits >> k) & 1u;
            } else if (!SPARSE) { if (k < nvalid) acc = Mon::apply(acc, z); }
            else if (pbits & (1u << k)) { acc = has ? Mon::apply(acc, z) : z; has = true; }
        }
        if (!SPARSE || has) {
            Mon::atomic(a.w + r, acc);
            if (SPARSE && a.wp) a.wp[r] = 1;
        }
      } while (0);
    }
    if (PROGRESS) red_report(a.progress, pend_lo, pend_hi, lane);     // whatever this warp still owes
    if (PROGRESS && a.stamps && lane == 0) { unsigned long long now; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now)); atomicMax(a.stamps + 1, now); }
}

// w := identity where absent (accumulating call) or everywhere with presence cleared (overwriting call)
template <class T>
__global__ void __launch_bounds__(256) k_red_prefill(T *w, uint8_t *wp, int64_t n, T ident, int overwrite) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (overwrite) { w[i] = ident; wp[i] = 0; }
        else if (!wp[i]) w[i] = ident;
    }
}

template <class T>
__global__ void k_gather_hubs(const T *u, const uint8_t *up, const uint32_t *hub_ids, int h, T *hub_v, uint8_t *hub_p) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= h) return;
    const uint32_t c = hub_ids[i];
    hub_v[i] = u[c];
    hub_p[i] = up ? up[c] : (uint8_t)1;
}

// ---- masked pull with early exit (ANY monoid): one thread per row, stop at the first hit ----
// A lane scans the first PULL_ANY_SERIAL entries of its row alone (on a BFS level that is worth pulling, most rows hit within
// two).  Rows still undecided after that are finished by the whole warp, 32 entries per step with one ballot: a single lane
// walking a 300-entry row costs 300 dependent L2 round trips, and the longest such row set the kernel's time (it is what made
// the pull over a rank's column block, where a row's frontier neighbours may all live on another rank, 3x slower than the
// pull over the whole matrix).  The first hit in row order is taken either way, as the oracle does.
constexpr int PULL_ANY_SERIAL = 8;
template <class SR>
__global__ void __launch_bounds__(256) k_pull_any(PullArgs<typename SR::type> a) {
    using T = typename SR::type;
    const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const bool full = (a.up == nullptr);
    const bool allowed = r < a.nrows && epi_allowed(a.epi, r);
    long long p = 0, e = 0, hit = -1;
    if (allowed) {
        p = a.rowptr[r]; e = a.rowptr[r + 1];
        const long long stop = p + PULL_ANY_SERIAL < e ? p + PULL_ANY_SERIAL : e;
        for (; p < stop; p++)
            if (full || __ldg(a.up + __ldg(a.col + p))) { hit = p; break; }
    }
    unsigned pending = __ballot_sync(0xffffffffu, hit < 0 && p < e);
    while (pending) {
        const int src = __ffs(pending) - 1;
        pending &= pending - 1;
        const long long ps = __shfl_sync(0xffffffffu, p, src), es = __shfl_sync(0xffffffffu, e, src);
        long long found = -1;
        for (long long base = ps; base < es; base += 32) {
            const long long q = base + lane;
            const bool h = q < es && (full || __ldg(a.up + __ldg(a.col + q)));
            const unsigned b = __ballot_sync(0xffffffffu, h);
            if (b) { found = base + (__ffs(b) - 1); break; }
        }
        if (lane == src) hit = found;
    }
    int my_count = 0;
    if (r < a.nrows) {
        T acc = (T)0;
        if (hit >= 0) {
            const uint32_t c = __ldg(a.col + hit);
            T av = SR::kUsesX ? a.av[a.a_iso ? 0 : hit] : (T)0;
            T x = SR::kUsesY ? a.u[c] : (T)0;
            acc = SR::mult(av, x);
        }
        my_count = epi_store(a.epi, r, allowed, hit >= 0, acc);
    }
    block_count_add(a.epi.count, my_count);
}

// ---- dense write-back w<m,replace> = accum(w, t) (after push, or when types had to be cast) ----
template <class T>
__global__ void __launch_bounds__(256) k_vec_writeback(EpiArgs<T> e, const T *t, const uint8_t *tp, int64_t n) {
    int my_count = 0;
    const bool need_old = epi_needs_old(e);
    const int64_t base = blockIdx.x * (int64_t)(256 * EW_VEC) + threadIdx.x;
    bool al[EW_VEC], th[EW_VEC], wh[EW_VEC]; T tv[EW_VEC], wo[EW_VEC];
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) {
        const int64_t i = base + k * 256;
        al[k] = th[k] = wh[k] = false; tv[k] = wo[k] = (T)0;
        if (i < n) {
            al[k] = epi_allowed(e, i); th[k] = tp[i] != 0; tv[k] = t[i];
            if (epi_reads_old(e, al[k])) { wh[k] = e.wp_in[i] != 0; if (need_old) wo[k] = e.w_in[i]; }
        }
    }
#pragma unroll
    for (int k = 0; k < EW_VEC; k++) {
        const int64_t i = base + k * 256;
        if (i < n) my_count += epi_apply(e, i, al[k], th[k], tv[k], wh[k], wo[k]);
    }
    block_count_add(e.count, my_count);
}

// ---- the same write-back, 16 positions per thread (no accumulator, in place, structural or 1-byte value mask) ----
// The byte-per-thread kernels above issue one 32-byte request per warp and array and spend ~85 instructions per position;
// a masked BFS level over 16.7 M vertices took 76 us (assign) and 70 us (write-back) for 34-50 MB of traffic.  Here a thread
// owns 16 consecutive positions: presence, mask and (1-byte) values move as one 16-byte word each, the mask logic runs on
// four packed bytes at a time, old presence is read only by threads that keep something, and presence is stored only by
// threads that change something.  SCALAR: t is the constant x, present everywhere (GrB_assign of a scalar).
template <class T, bool SCALAR>
__global__ void __launch_bounds__(256) k_writeback16(EpiArgs<T> e, const T *t, const uint8_t *tp, T x, int64_t n) {
    const int64_t base = (blockIdx.x * (int64_t)256 + threadIdx.x) * 16;
    int c = 0;
    if (base + 16 <= n) {
        uint32_t m[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu}, th[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
        if (e.mask_p) {
            const uint4 q = *reinterpret_cast<const uint4 *>(e.mask_p + base);
            m[0] = __vcmpne4(q.x, 0); m[1] = __vcmpne4(q.y, 0); m[2] = __vcmpne4(q.z, 0); m[3] = __vcmpne4(q.w, 0);
            if (e.mask_vsize) {
                const uint4 v = *reinterpret_cast<const uint4 *>((const uint8_t *)e.mask_v + base);
                m[0] &= __vcmpne4(v.x, 0); m[1] &= __vcmpne4(v.y, 0); m[2] &= __vcmpne4(v.z, 0); m[3] &= __vcmpne4(v.w, 0);
            }
        }
        if (e.comp) { m[0] = ~m[0]; m[1] = ~m[1]; m[2] = ~m[2]; m[3] = ~m[3]; }
        if (!SCALAR) {
            const uint4 q = *reinterpret_cast<const uint4 *>(tp + base);
            th[0] = __vcmpne4(q.x, 0); th[1] = __vcmpne4(q.y, 0); th[2] = __vcmpne4(q.z, 0); th[3] = __vcmpne4(q.w, 0);
        }
        const bool some_out = (m[0] & m[1] & m[2] & m[3]) != 0xffffffffu;       // a position the mask does not allow
        const bool some_in = (m[0] | m[1] | m[2] | m[3]) != 0;                   // a position it does
        uint32_t kept[4] = {0, 0, 0, 0};                                          // masked-out positions that keep an old entry
        if (!e.replace && some_out) {
            const uint4 q = *reinterpret_cast<const uint4 *>(e.wp_in + base);
            kept[0] = ~m[0] & __vcmpne4(q.x, 0); kept[1] = ~m[1] & __vcmpne4(q.y, 0);
            kept[2] = ~m[2] & __vcmpne4(q.z, 0); kept[3] = ~m[3] & __vcmpne4(q.w, 0);
        }
        uint32_t sel[4], wn[4];
#pragma unroll
        for (int k = 0; k < 4; k++) { sel[k] = m[k] & th[k]; wn[k] = (sel[k] | kept[k]) & 0x01010101u; c += __popc(wn[k]); }
        if (e.replace || some_in)       // otherwise every position keeps what it has
            *reinterpret_cast<uint4 *>(e.wp_out + base) = make_uint4(wn[0], wn[1], wn[2], wn[3]);
        if ((sel[0] | sel[1] | sel[2] | sel[3]) != 0) {
            if (sizeof(T) == 1 && (kept[0] | kept[1] | kept[2] | kept[3]) == 0) {
                // no old value survives in this group: the 16 values go out as one word (positions without an entry hold junk)
                uint4 v;
                if (SCALAR) { uint32_t b; { uint8_t xb; memcpy(&xb, &x, 1); b = xb * 0x01010101u; } v = make_uint4(b, b, b, b); }
                else v = *reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(t) + base);
                *reinterpret_cast<uint4 *>(reinterpret_cast<uint8_t *>(e.w_out) + base) = v;
            } else {
#pragma unroll
                for (int k = 0; k < 16; k++)
                    if ((sel[k >> 2] >> ((k & 3) * 8)) & 1u) e.w_out[base + k] = SCALAR ? x : t[base + k];
            }
        }
    } else if (base < n) {
        for (int64_t i = base; i < n; i++) {
            const bool al = epi_allowed(e, i);
            const bool has = SCALAR ? true : tp[i] != 0;
            c += epi_apply(e, i, al, has, SCALAR ? x : t[i], e.wp_in[i] != 0, e.w_in[i]);
        }
    }
    block_count_add(e.count, c);
}
template <class T>
static inline bool writeback16_ok(const EpiArgs<T> &e, const void *t, const void *tp, int64_t n) {
    auto al = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return n >= 16 && e.accum < 0 && e.w_out == e.w_in && e.wp_out == e.wp_in && (e.mask_vsize == 0 || e.mask_vsize == 1) &&
           al(e.wp_in) && (!e.mask_p || al(e.mask_p)) && (!e.mask_vsize || al(e.mask_v)) && (!tp || al(tp)) &&
           (sizeof(T) != 1 || (al(e.w_out) && (!t || al(t))));
}
static inline int writeback16_grid(int64_t n) { return (int)cdiv(cdiv(n, 16), 256); }
// launch the wide form when it applies, else the general kernel
template <class T>
static inline void launch_vec_writeback(const EpiArgs<T> &e, const T *t, const uint8_t *tp, int64_t n) {
    if (writeback16_ok(e, t, tp, n)) {
        auto wide = k_writeback16<T, false>;
        GRB_LAUNCH("vec_writeback16", wide, writeback16_grid(n), 256, 0, e, t, tp, (T)0, n);
    } else
        GRB_LAUNCH("vec_writeback", k_vec_writeback<T>, ew_grid4(n), 256, 0, e, t, tp, n);
}

// ---- push ----
template <class T>
__global__ void k_fill_identity(T *t, uint8_t *tp, int64_t n, T ident) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        t[i] = ident; tp[i] = 0;
    }
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
