// grb_peer.cu — multi-GPU exchange over NVLink peer memory for column-partitioned vector products.
//
// Partition: rank p owns the vertices [lo_p, hi_p) and the matching COLUMN block of the pull matrix, so a product on rank p
// gathers only from its own block of the vector and yields partial sums for ALL n rows.  Those partial sums live in a
// buffer that every rank allocates symmetrically and maps into every other rank's address space (NVLink P2P, and one
// multicast address when NVSwitch multicast is available).  What a reduce-scatter would do is done here by ONE light
// kernel per row chunk that runs beside the next chunk's product:
//   * announce "my partial sums of this chunk are complete" in every peer's flag words (release, system scope),
//   * wait until every peer has announced the same,
//   * out[i] = sum over ranks of partial_q[block offset + i]: one multimem.ld_reduce per 16 bytes when the multicast address
//     exists (the switch adds the 2/4/8 copies: the rank receives n/N floats, not (N-1)/N * n), otherwise P2P loads,
//   * optionally out_scaled[i] = out[i] / div[i]: the next iteration's w = r / d of PageRank, fused.
// Small CTAs without shared memory: the fused form (64 threads, 72 registers = 4.6 K registers) fits on an SM beside a 1024-thread
// k_spmv_red CTA (56 registers = 57 K of the 64 K, 194 KB shared), which an NCCL kernel does not -- that is why NCCL's reduce-scatter serialised behind the
// products (profiles/r02b_*).  No counterpart in the reference (python-graphblas is single-process).
#include "grb_internal.cuh"

namespace {

constexpr int PEER_MAX_WORLD = 16;
struct PeerArgs {
    const float *peer[PEER_MAX_WORLD];     // every rank's partial-sum buffer (P2P mapped), already offset to this chunk's block
    uint32_t *flags[PEER_MAX_WORLD];       // every rank's flag group for this (buffer, chunk): `world` words, one per source
    const float *mc;                       // multicast alias of the same block, or nullptr
    float *out, *out_scaled;
    const float *div;
    int64_t count;
    int world, rank;
    uint32_t seq;
    float *fill; int64_t fill_count; float fill_value;     // optional: refill a local buffer (the other generation) afterwards
};

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// a handshake that does not complete within 20 s is a protocol error (a peer died, or the caller broke the launch order):
// trap, so the process fails with a CUDA error instead of hanging the GPU
#define PEER_SPIN_GUARD(t0) do { if (globaltimer_ns() - (t0) > 20000000000ull) __trap(); } while (0)
// polling load: coherent at system scope but WITHOUT acquire semantics -- an acquire at every poll invalidates the SM's L1 each
// time, which slowed the product kernel running on the same SM 3x (profiles/r02g_*); one fence follows the successful poll
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_gpu_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float *p) {      // system-coherent, never from a stale L1 line
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_peer_f1(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_multimem_add_f4(const float *p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// SMALL: 64 threads, <= 40 registers, 4 loads in flight per thread: CTAs that fit on an SM BESIDE a running k_spmv_red CTA, for
// the chunked pipeline (exchange of row chunk c on a second stream while chunk c + 1 is being multiplied)
template <bool SMALL>
__global__ void __launch_bounds__(SMALL ? 64 : 256, SMALL ? 24 : 1) k_peer_reduce_scatter(PeerArgs a) {
    // ---- announce + wait (every CTA waits by itself: no grid-wide barrier, CTAs need not be co-resident)
    if (blockIdx.x == 0 && threadIdx.x < a.world) {
        __threadfence_system();
        st_release_sys(a.flags[threadIdx.x] + a.rank, a.seq);
    }
    if (threadIdx.x < a.world) {
        const uint32_t *f = a.flags[a.rank] + threadIdx.x;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(f) - a.seq) < 0) { __nanosleep(64); PEER_SPIN_GUARD(t0); }
    }
    __syncthreads();
    // ---- reduce this rank's block.  One multimem.ld_reduce (or N peer loads) takes microseconds to come back over the
    // switch, so each thread keeps UNROLL of them in flight (24 CTAs x 256 threads x 1 load: 25 GB/s measured; profiles/r02c_*)
    constexpr int UNROLL = SMALL ? 4 : 8;
    const int64_t n4 = a.count >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    for (; i + (UNROLL - 1) * stride < n4; i += UNROLL * stride) {
        float4 s[UNROLL];
        if (a.mc) {
#pragma unroll
            for (int u = 0; u < UNROLL; u++) s[u] = ld_multimem_add_f4(a.mc + 4 * (i + u * stride));
        } else {
#pragma unroll
            for (int u = 0; u < UNROLL; u++) s[u] = ld_peer_f4(a.peer[0] + 4 * (i + u * stride));
            for (int q = 1; q < a.world; q++) {
                float4 v[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; u++) v[u] = ld_peer_f4(a.peer[q] + 4 * (i + u * stride));
#pragma unroll
                for (int u = 0; u < UNROLL; u++) { s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w; }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int64_t j = i + u * stride;
            reinterpret_cast<float4 *>(a.out)[j] = s[u];
            if (a.out_scaled) {
                const float4 d = reinterpret_cast<const float4 *>(a.div)[j];
                reinterpret_cast<float4 *>(a.out_scaled)[j] = make_float4(s[u].x / d.x, s[u].y / d.y, s[u].z / d.z, s[u].w / d.w);
            }
        }
    }
    for (; i < n4; i += stride) {
        float4 s;
        if (a.mc) s = ld_multimem_add_f4(a.mc + 4 * i);
        else {
            s = ld_peer_f4(a.peer[0] + 4 * i);
            for (int q = 1; q < a.world; q++) {
                const float4 v = ld_peer_f4(a.peer[q] + 4 * i);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
        }
        reinterpret_cast<float4 *>(a.out)[i] = s;
        if (a.out_scaled) {
            const float4 d = reinterpret_cast<const float4 *>(a.div)[i];
            reinterpret_cast<float4 *>(a.out_scaled)[i] = make_float4(s.x / d.x, s.y / d.y, s.z / d.z, s.w / d.w);
        }
    }
    if (a.fill) {
        const float4 v = make_float4(a.fill_value, a.fill_value, a.fill_value, a.fill_value);
        const int64_t m4 = a.fill_count >> 2;
        for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < m4; j += stride) reinterpret_cast<float4 *>(a.fill)[j] = v;
    }
    for (int64_t t = (n4 << 2) + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < a.count; t += stride) {   // tail (< 4 elements)
        const int64_t i = t;
        float s = 0.f;
        for (int q = 0; q < a.world; q++) s += ld_peer_f1(a.peer[q] + i);
        a.out[i] = s;
        if (a.out_scaled) a.out_scaled[i] = s / a.div[i];
    }
}

// ---- the fused form: ONE exchange launch per product, running BESIDE the product kernel ------------------------------------
// k_spmv_red reports, per row block, how many of the tiles touching it are done (grb_spmv.cuh, RedArgs::progress).  CTA 0 of this
// kernel watches those counters and tells rank b "my partial sums for your rows are complete" the moment block b is; the other
// CTAs wait until every rank has said so about THIS rank's block, sum the N copies (multimem.ld_reduce or P2P loads), write
// r_p and the next w_p = r_p / d_p, and refill the generation of the local partial-sum buffer that the product after next
// will accumulate into (three generations: nobody reads generation k - 1 any more once every peer has reported for k).
// Rows are in natural order, so rank 0's block ships first and rank N-1's last: the pulls are staggered over the product's
// run time and only the last block's pull (1/N of the traffic) is exposed.
// phase stamps of the last fused exchange (globaltimer ns): [0] first CTA started, [1] first block announced, [2] last block
// announced (~ the product finished), [3] flags of this rank's block complete, [4] reduce done (max over CTAs), [5] refill done
static __device__ unsigned long long g_peer_stamps[8];
struct PeerFusedArgs {
    PeerArgs x;
    const unsigned long long *progress, *expected;
    unsigned long long product_seq;
    float *fill; int64_t fill_count; float fill_value;
};
// <= 64 registers: beside a k_spmv_red CTA (8 warps x 56 registers per SM sub-partition) 2048 registers per sub-partition are free,
// i.e. exactly one warp of 64 registers; with 72 the CTAs waited until the product had left the SM (no overlap, profiles/r02o_*)
__global__ void __launch_bounds__(64, 24) k_peer_exchange_fused(PeerFusedArgs f) {
    const PeerArgs &a = f.x;
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) { g_peer_stamps[0] = globaltimer_ns(); g_peer_stamps[4] = 0; g_peer_stamps[5] = 0; }
        if (threadIdx.x < a.world) {            // announcer: one thread per destination rank
            const int b = threadIdx.x;
            const unsigned long long target = f.expected[b] * f.product_seq;
            const unsigned long long t0 = globaltimer_ns();
            while (ld_relaxed_gpu_u64(f.progress + b) < target) { __nanosleep(500); PEER_SPIN_GUARD(t0); }
            __threadfence_system();
            st_release_sys(a.flags[b] + a.rank, a.seq);
            if (b == 0) g_peer_stamps[1] = globaltimer_ns();
            if (b == a.world - 1) g_peer_stamps[2] = globaltimer_ns();
        }
        return;
    }
    if (threadIdx.x < a.world) {
        const uint32_t *fl = a.flags[a.rank] + threadIdx.x;
        const unsigned long long t0 = globaltimer_ns();
        while ((int32_t)(ld_relaxed_sys(fl) - a.seq) < 0) { __nanosleep(1000); PEER_SPIN_GUARD(t0); }
        asm volatile("fence.acq_rel.sys;" ::: "memory");
    }
    __syncthreads();
    if (blockIdx.x == 1 && threadIdx.x == 0) g_peer_stamps[3] = globaltimer_ns();
    constexpr int UNROLL = 4;
    const int64_t n4 = a.count >> 2;
    const int64_t stride = (int64_t)(gridDim.x - 1) * blockDim.x;
    int64_t i = (blockIdx.x - 1) * (int64_t)blockDim.x + threadIdx.x;
    for (; i < n4; i += UNROLL * stride) {
        float4 s[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int64_t j = i + u * stride;
            if (j >= n4) { s[u] = make_float4(0.f, 0.f, 0.f, 0.f); continue; }
            if (a.mc) s[u] = ld_multimem_add_f4(a.mc + 4 * j);
            else s[u] = ld_peer_f4(a.peer[0] + 4 * j);
        }
        if (!a.mc) {
            for (int q = 1; q < a.world; q++) {
                float4 v[UNROLL];
#pragma unroll
                for (int u = 0; u < UNROLL; u++) { const int64_t j = i + u * stride; v[u] = (j < n4) ? ld_peer_f4(a.peer[q] + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f); }
#pragma unroll
                for (int u = 0; u < UNROLL; u++) { s[u].x += v[u].x; s[u].y += v[u].y; s[u].z += v[u].z; s[u].w += v[u].w; }
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int64_t j = i + u * stride;
            if (j >= n4) continue;
            reinterpret_cast<float4 *>(a.out)[j] = s[u];
            if (a.out_scaled) {
                const float4 d = reinterpret_cast<const float4 *>(a.div)[j];
                reinterpret_cast<float4 *>(a.out_scaled)[j] = make_float4(s[u].x / d.x, s[u].y / d.y, s[u].z / d.z, s[u].w / d.w);
            }
        }
    }
    if (threadIdx.x == 0) atomicMax(&g_peer_stamps[4], globaltimer_ns());
    if (f.fill) {
        const float4 v = make_float4(f.fill_value, f.fill_value, f.fill_value, f.fill_value);
        const int64_t m4 = f.fill_count >> 2;
        for (int64_t j = (blockIdx.x - 1) * (int64_t)blockDim.x + threadIdx.x; j < m4; j += stride) reinterpret_cast<float4 *>(f.fill)[j] = v;
    }
    if (threadIdx.x == 0) atomicMax(&g_peer_stamps[5], globaltimer_ns());
}

}  // namespace

extern "C" {

// A full vector whose value array is caller-owned device memory (e.g. a torch symmetric-memory buffer that peers map):
// products accumulate straight into memory the other GPUs can read.  The library never frees or swaps that array.
GrB_Info GxB_b200_Vector_wrap_device(GrB_Vector *v_, GrB_Type type, GrB_Index n, void *values) {
    if (!v_) return GrB_NULL_POINTER;
    *v_ = nullptr;
    GB_Object *o = nullptr;
    API_BEGIN
    if (!values) grb_throw(GrB_NULL_POINTER, "values is NULL");
    o = obj_new(K_VECTOR, type, n, 1);
    dfree(o->vals);
    o->vals = values;
    o->external_vals = true;
    CUDA_TRY(cudaMemsetAsync(o->present, 1, (size_t)n, g.stream));
    o->nvals = (int64_t)n; o->nvals_known = true; o->list_valid = false;
    *v_ = (GrB_Vector)o;
    API_END(nullptr)
}

GrB_Info GxB_b200_peer_reduce_scatter_FP32(float *out, float *out_scaled, const float *div, const uint64_t *peer_bases,
                                           uint64_t multicast_base, uint64_t data_offset_bytes, uint64_t count,
                                           uint64_t flag_offset_bytes, int world, int rank, uint32_t seq,
                                           float *fill, uint64_t fill_count, float fill_value, int ctas, void *stream) {
    if (g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return e.info; } }   // raw pointers: materialise first
    if ((fill_count & 3) || ((uintptr_t)fill & 15)) return GrB_INVALID_VALUE;
    if (!out || !peer_bases || (out_scaled && !div)) return GrB_NULL_POINTER;
    if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world) return GrB_INVALID_VALUE;
    if ((data_offset_bytes & 15) || (flag_offset_bytes & 3) || ((uintptr_t)out & 15) || (out_scaled && (((uintptr_t)out_scaled | (uintptr_t)div) & 15)))
        return GrB_INVALID_VALUE;
    PeerArgs a;
    for (int q = 0; q < world; q++) {
        if (!peer_bases[q]) return GrB_NULL_POINTER;
        a.peer[q] = (const float *)(peer_bases[q] + data_offset_bytes);
        a.flags[q] = (uint32_t *)(peer_bases[q] + flag_offset_bytes);
    }
    a.mc = multicast_base ? (const float *)(multicast_base + data_offset_bytes) : nullptr;
    a.out = out; a.out_scaled = out_scaled; a.div = div;
    a.count = (int64_t)count; a.world = world; a.rank = rank; a.seq = seq;
    a.fill = fill; a.fill_count = (int64_t)fill_count; a.fill_value = fill_value;
    const bool small = ctas < 0;           // negative: |ctas| small CTAs (64 threads) that co-reside with a product kernel
    if (small) ctas = -ctas;
    if (ctas < 1) ctas = 1;
    if (ctas > 1024) ctas = 1024;
    const int slot = g.timing ? ktime_begin("peer_reduce_scatter", (cudaStream_t)stream) : -1;
    if (small) {
        static int carveout_set = -1;      // an SM changes its shared-memory carve-out only when idle: ask for k_spmv_red's
        if (carveout_set != (int)g.peer_carveout) {
            cudaFuncSetAttribute(k_peer_reduce_scatter<true>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)g.peer_carveout);
            carveout_set = (int)g.peer_carveout;
        }
        k_peer_reduce_scatter<true><<<ctas, 64, 0, (cudaStream_t)stream>>>(a);
    } else k_peer_reduce_scatter<false><<<ctas, 256, 0, (cudaStream_t)stream>>>(a);
    if (slot >= 0) ktime_end(slot);
    g.launches++;
    g.last_kernel = "peer_reduce_scatter";
    return cudaGetLastError() == cudaSuccess ? GrB_SUCCESS : GrB_PANIC;
}

// The next product (GrB_mxv / GrB_vxm taking the unmasked k_spmv_red path) reports its progress per row block of
// `rows_per_block` output rows into counters[nblocks] (uint64, monotonic: never reset) and copies the number of tiles that
// touch each block into expected[nblocks].  One shot: consumed by the next product, successful or not.
GrB_Info GxB_b200_next_product_progress(uint64_t *counters, uint64_t *expected, int nblocks, GrB_Index rows_per_block) {
    if (!counters || !expected) return GrB_NULL_POINTER;
    if (nblocks < 1 || nblocks > PEER_MAX_WORLD || rows_per_block < 1) return GrB_INVALID_VALUE;
    g.progress_dev = (unsigned long long *)counters; g.progress_expected_dev = (unsigned long long *)expected;
    g.progress_blocks = nblocks; g.progress_rpb = (int64_t)rows_per_block;
    return GrB_SUCCESS;
}

// debugging aid: the phase stamps of the last fused exchange on the current device (8 x uint64 nanoseconds, see g_peer_stamps)
GrB_Info GxB_b200_peer_debug_stamps(uint64_t *out8) {
    if (!out8) return GrB_NULL_POINTER;
    return cudaMemcpyFromSymbol(out8, g_peer_stamps, sizeof(unsigned long long) * 8) == cudaSuccess ? GrB_SUCCESS : GrB_PANIC;
}

// count must be a multiple of 4 (whole float4s); the block of this rank is [data_offset_bytes, + 4 * count) in every buffer.
GrB_Info GxB_b200_peer_exchange_fused_FP32(float *out, float *out_scaled, const float *div, const uint64_t *peer_bases,
                                           uint64_t multicast_base, uint64_t data_offset_bytes, uint64_t count,
                                           uint64_t flag_offset_bytes, int world, int rank, uint32_t seq,
                                           const uint64_t *progress, const uint64_t *expected, uint64_t product_seq,
                                           float *fill, uint64_t fill_count, float fill_value, int ctas, void *stream) {
    if (g.pending_n) { try { pending_flush(); } catch (const GrbError &e) { return e.info; } }   // raw pointers: materialise first
    if (!out || !peer_bases || (out_scaled && !div) || !progress || !expected) return GrB_NULL_POINTER;
    if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world || (count & 3) || (fill_count & 3)) return GrB_INVALID_VALUE;
    if ((data_offset_bytes & 15) || (flag_offset_bytes & 3) || ((uintptr_t)out & 15) || ((uintptr_t)fill & 15) ||
        (out_scaled && (((uintptr_t)out_scaled | (uintptr_t)div) & 15)))
        return GrB_INVALID_VALUE;
    PeerFusedArgs f;
    PeerArgs &a = f.x;
    for (int q = 0; q < world; q++) {
        if (!peer_bases[q]) return GrB_NULL_POINTER;
        a.peer[q] = (const float *)(peer_bases[q] + data_offset_bytes);
        a.flags[q] = (uint32_t *)(peer_bases[q] + flag_offset_bytes);
    }
    a.mc = multicast_base ? (const float *)(multicast_base + data_offset_bytes) : nullptr;
    a.out = out; a.out_scaled = out_scaled; a.div = div;
    a.count = (int64_t)count; a.world = world; a.rank = rank; a.seq = seq;
    a.fill = nullptr; a.fill_count = 0; a.fill_value = 0.f;
    f.progress = (const unsigned long long *)progress; f.expected = (const unsigned long long *)expected; f.product_seq = product_seq;
    f.fill = fill; f.fill_count = (int64_t)fill_count; f.fill_value = fill_value;
    if (ctas < 2) ctas = 2;
    if (ctas > 1024) ctas = 1024;
    // Launched with programmatic stream serialization on the stream of the product it belongs to, directly behind it: it
    // starts when every CTA of k_spmv_red has executed griddepcontrol.launch_dependents, i.e. is resident, and then runs
    // BESIDE it.  (Launched freely on a second stream it could occupy the SMs first; k_spmv_red needs another shared-memory
    // carve-out and would wait for them to drain while this kernel waits for its progress: a deadlock, seen on B200.)
    cudaStream_t st = stream ? (cudaStream_t)stream : g.stream;
    // an SM changes its shared-memory carve-out only when idle: ask for the carve-out the product kernel runs with (196 KB of
    // 228 = 86 %; option peer_carveout), or the CTAs of this kernel wait until the product's have left (no overlap, measured)
    static int carveout_set = -1;
    if (carveout_set != (int)g.peer_carveout) {
        cudaFuncSetAttribute(k_peer_exchange_fused, cudaFuncAttributePreferredSharedMemoryCarveout, (int)g.peer_carveout);
        carveout_set = (int)g.peer_carveout;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)ctas); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, k_peer_exchange_fused, f) != cudaSuccess) { cudaGetLastError(); return GrB_PANIC; }
    g.launches++;
    g.last_kernel = "peer_exchange_fused";
    return cudaGetLastError() == cudaSuccess ? GrB_SUCCESS : GrB_PANIC;
}

}  // extern "C"
