// gather_bench5.cu — builds k_spmv_red up from the TMA skeleton of gather_bench4 one feature at a time, to see
// which feature costs what on B200.  Feature bits (template F):
//   1  fence.proxy.async before the stage is re-armed
//   2  third bulk copy per tile (128 B chunk_first) + its LDS
//   4  dependent row-id loads  nzrow[j0 .. j0+2]
//   8  segmented reduction between row-start flags + one RED per segment (else: plain sum, no output)
//  16  single-row-tile check (shfl + vote)
//  32  values are NOT staged (pattern-only semiring: half the stream)
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>
__device__ __host__ __forceinline__ uint32_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(x ^ (x >> 31));
}
__global__ void k_init(uint32_t *col, float *val, int64_t nnz, uint32_t n, int hub_pct, uint32_t hubs) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        uint32_t c = mix(p * 7 + 1) % n;
        if ((int)(mix(p) % 100) < hub_pct) c = 0x80000000u | (mix(p * 3 + 5) % hubs);
        if (p == 0 || mix(p * 11 + 3) % 16 == 0) c |= 0x40000000u;      // row start (mean row length 16)
        col[p] = c; val[p] = 1.0f;
    }
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void tma_load_1d(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
}
constexpr int TILE = 256;
constexpr uint32_t CMASK = 0x3fffffffu;
template <int WARPS, int F>
__global__ void __launch_bounds__(WARPS * 32, 1) k_red(const uint32_t *col, const float *val, const float *u, int64_t nnz, float *w,
                                                       const uint32_t *cfirst, const uint32_t *nzrow, int hubs) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr uint32_t STAGE = 2048 + 128;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float *hub = (float *)smem;
    unsigned char *stage_base = smem + (((size_t)hubs * 4 + 127) & ~(size_t)127);
    unsigned char *my = stage_base + (size_t)warp * STAGE;
    const uint32_t bar = smem_u32((uint64_t *)(stage_base + (size_t)WARPS * STAGE) + warp);
    const uint32_t my_a = smem_u32(my);
    for (int i = tid; i < hubs; i += WARPS * 32) hub[i] = u[i];
    if (lane == 0) mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const int64_t ntiles = nnz / TILE;   // ragged tail ignored in this benchmark
    const int64_t tstep = (int64_t)gridDim.x * WARPS;
    int64_t t = blockIdx.x * (int64_t)WARPS + warp;
    auto issue = [&](int64_t tt) {
        mbar_expect_tx(bar, 1024 + ((F & 32) ? 0 : 1024) + ((F & 2) ? 128 : 0));
        tma_load_1d(my_a, col + tt * TILE, 1024, bar);
        if (!(F & 32)) tma_load_1d(my_a + 1024, val + tt * TILE, 1024, bar);
        if (F & 2) tma_load_1d(my_a + 2048, cfirst + tt * 32, 128, bar);
    };
    if (lane == 0 && t < ntiles) issue(t);
    float sink = 0;
    uint32_t parity = 0;
    for (; t < ntiles; t += tstep) {
        mbar_wait(bar, parity); parity ^= 1;
        const uint4 *pc = (const uint4 *)my + lane * 2;
        const float4 *pv = (const float4 *)(my + 1024) + lane * 2;
        uint4 c0 = pc[0], c1 = pc[1];
        float4 v0 = make_float4(1, 1, 1, 1), v1 = v0;
        if (!(F & 32)) { v0 = pv[0]; v1 = pv[1]; }
        uint32_t j0 = (uint32_t)(t * 32 + lane);
        if (F & 2) j0 = ((const uint32_t *)(my + 2048))[lane];
        if (F & 1) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0 && t + tstep < ntiles) issue(t + tstep);
        uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        unsigned fbits = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) fbits |= ((c[k] >> 30) & 1u) << k;
        const int nseg = __popc(fbits >> 1);
        uint32_t r0 = j0, r1 = j0 + 1, r2 = j0 + 2;
        if (F & 4) { r0 = nzrow[j0]; if (nseg >= 1) r1 = nzrow[j0 + 1]; if (nseg >= 2) r2 = nzrow[j0 + 2]; }
        float x[8];
#pragma unroll
        for (int k = 0; k < 8; k++) { x[k] = 0; if ((int)c[k] >= 0) x[k] = __ldg(u + (c[k] & CMASK)); }
#pragma unroll
        for (int k = 0; k < 8; k++) if ((int)c[k] < 0) x[k] = hub[c[k] & CMASK];
        if (F & 16) {
            const uint32_t jl = __shfl_sync(0xffffffffu, j0, 0);
            if (__all_sync(0xffffffffu, (fbits >> 1) == 0 && j0 == jl)) {
                float a = 0;
#pragma unroll
                for (int k = 0; k < 8; k++) a += v[k] * x[k];
#pragma unroll
                for (int d = 16; d; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
                if (lane == 0) atomicAdd(w + r0, a);
                continue;
            }
        }
        if (F & 8) {
            float acc = v[0] * x[0];
            int m = 0; uint32_t r = r0;
#pragma unroll
            for (int k = 1; k < 8; k++) {
                const float z = v[k] * x[k];
                if (fbits & (1u << k)) {
                    atomicAdd(w + r, acc);
                    m++;
                    r = (m == 1) ? r1 : (m == 2) ? r2 : ((F & 4) ? nzrow[j0 + m] : j0 + m);
                    acc = z;
                } else acc += z;
            }
            atomicAdd(w + r, acc);
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) sink += v[k] * x[k];
            sink += (float)(r0 + r1 + r2);
        }
    }
    if (sink == 123.456f) w[0] = sink;
}
template <int WARPS, int F>
static void run(const uint32_t *col, const float *val, const float *u, int64_t nnz, float *w, const uint32_t *cf, const uint32_t *nzrow, int hubs) {
    const size_t smem = (((size_t)hubs * 4 + 127) & ~(size_t)127) + (size_t)WARPS * (2048 + 128) + WARPS * 8 + 128;
    cudaFuncSetAttribute(k_red<WARPS, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) k_red<WARPS, F><<<148, WARPS * 32, smem>>>(col, val, u, nnz, w, cf, nzrow, hubs);
    cudaEventRecord(a);
    for (int i = 0; i < 5; i++) k_red<WARPS, F><<<148, WARPS * 32, smem>>>(col, val, u, nnz, w, cf, nzrow, hubs);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 5;
    printf("  features=%2d warps=%2d smem=%6zu hubs=%5d  %8.1f us (%s)\n", F, WARPS, smem, hubs, ms * 1e3, cudaGetErrorString(cudaGetLastError()));
    fflush(stdout);
}
int main() {
    const int64_t nnz = 65244187; const uint32_t n = 1u << 22;
    uint32_t *col, *cf, *nzrow; float *val, *u, *w;
    cudaMalloc(&col, nnz * 4); cudaMalloc(&val, nnz * 4); cudaMalloc(&u, n * 4); cudaMalloc(&w, (size_t)n * 4 * 2);
    cudaMalloc(&cf, (nnz / 8 + 64) * 4); cudaMalloc(&nzrow, (size_t)(nnz / 8) * 4 + 1024);
    cudaMemset(u, 0, n * 4); cudaMemset(w, 0, (size_t)n * 8);
    const int hubs = 32768;
    for (int pct : {55, 100}) {
        k_init<<<148 * 8, 256>>>(col, val, nnz, n, pct, hubs); cudaDeviceSynchronize();
        // chunk_first on the host: index of the row that owns entry 8*i
        std::vector<uint32_t> hc(nnz), hcf(nnz / 8 + 64, 0);
        cudaMemcpy(hc.data(), col, nnz * 4, cudaMemcpyDeviceToHost);
        uint32_t rows = 0;
        for (int64_t p = 0; p < nnz; p++) {
            if (hc[p] & 0x40000000u) rows++;
            if ((p & 7) == 0) hcf[p >> 3] = rows - 1;
        }
        cudaMemcpy(cf, hcf.data(), hcf.size() * 4, cudaMemcpyHostToDevice);
        std::vector<uint32_t> hr(rows + 64);
        for (uint32_t i = 0; i < hr.size(); i++) hr[i] = i;        // rows are sorted: row id grows with the stream
        cudaMemcpy(nzrow, hr.data(), hr.size() * 4, cudaMemcpyHostToDevice);
        printf("hub share %d%%, rows %u\n", pct, rows);
        run<32, 0>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 1>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 3>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 7>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 15>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 31>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 8>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 9>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 11>(col, val, u, nnz, w, cf, nzrow, hubs);
        run<32, 63>(col, val, u, nnz, w, cf, nzrow, hubs);
    }
    return 0;
}
