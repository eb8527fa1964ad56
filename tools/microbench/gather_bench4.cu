// gather_bench4.cu — stream staged by per-warp TMA bulk copies (cp.async.bulk + mbarrier), hubs in shared memory,
// non-hub gathers through LSU: the structure of k_spmv_red2.  Reports time for STAGES = 1..3.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(x ^ (x >> 31));
}
__global__ void k_init(uint32_t *col, float *val, int64_t nnz, uint32_t n, int hub_pct, uint32_t hubs) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        uint32_t c = mix(p * 7 + 1) % n;
        if ((int)(mix(p) % 100) < hub_pct) c = 0x80000000u | (mix(p * 3 + 5) % hubs);
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
template <int WARPS, int STAGES>
__global__ void __launch_bounds__(WARPS * 32, 1) k_tma(const uint32_t *col, const float *val, const float *u, int64_t nnz, float *out, int hubs) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float *hub = (float *)smem;
    unsigned char *stage_base = smem + (((size_t)hubs * 4 + 127) & ~(size_t)127);
    // per warp: STAGES x (1 KB col + 1 KB val); barriers after all stages
    unsigned char *my = stage_base + (size_t)warp * STAGES * 2048;
    uint64_t *bars = (uint64_t *)(stage_base + (size_t)WARPS * STAGES * 2048) + warp * STAGES;
    for (int i = tid; i < hubs; i += WARPS * 32) hub[i] = u[i];
    if (lane == 0) for (int s = 0; s < STAGES; s++) mbar_init(smem_u32(bars + s), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const int64_t ntiles = nnz / TILE;   // ragged tail ignored in this benchmark
    const int64_t tstep = (int64_t)gridDim.x * WARPS;
    int64_t t = blockIdx.x * (int64_t)WARPS + warp;
    // prologue: fill the pipeline
    if (lane == 0) {
        for (int s = 0; s < STAGES; s++) {
            const int64_t tt = t + s * tstep;
            if (tt < ntiles) {
                mbar_expect_tx(smem_u32(bars + s), 2048);
                tma_load_1d(smem_u32(my + s * 2048), col + tt * TILE, 1024, smem_u32(bars + s));
                tma_load_1d(smem_u32(my + s * 2048 + 1024), val + tt * TILE, 1024, smem_u32(bars + s));
            }
        }
    }
    float acc = 0;
    int s = 0; uint32_t parity = 0;
    for (; t < ntiles; t += tstep) {
        mbar_wait(smem_u32(bars + s), parity);
        const uint4 *pc = (const uint4 *)(my + s * 2048) + lane * 2;
        const float4 *pv = (const float4 *)(my + s * 2048 + 1024) + lane * 2;
        uint4 c0 = pc[0], c1 = pc[1];
        float4 v0 = pv[0], v1 = pv[1];
        __syncwarp();
        // refill this stage with the tile STAGES ahead
        if (lane == 0) {
            const int64_t tt = t + STAGES * tstep;
            if (tt < ntiles) {
                mbar_expect_tx(smem_u32(bars + s), 2048);
                tma_load_1d(smem_u32(my + s * 2048), col + tt * TILE, 1024, smem_u32(bars + s));
                tma_load_1d(smem_u32(my + s * 2048 + 1024), val + tt * TILE, 1024, smem_u32(bars + s));
            }
        }
        uint32_t c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
        float x[8];
#pragma unroll
        for (int k = 0; k < 8; k++) { x[k] = 0; if ((int)c[k] >= 0) x[k] = __ldg(u + c[k]); }
#pragma unroll
        for (int k = 0; k < 8; k++) if ((int)c[k] < 0) x[k] = hub[c[k] & 0x7fffffffu];
#pragma unroll
        for (int k = 0; k < 8; k++) acc += v[k] * x[k];
        if (++s == STAGES) { s = 0; parity ^= 1; }
    }
    if (acc == 123.456f) out[0] = acc;
}
template <int WARPS, int STAGES>
static void run(const char *name, const uint32_t *col, const float *val, const float *u, int64_t nnz, float *out, int hubs) {
    const size_t smem = (((size_t)hubs * 4 + 127) & ~(size_t)127) + (size_t)WARPS * STAGES * 2048 + WARPS * STAGES * 8 + 128;
    cudaFuncSetAttribute(k_tma<WARPS, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) k_tma<WARPS, STAGES><<<148, WARPS * 32, smem>>>(col, val, u, nnz, out, hubs);
    cudaEventRecord(a);
    for (int i = 0; i < 5; i++) k_tma<WARPS, STAGES><<<148, WARPS * 32, smem>>>(col, val, u, nnz, out, hubs);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 5;
    printf("  %-36s warps=%2d stages=%d smem=%6zu hubs=%5d  %8.1f us (%s)\n", name, WARPS, STAGES, smem, hubs, ms * 1e3, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    const int64_t nnz = 65244187; const uint32_t n = 1u << 22;
    uint32_t *col; float *val, *u, *out;
    cudaMalloc(&col, nnz * 4); cudaMalloc(&val, nnz * 4); cudaMalloc(&u, n * 4); cudaMalloc(&out, 4);
    cudaMemset(u, 0, n * 4);
    for (int pct : {50, 0, 100}) {
        k_init<<<148 * 8, 256>>>(col, val, nnz, n, pct, 16384); cudaDeviceSynchronize();
        printf("hub share %d%%\n", pct);
        run<32, 1>("TMA stream + hubs + gather", col, val, u, nnz, out, 16384);
        run<32, 2>("TMA stream + hubs + gather", col, val, u, nnz, out, 16384);
        run<32, 3>("TMA stream + hubs + gather", col, val, u, nnz, out, 8192 * 0 + 4096);
        run<16, 4>("TMA stream + hubs + gather", col, val, u, nnz, out, 16384);
    }
    return 0;
}
