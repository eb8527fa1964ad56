// gather_bench3.cu — 50% of the gathers are "hubs": where should they be served from?
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(x ^ (x >> 31));
}
__global__ void k_init(uint32_t *col, int64_t nnz, uint32_t n, int hub_pct, uint32_t hubs) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        uint32_t c = mix(p * 7 + 1) % n;
        if ((int)(mix(p) % 100) < hub_pct) c = 0x80000000u | (mix(p * 3 + 5) % hubs);
        col[p] = c;
    }
}
// HUBSRC 0: ALU only (hub lanes do no memory access), 1: shared memory, 2: small global array (L1), 3: skip hub lanes entirely but keep smem allocated
template <int HUBSRC>
__global__ void k_gather(const uint32_t *col, const float *u, const float *hubg, int64_t nnz, float *out, int hubs) {
    extern __shared__ float hub[];
    if (HUBSRC == 1) { for (int i = threadIdx.x; i < hubs; i += blockDim.x) hub[i] = hubg[i]; __syncthreads(); }
    float acc = 0;
    const int EPT = 8;
    const int64_t tile = (int64_t)blockDim.x * EPT;
    for (int64_t base = blockIdx.x * tile; base < nnz; base += (int64_t)gridDim.x * tile) {
        uint32_t c[EPT]; float x[EPT];
#pragma unroll
        for (int k = 0; k < EPT; k++) { int64_t p = base + k * blockDim.x + threadIdx.x; c[k] = p < nnz ? __ldg(col + p) : 0; }
#pragma unroll
        for (int k = 0; k < EPT; k++) { x[k] = 0; if ((int)c[k] >= 0) x[k] = __ldg(u + c[k]); }
#pragma unroll
        for (int k = 0; k < EPT; k++) {
            if ((int)c[k] < 0) {
                const uint32_t h = c[k] & 0x7fffffffu;
                if (HUBSRC == 0 || HUBSRC == 3) x[k] = (float)h;
                else if (HUBSRC == 1) x[k] = hub[h];
                else x[k] = __ldg(hubg + h);
            }
        }
#pragma unroll
        for (int k = 0; k < EPT; k++) acc += x[k];
    }
    if (acc == 123.456f) out[0] = acc;
}
template <int HUBSRC>
static void run(const char *name, int grid, int block, size_t smem, const uint32_t *col, const float *u, const float *hubg, int64_t nnz, float *out, int hubs) {
    cudaFuncSetAttribute(k_gather<HUBSRC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) k_gather<HUBSRC><<<grid, block, smem>>>(col, u, hubg, nnz, out, hubs);
    cudaEventRecord(a);
    for (int i = 0; i < 5; i++) k_gather<HUBSRC><<<grid, block, smem>>>(col, u, hubg, nnz, out, hubs);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 5;
    printf("  %-52s grid=%4d block=%4d smem=%6zu  %8.1f us (%s)\n", name, grid, block, smem, ms * 1e3, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    const int64_t nnz = 65244187; const uint32_t n = 1u << 22;
    uint32_t *col; float *u, *hubg, *out;
    cudaMalloc(&col, nnz * 4); cudaMalloc(&u, n * 4); cudaMalloc(&hubg, 1 << 20); cudaMalloc(&out, 4);
    cudaMemset(u, 0, n * 4); cudaMemset(hubg, 0, 1 << 20);
    for (int hubs : {26624, 4096}) {
        k_init<<<148 * 8, 256>>>(col, nnz, n, 50, hubs); cudaDeviceSynchronize();
        printf("50%% hubs over %d entries\n", hubs);
        run<0>("hub lanes: ALU only, no smem", 148 * 8, 256, 0, col, u, hubg, nnz, out, hubs);
        run<3>("hub lanes: ALU only, 100 KB smem allocated x2 CTA", 148 * 2, 1024, 100 * 1024, col, u, hubg, nnz, out, hubs);
        run<3>("hub lanes: ALU only, 100 KB smem allocated x1 CTA", 148, 1024, 100 * 1024, col, u, hubg, nnz, out, hubs);
        if (hubs * 4 <= 24 * 1024) run<1>("hub lanes: shared memory, 256 thr, 8 CTA/SM", 148 * 8, 256, hubs * 4, col, u, hubg, nnz, out, hubs);
        run<1>("hub lanes: shared memory, 1024 thr x 1 CTA/SM", 148, 1024, hubs * 4, col, u, hubg, nnz, out, hubs);
        run<1>("hub lanes: shared memory, 1024 thr x 2 CTA/SM", 148 * 2, 1024, hubs * 4, col, u, hubg, nnz, out, hubs);
        run<2>("hub lanes: small global array (L1), 256 thr x8", 148 * 8, 256, 0, col, u, hubg, nnz, out, hubs);
    }
    return 0;
}
