// gather_bench.cu — what the B200 memory system gives a pull SpMV: stream-only, stream + random gather, and
// gather with a share of the indices served from shared memory.  Run: nvcc -arch=sm_100a -O3 gather_bench.cu && ./a.out
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(x ^ (x >> 31));
}
__global__ void k_init(uint32_t *col, float *val, int64_t nnz, uint32_t n, int hub_pct, uint32_t hubs) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        uint32_t h = mix(p);
        uint32_t c = mix(p * 7 + 1) % n;
        if ((int)(h % 100) < hub_pct) c = 0x80000000u | (mix(p * 3 + 5) % hubs);
        col[p] = c; val[p] = 1.0f;
    }
}
// EPT entries per thread, coalesced (k*blockDim + tid); MODE 0 stream only, 1 gather, 2 gather + smem hubs
template <int EPT, int MODE>
__global__ void k_bench(const uint32_t *col, const float *val, const float *u, int64_t nnz, float *out, int hubs) {
    extern __shared__ float hub[];
    if (MODE == 2) { for (int i = threadIdx.x; i < hubs; i += blockDim.x) hub[i] = u[i]; __syncthreads(); }
    float acc = 0;
    const int64_t tile = (int64_t)blockDim.x * EPT;
    for (int64_t base = blockIdx.x * tile; base < nnz; base += (int64_t)gridDim.x * tile) {
        uint32_t c[EPT]; float v[EPT], x[EPT];
#pragma unroll
        for (int k = 0; k < EPT; k++) { int64_t p = base + k * blockDim.x + threadIdx.x; c[k] = p < nnz ? __ldg(col + p) : 0; v[k] = p < nnz ? __ldg(val + p) : 0; }
#pragma unroll
        for (int k = 0; k < EPT; k++) {
            if (MODE == 0) x[k] = (float)(c[k] & 1);
            else if (MODE == 1) x[k] = __ldg(u + (c[k] & 0x3fffffu));
            else { x[k] = 0; if ((int)c[k] >= 0) x[k] = __ldg(u + c[k]); }
        }
        if (MODE == 2) {
#pragma unroll
            for (int k = 0; k < EPT; k++) if ((int)c[k] < 0) x[k] = hub[c[k] & 0x7fffffffu];
        }
#pragma unroll
        for (int k = 0; k < EPT; k++) acc += v[k] * x[k];
    }
    if (acc == 123.456f) out[0] = acc;
}
template <int EPT, int MODE>
static float run(const char *name, int grid, int block, size_t smem, const uint32_t *col, const float *val, const float *u, int64_t nnz, float *out, int hubs) {
    if (smem > 48 * 1024) cudaFuncSetAttribute(k_bench<EPT, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; i++) k_bench<EPT, MODE><<<grid, block, smem>>>(col, val, u, nnz, out, hubs);
    cudaEventRecord(a);
    for (int i = 0; i < 10; i++) k_bench<EPT, MODE><<<grid, block, smem>>>(col, val, u, nnz, out, hubs);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
    printf("%-44s grid=%5d block=%4d  %8.1f us  %7.1f Gedges/s  stream %6.0f GB/s  err=%s\n", name, grid, block, ms * 1e3, nnz / ms / 1e6, nnz * 8.0 / ms / 1e6,
           cudaGetErrorString(cudaGetLastError()));
    return ms;
}
int main() {
    const int64_t nnz = 65244187; const uint32_t n = 1u << 22;
    uint32_t *col; float *val, *u, *out;
    CK(cudaMalloc(&col, nnz * 4)); CK(cudaMalloc(&val, nnz * 4)); CK(cudaMalloc(&u, n * 4)); CK(cudaMalloc(&out, 4));
    CK(cudaMemset(u, 0, n * 4));
    const int hubs = 26624;
    for (int hub_pct : {0, 50}) {
        k_init<<<148 * 8, 256>>>(col, val, nnz, n, hub_pct, hubs);
        CK(cudaDeviceSynchronize());
        printf("---- hub share %d%% ----\n", hub_pct);
        if (hub_pct == 0) {
            run<8, 0>("stream only, 8/thread, 256 thr x 8 CTA/SM", 148 * 8, 256, 0, col, val, u, nnz, out, 0);
            run<8, 0>("stream only, 8/thread, full grid", (int)((nnz + 2047) / 2048), 256, 0, col, val, u, nnz, out, 0);
            run<8, 1>("gather, 8/thread, 256 thr x 8 CTA/SM", 148 * 8, 256, 0, col, val, u, nnz, out, 0);
            run<8, 1>("gather, 8/thread, full grid", (int)((nnz + 2047) / 2048), 256, 0, col, val, u, nnz, out, 0);
            run<4, 1>("gather, 4/thread, 256 thr x 8 CTA/SM", 148 * 8, 256, 0, col, val, u, nnz, out, 0);
            run<16, 1>("gather, 16/thread, 256 thr x 4 CTA/SM", 148 * 4, 256, 0, col, val, u, nnz, out, 0);
            run<8, 1>("gather, 8/thread, 1024 thr x 1 CTA/SM", 148, 1024, 0, col, val, u, nnz, out, 0);
            run<8, 1>("gather, 8/thread, 1024 thr x 2 CTA/SM", 148 * 2, 1024, 0, col, val, u, nnz, out, 0);
        } else {
            run<8, 2>("gather+smem hubs, 8/thr, 1024 thr x 1 CTA/SM", 148, 1024, hubs * 4, col, val, u, nnz, out, hubs);
            run<8, 2>("gather+smem hubs, 8/thr, 1024 thr x 2 CTA/SM", 148 * 2, 1024, hubs * 4, col, val, u, nnz, out, hubs);
            run<8, 2>("gather+smem hubs, 8/thr, 768 thr x 2 CTA/SM", 148 * 2, 768, hubs * 4, col, val, u, nnz, out, hubs);
            run<16, 2>("gather+smem hubs, 16/thr, 512 thr x 2 CTA/SM", 148 * 2, 512, hubs * 4, col, val, u, nnz, out, hubs);
        }
    }
    return 0;
}
