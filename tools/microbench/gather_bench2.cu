// gather_bench2.cu — how the cost of a divergent gather depends on load flavour, active lanes, sector sharing, footprint
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t mix(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(x ^ (x >> 31));
}
// idx pattern: mode 0 uniform over n; mode 1: warp-clustered (each group of `share` consecutive lanes hits one 32B sector)
__global__ void k_init(uint32_t *col, int64_t nnz, uint32_t n, int share) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < nnz; p += (int64_t)gridDim.x * blockDim.x) {
        if (share <= 1) col[p] = mix(p * 7 + 1) % n;
        else { int64_t g = p / share; col[p] = ((mix(g * 7 + 1) % (n / 8)) * 8 + (p % share) % 8); }
    }
}
template <int FL> __device__ __forceinline__ float ldf(const float *p) {
    float v;
    if (FL == 0) v = __ldg(p);
    else if (FL == 1) v = __ldcg(p);
    else if (FL == 2) v = *(volatile const float *)p;
    else if (FL == 3) asm volatile("ld.global.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (FL == 4) v = __ldcs(p);
    else asm volatile("ld.global.nc.L1::evict_last.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
template <int FL, int HALF>
__global__ void k_gather(const uint32_t *col, const float *u, int64_t nnz, float *out) {
    float acc = 0;
    const int EPT = 8;
    const int64_t tile = (int64_t)blockDim.x * EPT;
    for (int64_t base = blockIdx.x * tile; base < nnz; base += (int64_t)gridDim.x * tile) {
        uint32_t c[EPT]; float x[EPT];
#pragma unroll
        for (int k = 0; k < EPT; k++) { int64_t p = base + k * blockDim.x + threadIdx.x; c[k] = p < nnz ? __ldg(col + p) : 0; }
#pragma unroll
        for (int k = 0; k < EPT; k++) { x[k] = 0; if (!HALF || (threadIdx.x & 1)) x[k] = ldf<FL>(u + c[k]); }
#pragma unroll
        for (int k = 0; k < EPT; k++) acc += x[k];
    }
    if (acc == 123.456f) out[0] = acc;
}
template <int FL, int HALF>
static void run(const char *name, const uint32_t *col, const float *u, int64_t nnz, float *out) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; i++) k_gather<FL, HALF><<<148 * 8, 256>>>(col, u, nnz, out);
    cudaEventRecord(a);
    for (int i = 0; i < 5; i++) k_gather<FL, HALF><<<148 * 8, 256>>>(col, u, nnz, out);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 5;
    printf("  %-40s %8.1f us  %7.1f Ggathers/s (%s)\n", name, ms * 1e3, nnz / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    const int64_t nnz = 65244187;
    uint32_t *col; float *u, *out;
    cudaMalloc(&col, nnz * 4); cudaMalloc(&u, (64u << 20) * 4); cudaMalloc(&out, 4);
    cudaMemset(u, 0, (64u << 20) * 4);
    for (uint32_t n : {1u << 15, 1u << 18, 1u << 22, 1u << 24, 1u << 26}) {
        k_init<<<148 * 8, 256>>>(col, nnz, n, 1); cudaDeviceSynchronize();
        printf("footprint %u floats (%.1f MB), uniform random\n", n, n * 4.0 / 1e6);
        run<0, 0>("__ldg", col, u, nnz, out);
        if (n == (1u << 22)) {
            run<1, 0>("__ldcg", col, u, nnz, out);
            run<2, 0>("volatile", col, u, nnz, out);
            run<3, 0>("L1::no_allocate", col, u, nnz, out);
            run<4, 0>("__ldcs", col, u, nnz, out);
            run<5, 0>("nc.L1::evict_last", col, u, nnz, out);
            run<0, 1>("__ldg, odd lanes only", col, u, nnz, out);
        }
    }
    for (int share : {2, 4, 8, 32}) {
        k_init<<<148 * 8, 256>>>(col, nnz, 1u << 22, share); cudaDeviceSynchronize();
        printf("footprint 4M floats, %d consecutive lanes share a sector\n", share);
        run<0, 0>("__ldg", col, u, nnz, out);
    }
    return 0;
}
