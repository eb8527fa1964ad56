// grb_gemm_launch.cuh — type-erased launchers of the SpGEMM kernels (grb_spgemm.cuh); the numeric kernels are
// instantiated per monoid in grb_gemm_inst_<MON>.cu.  The symbolic (pattern-only) kernels do not depend on the
// semiring and are instantiated once in grb_mxm.cu.
#pragma once
#include "grb_internal.cuh"

enum GemmKernel { GK_WARP = 0, GK_BLOCK = 1, GK_DENSE = 2, GK_PANEL = 3 };
struct GemmLaunch {
    int kernel;          // GemmKernel
    int grid, threads;
    int table;           // GK_BLOCK: hash slots (power of two)
    size_t smem;         // dynamic shared memory bytes
    const void *args;    // GemmArgs<T> of the semiring's type
};
// Return false if (mon, mul, type) has no kernel.
bool gemm_launch_numeric(int mon, int mul, int tcode, const GemmLaunch &L);
bool gemm_launch_dot(int mon, int mul, int tcode, const void *dot_args, int64_t mnnz);

#define GEMM_DECL(MON)                                                          \
    bool gemm_numeric_##MON(int mul, int tcode, const GemmLaunch &L);           \
    bool gemm_dot_##MON(int mul, int tcode, const void *dot_args, int64_t mnnz);
GEMM_DECL(PLUS) GEMM_DECL(MIN) GEMM_DECL(MAX) GEMM_DECL(ANY) GEMM_DECL(BOOLMON)
#undef GEMM_DECL
