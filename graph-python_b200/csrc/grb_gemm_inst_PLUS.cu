// SpGEMM numeric + masked-dot kernels with the PLUS monoid (instantiated at build time; see grb_spgemm.cuh)
#include "grb_gemm_inst.inc"
bool gemm_numeric_PLUS(int mul, int tcode, const GemmLaunch &L) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_gemm_numeric_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(L); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
#undef X
    return false;
}
bool gemm_dot_PLUS(int mul, int tcode, const void *da, int64_t mnnz) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_dot_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(da, mnnz); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
#undef X
    return false;
}
