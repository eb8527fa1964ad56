// SpGEMM numeric + masked-dot kernels with the boolean monoids LOR / LAND / LXOR / LXNOR
#include "grb_gemm_inst.inc"
bool gemm_numeric_BOOLMON(int mul, int tcode, const GemmLaunch &L) {
    // the monoid arrives encoded in the upper bits of mul (mul | mon << 8)
    const int mon = mul >> 8; mul &= 0xff;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_gemm_numeric_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(L); return true; }
#define CUR OP_LOR
    GRB_SEMIRINGS_OF_LOR(X)
#undef CUR
#define CUR OP_LAND
    GRB_SEMIRINGS_OF_LAND(X)
#undef CUR
#define CUR OP_LXOR
    GRB_SEMIRINGS_OF_LXOR(X)
#undef CUR
#define CUR OP_LXNOR
    GRB_SEMIRINGS_OF_LXNOR(X)
#undef CUR
#undef X
    return false;
}
bool gemm_dot_BOOLMON(int mul, int tcode, const void *da, int64_t mnnz) {
    const int mon = mul >> 8; mul &= 0xff;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_dot_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(da, mnnz); return true; }
#define CUR OP_LOR
    GRB_SEMIRINGS_OF_LOR(X)
#undef CUR
#define CUR OP_LAND
    GRB_SEMIRINGS_OF_LAND(X)
#undef CUR
#define CUR OP_LXOR
    GRB_SEMIRINGS_OF_LXOR(X)
#undef CUR
#define CUR OP_LXNOR
    GRB_SEMIRINGS_OF_LXNOR(X)
#undef CUR
#undef X
    return false;
}
