// semiring kernels with the boolean monoids LOR / LAND / LXOR / LXNOR
#include "grb_spmv_inst.inc"
static int g_mon;
bool spmv_pull_BOOLMON(int mul, int tcode, const void *pa, int mode) {
    // the monoid arrives encoded in the upper bits of mul (mul | mon << 8)
    int mon = mul >> 8; mul &= 0xff; (void)g_mon;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_pull_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(pa, mode); return true; }
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
bool spmv_push_BOOLMON(int mul, int tcode, const void *pa) {
    int mon = mul >> 8; mul &= 0xff;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_push_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(pa); return true; }
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
bool spmv_stream_BOOLMON(int mul, int tcode, const void *sa) {
    int mon = mul >> 8; mul &= 0xff;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_stream_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(sa); return true; }
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
bool spmv_red_BOOLMON(int mul, int tcode, const void *ra) {
    int mon = mul >> 8; mul &= 0xff;
#define X(MUL, TYPE) if (mon == CUR && mul == OP_##MUL && tcode == TC_##TYPE) { launch_red_t<Semiring<CUR, OP_##MUL, CT_##TYPE>>(ra); return true; }
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
