// semiring kernels with the PLUS monoid (instantiated at build time; see grb_spmv.cuh)
#include "grb_spmv_inst.inc"
// plus_div under vxm: the kernels see (A entry, u entry), the semiring wants u / A -> reversed divide, every plus_div type
#define GRB_PLUS_RDIV_INTERNAL(X) X(RDIV, INT32) X(RDIV, INT64) X(RDIV, UINT32) X(RDIV, UINT64) X(RDIV, FP32) X(RDIV, FP64) \
    X(RDIV_N1, INT32) X(RDIV_N2, INT32) X(RDIV_N3, UINT32) X(RDIV_N4, UINT32)     /* narrow-integer plus_div under vxm */
bool spmv_pull_PLUS(int mul, int tcode, const void *pa, int mode) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_pull_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(pa, mode); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
    GRB_PLUS_RDIV_INTERNAL(X)
#undef X
    return false;
}
bool spmv_push_PLUS(int mul, int tcode, const void *pa) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_push_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(pa); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
    GRB_PLUS_RDIV_INTERNAL(X)
#undef X
    return false;
}
bool spmv_stream_PLUS(int mul, int tcode, const void *sa) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_stream_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(sa); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
    GRB_PLUS_RDIV_INTERNAL(X)
#undef X
    return false;
}
bool spmv_red_PLUS(int mul, int tcode, const void *ra) {
#define X(MUL, TYPE) if (mul == OP_##MUL && tcode == TC_##TYPE) { launch_red_t<Semiring<OP_PLUS, OP_##MUL, CT_##TYPE>>(ra); return true; }
    GRB_SEMIRINGS_OF_PLUS(X)
    GRB_PLUS_RDIV_INTERNAL(X)
#undef X
    return false;
}
