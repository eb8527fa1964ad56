// grb_todo_stubs.cu — entry points declared in include/GraphBLAS_b200.h whose kernels are not written yet.
// They fail loudly with GrB_NOT_IMPLEMENTED (python: NotImplementedException); there is no CPU fallback.
// Each stub disappears from this file when its implementation lands.
#include "grb_internal.cuh"
#include "grb_abi_lists.inc"

#define NOT_YET(obj, what)                                                        \
    do {                                                                          \
        GB_Object *_o = (GB_Object *)(obj);                                       \
        if (_o && _o->magic == GRB_MAGIC) _o->err = what " is not implemented in libgrb_b200 yet"; \
        return GrB_NOT_IMPLEMENTED;                                               \
    } while (0)

extern "C" {
#ifndef HAVE_MXM
GrB_Info GrB_mxm(GrB_Matrix C, const GrB_Matrix, const GrB_BinaryOp, const GrB_Semiring, const GrB_Matrix,
                 const GrB_Matrix, const GrB_Descriptor) { NOT_YET(C, "GrB_mxm"); }
GrB_Info GrB_transpose(GrB_Matrix C, const GrB_Matrix, const GrB_BinaryOp, const GrB_Matrix, const GrB_Descriptor) { NOT_YET(C, "GrB_transpose"); }
GrB_Info GrB_Matrix_eWiseMult_BinaryOp(GrB_Matrix C, const GrB_Matrix, const GrB_BinaryOp, const GrB_BinaryOp,
                                       const GrB_Matrix, const GrB_Matrix, const GrB_Descriptor) { NOT_YET(C, "GrB_Matrix_eWiseMult"); }
GrB_Info GrB_Matrix_eWiseAdd_BinaryOp(GrB_Matrix C, const GrB_Matrix, const GrB_BinaryOp, const GrB_BinaryOp,
                                      const GrB_Matrix, const GrB_Matrix, const GrB_Descriptor) { NOT_YET(C, "GrB_Matrix_eWiseAdd"); }
GrB_Info GrB_Matrix_select_INT64(GrB_Matrix C, const GrB_Matrix, const GrB_BinaryOp, const GrB_IndexUnaryOp,
                                 const GrB_Matrix, int64_t, const GrB_Descriptor) { NOT_YET(C, "GrB_Matrix_select"); }
GrB_Info GxB_b200_Matrix_rowblock(GrB_Matrix *, const GrB_Matrix A, GrB_Index, GrB_Index) { NOT_YET(A, "GxB_b200_Matrix_rowblock"); }
#endif
#ifndef HAVE_RMAT
GrB_Info GxB_b200_Matrix_rmat(GrB_Matrix *, GrB_Type, int, int, uint64_t, int, int) { return GrB_NOT_IMPLEMENTED; }
GrB_Info GxB_b200_rmat_edges_host(uint64_t *, uint64_t *, int, uint64_t, uint64_t) { return GrB_NOT_IMPLEMENTED; }
#endif
}
