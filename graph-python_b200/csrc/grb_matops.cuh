// grb_matops.cuh — matrix write-back shared by GrB_mxm / GrB_transpose / matrix eWise / select (grb_matops.cu)
#pragma once
#include "grb_internal.cuh"

// C<M,replace> = accum(C, T) with C API 2.0 semantics (SURVEY 8c "What a CPU restatement must follow").
// T holds values of type tcode and is consumed (freed or adopted).  t_in_mask: T's pattern is already a subset of
// the allowed positions (masked dot product), which enables the no-merge fast path.
void mat_writeback(GB_Object *C, GB_Object *mask, const DescFlags &f, int accum, Csr &T, int tcode, bool t_in_mask);
int accum_code_of(const GrB_BinaryOp accum);
// exclusive scan of int64 counts[n+1] (counts[n] ignored/zeroed) into out[n+1]; returns out[n] (synchronises)
int64_t scan_counts(int64_t *counts, int64_t *out, int64_t n);
