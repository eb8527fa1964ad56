#!/usr/bin/env python3
"""Write tests/golden/reference_kats.json: the literal known-answer vectors the reference's own
tests hold for GrB_mxm / GrB_mxv / GrB_vxm.

The reference cannot be executed here (its arithmetic is in SuiteSparse:GraphBLAS, absent from the
image), so these are transcriptions of the literals in /root/reference/graphblas/tests, each with
its file:line.  They are data (COO triples), not code.  All values INT64 unless stated.
"""
import json
import os

A = dict(  # graphblas/tests/test_matrix.py:34-49 (same fixture test_vector.py:34-41)
    rows=[3, 0, 3, 5, 6, 0, 6, 1, 6, 2, 4, 1],
    cols=[0, 1, 2, 2, 2, 3, 3, 4, 4, 5, 5, 6],
    vals=[3, 2, 3, 1, 5, 3, 7, 8, 3, 1, 7, 4],
    nrows=7, ncols=7,
)
v = dict(idx=[1, 3, 4, 6], vals=[1, 1, 2, 0], size=7)  # test_matrix.py:52-55, test_vector.py:44-47

KATS = [
    dict(name="test_mxm", cite="graphblas/tests/test_matrix.py:307-314", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C=None,
         expect=dict(rows=[0, 0, 0, 0, 1, 1, 1, 1, 2, 3, 3, 3, 4, 5, 6, 6, 6],
                     cols=[0, 2, 4, 6, 2, 3, 4, 5, 2, 1, 3, 5, 2, 5, 0, 2, 5],
                     vals=[9, 9, 16, 8, 20, 28, 12, 56, 1, 6, 9, 3, 7, 1, 21, 21, 26])),
    dict(name="test_mxm_transpose_T1", cite="graphblas/tests/test_matrix.py:317-325", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C="A", tran1=True,
         expect=dict(rows=[0, 0, 1, 1, 2, 2, 3, 3, 3, 4, 4, 5, 5, 5, 6, 6, 6, 6, 6],
                     cols=[0, 6, 1, 6, 2, 4, 3, 5, 6, 2, 4, 3, 5, 6, 0, 1, 3, 5, 6],
                     vals=[13, 21, 80, 24, 1, 7, 18, 3, 15, 7, 49, 3, 1, 5, 21, 24, 15, 5, 83])),
    dict(name="test_mxm_transpose_T0", cite="graphblas/tests/test_matrix.py:326-332", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C="A", tran0=True,
         expect=dict(rows=[0, 0, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 6, 6],
                     cols=[0, 2, 1, 3, 0, 2, 3, 4, 1, 2, 3, 4, 2, 3, 4, 6, 5, 4, 6],
                     vals=[9, 9, 4, 6, 9, 35, 35, 15, 6, 35, 58, 21, 15, 21, 73, 32, 50, 32, 16])),
    dict(name="test_mxm_nonsquare", cite="graphblas/tests/test_matrix.py:335-341", op="mxm",
         semiring="MAX_PLUS",
         A=dict(rows=[0, 0, 0], cols=[0, 2, 4], vals=[1, 2, 3], nrows=1, ncols=5),
         B=dict(rows=[0, 2, 4], cols=[0, 0, 0], vals=[10, 20, 30], nrows=5, ncols=1), C=None,
         expect=dict(rows=[0], cols=[0], vals=[33])),
    dict(name="test_mxm_mask_value", cite="graphblas/tests/test_matrix.py:348-359", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C="A",
         mask=dict(rows=[0, 3, 4], cols=[2, 3, 2], vals=[1, 1, 1], dtype="BOOL"),
         expect=dict(rows=[0, 0, 0, 1, 1, 2, 3, 3, 3, 4, 4, 5, 6, 6, 6],
                     cols=[1, 2, 3, 4, 6, 5, 0, 2, 3, 2, 5, 2, 2, 3, 4],
                     vals=[2, 9, 3, 8, 4, 1, 3, 3, 9, 7, 7, 1, 5, 7, 3])),
    dict(name="test_mxm_mask_value_comp", cite="graphblas/tests/test_matrix.py:360-367", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C="A", comp=True,
         mask=dict(rows=[0, 3, 4], cols=[2, 3, 2], vals=[1, 1, 1], dtype="BOOL"),
         expect=dict(rows=[0, 0, 0, 1, 1, 1, 1, 2, 3, 3, 5, 6, 6, 6],
                     cols=[0, 4, 6, 2, 3, 4, 5, 2, 1, 5, 5, 0, 2, 5],
                     vals=[9, 16, 8, 20, 28, 12, 56, 1, 6, 3, 1, 21, 21, 26])),
    dict(name="test_mxm_mask_struct_replace", cite="graphblas/tests/test_matrix.py:349,368-371",
         op="mxm", semiring="PLUS_TIMES", A="A", B="A", C="A", structure=True, replace=True,
         mask=dict(rows=[0, 3, 4], cols=[2, 3, 2], vals=[1, 0, 0], dtype="INT64"),
         expect=dict(rows=[0, 3, 4], cols=[2, 3, 2], vals=[9, 9, 7])),
    dict(name="test_mxm_accum", cite="graphblas/tests/test_matrix.py:377-386", op="mxm",
         semiring="PLUS_TIMES", A="A", B="A", C="A", accum="PLUS",
         expect=dict(
             rows=[0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 2, 2, 3, 3, 3, 3, 3, 4, 4, 5, 5, 6, 6, 6, 6, 6],
             cols=[0, 1, 2, 3, 4, 6, 2, 3, 4, 5, 6, 2, 5, 0, 1, 2, 3, 5, 2, 5, 2, 5, 0, 2, 3, 4, 5],
             vals=[9, 2, 9, 3, 16, 8, 20, 28, 20, 56, 4, 1, 1, 3, 6, 3, 9, 3, 7, 7, 1, 1, 21, 26, 7,
                   3, 26])),
    dict(name="test_mxv", cite="graphblas/tests/test_matrix.py:389-392", op="mxv",
         semiring="PLUS_TIMES", A="A", u="v", w=None, expect=dict(idx=[0, 1, 6], vals=[5, 16, 13])),
    dict(name="test_vxm", cite="graphblas/tests/test_vector.py:297-300", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w=None,
         expect=dict(idx=[0, 2, 3, 4, 5, 6], vals=[3, 3, 0, 8, 14, 4])),
    dict(name="test_vxm_transpose", cite="graphblas/tests/test_vector.py:303-306", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w=None, tran1=True,
         expect=dict(idx=[0, 1, 6], vals=[5, 16, 13])),
    dict(name="test_vxm_nonsquare", cite="graphblas/tests/test_vector.py:309-316", op="vxm",
         semiring="MIN_PLUS", A=dict(rows=[0, 3], cols=[0, 1], vals=[10, 20], nrows=7, ncols=2),
         u="v", w=dict(idx=[], vals=[], size=2), expect=dict(idx=[1], vals=[21])),
    dict(name="test_vxm_mask_struct", cite="graphblas/tests/test_vector.py:324-330", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w="v", structure=True,
         mask=dict(idx=[0, 3, 4], vals=[0, 0, 0], dtype="BOOL"),
         expect=dict(idx=[0, 1, 3, 4, 6], vals=[3, 1, 0, 8, 0])),
    dict(name="test_vxm_mask_struct_comp", cite="graphblas/tests/test_vector.py:334-337", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w="v", structure=True, comp=True,
         mask=dict(idx=[0, 3, 4], vals=[0, 0, 0], dtype="BOOL"),
         expect=dict(idx=[2, 3, 4, 5, 6], vals=[3, 1, 2, 14, 4])),
    dict(name="test_vxm_mask_value_replace", cite="graphblas/tests/test_vector.py:338-346", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w="v", replace=True,
         mask=dict(idx=[0, 1, 2, 3, 4], vals=[1, 0, 0, 1, 1], dtype="BOOL"),
         expect=dict(idx=[0, 3, 4], vals=[3, 0, 8])),
    dict(name="test_vxm_accum", cite="graphblas/tests/test_vector.py:349-353", op="vxm",
         semiring="PLUS_TIMES", A="A", u="v", w="v", accum="PLUS",
         expect=dict(idx=[0, 1, 2, 3, 4, 5, 6], vals=[3, 1, 3, 1, 10, 14, 4])),
    dict(name="test_inner", cite="graphblas/tests/test_vector.py:1479-1491", op="inner",
         semiring="PLUS_TIMES", u="v", expect=dict(value=6)),
]

if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "reference_kats.json"), "w") as f:
        json.dump(dict(fixtures=dict(A=A, v=v), kats=KATS), f, indent=1)
    print(len(KATS), "KATs written")
