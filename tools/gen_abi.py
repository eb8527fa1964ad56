#!/usr/bin/env python3
"""Generate the C-ABI surface of libgrb_b200.so from one set of lists.

Outputs
  include/GraphBLAS_b200.h                  fully expanded declarations (cffi can parse the
                                            region between the cffi-begin/cffi-end markers)
  graph-python_b200/csrc/grb_abi_lists.inc  X-macro lists used by the C++/CUDA side

Symbol names follow the GraphBLAS C API 2.0 and the SuiteSparse GxB_ names, because
python-graphblas discovers operators by regex-scanning dir(lib)
(reference graphblas/core/operator/base.py:291, semiring regexes operator/semiring.py:170-204).
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# (suffix, C type, type code)
TYPES = [
    ("BOOL", "bool", 0),
    ("INT8", "int8_t", 1),
    ("INT16", "int16_t", 2),
    ("INT32", "int32_t", 3),
    ("INT64", "int64_t", 4),
    ("UINT8", "uint8_t", 5),
    ("UINT16", "uint16_t", 6),
    ("UINT32", "uint32_t", 7),
    ("UINT64", "uint64_t", 8),
    ("FP32", "float", 9),
    ("FP64", "double", 10),
]
NUMERIC = [t for t in TYPES if t[0] != "BOOL"]
# types that have semiring kernels instantiated (round 1)
KERNEL_NUMERIC = ["INT32", "INT64", "UINT32", "UINT64", "FP32", "FP64"]

# binary ops: (NAME, prefix) typed over all 11 types, z type = x type
BINOPS_T = [
    ("FIRST", "GrB"), ("SECOND", "GrB"), ("ONEB", "GrB"), ("PAIR", "GxB"), ("ANY", "GxB"),
    ("PLUS", "GrB"), ("MINUS", "GrB"), ("TIMES", "GrB"), ("DIV", "GrB"),
    ("MIN", "GrB"), ("MAX", "GrB"),
]
# comparison ops: typed over all 11, z = bool
CMPOPS_T = [("EQ", "GrB"), ("NE", "GrB"), ("GT", "GrB"), ("LT", "GrB"), ("GE", "GrB"), ("LE", "GrB")]
# bool-only logical ops (no suffix in the C API)
BOOLOPS = ["LOR", "LAND", "LXOR", "LXNOR"]

# monoids over numeric types
MONOIDS_NUM = ["PLUS", "TIMES", "MIN", "MAX"]
MONOIDS_BOOL = ["LOR", "LAND", "LXOR", "LXNOR"]

# semirings (monoid, multiply) for numeric kernel types; official GrB names where the C API has them
GRB_SEMIRINGS = {
    ("PLUS", "TIMES"), ("PLUS", "MIN"), ("MIN", "PLUS"), ("MIN", "TIMES"), ("MIN", "FIRST"),
    ("MIN", "SECOND"), ("MIN", "MAX"), ("MAX", "PLUS"), ("MAX", "TIMES"), ("MAX", "FIRST"),
    ("MAX", "SECOND"), ("MAX", "MIN"),
}
SEMIRINGS_NUM = [
    ("PLUS", "TIMES"), ("PLUS", "PLUS"), ("PLUS", "FIRST"), ("PLUS", "SECOND"), ("PLUS", "PAIR"),
    ("PLUS", "MIN"), ("PLUS", "MAX"), ("PLUS", "DIV"),
    ("MIN", "PLUS"), ("MIN", "TIMES"), ("MIN", "FIRST"), ("MIN", "SECOND"), ("MIN", "MAX"),
    ("MAX", "PLUS"), ("MAX", "TIMES"), ("MAX", "FIRST"), ("MAX", "SECOND"), ("MAX", "MIN"),
    ("ANY", "PAIR"), ("ANY", "FIRST"), ("ANY", "SECOND"), ("ANY", "TIMES"), ("ANY", "PLUS"),
]
GRB_SEMIRINGS_BOOL = {("LOR", "LAND"), ("LAND", "LOR"), ("LXOR", "LAND"), ("LXNOR", "LOR")}
SEMIRINGS_BOOL = [
    ("LOR", "LAND"), ("LAND", "LOR"), ("LXOR", "LAND"), ("LXNOR", "LOR"),
    ("ANY", "PAIR"), ("ANY", "FIRST"), ("ANY", "SECOND"), ("ANY", "LAND"), ("LOR", "PAIR"),
]

DESC = [
    "T1", "T0", "T0T1", "C", "S", "CT1", "ST1", "CT0", "ST0", "CT0T1", "ST0T1", "SC", "SCT1",
    "SCT0", "SCT0T1", "R", "RT1", "RT0", "RT0T1", "RC", "RS", "RCT1", "RST1", "RCT0", "RST0",
    "RCT0T1", "RST0T1", "RSC", "RSCT1", "RSCT0", "RSCT0T1",
]


def semiring_name(mon, mul, t):
    if t == "BOOL":
        if (mon, mul) in GRB_SEMIRINGS_BOOL:
            return f"GrB_{mon}_{mul}_SEMIRING_BOOL"
        return f"GxB_{mon}_{mul}_BOOL"
    if (mon, mul) in GRB_SEMIRINGS:
        return f"GrB_{mon}_{mul}_SEMIRING_{t}"
    return f"GxB_{mon}_{mul}_{t}"


INTS = ["INT8", "INT16", "INT32", "INT64", "UINT8", "UINT16", "UINT32", "UINT64"]
UINTS = ["UINT8", "UINT16", "UINT32", "UINT64"]
NUM10 = [t for t, _, _ in NUMERIC]
ALL11 = [t for t, _, _ in TYPES]
NARROW = ["INT8", "INT16", "UINT8", "UINT16"]   # semirings computed in the 32-bit type, results wrapped to the narrow one


def all_semirings():
    """Semirings that have sm_100a kernels: (symbol, MON, MUL, TYPE).  Both spellings SuiteSparse exports
    (GrB_<MON>_<MUL>_SEMIRING_<T> and GxB_<MON>_<MUL>_<T>) point at the same kernels."""
    out = []
    for mon, mul in SEMIRINGS_NUM:
        for t in KERNEL_NUMERIC + NARROW:
            out.append((semiring_name(mon, mul, t), mon, mul, t))
            if (mon, mul) in GRB_SEMIRINGS:
                out.append((f"GxB_{mon}_{mul}_{t}", mon, mul, t))
    for mon, mul in SEMIRINGS_BOOL:
        out.append((semiring_name(mon, mul, "BOOL"), mon, mul, "BOOL"))
        if (mon, mul) in GRB_SEMIRINGS_BOOL:
            out.append((f"GxB_{mon}_{mul}_BOOL", mon, mul, "BOOL"))
    return out


def kernel_semirings():
    """(MON, MUL, TYPE) triples the kernel translation units instantiate (wide types + BOOL)."""
    seen, out = set(), []
    for _, mon, mul, t in all_semirings():
        if t in NARROW or (mon, mul, t) in seen:
            continue
        seen.add((mon, mul, t))
        out.append((mon, mul, t))
    # multiplies that wrap to a narrow integer type inside the 32-bit kernels (grb_internal.cuh: narrow_mul);
    # N1..N4 = INT8, INT16, UINT8, UINT16
    wide = {1: "INT32", 2: "INT32", 3: "UINT32", 4: "UINT32"}
    for nid, t in wide.items():
        for mon in ("MIN", "MAX"):
            for mul in ("PLUS", "TIMES"):
                out.append((mon, f"{mul}_N{nid}", t))
        out.append(("PLUS", f"DIV_N{nid}", t))
    return out


def registry_only_semirings():
    """Handles python-graphblas's registry finds by regex (reference graphblas/core/operator/semiring.py:170-204) that
    have no kernel here: using one returns GrB_NOT_IMPLEMENTED (north_star: raise, no CPU fallback)."""
    have = {n for n, *_ in all_semirings()}
    out = []
    muls = ["FIRST", "SECOND", "PAIR", "MIN", "MAX", "PLUS", "MINUS", "RMINUS", "TIMES", "DIV", "RDIV", "ISEQ", "ISNE",
            "ISGT", "ISLT", "ISGE", "ISLE", "LOR", "LAND", "LXOR"]
    for mon in ["MIN", "MAX", "PLUS", "TIMES", "ANY"]:
        for mul in muls:
            out += [f"GxB_{mon}_{mul}_{t}" for t in NUM10]
        for mul in ["FIRSTI1", "FIRSTI", "FIRSTJ1", "FIRSTJ", "SECONDI1", "SECONDI", "SECONDJ1", "SECONDJ"]:
            out += [f"GxB_{mon}_{mul}_{t}" for t in ("INT32", "INT64")]
    for mon, mul in sorted(GRB_SEMIRINGS):
        out += [f"GrB_{mon}_{mul}_SEMIRING_{t}" for t in NUM10]
    for mon in ["LOR", "LAND", "LXOR", "EQ", "ANY"]:
        for mul in ["EQ", "NE", "GT", "LT", "GE", "LE"]:
            out += [f"GxB_{mon}_{mul}_{t}" for t in NUM10]
        for mul in ["FIRST", "SECOND", "PAIR", "LOR", "LAND", "LXOR", "EQ", "GT", "LT", "GE", "LE"]:
            out.append(f"GxB_{mon}_{mul}_BOOL")
    for mon in ["BOR", "BAND", "BXOR", "BXNOR"]:
        for mul in ["BOR", "BAND", "BXOR", "BXNOR"]:
            out += [f"GxB_{mon}_{mul}_{t}" for t in UINTS]
    seen, res = set(), []
    for n in out:
        if n not in have and n not in seen:
            seen.add(n)
            res.append(n)
    return res


def all_monoids():
    out = []
    for mon in MONOIDS_NUM:
        for t, _, _ in NUMERIC:
            out.append((f"GrB_{mon}_MONOID_{t}", mon, t))
    for t, _, _ in TYPES:
        out.append((f"GxB_ANY_{t}_MONOID", "ANY", t))
    for mon in MONOIDS_BOOL:
        out.append((f"GrB_{mon}_MONOID_BOOL", mon, "BOOL"))
    out.append(("GxB_EQ_BOOL_MONOID", "LXNOR", "BOOL"))      # eq on bool is lxnor
    return out


def registry_only_monoids():
    """reference graphblas/core/operator/monoid.py:179-195"""
    return [f"GxB_{m}_{t}_MONOID" for m in ("BOR", "BAND", "BXOR", "BXNOR") for t in UINTS]


UNOPS_T = [("IDENTITY", "GrB"), ("AINV", "GrB"), ("MINV", "GrB"), ("ABS", "GrB"), ("ONE", "GxB")]


def all_unops():
    out = []  # (symbol, OP, type)
    for op, pre in UNOPS_T:
        for t, _, _ in TYPES:
            out.append((f"{pre}_{op}_{t}", op, t))
    for t, _, _ in TYPES:
        if "INT" in t:
            out.append((f"GrB_BNOT_{t}", "BNOT", t))
    out.append(("GrB_LNOT", "LNOT", "BOOL"))
    return out


def registry_only_unops():
    """reference graphblas/core/operator/unary.py:142-167 (the GxB math set UnaryOp._initialize walks, :395-425)"""
    out = [f"GxB_LNOT_{t}" for t in ALL11]
    out += [f"GxB_{op}_{t}" for op in ("POSITIONI1", "POSITIONI", "POSITIONJ1", "POSITIONJ") for t in ("INT32", "INT64")]
    math = ["SQRT", "LOG", "EXP", "LOG2", "SIN", "COS", "TAN", "ACOS", "ASIN", "ATAN", "SINH", "COSH", "TANH", "ACOSH",
            "ASINH", "ATANH", "SIGNUM", "CEIL", "FLOOR", "ROUND", "TRUNC", "EXP2", "EXPM1", "LOG10", "LOG1P",
            "LGAMMA", "TGAMMA", "ERF", "ERFC", "FREXPX", "FREXPE", "CBRT", "ISINF", "ISNAN", "ISFINITE"]
    out += [f"GxB_{op}_{t}" for op in math for t in ("FP32", "FP64")]
    return out


def all_binops():
    out = []  # (symbol, OP, xtype, ztype)
    for op, pre in BINOPS_T:
        for t, _, _ in TYPES:
            out.append((f"{pre}_{op}_{t}", "ONEB" if op == "PAIR" else op, t, t))
    for op, pre in CMPOPS_T:
        for t, _, _ in TYPES:
            out.append((f"{pre}_{op}_{t}", op, t, "BOOL"))
    for op in BOOLOPS:
        out.append((f"GrB_{op}", op, "BOOL", "BOOL"))
    # SuiteSparse's typed logical ops (coerce to 0/1 in the operand type) and the reversed divide
    for op in ("LOR", "LAND", "LXOR"):
        for t, _, _ in TYPES:
            out.append((f"GxB_{op}_{t}", op, t, t))
    for t, _, _ in TYPES:
        out.append((f"GxB_RDIV_{t}", "RDIV", t, t))
    return out


def registry_only_binops():
    """reference graphblas/core/operator/binary.py:332-369"""
    out = [f"GrB_{op}_{t}" for op in ("BOR", "BAND", "BXOR", "BXNOR") for t in INTS]
    out += [f"GxB_{op}_{t}" for op in ("POW", "RMINUS", "ISEQ", "ISNE", "ISGT", "ISLT", "ISGE", "ISLE") for t in ALL11]
    out += [f"GxB_{op}_{t}" for op in ("ATAN2", "HYPOT", "FMOD", "REMAINDER", "LDEXP", "COPYSIGN") for t in ("FP32", "FP64")]
    out += [f"GxB_{op}_{t}" for op in ("BGET", "BSET", "BCLR", "BSHIFT") for t in INTS]
    out += [f"GxB_{op}_{t}" for op in ("FIRSTI1", "FIRSTI", "FIRSTJ1", "FIRSTJ", "SECONDI1", "SECONDI", "SECONDJ1", "SECONDJ")
            for t in ("INT32", "INT64")]
    return out


IDXOPS = ["TRIL", "TRIU", "DIAG", "OFFDIAG"]


def registry_only_idxops():
    """reference graphblas/core/operator/indexunary.py:98-113 (select.valuene is imported by core/mask.py:4)"""
    out = [f"GrB_{op}_{t}" for op in ("ROWINDEX", "COLINDEX", "DIAGINDEX") for t in ("INT32", "INT64")]
    out += [f"GrB_{op}" for op in ("COLLE", "COLGT", "ROWLE", "ROWGT")]
    out += [f"GrB_{op}_{t}" for op in ("VALUEEQ", "VALUENE", "VALUEGT", "VALUEGE", "VALUELT", "VALUELE") for t in ALL11]
    return out


HEADER_TOP = r'''/* GraphBLAS_b200.h — C-ABI of libgrb_b200.so (GENERATED by tools/gen_abi.py; do not edit by hand).
 *
 * A GraphBLAS C-API-2.0 subset: exactly the entry points python-graphblas's cffi binding calls on
 * the semiring hot path, plus the container calls needed to hand CSR/CSC/bitmap/full storage across
 * the FFI. All signatures are plain C (pointers, sizes, opaque handles); no torch types.
 *
 * Reference interfaces replaced (paths relative to the python-graphblas tree):
 *   GrB_mxm  <- graphblas/core/matrix.py:2206-2251  (cfunc_name "GrB_mxm",  args assembled base.py:476-483)
 *   GrB_mxv  <- graphblas/core/matrix.py:2163-2204  (cfunc_name "GrB_mxv")
 *   GrB_vxm  <- graphblas/core/vector.py:1259-1306  (cfunc_name "GrB_vxm"); Vector.inner/outer :1609-1693
 *   GrB_Info codes          <- graphblas/exceptions.py:93-117
 *   GrB_DESC_* / Descriptor <- graphblas/core/descriptor.py:51-156, core/ss/descriptor.py:24-105
 *   GrB_<TYPE> handles      <- graphblas/core/dtypes.py:154-245
 *   operator globals        <- graphblas/core/operator/base.py:291,410-486; semiring.py:170-204
 *   new/free/dup/nvals/...  <- graphblas/core/matrix.py:185,213 ; core/vector.py (same pattern)
 *   build / extractTuples   <- graphblas/core/matrix.py:643-697, 885
 *   import/export (CSR/CSC/COO, copy semantics) <- graphblas/core/matrix.py:1057-1133, 1658-1702
 *   GxB pack/unpack (ownership transfer)        <- graphblas/core/ss/matrix.py:744-1162,1282-1352,
 *                                                  graphblas/core/ss/vector.py:494-636,946-1016
 *   *_error strings         <- graphblas/exceptions.py:143-154
 * The loader seam is `from suitesparse_graphblas import ffi, lib, initialize, is_initialized`
 * (graphblas/__init__.py:141); see INTEGRATION.md for the binding a maintainer adds.
 */
#ifndef GRAPHBLAS_B200_H
#define GRAPHBLAS_B200_H
#include <stdint.h>
#include <stdbool.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
/* cffi-begin */
typedef uint64_t GrB_Index;

typedef enum {
    GrB_SUCCESS = 0,
    GrB_NO_VALUE = 1,
    GxB_EXHAUSTED = 7089,
    GrB_UNINITIALIZED_OBJECT = -1,
    GrB_NULL_POINTER = -2,
    GrB_INVALID_VALUE = -3,
    GrB_INVALID_INDEX = -4,
    GrB_DOMAIN_MISMATCH = -5,
    GrB_DIMENSION_MISMATCH = -6,
    GrB_OUTPUT_NOT_EMPTY = -7,
    GrB_NOT_IMPLEMENTED = -8,
    GrB_PANIC = -101,
    GrB_OUT_OF_MEMORY = -102,
    GrB_INSUFFICIENT_SPACE = -103,
    GrB_INVALID_OBJECT = -104,
    GrB_INDEX_OUT_OF_BOUNDS = -105,
    GrB_EMPTY_OBJECT = -106
} GrB_Info;

typedef enum { GrB_NONBLOCKING = 0, GrB_BLOCKING = 1 } GrB_Mode;
typedef enum { GrB_COMPLETE = 0, GrB_MATERIALIZE = 1 } GrB_WaitMode;
typedef enum { GrB_CSR_FORMAT = 0, GrB_CSC_FORMAT = 1, GrB_COO_FORMAT = 2 } GrB_Format;

typedef enum {
    GrB_OUTP = 0, GrB_MASK = 1, GrB_INP0 = 2, GrB_INP1 = 3,
    GxB_DESCRIPTOR_NTHREADS = 5, GxB_DESCRIPTOR_CHUNK = 7, GxB_DESCRIPTOR_GPU_ID = 21,
    GxB_AxB_METHOD = 1000, GxB_SORT = 35, GxB_COMPRESSION = 36, GxB_IMPORT = 37
} GrB_Desc_Field;

typedef enum {
    GxB_DEFAULT = 0, GrB_REPLACE = 1, GrB_COMP = 2, GrB_TRAN = 3, GrB_STRUCTURE = 4,
    GxB_AxB_GUSTAVSON = 1001, GxB_AxB_DOT = 1003, GxB_AxB_HASH = 1004, GxB_AxB_SAXPY = 1005
} GrB_Desc_Value;

/* object / global options queried by the reference (graphblas/__init__.py:158, core/ss/matrix.py:216-220) */
typedef enum {
    GxB_HYPER_SWITCH = 0, GxB_FORMAT = 1, GxB_MODE = 2, GxB_GLOBAL_NTHREADS = 5, GxB_GLOBAL_CHUNK = 7,
    GxB_LIBRARY_NAME = 8, GxB_LIBRARY_VERSION = 9, GxB_LIBRARY_DATE = 10, GxB_LIBRARY_ABOUT = 11, GxB_LIBRARY_URL = 12,
    GxB_LIBRARY_LICENSE = 13, GxB_LIBRARY_COMPILE_DATE = 14, GxB_LIBRARY_COMPILE_TIME = 15, GxB_API_VERSION = 16,
    GxB_API_DATE = 17, GxB_API_ABOUT = 18, GxB_API_URL = 19, GxB_GLOBAL_GPU_CONTROL = 21, GxB_GLOBAL_GPU_CHUNK = 22,
    GxB_COMPILER_VERSION = 23, GxB_COMPILER_NAME = 24, GxB_LIBRARY_OPENMP = 25,
    GxB_SPARSITY_CONTROL = 32, GxB_SPARSITY_STATUS = 33, GxB_BITMAP_SWITCH = 34,
    GxB_BURBLE = 99, GxB_PRINTF = 101, GxB_FLUSH = 102, GxB_MEMORY_POOL = 103, GxB_PRINT_1BASED = 104
} GxB_Option_Field;
typedef enum { GxB_BY_ROW = 0, GxB_BY_COL = 1, GxB_NO_FORMAT = -1 } GxB_Format_Value;
enum { GxB_HYPERSPARSE = 1, GxB_SPARSE = 2, GxB_BITMAP = 4, GxB_FULL = 8, GxB_AUTO_SPARSITY = 15 };
/* constants python-graphblas reads at import (graphblas/core/ss/descriptor.py:9-16, core/ss/config.py, core/ss/matrix.py) */
enum { GxB_COMPRESSION_NONE = -1, GxB_COMPRESSION_DEFAULT = 0, GxB_COMPRESSION_LZ4 = 1000, GxB_COMPRESSION_LZ4HC = 2000,
       GxB_COMPRESSION_ZSTD = 3000 };
enum { GxB_FAST_IMPORT = 0, GxB_SECURE_IMPORT = 502 };
enum { GxB_RANGE = 0, GxB_STRIDE = 1, GxB_BACKWARDS = 2, GxB_BEGIN = 0, GxB_END = 1, GxB_INC = 2 };
enum { GxB_NBITMAP_SWITCH = 8, GxB_MAX_NAME_LEN = 128, GxB_FORMAT_DEFAULT = 0 };
enum { GxB_GPU_ALWAYS = 2001, GxB_GPU_NEVER = 2002 };
extern double GxB_HYPER_DEFAULT, GxB_ALWAYS_HYPER, GxB_NEVER_HYPER;   /* plain globals: cffi ABI mode cannot read const doubles */

typedef struct GB_Type_opaque *GrB_Type;
typedef struct GB_UnaryOp_opaque *GrB_UnaryOp;
typedef struct GB_IndexUnaryOp_opaque *GrB_IndexUnaryOp;
typedef struct GB_BinaryOp_opaque *GrB_BinaryOp;
typedef struct GB_Monoid_opaque *GrB_Monoid;
typedef struct GB_Semiring_opaque *GrB_Semiring;
typedef struct GB_Descriptor_opaque *GrB_Descriptor;
typedef struct GB_Scalar_opaque *GrB_Scalar;
typedef struct GB_Vector_opaque *GrB_Vector;
typedef struct GB_Matrix_opaque *GrB_Matrix;

extern const uint64_t *GrB_ALL;

/* ---- context ---- */
GrB_Info GrB_init(GrB_Mode mode);
GrB_Info GxB_init(GrB_Mode mode, void *(*user_malloc)(size_t), void *(*user_calloc)(size_t, size_t),
                  void *(*user_realloc)(void *, size_t), void (*user_free)(void *));
GrB_Info GrB_finalize(void);
GrB_Info GrB_getVersion(unsigned int *version, unsigned int *subversion);

/* ---- the hot path ---- */
GrB_Info GrB_mxm(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum, const GrB_Semiring op,
                 const GrB_Matrix A, const GrB_Matrix B, const GrB_Descriptor desc);
GrB_Info GrB_mxv(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Semiring op,
                 const GrB_Matrix A, const GrB_Vector u, const GrB_Descriptor desc);
GrB_Info GrB_vxm(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Semiring op,
                 const GrB_Vector u, const GrB_Matrix A, const GrB_Descriptor desc);

/* ---- descriptors ---- */
GrB_Info GrB_Descriptor_new(GrB_Descriptor *desc);
GrB_Info GrB_Descriptor_set(GrB_Descriptor desc, GrB_Desc_Field field, GrB_Desc_Value val);
GrB_Info GxB_Desc_set_INT32(GrB_Descriptor desc, GrB_Desc_Field field, int32_t val);
GrB_Info GrB_Descriptor_free(GrB_Descriptor *desc);
GrB_Info GrB_Descriptor_error(const char **error, const GrB_Descriptor desc);

/* ---- algebra objects.  GrB_Semiring_new composes built-in monoids and multiplies (the reference's registry does so at
 *      import, graphblas/core/operator/semiring.py:205-243); a triple without an instantiated kernel, and every
 *      user-defined function, is refused with GrB_NOT_IMPLEMENTED when used (north_star: raise, no CPU fallback) ---- */
GrB_Info GrB_Semiring_new(GrB_Semiring *semiring, GrB_Monoid add, GrB_BinaryOp multiply);
GrB_Info GrB_Semiring_free(GrB_Semiring *semiring);
GrB_Info GrB_UnaryOp_new(GrB_UnaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype);
GrB_Info GrB_BinaryOp_new(GrB_BinaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype, GrB_Type ytype);
GrB_Info GrB_IndexUnaryOp_new(GrB_IndexUnaryOp *op, void *function, GrB_Type ztype, GrB_Type xtype, GrB_Type ytype);
GrB_Info GrB_UnaryOp_free(GrB_UnaryOp *op);
GrB_Info GrB_BinaryOp_free(GrB_BinaryOp *op);
GrB_Info GrB_IndexUnaryOp_free(GrB_IndexUnaryOp *op);
GrB_Info GrB_Monoid_free(GrB_Monoid *op);
GrB_Info GrB_Semiring_error(const char **error, const GrB_Semiring op);
GrB_Info GrB_Monoid_error(const char **error, const GrB_Monoid op);
GrB_Info GrB_BinaryOp_error(const char **error, const GrB_BinaryOp op);
GrB_Info GrB_UnaryOp_error(const char **error, const GrB_UnaryOp op);
GrB_Info GrB_IndexUnaryOp_error(const char **error, const GrB_IndexUnaryOp op);
GrB_Info GrB_Type_error(const char **error, const GrB_Type t);

/* ---- Matrix ---- */
GrB_Info GrB_Matrix_new(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols);
GrB_Info GrB_Matrix_dup(GrB_Matrix *C, const GrB_Matrix A);
GrB_Info GrB_Matrix_free(GrB_Matrix *A);
GrB_Info GrB_Matrix_clear(GrB_Matrix A);
GrB_Info GrB_Matrix_nrows(GrB_Index *nrows, const GrB_Matrix A);
GrB_Info GrB_Matrix_ncols(GrB_Index *ncols, const GrB_Matrix A);
GrB_Info GrB_Matrix_nvals(GrB_Index *nvals, const GrB_Matrix A);
GrB_Info GrB_Matrix_wait(GrB_Matrix A, GrB_WaitMode mode);
GrB_Info GrB_Matrix_error(const char **error, const GrB_Matrix A);
GrB_Info GxB_Matrix_type(GrB_Type *type, const GrB_Matrix A);
GrB_Info GrB_Matrix_removeElement(GrB_Matrix C, GrB_Index i, GrB_Index j);
GrB_Info GrB_Matrix_exportSize(GrB_Index *Ap_len, GrB_Index *Ai_len, GrB_Index *Ax_len,
                               GrB_Format format, const GrB_Matrix A);
GrB_Info GrB_Matrix_exportHint(GrB_Format *format, const GrB_Matrix A);
GrB_Info GrB_transpose(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                       const GrB_Matrix A, const GrB_Descriptor desc);

/* ---- Vector ---- */
GrB_Info GrB_Vector_new(GrB_Vector *v, GrB_Type type, GrB_Index n);
GrB_Info GrB_Vector_dup(GrB_Vector *w, const GrB_Vector u);
GrB_Info GrB_Vector_free(GrB_Vector *v);
GrB_Info GrB_Vector_clear(GrB_Vector v);
GrB_Info GrB_Vector_size(GrB_Index *n, const GrB_Vector v);
GrB_Info GrB_Vector_nvals(GrB_Index *nvals, const GrB_Vector v);
GrB_Info GrB_Vector_wait(GrB_Vector v, GrB_WaitMode mode);
GrB_Info GrB_Vector_error(const char **error, const GrB_Vector v);
GrB_Info GxB_Vector_type(GrB_Type *type, const GrB_Vector v);
GrB_Info GrB_Vector_removeElement(GrB_Vector w, GrB_Index i);

/* ---- Scalar ---- */
GrB_Info GrB_Scalar_new(GrB_Scalar *s, GrB_Type type);
GrB_Info GrB_Scalar_dup(GrB_Scalar *s, const GrB_Scalar t);
GrB_Info GrB_Scalar_free(GrB_Scalar *s);
GrB_Info GrB_Scalar_clear(GrB_Scalar s);
GrB_Info GrB_Scalar_nvals(GrB_Index *nvals, const GrB_Scalar s);
GrB_Info GrB_Scalar_wait(GrB_Scalar s, GrB_WaitMode mode);
GrB_Info GrB_Scalar_error(const char **error, const GrB_Scalar s);

/* ---- loop-body operations around the products (SURVEY 8f rank 1) ---- */
GrB_Info GrB_Vector_eWiseMult_BinaryOp(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                                       const GrB_BinaryOp op, const GrB_Vector u, const GrB_Vector v,
                                       const GrB_Descriptor desc);
GrB_Info GrB_Vector_eWiseAdd_BinaryOp(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                                      const GrB_BinaryOp op, const GrB_Vector u, const GrB_Vector v,
                                      const GrB_Descriptor desc);
GrB_Info GrB_Matrix_eWiseMult_BinaryOp(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                                       const GrB_BinaryOp op, const GrB_Matrix A, const GrB_Matrix B,
                                       const GrB_Descriptor desc);
GrB_Info GrB_Matrix_eWiseAdd_BinaryOp(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                                      const GrB_BinaryOp op, const GrB_Matrix A, const GrB_Matrix B,
                                      const GrB_Descriptor desc);
GrB_Info GrB_Vector_assign(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                           const GrB_Vector u, const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc);
GrB_Info GrB_Vector_apply(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_UnaryOp op,
                          const GrB_Vector u, const GrB_Descriptor desc);
GrB_Info GrB_Vector_extract(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Vector u,
                            const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc);
GrB_Info GrB_Matrix_reduce_Monoid(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum,
                                  const GrB_Monoid monoid, const GrB_Matrix A, const GrB_Descriptor desc);
GrB_Info GxB_Global_Option_get_INT32(GxB_Option_Field field, int32_t *value);
GrB_Info GxB_Matrix_Option_get_INT32(GrB_Matrix A, GxB_Option_Field field, int32_t *value);
GrB_Info GxB_Vector_Option_get_INT32(GrB_Vector v, GxB_Option_Field field, int32_t *value);
GrB_Info GrB_Matrix_select_INT64(GrB_Matrix C, const GrB_Matrix Mask, const GrB_BinaryOp accum,
                                 const GrB_IndexUnaryOp op, const GrB_Matrix A, int64_t y,
                                 const GrB_Descriptor desc);

/* ---- B200 extensions (GxB_b200_*): stream plumbing, introspection, bench utilities ---- */
GrB_Info GxB_b200_set_stream(void *cuda_stream);
GrB_Info GxB_b200_sync(void);
GrB_Info GxB_b200_device_info(int *device, int *sm_count, size_t *free_bytes, size_t *total_bytes);
uint64_t GxB_b200_launch_count(void);
const char *GxB_b200_last_kernel(void);
GrB_Info GxB_b200_kernel_time(const char *name, double *total_ms, uint64_t *launches);
GrB_Info GxB_b200_kernel_time_reset(void);
GrB_Info GxB_b200_kernel_time_dump(void);
GrB_Info GxB_b200_set_option(const char *name, int64_t value);
GrB_Info GxB_b200_Vector_device_ptrs(void **values, void **present, const GrB_Vector v);
GrB_Info GxB_b200_Vector_set_dense_nvals(GrB_Vector v, int64_t nvals);
GrB_Info GxB_b200_Vector_isequal(bool *result, const GrB_Vector u, const GrB_Vector v);
GrB_Info GxB_b200_Matrix_device_ptrs(void **rowptr, void **colidx, void **values, const GrB_Matrix A);
GrB_Info GxB_b200_Matrix_prepare(GrB_Matrix A, int want_transpose);
GrB_Info GxB_b200_Matrix_rmat(GrB_Matrix *A, GrB_Type type, int scale, int edgefactor, uint64_t seed,
                              int symmetrize, int value_mode);
GrB_Info GxB_b200_rmat_edges_host(uint64_t *src, uint64_t *dst, int scale, uint64_t nedges, uint64_t seed);
GrB_Info GxB_b200_Matrix_rowblock(GrB_Matrix *C, const GrB_Matrix A, GrB_Index row_begin, GrB_Index row_end);
/* index lists that already live in HBM: I = the values of a full integer vector (as SuiteSparse 9's GxB_Vector_assign_Vector /
 * GxB_Vector_extract_Vector).  w<mask>(I) = accum(w(I), u), duplicates folded under MIN/MAX/PLUS; w<mask> = accum(w, u(I)).
 * FastSV's `f(min)[I] << mngp` and `gp << f[I]` (reference notebooks/Connected Components -- FastSV.ipynb cell 30) with I = f. */
GrB_Info GxB_b200_Vector_assign_Vector(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Vector u,
                                       const GrB_Vector I, const GrB_Descriptor desc);
GrB_Info GxB_b200_Vector_extract_Vector(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, const GrB_Vector u,
                                        const GrB_Vector I, const GrB_Descriptor desc);
/* multi-GPU exchange over NVLink peer memory (csrc/grb_peer.cu; no reference counterpart: python-graphblas is single-process).
 * Vector_wrap_device: a full vector over caller-owned device memory (a symmetric buffer the peers map), so a product's
 *   partial sums land where the other GPUs can read them.
 * peer_reduce_scatter_FP32: out[i] = sum over ranks q of ((float *)(peer_bases[q] + data_offset_bytes))[i], i < count, after a
 *   flag handshake with every peer (flag group of `world` uint32 at flag_offset_bytes in every buffer, value seq, compared
 *   modulo 2^32); uses multimem.ld_reduce through multicast_base when it is non-zero; optionally out_scaled = out / div and,
 *   afterwards, fill[0 .. fill_count) = fill_value (the other generation of the local partial-sum buffer: once every rank has
 *   reported for iteration k nobody reads generation k - 1 any more).  ctas < 0: |ctas| small CTAs that fit beside a product.
 *   peer_bases is a HOST array of `world` device addresses.  Runs on `stream`, beside the next chunk's product. */
GrB_Info GxB_b200_Vector_wrap_device(GrB_Vector *v, GrB_Type type, GrB_Index n, void *values);
GrB_Info GxB_b200_peer_reduce_scatter_FP32(float *out, float *out_scaled, const float *div, const uint64_t *peer_bases,
                                           uint64_t multicast_base, uint64_t data_offset_bytes, uint64_t count,
                                           uint64_t flag_offset_bytes, int world, int rank, uint32_t seq,
                                           float *fill, uint64_t fill_count, float fill_value, int ctas, void *stream);
/* the fused form: GxB_b200_next_product_progress makes the NEXT unmasked pull product count finished tiles per row block
 * (counters: uint64[nblocks + 2] in device memory, monotonic, the last two words receive debug time stamps; expected[b] receives the tiles touching block b);
 * GxB_b200_peer_exchange_fused_FP32 runs beside that product on another stream: it announces every block to its owner as
 * soon as the local product has passed it, waits for all ranks' announcements about this rank's block, sums the N copies
 * (as peer_reduce_scatter) and optionally refills `fill[fill_count]` with fill_value (the generation of the partial-sum
 * buffer the product after next accumulates into).  product_seq = how many products have reported into `counters` so far,
 * this one included. */
GrB_Info GxB_b200_next_product_progress(uint64_t *counters, uint64_t *expected, int nblocks, GrB_Index rows_per_block);
GrB_Info GxB_b200_peer_debug_stamps(uint64_t *out8);   /* phase stamps (ns) of the last fused exchange: start, first/last announce, flags, reduce, refill */
GrB_Info GxB_b200_peer_exchange_fused_FP32(float *out, float *out_scaled, const float *div, const uint64_t *peer_bases,
                                           uint64_t multicast_base, uint64_t data_offset_bytes, uint64_t count,
                                           uint64_t flag_offset_bytes, int world, int rank, uint32_t seq,
                                           const uint64_t *progress, const uint64_t *expected, uint64_t product_seq,
                                           float *fill, uint64_t fill_count, float fill_value, int ctas, void *stream);

/* ---- allocation-free product.  No counterpart in the reference: added at the request of the run's operators for callers
 *      that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 *      w = A (+).(x) u, plus_times FP32, u full: a variant of the tiled pull kernel behind GrB_mxv whose scratch (tile table
 *      and per-tile carries) is a file-scope static __device__ array instead of pool memory.  All pointers are DEVICE
 *      pointers: rowptr int64[nrows+1], colidx uint32[nvals], values float[nvals], u float[ncols]; outputs w float[nrows]
 *      and presence bytes w_present[nrows] (rows without entries are absent).  w_nvals (host, may be NULL) receives the
 *      number of entries of w and makes the call synchronise.  Limit: nvals <= GXB_B200_NOALLOC_MAX_NVALS (2^26), else
 *      GrB_INSUFFICIENT_SPACE and *error (host, may be NULL) points at a static message.  It calls no cudaMalloc/cudaFree
 *      and no allocating library routine.  Every call shares the one scratch array (3.5 MB per GPU that loads the
 *      library, until exit): calls that overlap on the GPU race; calls on the library's stream are ordered. ---- */
#define GXB_B200_NOALLOC_MAX_NVALS 67108864
GrB_Info GxB_b200_mxv_plus_times_FP32_noalloc(float *w, uint8_t *w_present, GrB_Index *w_nvals, const int64_t *rowptr,
                                              const uint32_t *colidx, const float *values, const float *u,
                                              GrB_Index nrows, GrB_Index ncols, GrB_Index nvals, const char **error);
'''

HEADER_BOTTOM = r'''/* cffi-end */
#ifdef __cplusplus
}
#endif
#endif /* GRAPHBLAS_B200_H */
'''


def typed_functions():
    """Per-type entry points (suffix = GraphBLAS type name)."""
    out = []
    for t, c, _ in TYPES:
        out += [
            f"GrB_Info GrB_Matrix_build_{t}(GrB_Matrix C, const GrB_Index *I, const GrB_Index *J, "
            f"const {c} *X, GrB_Index nvals, const GrB_BinaryOp dup);",
            f"GrB_Info GrB_Matrix_extractTuples_{t}(GrB_Index *I, GrB_Index *J, {c} *X, GrB_Index *nvals, "
            f"const GrB_Matrix A);",
            f"GrB_Info GrB_Matrix_setElement_{t}(GrB_Matrix C, {c} x, GrB_Index i, GrB_Index j);",
            f"GrB_Info GrB_Matrix_extractElement_{t}({c} *x, const GrB_Matrix A, GrB_Index i, GrB_Index j);",
            f"GrB_Info GrB_Matrix_import_{t}(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, "
            f"const GrB_Index *Ap, const GrB_Index *Ai, const {c} *Ax, GrB_Index Ap_len, GrB_Index Ai_len, "
            f"GrB_Index Ax_len, GrB_Format format);",
            f"GrB_Info GrB_Matrix_export_{t}(GrB_Index *Ap, GrB_Index *Ai, {c} *Ax, GrB_Index *Ap_len, "
            f"GrB_Index *Ai_len, GrB_Index *Ax_len, GrB_Format format, const GrB_Matrix A);",
            f"GrB_Info GrB_Matrix_reduce_{t}({c} *c, const GrB_BinaryOp accum, const GrB_Monoid monoid, "
            f"const GrB_Matrix A, const GrB_Descriptor desc);",
            f"GrB_Info GrB_Vector_build_{t}(GrB_Vector w, const GrB_Index *I, const {c} *X, GrB_Index nvals, "
            f"const GrB_BinaryOp dup);",
            f"GrB_Info GrB_Vector_extractTuples_{t}(GrB_Index *I, {c} *X, GrB_Index *nvals, const GrB_Vector v);",
            f"GrB_Info GrB_Vector_setElement_{t}(GrB_Vector w, {c} x, GrB_Index i);",
            f"GrB_Info GrB_Vector_extractElement_{t}({c} *x, const GrB_Vector v, GrB_Index i);",
            f"GrB_Info GrB_Vector_reduce_{t}({c} *c, const GrB_BinaryOp accum, const GrB_Monoid monoid, "
            f"const GrB_Vector u, const GrB_Descriptor desc);",
            f"GrB_Info GrB_Vector_assign_{t}(GrB_Vector w, const GrB_Vector mask, const GrB_BinaryOp accum, "
            f"{c} x, const GrB_Index *I, GrB_Index ni, const GrB_Descriptor desc);",
            f"GrB_Info GrB_Vector_apply_BinaryOp2nd_{t}(GrB_Vector w, const GrB_Vector mask, "
            f"const GrB_BinaryOp accum, const GrB_BinaryOp op, const GrB_Vector u, {c} y, "
            f"const GrB_Descriptor desc);",
            f"GrB_Info GrB_Scalar_setElement_{t}(GrB_Scalar s, {c} x);",
            f"GrB_Info GrB_Scalar_extractElement_{t}({c} *x, const GrB_Scalar s);",
        ]
    return out


UNTYPED_STORAGE = r'''
/* ---- SuiteSparse-style hand-off with ownership transfer (buffers from the allocator given to GxB_init).
 *      pack: the library copies host->HBM, frees the host buffer, sets the caller's pointer to NULL.
 *      unpack: the library allocates host buffers with user_malloc, copies HBM->host, empties the object. */
GrB_Info GxB_Matrix_pack_CSR(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Aj, void **Ax, GrB_Index Ap_size,
                             GrB_Index Aj_size, GrB_Index Ax_size, bool iso, bool jumbled,
                             const GrB_Descriptor desc);
GrB_Info GxB_Matrix_pack_CSC(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index Ap_size,
                             GrB_Index Ai_size, GrB_Index Ax_size, bool iso, bool jumbled,
                             const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_CSR(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Aj, void **Ax, GrB_Index *Ap_size,
                               GrB_Index *Aj_size, GrB_Index *Ax_size, bool *iso, bool *jumbled,
                               const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_CSC(GrB_Matrix A, GrB_Index **Ap, GrB_Index **Ai, void **Ax, GrB_Index *Ap_size,
                               GrB_Index *Ai_size, GrB_Index *Ax_size, bool *iso, bool *jumbled,
                               const GrB_Descriptor desc);
GrB_Info GxB_Vector_pack_Full(GrB_Vector v, void **vx, GrB_Index vx_size, bool iso, const GrB_Descriptor desc);
GrB_Info GxB_Vector_unpack_Full(GrB_Vector v, void **vx, GrB_Index *vx_size, bool *iso,
                                const GrB_Descriptor desc);
GrB_Info GxB_Vector_pack_Bitmap(GrB_Vector v, int8_t **vb, void **vx, GrB_Index vb_size, GrB_Index vx_size,
                                bool iso, GrB_Index nvals, const GrB_Descriptor desc);
GrB_Info GxB_Vector_unpack_Bitmap(GrB_Vector v, int8_t **vb, void **vx, GrB_Index *vb_size,
                                  GrB_Index *vx_size, bool *iso, GrB_Index *nvals, const GrB_Descriptor desc);
GrB_Info GxB_Vector_pack_CSC(GrB_Vector v, GrB_Index **vi, void **vx, GrB_Index vi_size, GrB_Index vx_size, bool iso,
                             GrB_Index nvals, bool jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Vector_unpack_CSC(GrB_Vector v, GrB_Index **vi, void **vx, GrB_Index *vi_size, GrB_Index *vx_size, bool *iso,
                               GrB_Index *nvals, bool *jumbled, const GrB_Descriptor desc);
/* bitmap / full matrices (R = row-major, C = column-major host arrays); converted to CSR on the device, <= 2^31 cells */
GrB_Info GxB_Matrix_pack_BitmapR(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index Ab_size, GrB_Index Ax_size, bool iso,
                                 GrB_Index nvals, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_pack_BitmapC(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index Ab_size, GrB_Index Ax_size, bool iso,
                                 GrB_Index nvals, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_BitmapR(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso,
                                   GrB_Index *nvals, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_BitmapC(GrB_Matrix A, int8_t **Ab, void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso,
                                   GrB_Index *nvals, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_pack_FullR(GrB_Matrix A, void **Ax, GrB_Index Ax_size, bool iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_pack_FullC(GrB_Matrix A, void **Ax, GrB_Index Ax_size, bool iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_FullR(GrB_Matrix A, void **Ax, GrB_Index *Ax_size, bool *iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_unpack_FullC(GrB_Matrix A, void **Ax, GrB_Index *Ax_size, bool *iso, const GrB_Descriptor desc);
/* import = new + pack, export = unpack + free: what Matrix.ss.import_csr / Vector.ss.import_sparse / .ss.export call
 * (graphblas/core/ss/matrix.py:879,920,1067-1069,1123-1125,1327,1517; core/ss/vector.py:530,568,601,991,1179,1341) */
GrB_Info GxB_Matrix_import_CSR(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, GrB_Index **Ap,
                               GrB_Index **Aj, void **Ax, GrB_Index Ap_size, GrB_Index Aj_size, GrB_Index Ax_size,
                               bool iso, bool jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_import_CSC(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, GrB_Index **Ap,
                               GrB_Index **Ai, void **Ax, GrB_Index Ap_size, GrB_Index Ai_size, GrB_Index Ax_size,
                               bool iso, bool jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_CSR(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, GrB_Index **Ap,
                               GrB_Index **Aj, void **Ax, GrB_Index *Ap_size, GrB_Index *Aj_size, GrB_Index *Ax_size,
                               bool *iso, bool *jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_CSC(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, GrB_Index **Ap,
                               GrB_Index **Ai, void **Ax, GrB_Index *Ap_size, GrB_Index *Ai_size, GrB_Index *Ax_size,
                               bool *iso, bool *jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_import_BitmapR(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, int8_t **Ab, void **Ax,
                                   GrB_Index Ab_size, GrB_Index Ax_size, bool iso, GrB_Index nvals,
                                   const GrB_Descriptor desc);
GrB_Info GxB_Matrix_import_BitmapC(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, int8_t **Ab, void **Ax,
                                   GrB_Index Ab_size, GrB_Index Ax_size, bool iso, GrB_Index nvals,
                                   const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_BitmapR(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, int8_t **Ab,
                                   void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso, GrB_Index *nvals,
                                   const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_BitmapC(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, int8_t **Ab,
                                   void **Ax, GrB_Index *Ab_size, GrB_Index *Ax_size, bool *iso, GrB_Index *nvals,
                                   const GrB_Descriptor desc);
GrB_Info GxB_Matrix_import_FullR(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, void **Ax,
                                 GrB_Index Ax_size, bool iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_import_FullC(GrB_Matrix *A, GrB_Type type, GrB_Index nrows, GrB_Index ncols, void **Ax,
                                 GrB_Index Ax_size, bool iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_FullR(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, void **Ax,
                                 GrB_Index *Ax_size, bool *iso, const GrB_Descriptor desc);
GrB_Info GxB_Matrix_export_FullC(GrB_Matrix *A, GrB_Type *type, GrB_Index *nrows, GrB_Index *ncols, void **Ax,
                                 GrB_Index *Ax_size, bool *iso, const GrB_Descriptor desc);
GrB_Info GxB_Vector_import_CSC(GrB_Vector *v, GrB_Type type, GrB_Index n, GrB_Index **vi, void **vx, GrB_Index vi_size,
                               GrB_Index vx_size, bool iso, GrB_Index nvals, bool jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Vector_export_CSC(GrB_Vector *v, GrB_Type *type, GrB_Index *n, GrB_Index **vi, void **vx, GrB_Index *vi_size,
                               GrB_Index *vx_size, bool *iso, GrB_Index *nvals, bool *jumbled, const GrB_Descriptor desc);
GrB_Info GxB_Vector_import_Bitmap(GrB_Vector *v, GrB_Type type, GrB_Index n, int8_t **vb, void **vx, GrB_Index vb_size,
                                  GrB_Index vx_size, bool iso, GrB_Index nvals, const GrB_Descriptor desc);
GrB_Info GxB_Vector_export_Bitmap(GrB_Vector *v, GrB_Type *type, GrB_Index *n, int8_t **vb, void **vx, GrB_Index *vb_size,
                                  GrB_Index *vx_size, bool *iso, GrB_Index *nvals, const GrB_Descriptor desc);
GrB_Info GxB_Vector_import_Full(GrB_Vector *v, GrB_Type type, GrB_Index n, void **vx, GrB_Index vx_size, bool iso,
                                const GrB_Descriptor desc);
GrB_Info GxB_Vector_export_Full(GrB_Vector *v, GrB_Type *type, GrB_Index *n, void **vx, GrB_Index *vx_size, bool *iso,
                                const GrB_Descriptor desc);
GrB_Info GxB_Matrix_iso(bool *iso, const GrB_Matrix A);
GrB_Info GxB_Vector_iso(bool *iso, const GrB_Vector v);
/* copy-semantics dense hand-off used by the host mirror and the benchmark (host buffers stay the caller's) */
GrB_Info GxB_b200_Vector_import_dense(GrB_Vector v, const void *values, const int8_t *present, GrB_Index n);
GrB_Info GxB_b200_Vector_export_dense(void *values, int8_t *present, GrB_Index n, const GrB_Vector v);
'''


def emit_header():
    L = [HEADER_TOP, UNTYPED_STORAGE, "\n/* ---- per-type entry points ---- */"]
    L += typed_functions()
    L.append("\n/* ---- built-in types (graphblas/core/dtypes.py:154-245) ---- */")
    L += [f"extern GrB_Type GrB_{t};" for t, _, _ in TYPES]
    L.append("\n/* ---- built-in binary ops (accum / dup / eWise) ---- */")
    L += [f"extern GrB_BinaryOp {n};" for n, *_ in all_binops()]
    L.append("\n/* ---- built-in unary ops (apply) ---- */")
    L += [f"extern GrB_UnaryOp {n};" for n, *_ in all_unops()]
    L.append("\n/* ---- index-unary ops for select (tril/triu build the triangle-count input) ---- */")
    L += [f"extern GrB_IndexUnaryOp GrB_{n};" for n in IDXOPS]
    L.append("\n/* ---- built-in monoids ---- */")
    L += [f"extern GrB_Monoid {n};" for n, *_ in all_monoids()]
    L.append("\n/* ---- built-in semirings with sm_100a kernels (semiring.py:170-204 name patterns) ---- */")
    L += [f"extern GrB_Semiring {n};" for n, *_ in all_semirings()]
    L.append("\n/* ---- operator handles the reference's registry discovers by regex over dir(lib)")
    L.append(" *      (graphblas/core/operator/base.py:291,410-486) that have NO kernel in this library: they exist so that")
    L.append(" *      `import graphblas` builds its namespaces; using one in a call returns GrB_NOT_IMPLEMENTED. ---- */")
    L += [f"extern GrB_UnaryOp {n};" for n in registry_only_unops()]
    L += [f"extern GrB_BinaryOp {n};" for n in registry_only_binops()]
    L += [f"extern GrB_IndexUnaryOp {n};" for n in registry_only_idxops()]
    L += [f"extern GrB_Monoid {n};" for n in registry_only_monoids()]
    L += [f"extern GrB_Semiring {n};" for n in registry_only_semirings()]
    L.append("\n/* ---- predefined descriptors (graphblas/core/descriptor.py:51-87) ---- */")
    L += [f"extern GrB_Descriptor GrB_DESC_{d};" for d in DESC]
    L.append(HEADER_BOTTOM)
    return "\n".join(L)


def emit_lists():
    L = ["// GENERATED by tools/gen_abi.py — X-macro lists for the C++/CUDA side.", ""]
    L.append("#define GRB_TYPE_LIST(X) \\")
    L += [f"    X({t}, {c}, {code}) \\" for t, c, code in TYPES]
    L.append("")
    L.append("#define GRB_BINOP_LIST(X) \\")
    L += [f"    X({n}, {op}, {x}, {z}) \\" for n, op, x, z in all_binops()]
    L.append("")
    L.append("#define GRB_UNOP_LIST(X) \\")
    L += [f"    X({n}, {op}, {t}) \\" for n, op, t in all_unops()]
    L.append("")
    L.append("#define GRB_MONOID_LIST(X) \\")
    L += [f"    X({n}, {m}, {t}) \\" for n, m, t in all_monoids()]
    L.append("")
    L.append("#define GRB_SEMIRING_LIST(X) \\")
    L += [f"    X({n}, {m}, {u}, {t}) \\" for n, m, u, t in all_semirings()]
    L.append("")
    L.append("// (name, replace, comp, structure, tran0, tran1)")
    L.append("#define GRB_DESC_LIST(X) \\")
    for d in DESC:
        r = int(d.startswith("R"))
        rest = d[1:] if r else d
        s = int("S" in rest)
        c = int("C" in rest)
        t0 = int("T0" in rest)
        t1 = int("T1" in rest)
        L.append(f"    X({d}, {r}, {c}, {s}, {t0}, {t1}) \\")
    L.append("")
    # per-monoid semiring lists for the kernel instantiation units: X(MUL, TYPE)
    mons = sorted({m for m, _, _ in kernel_semirings()})
    for mon in mons:
        L.append(f"#define GRB_SEMIRINGS_OF_{mon}(X) \\")
        L += [f"    X({u}, {t}) \\" for m, u, t in kernel_semirings() if m == mon]
        L.append("")
    for name, fn in (("UNOP", registry_only_unops), ("BINOP", registry_only_binops), ("IDXOP", registry_only_idxops),
                     ("MONOID", registry_only_monoids), ("SEMIRING", registry_only_semirings)):
        L.append(f"#define GRB_{name}_STUB_LIST(X) \\")
        L += [f"    X({n}) \\" for n in fn()]
        L.append("")
    return "\n".join(L) + "\n"


if __name__ == "__main__":
    with open(os.path.join(ROOT, "include", "GraphBLAS_b200.h"), "w") as f:
        f.write(emit_header())
    with open(os.path.join(ROOT, "graph-python_b200", "csrc", "grb_abi_lists.inc"), "w") as f:
        f.write(emit_lists())
    print("semirings:", len(all_semirings()), "monoids:", len(all_monoids()), "binops:", len(all_binops()))
