// grb_spmv_launch.cuh — type-erased launchers; the semiring kernels are instantiated per monoid in
// grb_spmv_inst_<MON>.cu (one translation unit each so the build parallelises).
#pragma once
#include "grb_internal.cuh"

enum PullMode { PULL_STREAM = 0, PULL_ANY = 1 };

// args point to PullArgs<T> / PushArgs<T> of the semiring's type.  Return false if (mon, mul, type) has no kernel.
bool spmv_launch_pull(int mon, int mul, int tcode, const void *pull_args, int mode);
bool spmv_launch_push(int mon, int mul, int tcode, const void *push_args);
bool spmv_launch_stream(int mon, int mul, int tcode, const void *stream_args);
bool spmv_launch_red(int mon, int mul, int tcode, const void *red_args);         // RedArgs<T>: atomic-segment pull   // StreamArgs<T>: u full, stream form cached

#define SPMV_DECL(MON)                                                                   \
    bool spmv_pull_##MON(int mul, int tcode, const void *pull_args, int mode);           \
    bool spmv_push_##MON(int mul, int tcode, const void *push_args);           \
    bool spmv_stream_##MON(int mul, int tcode, const void *stream_args);       \
    bool spmv_red_##MON(int mul, int tcode, const void *red_args);
SPMV_DECL(PLUS) SPMV_DECL(MIN) SPMV_DECL(MAX) SPMV_DECL(ANY) SPMV_DECL(BOOLMON)
#undef SPMV_DECL

#define CT_BOOL bool
#define CT_INT8 int8_t
#define CT_INT16 int16_t
#define CT_INT32 int32_t
#define CT_INT64 int64_t
#define CT_UINT8 uint8_t
#define CT_UINT16 uint16_t
#define CT_UINT32 uint32_t
#define CT_UINT64 uint64_t
#define CT_FP32 float
#define CT_FP64 double
#define OP_PAIR OP_ONEB
