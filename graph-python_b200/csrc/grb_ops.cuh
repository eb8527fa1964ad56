// grb_ops.cuh — built-in operators as device functors.
//   BinOp<OP,T>::apply     compile-time operator (semiring multiply / monoid), instantiated at build time
//   binop_rt<T>(op,x,y)    run-time switch (accumulators, dup ops, eWise ops: once per output entry)
//   Monoid<OP,T>           identity + atomic combine for scatter (push) kernels
// Integer arithmetic wraps exactly like C on two's complement (computed in the unsigned type), which is
// what makes the INT64 power-chain test (reference tests/test_matrix.py:4367-4395) bit-exact.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <limits>
#include <type_traits>

#include "grb_internal.cuh"

#define HD __host__ __device__ __forceinline__

template <class T> struct UnsignedOf { using type = typename std::make_unsigned<T>::type; };
template <> struct UnsignedOf<bool> { using type = bool; };
template <> struct UnsignedOf<float> { using type = float; };
template <> struct UnsignedOf<double> { using type = double; };

template <class T> HD T op_plus(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x || y;
    else if constexpr (std::is_floating_point<T>::value) return x + y;
    else { using U = typename UnsignedOf<T>::type; return (T)((U)x + (U)y); }
}
template <class T> HD T op_minus(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x != y;
    else if constexpr (std::is_floating_point<T>::value) return x - y;
    else { using U = typename UnsignedOf<T>::type; return (T)((U)x - (U)y); }
}
template <class T> HD T op_times(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x && y;
    else if constexpr (std::is_floating_point<T>::value) return x * y;
    else { using U = typename UnsignedOf<T>::type; return (T)((U)x * (U)y); }
}
template <class T> HD T op_min(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x && y;
    else if constexpr (std::is_same<T, float>::value) return fminf(x, y);
    else if constexpr (std::is_same<T, double>::value) return fmin(x, y);
    else return x < y ? x : y;
}
template <class T> HD T op_max(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x || y;
    else if constexpr (std::is_same<T, float>::value) return fmaxf(x, y);
    else if constexpr (std::is_same<T, double>::value) return fmax(x, y);
    else return x > y ? x : y;
}
template <class T> HD T op_div(T x, T y) {
    if constexpr (std::is_same<T, bool>::value) return x;
    else if constexpr (std::is_floating_point<T>::value) return x / y;
    else if constexpr (std::is_signed<T>::value) {
        // SuiteSparse integer division: x/0 saturates by sign, MIN/-1 wraps; never traps
        if (y == 0) return x == 0 ? (T)0 : (x < 0 ? std::numeric_limits<T>::min() : std::numeric_limits<T>::max());
        if (y == (T)-1) { using U = typename UnsignedOf<T>::type; return (T)((U)0 - (U)x); }
        return (T)(x / y);
    } else {
        if (y == 0) return x == 0 ? (T)0 : std::numeric_limits<T>::max();
        return (T)(x / y);
    }
}

template <int ID> struct NarrowType;
template <> struct NarrowType<NARROW_INT8> { using type = int8_t; };
template <> struct NarrowType<NARROW_INT16> { using type = int16_t; };
template <> struct NarrowType<NARROW_UINT8> { using type = uint8_t; };
template <> struct NarrowType<NARROW_UINT16> { using type = uint16_t; };
// primary template: a multiply that wraps to a narrow integer type (code = op | narrow id << 8, see grb_internal.cuh);
// the built-in operators are the specialisations below
template <int OP, class T> struct BinOp {
    static_assert(OP >= 256, "unknown operator code");
    using N = typename NarrowType<(OP >> 8)>::type;
    static HD T apply(T x, T y) { return (T)BinOp<(OP & 255), N>::apply((N)x, (N)y); }
};
template <class T> struct BinOp<OP_FIRST, T> { static HD T apply(T x, T) { return x; } };
template <class T> struct BinOp<OP_SECOND, T> { static HD T apply(T, T y) { return y; } };
template <class T> struct BinOp<OP_ANY, T> { static HD T apply(T x, T) { return x; } };
template <class T> struct BinOp<OP_ONEB, T> { static HD T apply(T, T) { return (T)1; } };
template <class T> struct BinOp<OP_PLUS, T> { static HD T apply(T x, T y) { return op_plus(x, y); } };
template <class T> struct BinOp<OP_MINUS, T> { static HD T apply(T x, T y) { return op_minus(x, y); } };
template <class T> struct BinOp<OP_TIMES, T> { static HD T apply(T x, T y) { return op_times(x, y); } };
template <class T> struct BinOp<OP_DIV, T> { static HD T apply(T x, T y) { return op_div(x, y); } };
template <class T> struct BinOp<OP_RDIV, T> { static HD T apply(T x, T y) { return op_div(y, x); } };
template <class T> struct BinOp<OP_MIN, T> { static HD T apply(T x, T y) { return op_min(x, y); } };
template <class T> struct BinOp<OP_MAX, T> { static HD T apply(T x, T y) { return op_max(x, y); } };
template <class T> struct BinOp<OP_LOR, T> { static HD T apply(T x, T y) { return (T)((x != (T)0) || (y != (T)0)); } };
template <class T> struct BinOp<OP_LAND, T> { static HD T apply(T x, T y) { return (T)((x != (T)0) && (y != (T)0)); } };
template <class T> struct BinOp<OP_LXOR, T> { static HD T apply(T x, T y) { return (T)((x != (T)0) != (y != (T)0)); } };
template <class T> struct BinOp<OP_LXNOR, T> { static HD T apply(T x, T y) { return (T)((x != (T)0) == (y != (T)0)); } };

// run-time operator (z type == x type); comparison ops are handled by cmp_rt below
template <class T> HD T binop_rt(int op, T x, T y) {
    switch (op) {
    case OP_FIRST: return x;
    case OP_SECOND: return y;
    case OP_ANY: return x;
    case OP_ONEB: return (T)1;
    case OP_PLUS: return op_plus(x, y);
    case OP_MINUS: return op_minus(x, y);
    case OP_TIMES: return op_times(x, y);
    case OP_DIV: return op_div(x, y);
    case OP_RDIV: return op_div(y, x);
    case OP_MIN: return op_min(x, y);
    case OP_MAX: return op_max(x, y);
    case OP_LOR: return (T)((x != (T)0) || (y != (T)0));
    case OP_LAND: return (T)((x != (T)0) && (y != (T)0));
    case OP_LXOR: return (T)((x != (T)0) != (y != (T)0));
    case OP_LXNOR: return (T)((x != (T)0) == (y != (T)0));
    default: return y;
    }
}
template <class T> HD bool cmp_rt(int op, T x, T y) {
    switch (op) {
    case OP_EQ: return x == y;
    case OP_NE: return x != y;
    case OP_GT: return x > y;
    case OP_LT: return x < y;
    case OP_GE: return x >= y;
    case OP_LE: return x <= y;
    default: return false;
    }
}

// ---- monoids: identity, terminal ("any" stops at the first product), atomic combine ----
template <class T> __device__ __forceinline__ void atomic_cas_combine_min(T *p, T v) {
    if constexpr (sizeof(T) == 4) {
        unsigned int *a = (unsigned int *)p, old = *a, assumed;
        do {
            assumed = old;
            T cur = __uint_as_float(assumed);
            T nv = op_min(cur, v);
            if (__float_as_uint(nv) == assumed) break;
            old = atomicCAS(a, assumed, __float_as_uint(nv));
        } while (old != assumed);
    } else {
        unsigned long long *a = (unsigned long long *)p, old = *a, assumed;
        do {
            assumed = old;
            T cur = __longlong_as_double((long long)assumed);
            T nv = op_min(cur, v);
            if ((unsigned long long)__double_as_longlong(nv) == assumed) break;
            old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(nv));
        } while (old != assumed);
    }
}
template <class T> __device__ __forceinline__ void atomic_cas_combine_max(T *p, T v) {
    if constexpr (sizeof(T) == 4) {
        unsigned int *a = (unsigned int *)p, old = *a, assumed;
        do {
            assumed = old;
            T cur = __uint_as_float(assumed);
            T nv = op_max(cur, v);
            if (__float_as_uint(nv) == assumed) break;
            old = atomicCAS(a, assumed, __float_as_uint(nv));
        } while (old != assumed);
    } else {
        unsigned long long *a = (unsigned long long *)p, old = *a, assumed;
        do {
            assumed = old;
            T cur = __longlong_as_double((long long)assumed);
            T nv = op_max(cur, v);
            if ((unsigned long long)__double_as_longlong(nv) == assumed) break;
            old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(nv));
        } while (old != assumed);
    }
}

// xor one byte inside its aligned 32-bit word (bool LXOR/LXNOR scatter)
__device__ __forceinline__ void atomic_xor_byte(bool *p, bool v) {
    if (!v) return;
    uintptr_t a = (uintptr_t)p;
    unsigned int *w = (unsigned int *)(a & ~(uintptr_t)3);
    atomicXor(w, 1u << (8 * (a & 3)));
}

template <int OP, class T> struct Monoid;
template <class T> struct Monoid<OP_PLUS, T> {
    static constexpr bool kAny = false;
    static HD T identity() { return (T)0; }
    static HD T apply(T x, T y) { return op_plus(x, y); }
    static __device__ __forceinline__ void atomic(T *p, T v) {
        if constexpr (std::is_same<T, int64_t>::value || std::is_same<T, uint64_t>::value)
            atomicAdd((unsigned long long *)p, (unsigned long long)v);
        else if constexpr (std::is_same<T, int32_t>::value) atomicAdd((int *)p, v);
        else if constexpr (std::is_same<T, uint32_t>::value) atomicAdd((unsigned int *)p, v);
        else atomicAdd(p, v);
    }
};
template <class T> struct Monoid<OP_MIN, T> {
    static constexpr bool kAny = false;
    static HD T identity() {
        if constexpr (std::is_floating_point<T>::value) return std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::max();
    }
    static HD T apply(T x, T y) { return op_min(x, y); }
    static __device__ __forceinline__ void atomic(T *p, T v) {
        if constexpr (std::is_floating_point<T>::value) atomic_cas_combine_min(p, v);
        else if constexpr (std::is_same<T, int64_t>::value) atomicMin((long long *)p, (long long)v);
        else if constexpr (std::is_same<T, uint64_t>::value) atomicMin((unsigned long long *)p, (unsigned long long)v);
        else if constexpr (std::is_same<T, int32_t>::value) atomicMin((int *)p, v);
        else atomicMin((unsigned int *)p, v);
    }
};
template <class T> struct Monoid<OP_MAX, T> {
    static constexpr bool kAny = false;
    static HD T identity() {
        if constexpr (std::is_floating_point<T>::value) return -std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::lowest();
    }
    static HD T apply(T x, T y) { return op_max(x, y); }
    static __device__ __forceinline__ void atomic(T *p, T v) {
        if constexpr (std::is_floating_point<T>::value) atomic_cas_combine_max(p, v);
        else if constexpr (std::is_same<T, int64_t>::value) atomicMax((long long *)p, (long long)v);
        else if constexpr (std::is_same<T, uint64_t>::value) atomicMax((unsigned long long *)p, (unsigned long long)v);
        else if constexpr (std::is_same<T, int32_t>::value) atomicMax((int *)p, v);
        else atomicMax((unsigned int *)p, v);
    }
};
template <class T> struct Monoid<OP_ANY, T> {
    static constexpr bool kAny = true;
    static HD T identity() { return (T)0; }
    static HD T apply(T x, T) { return x; }
    static __device__ __forceinline__ void atomic(T *p, T v) { *p = v; }  // any writer wins
};
template <> struct Monoid<OP_LOR, bool> {
    static constexpr bool kAny = false;
    static HD bool identity() { return false; }
    static HD bool apply(bool x, bool y) { return x || y; }
    static __device__ __forceinline__ void atomic(bool *p, bool v) { if (v) *p = true; }
};
template <> struct Monoid<OP_LAND, bool> {
    static constexpr bool kAny = false;
    static HD bool identity() { return true; }
    static HD bool apply(bool x, bool y) { return x && y; }
    static __device__ __forceinline__ void atomic(bool *p, bool v) { if (!v) *p = false; }
};
template <> struct Monoid<OP_LXOR, bool> {
    static constexpr bool kAny = false;
    static HD bool identity() { return false; }
    static HD bool apply(bool x, bool y) { return x != y; }
    static __device__ __forceinline__ void atomic(bool *p, bool v) { atomic_xor_byte(p, v); }
};
template <> struct Monoid<OP_LXNOR, bool> {
    static constexpr bool kAny = false;
    static HD bool identity() { return true; }
    static HD bool apply(bool x, bool y) { return x == y; }
    static __device__ __forceinline__ void atomic(bool *p, bool v) { atomic_xor_byte(p, !v); }
};

// ---- semiring = (monoid, multiply, type); which operand values the multiply reads decides the traffic ----
template <int MON, int MUL, class T> struct Semiring {
    using type = T;
    using Mon = Monoid<MON, T>;
    static constexpr int kMon = MON, kMul = MUL;
    static constexpr bool kUsesX = !(MUL == OP_SECOND || MUL == OP_ONEB);  // first operand's value
    static constexpr bool kUsesY = !(MUL == OP_FIRST || MUL == OP_ONEB);   // second operand's value
    static HD T mult(T x, T y) { return BinOp<MUL, T>::apply(x, y); }
};

// run-time monoid identity (host side: pre-filling dense accumulators)
template <class T> inline T monoid_identity_rt(int mon) {
    switch (mon) {
    case OP_MIN:
        if constexpr (std::is_same<T, bool>::value) return true;
        else if constexpr (std::is_floating_point<T>::value) return std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::max();
    case OP_MAX:
        if constexpr (std::is_same<T, bool>::value) return false;
        else if constexpr (std::is_floating_point<T>::value) return -std::numeric_limits<T>::infinity();
        else return std::numeric_limits<T>::lowest();
    case OP_LAND: case OP_LXNOR: case OP_TIMES: return (T)1;
    default: return (T)0;
    }
}
