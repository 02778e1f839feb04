// Runtime (dtype, io dtype, op, vector/scalar) -> template instantiation.
//
// Capability parity: the dispatch switch of handle_reduce (allreduce_over_mpi/mpi_mod.hpp:825-874).
// Differences: unknown combinations return false instead of exit(1); signed integers reuse the
// unsigned instantiation for SUM/BAND/BOR (identical bits under two's complement); bool is handled
// as a 0/1 byte with SUM == logical OR.
//
// GROUPS lets a translation unit compile only a slice of the type matrix so the CUDA build can be
// spread over several files:  bit0 fp32 | bit1 fp16+bf16 (+fp32 user I/O) | bit2 64-bit | bit3 32-bit
// ints | bit4 8/16-bit ints + bool.
#pragma once
#include <stdint.h>

#include "types.h"

namespace ft {

enum TypeGroup { TG_F32 = 1, TG_F16 = 2, TG_64 = 4, TG_I32 = 8, TG_SMALL = 16, TG_ALL = 31 };

inline int dtype_group(int dt) {
  switch (dt) {
    case DT_FLOAT: return TG_F32;
    case DT_HALF: case DT_BFLOAT16: case DT_FP8_BLOCK: return TG_F16;
    case DT_DOUBLE: case DT_INT64: case DT_UINT64: case DT_LONG_LONG: return TG_64;
    case DT_INT32: case DT_UINT32: return TG_I32;
    default: return TG_SMALL;
  }
}

// F is `template <typename T, typename IO, class Op, int VE> struct F { static void run(Args...); }`.
template <template <typename, typename, class, int> class F, int GROUPS, typename... Args>
struct Dispatcher {
  template <typename T, typename IO, class Op>
  static bool vec(bool vectorized, Args... args) {
    if (vectorized) F<T, IO, Op, VecElems<T>::value>::run(args...);
    else F<T, IO, Op, 1>::run(args...);
    return true;
  }
  template <typename T, typename IO>
  static bool float_ops(int op, bool v, Args... args) {
    switch (op) {
      case OP_SUM: return vec<T, IO, OpSum>(v, args...);
      case OP_MAX: return vec<T, IO, OpMax>(v, args...);
      case OP_MIN: return vec<T, IO, OpMin>(v, args...);
    }
    return false;
  }
  template <typename S, typename U>
  static bool int_ops(int op, bool v, Args... args) {
    switch (op) {
      case OP_SUM: return vec<U, U, OpSum>(v, args...);
      case OP_BAND: return vec<U, U, OpBand>(v, args...);
      case OP_BOR: return vec<U, U, OpBor>(v, args...);
      case OP_MAX: return vec<S, S, OpMax>(v, args...);
      case OP_MIN: return vec<S, S, OpMin>(v, args...);
    }
    return false;
  }
  static bool run(int dtype, int io_dtype, int op, bool v, Args... args) {
    if (!op_supported(op, dtype)) return false;
    if (io_dtype != dtype) {
      // only "fp32 user buffers, 16-bit wire" is defined (gradient compression on the DDP path)
      if constexpr ((GROUPS & TG_F16) != 0) {
        if (io_dtype == DT_FLOAT && dtype == DT_BFLOAT16) return float_ops<bf16_t, float>(op, v, args...);
        if (io_dtype == DT_FLOAT && dtype == DT_HALF) return float_ops<f16_t, float>(op, v, args...);
        // block-scaled fp8 wire: SUM only, vector path only (a pack is 16 values + a scale byte)
        if (dtype == DT_FP8_BLOCK && op == OP_SUM && v) {
          if (io_dtype == DT_FLOAT) { F<fp8_t, float, OpSum, 16>::run(args...); return true; }
          if (io_dtype == DT_BFLOAT16) { F<fp8_t, bf16_t, OpSum, 16>::run(args...); return true; }
        }
      }
      return false;
    }
    if constexpr ((GROUPS & TG_F32) != 0) {
      if (dtype == DT_FLOAT) return float_ops<float, float>(op, v, args...);
    }
    if constexpr ((GROUPS & TG_F16) != 0) {
      if (dtype == DT_BFLOAT16) return float_ops<bf16_t, bf16_t>(op, v, args...);
      if (dtype == DT_HALF) return float_ops<f16_t, f16_t>(op, v, args...);
    }
    if constexpr ((GROUPS & TG_64) != 0) {
      if (dtype == DT_DOUBLE) return float_ops<double, double>(op, v, args...);
      if (dtype == DT_INT64 || dtype == DT_LONG_LONG) return int_ops<int64_t, uint64_t>(op, v, args...);
      if (dtype == DT_UINT64) return int_ops<uint64_t, uint64_t>(op, v, args...);
    }
    if constexpr ((GROUPS & TG_I32) != 0) {
      if (dtype == DT_INT32) return int_ops<int32_t, uint32_t>(op, v, args...);
      if (dtype == DT_UINT32) return int_ops<uint32_t, uint32_t>(op, v, args...);
    }
    if constexpr ((GROUPS & TG_SMALL) != 0) {
      if (dtype == DT_INT8) return int_ops<int8_t, uint8_t>(op, v, args...);
      if (dtype == DT_UINT8) return int_ops<uint8_t, uint8_t>(op, v, args...);
      if (dtype == DT_INT16) return int_ops<int16_t, uint16_t>(op, v, args...);
      if (dtype == DT_UINT16) return int_ops<uint16_t, uint16_t>(op, v, args...);
      if (dtype == DT_BOOL) {
        switch (op) {  // 0/1 bytes: SUM, MAX, BOR == any ; BAND, MIN == all
          case OP_SUM: case OP_MAX: case OP_BOR: return vec<bool8_t, bool8_t, OpBoolOr>(v, args...);
          case OP_BAND: case OP_MIN: return vec<bool8_t, bool8_t, OpBand>(v, args...);
        }
      }
    }
    return false;
  }
};

}  // namespace ft
