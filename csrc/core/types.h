// Element types, reduction operators and 16-byte packs shared by the CPU model and the CUDA kernels.
//
// Capability parity: the (op x dtype) dispatch of handle_reduce and the k-way reducers
// reduce_sum<T> / reduce_band<T> (allreduce_over_mpi/mpi_mod.hpp:246-660, :825-874). The reference
// supports SUM over 10 dtypes and BAND over the integer ones and exits on anything else; this
// version adds MAX / MIN / BOR and fp16 / bf16 / uint32 / uint64, and accumulates 16-bit floats in
// fp32 inside a stage.
#pragma once
#include <stdint.h>
#include <string.h>

#include "plan.h"

#if defined(__CUDACC__)
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#endif

namespace ft {

// Numeric values are part of the C ABI (include/flextree.h).
enum DType {
  DT_UINT8 = 0, DT_INT8 = 1, DT_UINT16 = 2, DT_INT16 = 3, DT_INT32 = 4, DT_INT64 = 5,
  DT_FLOAT = 6, DT_DOUBLE = 7, DT_BOOL = 8, DT_LONG_LONG = 9,
  DT_UINT32 = 10, DT_UINT64 = 11, DT_HALF = 12, DT_BFLOAT16 = 13,
  DT_FP8_BLOCK = 14,  // wire-only: e4m3 values + one power-of-two scale byte per 16 elements (opt-in)
  DT_COUNT = 15
};
enum RedOp { OP_SUM = 0, OP_BAND = 1, OP_MAX = 2, OP_MIN = 3, OP_BOR = 4, OP_COUNT = 5 };

inline int dtype_size(int dt) {
  switch (dt) {
    case DT_UINT8: case DT_INT8: case DT_BOOL: case DT_FP8_BLOCK: return 1;
    case DT_UINT16: case DT_INT16: case DT_HALF: case DT_BFLOAT16: return 2;
    case DT_INT32: case DT_UINT32: case DT_FLOAT: return 4;
    case DT_INT64: case DT_UINT64: case DT_DOUBLE: case DT_LONG_LONG: return 8;
  }
  return 0;
}
inline bool dtype_is_float(int dt) {
  return dt == DT_FLOAT || dt == DT_DOUBLE || dt == DT_HALF || dt == DT_BFLOAT16 || dt == DT_FP8_BLOCK;
}
inline const char* dtype_name(int dt) {
  static const char* n[] = {"uint8", "int8", "uint16", "int16", "int32", "int64", "float32",
                            "float64", "bool", "longlong", "uint32", "uint64", "float16", "bfloat16", "fp8-block16"};
  return (dt >= 0 && dt < DT_COUNT) ? n[dt] : "?";
}
inline const char* op_name(int op) {
  static const char* n[] = {"sum", "band", "max", "min", "bor"};
  return (op >= 0 && op < OP_COUNT) ? n[op] : "?";
}
// Is (op, dtype) defined? Bitwise ops need integers (as in the reference, mpi_mod.hpp:851-858).
inline bool op_supported(int op, int dt) {
  if (op < 0 || op >= OP_COUNT || dt < 0 || dt >= DT_COUNT) return false;
  if ((op == OP_BAND || op == OP_BOR) && dtype_is_float(dt)) return false;
  return true;
}

// ---- 16-bit float storage types usable from plain C++ -------------------------------------------
struct bf16_t { uint16_t v; };
struct f16_t { uint16_t v; };
struct bool8_t { uint8_t v; };  // MPI_C_BOOL: SUM behaves as logical OR, BAND as logical AND

FT_HD float bits_to_float(uint32_t u) {
#if defined(__CUDA_ARCH__)
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}
FT_HD uint32_t float_to_bits(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_uint(f);
#else
  uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}

FT_HD float to_float(bf16_t x) { return bits_to_float((uint32_t)x.v << 16); }
FT_HD bf16_t bf16_from_float(float f) {
  bf16_t r;
#if defined(__CUDA_ARCH__)
  r.v = __bfloat16_as_ushort(__float2bfloat16_rn(f));
#else
  uint32_t u = float_to_bits(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) { r.v = (uint16_t)((u >> 16) | 0x40); return r; }  // quiet NaN
  u += 0x7fffu + ((u >> 16) & 1u);  // round to nearest even
  r.v = (uint16_t)(u >> 16);
#endif
  return r;
}

FT_HD float to_float(f16_t x) {
#if defined(__CUDA_ARCH__)
  return __half2float(__ushort_as_half(x.v));
#else
  uint32_t s = (uint32_t)(x.v & 0x8000u) << 16, e = (x.v >> 10) & 0x1f, m = x.v & 0x3ffu;
  if (e == 0) {
    if (m == 0) return bits_to_float(s);
    int sh = 0;  // subnormal: normalise
    while (!(m & 0x400u)) { m <<= 1; ++sh; }
    m &= 0x3ffu;
    return bits_to_float(s | ((uint32_t)(127 - 15 - sh + 1) << 23) | (m << 13));
  }
  if (e == 31) return bits_to_float(s | 0x7f800000u | (m << 13));
  return bits_to_float(s | ((e + 112) << 23) | (m << 13));
#endif
}
FT_HD f16_t f16_from_float(float f) {
  f16_t r;
#if defined(__CUDA_ARCH__)
  r.v = __half_as_ushort(__float2half_rn(f));
#else
  uint32_t u = float_to_bits(f), s = (u >> 16) & 0x8000u, a = u & 0x7fffffffu;
  if (a > 0x7f800000u) { r.v = (uint16_t)(s | 0x7e00u); return r; }       // NaN
  if (a >= 0x47800000u) { r.v = (uint16_t)(s | 0x7c00u); return r; }      // overflow -> inf (incl. inf)
  if (a < 0x33000001u) { r.v = (uint16_t)s; return r; }                   // underflow -> 0
  int e = (int)(a >> 23) - 127;
  uint32_t m = (a & 0x7fffffu) | 0x800000u;
  int shift = e < -14 ? (13 + (-14 - e)) : 13;   // bits dropped from the 24-bit significand
  uint32_t half_m = m >> shift, rem = m & ((1u << shift) - 1), halfway = 1u << (shift - 1);
  if (rem > halfway || (rem == halfway && (half_m & 1u))) ++half_m;
  uint32_t he = e < -14 ? 0 : (uint32_t)(e + 15);
  // half_m includes the implicit bit for normals; adding handles mantissa carry into the exponent
  uint32_t out = e < -14 ? half_m : ((he << 10) + (half_m - 0x400u));
  r.v = (uint16_t)(s | out);
#endif
  return r;
}

// ---- block-scaled fp8 wire format (opt-in) ------------------------------------------------------------------
// A 16-byte pack holds 16 e4m3 values; their common scale 2^(s-127) travels as one byte s in a separate
// "scale plane" of the window (1/16 of the payload). Every reduce-scatter stage dequantises, sums in
// fp32 and re-quantises with a fresh scale, so the error grows with the number of stages (measured
// bounds in the tests). Conversions are plain integer code so that CPU model and GPU agree bit for bit.
struct fp8_t { uint8_t v; };

FT_HD float e4m3_to_float(uint8_t b) {
  const uint32_t sign = (uint32_t)(b & 0x80) << 24, e = (b >> 3) & 0xf, m = b & 7;
  if (e == 0xf && m == 7) return bits_to_float(sign | 0x7fc00000u);  // NaN (e4m3 has no inf)
  if (e == 0) {                                                          // subnormal: m * 2^-9
    float f = (float)m * 0.001953125f;
    return sign ? -f : f;
  }
  return bits_to_float(sign | ((e + 120) << 23) | (m << 20));
}
// round to nearest even, saturate to +-448, NaN -> 0x7f
FT_HD uint8_t float_to_e4m3(float x) {
  const uint32_t u = float_to_bits(x), sign = (u >> 24) & 0x80, a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint8_t)(sign | 0x7f);
  if (a >= 0x43e00000u) return (uint8_t)(sign | 0x7e);  // >= 448 (also inf): saturate
  if (a < 0x3a800000u) return (uint8_t)sign;             // < 2^-10: below half of the smallest subnormal (2^-9)
  int e = (int)(a >> 23) - 127;                          // unbiased exponent
  uint32_t m = (a & 0x7fffffu) | 0x800000u;              // 24-bit significand
  int shift = e < -6 ? (20 + (-6 - e)) : 20;             // bits dropped (subnormal results drop more)
  uint32_t q = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
  if (rem > half || (rem == half && (q & 1u))) ++q;
  uint32_t out = e < -6 ? q : (((uint32_t)(e + 7) << 3) + (q - 8u));  // q carries the implicit bit for normals
  if (out > 0x7e) out = 0x7e;
  return (uint8_t)(sign | out);
}
// scale byte for a block whose largest magnitude is amax: smallest power of two with amax / 2^e <= 448
FT_HD uint8_t fp8_block_scale(float amax) {
  const uint32_t a = float_to_bits(amax) & 0x7fffffffu;
  if (a == 0 || a >= 0x7f800000u) return a == 0 ? 0 : 254;
  int ex = (int)(a >> 23) - 126;                  // amax = m * 2^ex, m in [0.5, 1)
  int e = ((a & 0x7fffffu) <= 0x600000u) ? ex - 9 : ex - 8;  // 448 = 0.875 * 2^9
  int s = e + 127;
  return (uint8_t)(s < 0 ? 0 : (s > 254 ? 254 : s));
}
FT_HD float fp8_scale_value(uint8_t s) { return bits_to_float(s == 0 ? 0x00400000u : ((uint32_t)s << 23)); }      // 2^(s-127)
FT_HD float fp8_scale_inverse(uint8_t s) { return bits_to_float(s >= 254 ? 0x00400000u : ((uint32_t)(254 - s) << 23)); }  // 2^(127-s)

#if defined(__CUDA_ARCH__)
// Hardware conversions for the GPU side of the fp8 wire (sm_100a: F2FP.SATFINITE.E4M3 / F2FP.F16.E4M3 in SASS). They
// agree bit for bit with the integer code above, which the CPU back end keeps using (round to nearest even, saturate
// to +-448, NaN -> 0x7f; e4m3 -> f16 is exact), so CPU model and GPU still produce identical bytes.
__device__ __forceinline__ void e4m3x4_to_float4(uint32_t w, float* f) {
  uint32_t h01, h23;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h01) : "h"((unsigned short)(w & 0xffffu)));
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h23) : "h"((unsigned short)(w >> 16)));
  const __half2 a = *reinterpret_cast<const __half2*>(&h01), b = *reinterpret_cast<const __half2*>(&h23);
  f[0] = __low2float(a);
  f[1] = __high2float(a);
  f[2] = __low2float(b);
  f[3] = __high2float(b);
}
__device__ __forceinline__ uint32_t float4_to_e4m3x4(float f0, float f1, float f2, float f3) {
  unsigned short lo, hi;  // cvt packs its FIRST source into the upper byte
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(lo) : "f"(f1), "f"(f0));
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi) : "f"(f3), "f"(f2));
  return (uint32_t)lo | ((uint32_t)hi << 16);
}
#endif

// ---- accumulator traits ---------------------------------------------------------------------------
template <typename T> struct Elem {  // default: integers and fp32/fp64 accumulate in themselves
  using acc = T;
  static FT_HD acc up(T x) { return x; }
  static FT_HD T down(acc a) { return a; }
  static constexpr bool is_float = false;
};
template <> struct Elem<float> {
  using acc = float;
  static FT_HD acc up(float x) { return x; }
  static FT_HD float down(acc a) { return a; }
  static constexpr bool is_float = true;
};
template <> struct Elem<double> {
  using acc = double;
  static FT_HD acc up(double x) { return x; }
  static FT_HD double down(acc a) { return a; }
  static constexpr bool is_float = true;
};
template <> struct Elem<bf16_t> {
  using acc = float;
  static FT_HD acc up(bf16_t x) { return to_float(x); }
  static FT_HD bf16_t down(acc a) { return bf16_from_float(a); }
  static constexpr bool is_float = true;
};
template <> struct Elem<f16_t> {
  using acc = float;
  static FT_HD acc up(f16_t x) { return to_float(x); }
  static FT_HD f16_t down(acc a) { return f16_from_float(a); }
  static constexpr bool is_float = true;
};
template <> struct Elem<fp8_t> {  // per-element view WITHOUT the block scale: only the quantised tile functions use it
  using acc = float;
  static FT_HD acc up(fp8_t x) { return e4m3_to_float(x.v); }
  static FT_HD fp8_t down(acc a) { fp8_t r; r.v = float_to_e4m3(a); return r; }
  static constexpr bool is_float = true;
};
template <> struct Elem<bool8_t> {
  using acc = uint8_t;
  static FT_HD acc up(bool8_t x) { return x.v ? 1 : 0; }
  static FT_HD bool8_t down(acc a) { bool8_t r; r.v = a ? 1 : 0; return r; }
  static constexpr bool is_float = false;
};

// ---- operators (applied on accumulators) ----------------------------------------------------------
struct OpSum { template <typename A> static FT_HD A apply(A a, A b) { return (A)(a + b); } };
struct OpMax { template <typename A> static FT_HD A apply(A a, A b) { return a > b ? a : (b > a ? b : (a == a ? a : b)); } };
struct OpMin { template <typename A> static FT_HD A apply(A a, A b) { return a < b ? a : (b < a ? b : (a == a ? a : b)); } };
struct OpBand { template <typename A> static FT_HD A apply(A a, A b) { return (A)(a & b); } };
struct OpBor { template <typename A> static FT_HD A apply(A a, A b) { return (A)(a | b); } };
// bool: SUM == OR (any), BAND == AND; on the 0/1 accumulator '+' could overflow to 2, so map it.
struct OpBoolOr { template <typename A> static FT_HD A apply(A a, A b) { return (A)((a | b) ? 1 : 0); } };

// ---- packs ----------------------------------------------------------------------------------------
// A Pack is what one thread moves per memory instruction: 16 bytes on the vector path, one element
// on the scalar (unaligned user buffer) path.
template <typename T, int VE> struct alignas(sizeof(T) * VE >= 16 ? 16 : sizeof(T) * VE) Pack { T e[VE]; };

template <typename T> struct VecElems { static constexpr int value = 16 / (int)sizeof(T); };

}  // namespace ft
