// CUDA back end of the algorithm templates (core/staged_core.h): one instance per CTA.
//
// Everything that touches another GPU's memory is here, as inline PTX for sm_100a:
//   peer payload loads   ld.relaxed.sys.global.L1::no_allocate.v4  -> LDG.E.NA.128.STRONG.SYS
//   flag publish         fence.acq_rel.sys ; st.relaxed.sys.u64   -> MEMBAR.ALL.SYS + CCTL.IVALL ; STG.STRONG.SYS
//   flag poll            ld.relaxed.sys.u64 (+ fence.acq_rel.sys once it is seen)
//   NVLS                 multimem.ld_reduce.relaxed.sys.add[.acc::f32].v4  -> LDGMC.E.ADD.* ; multimem.st
//   LL lines             st/ld.relaxed.sys.v4.u32 (two {data,flag} halves)
// These replace MPI_Isend/Irecv/Waitall/Barrier of the reference data plane
// (allreduce_over_mpi/mpi_mod.hpp:663-765, :1003, :1027-1028).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../core/comm_dev.h"
#include "../core/types.h"
#include "tma_tiles.cuh"
#include "nvls_ws.cuh"

namespace ft {

struct CudaBE {
  static constexpr bool kTma = true;
  static constexpr bool kBulkCopyIn = false;  // CPU back end only (host/cpu_backend.h)
  int* abort_;
  TmaRing ring_;
  __device__ __forceinline__ CudaBE(int* abort_smem, const TmaRing& ring) : abort_(abort_smem), ring_(ring) {}
  template <typename T, typename IO, class Op, int VE, class Cta>
  __device__ __forceinline__ void rs_tile_tma(Cta& c, int k, int t) { rs_tile_tma_impl<T, IO, Op, VE>(c, ring_, k, t); }
  template <typename T, typename IO, class Op, int VE, class Cta>
  __device__ __forceinline__ void ag_tile_tma(Cta& c, int k, int t) { ag_tile_tma_impl<T, IO, Op, VE>(c, ring_, k, t); }
  template <typename T, typename IO, class Op, int VE, bool EMU, class Cta>
  __device__ __forceinline__ void nvls_tile_ws(Cta& c, int t) { nvls_tile_ws_impl<T, IO, Op, VE, EMU>(c, ring_.ring, t); }
  __device__ __forceinline__ int tid() const { return (int)threadIdx.x; }
  __device__ __forceinline__ int nthr() const { return (int)blockDim.x; }
  __device__ __forceinline__ void sync() { __syncthreads(); }
  __device__ __forceinline__ int* abort_ptr() { return abort_; }

  // ---- payload ------------------------------------------------------------------------------------
  template <int B>
  static __device__ __forceinline__ void ld(void* dst, const void* src, bool remote) {
    if constexpr (B > 16) {
      static_assert(B % 16 == 0, "unsupported access width");
#pragma unroll
      for (int i = 0; i < B / 16; ++i) ld<16>((char*)dst + 16 * i, (const char*)src + 16 * i, remote);
    } else if constexpr (B == 16) {
      uint4 v;
      if (remote)
        asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src) : "memory");
      else
        v = *reinterpret_cast<const uint4*>(src);
      *reinterpret_cast<uint4*>(dst) = v;
    } else if constexpr (B == 8) {
      uint2 v;
      if (remote)
        asm volatile("ld.relaxed.sys.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
                     : "=r"(v.x), "=r"(v.y) : "l"(src) : "memory");
      else
        v = *reinterpret_cast<const uint2*>(src);
      *reinterpret_cast<uint2*>(dst) = v;
    } else if constexpr (B == 4) {
      uint32_t v;
      if (remote) asm volatile("ld.relaxed.sys.global.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
      else v = *reinterpret_cast<const uint32_t*>(src);
      *reinterpret_cast<uint32_t*>(dst) = v;
    } else if constexpr (B == 2) {
      uint16_t v;
      if (remote) asm volatile("ld.relaxed.sys.global.L1::no_allocate.u16 %0, [%1];" : "=h"(v) : "l"(src) : "memory");
      else v = *reinterpret_cast<const uint16_t*>(src);
      *reinterpret_cast<uint16_t*>(dst) = v;
    } else {
      static_assert(B == 1, "unsupported access width");
      uint32_t v;
      if (remote) asm volatile("ld.relaxed.sys.global.L1::no_allocate.u8 %0, [%1];" : "=r"(v) : "l"(src) : "memory");
      else v = *reinterpret_cast<const uint8_t*>(src);
      *reinterpret_cast<uint8_t*>(dst) = (uint8_t)v;
    }
  }
  template <int B>
  static __device__ __forceinline__ void st(void* dst, const void* src) {
    if constexpr (B > 16) {
      static_assert(B % 16 == 0, "unsupported access width");
#pragma unroll
      for (int i = 0; i < B / 16; ++i) st<16>((char*)dst + 16 * i, (const char*)src + 16 * i);
    } else if constexpr (B == 16) {
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
    } else if constexpr (B == 8) {
      *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(src);
    } else if constexpr (B == 4) {
      *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(src);
    } else if constexpr (B == 2) {
      *reinterpret_cast<uint16_t*>(dst) = *reinterpret_cast<const uint16_t*>(src);
    } else {
      static_assert(B == 1, "unsupported access width");
      *reinterpret_cast<uint8_t*>(dst) = *reinterpret_cast<const uint8_t*>(src);
    }
  }

  // ---- flags --------------------------------------------------------------------------------------
  __device__ __forceinline__ uint64_t flag_ld(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
  }
  __device__ __forceinline__ void flag_st(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
  }
  __device__ __forceinline__ void fence() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
  // Acquire side of a flag hop: after the relaxed spin has seen the value, one ld.acquire.sys of the same
  // word orders everything after it (LDG.STRONG.SYS + CCTL.IVALL, no MEMBAR.SYS - measurably cheaper than a
  // full fence on every hop).
  __device__ __forceinline__ void acquire(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    (void)v;
  }
  __device__ __forceinline__ uint64_t load_epoch(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
  }
  __device__ __forceinline__ unsigned err_ld(const unsigned* p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
  }
  // First error wins. The winner also writes the mirror in mapped host memory, which the host checks at the
  // start of the next call without synchronising anything.
  __device__ __forceinline__ void err_set(const CommDev& cd, unsigned v) {
    if (atomicCAS(cd.err, 0u, v) == 0u && cd.err_host) {
      *(volatile unsigned*)cd.err_host = v;
      __threadfence_system();
    }
  }
  __device__ __forceinline__ uint64_t now_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
  }
  __device__ __forceinline__ void pause() {}

  // Last CTA of the launch advances the communicator's epoch (device-resident: graph replays need no
  // host-side reset).
  __device__ __forceinline__ void finish(const CommDev& cd, int ctas) {
    __syncthreads();
    if (threadIdx.x == 0) {
      uint64_t epoch = load_epoch(cd.epoch);
      __threadfence();
      unsigned old = atomicAdd(cd.done, 1u);
      if (old == (unsigned)ctas - 1u) {
        atomicExch(cd.done, 0u);
        __threadfence();
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(cd.epoch), "l"(epoch + 1) : "memory");
      }
    }
  }

  // ---- NVLS ---------------------------------------------------------------------------------------
  // No multicast mapping (single-GPU emulation, where every "peer" is a slice of this GPU): the switch is played by
  // a loop over the N windows - sum in rank order, store to all - so that the NVLS control flow (flags, tiles,
  // inner-loop variants, graph replay) is testable on one GPU. Never taken on the multi-GPU path (rc.mc is set).
  template <typename T, class Op, int VE>
  __device__ __forceinline__ Pack<T, VE> mc_ld_reduce_emulated(const RankCall& rc, long long off, int nranks) {
    using EL = Elem<T>;
    typename EL::acc acc[VE];
    for (int p = 0; p < nranks; ++p) {
      Pack<T, VE> v;
      ld<(int)sizeof(Pack<T, VE>)>(&v, rc.win[p] + off, true);
#pragma unroll
      for (int i = 0; i < VE; ++i) acc[i] = p ? Op::apply(acc[i], EL::up(v.e[i])) : EL::up(v.e[i]);
    }
    Pack<T, VE> o;
#pragma unroll
    for (int i = 0; i < VE; ++i) o.e[i] = EL::down(acc[i]);
    return o;
  }
  template <typename T, class Op, int VE, bool EMU = false>
  __device__ __forceinline__ Pack<T, VE> mc_ld_reduce(const RankCall& rc, long long off, int nranks) {
    Pack<T, VE> o;
    if constexpr (EMU) return mc_ld_reduce_emulated<T, Op, VE>(rc, off, nranks);
    const char* p = rc.mc + off;
    if constexpr (sizeof(Pack<T, VE>) == 16 && SameT<T, float>::v && SameT<Op, OpSum>::v) {
      uint4 v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
      *reinterpret_cast<uint4*>(&o) = v;
    } else if constexpr (sizeof(Pack<T, VE>) == 16 && SameT<T, bf16_t>::v && SameT<Op, OpSum>::v) {
      uint4 v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
      *reinterpret_cast<uint4*>(&o) = v;
    } else if constexpr (sizeof(Pack<T, VE>) == 16 && SameT<T, f16_t>::v && SameT<Op, OpSum>::v) {
      uint4 v;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                   : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
      *reinterpret_cast<uint4*>(&o) = v;
    } else {
      __trap();  // host never launches NVLS for other (type, op, width) combinations
    }
    return o;
  }
  template <typename T, int VE, bool EMU = false>
  __device__ __forceinline__ void mc_st(const RankCall& rc, long long off, const Pack<T, VE>& v, int nranks) {
    if constexpr (EMU) {  // emulation, see above
      for (int p = 0; p < nranks; ++p) st<(int)sizeof(Pack<T, VE>)>(rc.win[p] + off, &v);
      return;
    }
    if constexpr (sizeof(Pack<T, VE>) == 16) {
      const uint4 u = *reinterpret_cast<const uint4*>(&v);
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(rc.mc + off), "r"(u.x),
                   "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
    } else {
      __trap();
    }
  }

  // ---- LL lines -----------------------------------------------------------------------------------
  __device__ __forceinline__ void ll_st(char* line, uint32_t d0, uint32_t d1, uint32_t tag) {
    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(line), "r"(d0), "r"(tag), "r"(d1),
                 "r"(tag) : "memory");
  }
  __device__ __forceinline__ bool ll_ld(const char* line, uint32_t tag, uint32_t* data) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(line) : "memory");
    data[0] = v.x;
    data[1] = v.z;
    return v.y == tag && v.w == tag;
  }

  template <typename X, typename Y> struct SameT { static constexpr bool v = false; };
  template <typename X> struct SameT<X, X> { static constexpr bool v = true; };
};

// Which (type, op) the switch can reduce (multimem.ld_reduce); everything else uses the P2P stages.
template <typename T, class Op> struct NvlsOk { static constexpr bool value = false; };
template <> struct NvlsOk<float, OpSum> { static constexpr bool value = true; };
template <> struct NvlsOk<bf16_t, OpSum> { static constexpr bool value = true; };
template <> struct NvlsOk<f16_t, OpSum> { static constexpr bool value = true; };

}  // namespace ft
