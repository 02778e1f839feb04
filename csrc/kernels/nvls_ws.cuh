// Warp-specialised in-switch reduce of one tile (CUDA only; the plain loops live in core/staged_core.h).
//
// In the plain loops one thread issues a batch of multimem.ld_reduce, waits for the switch to return the sums and
// only then issues the multimem.st that broadcasts them: per thread the inbound and the outbound direction of the
// links alternate. Here the two halves are decoupled: the LOWER half of the CTA's warps only pulls
// (multimem.ld_reduce -> registers -> shared-memory ring), the UPPER half only pushes (ring -> registers ->
// multimem.st), so both instruction streams are always in flight. Hand-off is by named barriers, one "full" and one
// "empty" per ring stage; bar.sync / bar.arrive order the shared-memory accesses of the participating threads.
//   loaders : [wait empty[s]]  ld_reduce x UNR  st.shared          arrive full[s]
//   storers :  wait full[s]    ld.shared  (scale)  multimem.st     [arrive empty[s]]
// An "empty" is only signalled when a later iteration will wait for it, so every barrier is back in its initial
// state when the tile ends (tiles follow each other in the same CTA).
// No counterpart in the reference (its reduction is a CPU loop after MPI_Waitall, mpi_mod.hpp:1003-1017).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../core/staged_core.h"
#include "launch.h"

namespace ft {

constexpr int kNvlsWsStages = 2;
constexpr int kNvlsWsUnroll = 16;  // a loader has one batch in flight at a time (it must see the data to park it)
// ring bytes = stages x (threads / 2) x unroll x 16 B  = 64 KiB at the 256 threads of the NVLS kernel
constexpr int kNvlsWsSmemBytes = kNvlsWsStages * (FT_NVLS_THREADS / 2) * kNvlsWsUnroll * 16;

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

template <typename T, typename IO, class Op, int VE, bool EMU, class Cta>
__device__ __forceinline__ void nvls_tile_ws_impl(Cta& c, char* smem, int t) {
  using EL = Elem<T>;
  using A = typename EL::acc;
  using PackT = Pack<T, VE>;
  constexpr int S = kNvlsWsStages, UNR = kNvlsWsUnroll;
  long long start;
  const long long n = c.tile_valid(c.r, t, &start);
  if (n <= 0) return;  // same decision in every thread of the CTA
  const long long nvec = (n + VE - 1) / VE;
  constexpr int nth = FT_NVLS_THREADS, half = nth >> 1;  // the dedicated kernel always runs FT_NVLS_THREADS threads
  const bool loader = (int)threadIdx.x < half;
  const int ltid = loader ? (int)threadIdx.x : (int)threadIdx.x - half;
  constexpr long long step = (long long)half * UNR;
  constexpr long long kStride = (long long)half * 16;  // bytes between a thread's vectors of one batch
  const long long iters = (nvec + step - 1) / step;
  const long long full = nvec / step;                  // batches without a ragged end
  const long long base = c.woff + start * (long long)sizeof(T) + (long long)ltid * 16;
  const int nranks = c.a.plan.nranks;
  // barrier ids: 0 is __syncthreads; full[s] = 1 + s, empty[s] = 1 + S + s
  // Addresses: one base per batch plus compile-time offsets, see StagedCta::nvls_tile_stream for why.
  if (loader) {
    for (long long it = 0; it < iters; ++it) {
      const int s = (int)(it % S);
      if (it >= S) named_bar_sync(1 + S + s, nth);  // the storers have drained this stage
      PackT v[UNR];
      const long long o = base + it * step * 16;
      if (it < full) {
#pragma unroll
        for (int u = 0; u < UNR; ++u) v[u] = c.be.template mc_ld_reduce<T, Op, VE, EMU>(c.rc, o + u * kStride, nranks);
      } else {
#pragma unroll
        for (int u = 0; u < UNR; ++u)
          if (it * step + (long long)u * half + ltid < nvec) v[u] = c.be.template mc_ld_reduce<T, Op, VE, EMU>(c.rc, o + u * kStride, nranks);
      }
      char* slot = smem + (size_t)s * (size_t)half * UNR * 16;
#pragma unroll
      for (int u = 0; u < UNR; ++u) *reinterpret_cast<PackT*>(slot + ((size_t)u * half + ltid) * 16) = v[u];
      named_bar_arrive(1 + s, nth);
    }
  } else {
    for (long long it = 0; it < iters; ++it) {
      const int s = (int)(it % S);
      named_bar_sync(1 + s, nth);
      const char* slot = smem + (size_t)s * (size_t)half * UNR * 16;
      PackT v[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) v[u] = *reinterpret_cast<const PackT*>(slot + ((size_t)u * half + ltid) * 16);
      if (it + S < iters) named_bar_arrive(1 + S + s, nth);  // only if a later iteration waits for it
      const long long o = base + it * step * 16;
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if (it >= full && it * step + (long long)u * half + ltid >= nvec) continue;
        c.scale_pack(v[u]);
        c.be.template mc_st<T, VE, EMU>(c.rc, o + u * kStride, v[u], nranks);
      }
    }
  }
}

}  // namespace ft
