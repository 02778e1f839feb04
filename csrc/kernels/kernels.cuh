// The sm_100a kernels: one __global__ template, instantiated per (wire type, user type, op, vector
// width, algorithm). Included by kernels_g*.cu, each compiling one slice of the type matrix.
//
// Kernel inventory: K1/K2/K3/K5/K6 -> ALGO_STAGED (fused pull + k-way reduce + flags +
// scale/cast epilogue), K7 -> ALGO_RING, K10 -> ALGO_ONESHOT, K11 -> ALGO_NVLS, K13 -> epilogue of all.
#pragma once
#include <cuda_runtime.h>

#include "../core/dispatch.h"
#include "../core/staged_core.h"
#include "../planner/cost_model.h"
#include "backend.cuh"
#include "launch.h"

namespace ft {

template <typename T, typename IO, class Op, int VE, int ALGO>
__global__ void __launch_bounds__(FT_MAX_THREADS, 1) ft_allreduce_kernel(const __grid_constant__ StagedArgs a) {
  extern __shared__ __align__(128) unsigned char ft_dyn_smem[];
  __shared__ int s_abort;
  if (threadIdx.x == 0) s_abort = 0;
  __syncthreads();
  const int lr = (int)blockIdx.x / a.ctas;
  const int cta = (int)blockIdx.x - lr * a.ctas;
  TmaRing ring;
  ring.ring = (char*)ft_dyn_smem;
  ring.full = (uint64_t*)(ft_dyn_smem + kTmaRingBytes);
  ring.empty = ring.full + kTmaStages;
  if constexpr (ALGO == ALGO_STAGED_TMA) tma_ring_init(ring, ((int)blockDim.x >> 5) - 1);
  CudaBE be(&s_abort, ring);
  if (a.fault_rank_plus1 && a.rc[lr].comm->rank + 1 == a.fault_rank_plus1) {
    // fault injection (ft_comm_inject_fault): this rank plays dead - it touches no buffer and signals nobody, so
    // its peers run into the watchdog. The epoch still advances, which keeps the communicator usable afterwards.
    be.finish(*a.rc[lr].comm, a.ctas);
    return;
  }
  StagedCta<CudaBE, T, IO, Op, VE> k(be, a, a.rc[lr], cta);
  if constexpr (ALGO == ALGO_STAGED) k.template run<false>();
  else if constexpr (ALGO == ALGO_STAGED_TMA) k.template run<true>();
  else if constexpr (ALGO == ALGO_NVLS) k.run_nvls();
  else if constexpr (ALGO == ALGO_ONESHOT) k.run_oneshot();
  else if constexpr (ALGO == ALGO_TWOSHOT) k.run_twoshot();
  else k.run_ring();
}

template <typename T, typename IO, class Op, int VE>
struct CudaLaunch {
  template <int ALGO>
  static cudaError_t go(const LaunchCtx* c) {
    auto kern = ft_allreduce_kernel<T, IO, Op, VE, ALGO>;
    dim3 grid((unsigned)(c->args->nlocal * c->args->ctas)), block((unsigned)c->threads);
    size_t smem = 0;
    if (ALGO == ALGO_STAGED_TMA) {
      smem = kTmaSmemBytes;
      static bool attr_set[64] = {};  // per instantiation, per device
      int dev = 0;
      cudaGetDevice(&dev);
      if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        attr_set[dev] = true;
      }
    }
    if (c->cooperative) {
      void* params[] = {(void*)c->args};
      return cudaLaunchCooperativeKernel((const void*)kern, grid, block, params, smem, c->stream);
    }
    kern<<<grid, block, smem, c->stream>>>(*c->args);
    return cudaGetLastError();
  }
  static void run(LaunchCtx* c) {
    c->supported = true;
    if constexpr (SameType<T, fp8_t>::value) {  // block-scaled wire: register-staged P2P stages only
      if (c->algo == ALGO_STAGED) c->err = go<ALGO_STAGED>(c);
      else c->supported = false;
    } else {
      switch (c->algo) {
        case ALGO_STAGED: c->err = go<ALGO_STAGED>(c); break;
        case ALGO_STAGED_TMA:
          if constexpr (VE > 1) c->err = go<ALGO_STAGED_TMA>(c);
          else c->supported = false;
          break;
        case ALGO_ONESHOT: c->err = go<ALGO_ONESHOT>(c); break;
        case ALGO_TWOSHOT: c->err = go<ALGO_TWOSHOT>(c); break;
        case ALGO_RING: c->err = go<ALGO_RING>(c); break;
        case ALGO_NVLS:
          if constexpr (NvlsOk<T, Op>::value && VE > 1) c->err = go<ALGO_NVLS>(c);
          else c->supported = false;
          break;
        default: c->supported = false;
      }
    }
  }
};

}  // namespace ft

#define FT_DEFINE_GROUP_LAUNCH(GROUP)                                                             \
  namespace ft {                                                                                  \
  bool FT_CAT(cuda_launch_g, GROUP)(int dtype, int io_dtype, int op, bool vec, LaunchCtx* ctx) {  \
    return Dispatcher<CudaLaunch, GROUP, LaunchCtx*>::run(dtype, io_dtype, op, vec, ctx);         \
  }                                                                                               \
  }
#define FT_CAT2(a, b) a##b
#define FT_CAT(a, b) FT_CAT2(a, b)
