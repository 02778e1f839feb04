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
  __shared__ int s_abort;
  if (threadIdx.x == 0) s_abort = 0;
  __syncthreads();
  const int lr = (int)blockIdx.x / a.ctas;
  const int cta = (int)blockIdx.x - lr * a.ctas;
  CudaBE be(&s_abort);
  StagedCta<CudaBE, T, IO, Op, VE> k(be, a, a.rc[lr], cta);
  if constexpr (ALGO == ALGO_STAGED) k.run();
  else if constexpr (ALGO == ALGO_NVLS) k.run_nvls();
  else if constexpr (ALGO == ALGO_ONESHOT) k.run_oneshot();
  else k.run_ring();
}

template <typename T, typename IO, class Op, int VE>
struct CudaLaunch {
  template <int ALGO>
  static cudaError_t go(const LaunchCtx* c) {
    auto kern = ft_allreduce_kernel<T, IO, Op, VE, ALGO>;
    dim3 grid((unsigned)(c->args->nlocal * c->args->ctas)), block((unsigned)c->threads);
    if (c->cooperative) {
      void* params[] = {(void*)c->args};
      return cudaLaunchCooperativeKernel((const void*)kern, grid, block, params, 0, c->stream);
    }
    kern<<<grid, block, 0, c->stream>>>(*c->args);
    return cudaGetLastError();
  }
  static void run(LaunchCtx* c) {
    c->supported = true;
    switch (c->algo) {
      case ALGO_STAGED: c->err = go<ALGO_STAGED>(c); break;
      case ALGO_ONESHOT: c->err = go<ALGO_ONESHOT>(c); break;
      case ALGO_RING: c->err = go<ALGO_RING>(c); break;
      case ALGO_NVLS:
        if constexpr (NvlsOk<T, Op>::value && VE > 1) c->err = go<ALGO_NVLS>(c);
        else c->supported = false;
        break;
      default: c->supported = false;
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
