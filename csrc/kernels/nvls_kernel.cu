// The large-message hot path: NVLS allreduce IN PLACE on registered (symmetric, multicast-bound) buffers of the
// wire type - what bench.py, FlexTreeDataParallel and ft_allreduce_host run on 4 and 8 GPUs.
//
// It has a kernel of its own (instead of being one more branch of ft_allreduce_kernel) for two reasons:
//  * registers: the in-switch loop wants 8-16 sixteen-byte multimem operations in flight per thread. Inside the
//    general kernel, at the 128-register cap of 512-thread CTAs and next to the private-buffer roles, that loop
//    spilled (round 1: 109 LDL/STL for fp32, 179 for bf16). Here: 256-thread CTAs (255 registers available), nothing
//    but the loop and its three flag hops; every inner-loop variant is a separate instantiation, so each gets its
//    own register allocation and its own SASS listing (profiles/sass/).
//  * the switch path does not care whether 4096 threads sit in 8 or in 16 CTAs (profiles/tune_n8.md).
// Protocol: StagedCta::run_nvls_registered (core/staged_core.h). Replaces the reference's reduce_sum<T> after
// MPI_Waitall plus the all-gather sends (allreduce_over_mpi/mpi_mod.hpp:246-452, :1003-1017, :1050-1060).
#include <cuda_runtime.h>

#include "../core/staged_core.h"
#include "../planner/cost_model.h"
#include "backend.cuh"
#include "launch.h"

namespace ft {

template <typename T, class Op, int VARIANT, bool EMU>
__global__ void __launch_bounds__(FT_NVLS_THREADS, 1) ft_nvls_kernel(const __grid_constant__ StagedArgs a) {
  extern __shared__ __align__(128) unsigned char ft_dyn_smem[];
  __shared__ int s_abort;
  if (threadIdx.x == 0) s_abort = 0;
  __syncthreads();
  const int lr = (int)blockIdx.x / a.ctas;
  const int cta = (int)blockIdx.x - lr * a.ctas;
  TmaRing ring;
  ring.ring = (char*)ft_dyn_smem;  // only the warp-specialised variant uses shared memory (as a plain ring)
  ring.full = ring.empty = nullptr;
  CudaBE be(&s_abort, ring);
  if (a.fault_rank_plus1 && a.rc[lr].comm->rank + 1 == a.fault_rank_plus1) {  // fault injection, see kernels.cuh
    be.finish(*a.rc[lr].comm, a.ctas);
    return;
  }
  StagedCta<CudaBE, T, T, Op, VecElems<T>::value> k(be, a, a.rc[lr], cta);
  k.template run_nvls_registered<VARIANT, FT_NVLS_THREADS, EMU>();
}

namespace {
template <typename T, class Op, int VARIANT, bool EMU>
cudaError_t go(const LaunchCtx* c) {
  auto kern = ft_nvls_kernel<T, Op, VARIANT, EMU>;
  dim3 grid((unsigned)(c->args->nlocal * c->args->ctas)), block((unsigned)c->threads);
  const size_t smem = VARIANT == 5 ? (size_t)kNvlsWsSmemBytes : 0;
  if (smem > (48u << 10)) {
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
template <typename T, class Op, bool EMU>
cudaError_t by_variant(const LaunchCtx* c) {
  switch (c->args->nvls_variant) {
    case 1: return go<T, Op, 1, EMU>(c);
    case 2: return go<T, Op, 2, EMU>(c);
    case 3: return go<T, Op, 3, EMU>(c);
    case 4: return go<T, Op, 4, EMU>(c);
    case 5: return go<T, Op, 5, EMU>(c);
    case 6: return go<T, Op, 6, EMU>(c);
    case 7: return go<T, Op, 7, EMU>(c);
    case 8: return go<T, Op, 8, EMU>(c);
    default: return go<T, Op, 0, EMU>(c);
  }
}
template <typename T>
cudaError_t by_emu(const LaunchCtx* c) {
  return c->args->rc[0].mc ? by_variant<T, OpSum, false>(c) : by_variant<T, OpSum, true>(c);
}
}  // namespace

// Registered buffers, user type == wire type, fp32 / bf16 / fp16 SUM, whole 16-byte vectors. false = not my case.
bool cuda_launch_nvls_registered(int dtype, int op, LaunchCtx* ctx) {
  if (op != OP_SUM || ctx->threads != FT_NVLS_THREADS) return false;  // the streaming loops are compiled for that CTA size
  ctx->supported = true;
  switch (dtype) {
    case DT_FLOAT: ctx->err = by_emu<float>(ctx); return true;
    case DT_BFLOAT16: ctx->err = by_emu<bf16_t>(ctx); return true;
    case DT_HALF: ctx->err = by_emu<f16_t>(ctx); return true;
  }
  ctx->supported = false;
  return false;
}

}  // namespace ft
