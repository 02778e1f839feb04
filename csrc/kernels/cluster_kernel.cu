// Kernel + launcher of the thread-block-cluster form of the P2P stages (cluster_tiles.cuh): clusters of two CTAs,
// 512 threads each, one cluster per tile. Float types only (fp32, bf16, fp16; SUM / MAX / MIN): this variant exists
// to be measured against the one-CTA-per-tile TMA kernel (profiles/clusters.md), not to cover the type matrix.
#include <cuda_runtime.h>

#include "../core/staged_core.h"
#include "../planner/cost_model.h"
#include "cluster_tiles.cuh"
#include "launch.h"

namespace ft {

template <typename T, typename IO, class Op>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FT_MAX_THREADS, 1) ft_cluster_kernel(const __grid_constant__ StagedArgs a) {
  extern __shared__ __align__(128) unsigned char ft_dyn_smem[];
  __shared__ int s_abort;
  if (threadIdx.x == 0) s_abort = 0;
  __syncthreads();
  const int cluster_id = (int)blockIdx.x >> 1;
  const int lr = cluster_id / a.ctas;
  const int cl = cluster_id - lr * a.ctas;
  ClusterRing R;
  R.ring = (char*)ft_dyn_smem;
  R.xch = R.ring + kClRingBytes;
  R.full = (uint64_t*)(R.xch + 2 * kClXchBytes);
  R.empty = R.full + kClStages;
  R.xbar = R.empty + kClStages;
  cluster_ring_init(R, ((int)blockDim.x >> 5) - 1, (int)blockDim.x - 32);
  CudaClusterBE be(&s_abort, R);
  if (a.fault_rank_plus1 && a.rc[lr].comm->rank + 1 == a.fault_rank_plus1) {  // fault injection, see kernels.cuh
    be.finish(*a.rc[lr].comm, a.ctas);
    return;
  }
  StagedCta<CudaClusterBE, T, IO, Op, VecElems<T>::value> k(be, a, a.rc[lr], cl);
  k.template run<true>();
}

namespace {
template <typename T, typename IO, class Op>
cudaError_t go(const LaunchCtx* c) {
  auto kern = ft_cluster_kernel<T, IO, Op>;
  static bool attr_set[64] = {};  // per instantiation, per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaError_t e = cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kClSmemBytes);
    if (e != cudaSuccess) return e;
    attr_set[dev] = true;
  }
  // (no cooperative launch: a grid of <= 148 CTAs with one CTA per SM is resident as a whole on an idle GPU, and the
  // watchdog turns the pathological case into an error instead of a hang)
  dim3 grid((unsigned)(c->args->nlocal * c->args->ctas * 2)), block((unsigned)c->threads);
  kern<<<grid, block, kClSmemBytes, c->stream>>>(*c->args);
  return cudaGetLastError();
}
template <typename T, typename IO>
bool by_op(int op, LaunchCtx* c) {
  switch (op) {
    case OP_SUM: c->err = go<T, IO, OpSum>(c); return true;
    case OP_MAX: c->err = go<T, IO, OpMax>(c); return true;
    case OP_MIN: c->err = go<T, IO, OpMin>(c); return true;
  }
  return false;
}
}  // namespace

// P2P stages with a cluster of two CTAs per tile. false = not available for this (dtype, io dtype, op, shape).
bool cuda_launch_cluster(int dtype, int io_dtype, int op, bool vec, LaunchCtx* ctx) {
  if (!vec || ctx->threads != FT_MAX_THREADS) return false;
  ctx->supported = true;
  ctx->err = cudaSuccess;
  bool ok = false;
  if (dtype == DT_FLOAT && io_dtype == DT_FLOAT) ok = by_op<float, float>(op, ctx);
  else if (dtype == DT_BFLOAT16 && io_dtype == DT_BFLOAT16) ok = by_op<bf16_t, bf16_t>(op, ctx);
  else if (dtype == DT_HALF && io_dtype == DT_HALF) ok = by_op<f16_t, f16_t>(op, ctx);
  else if (dtype == DT_BFLOAT16 && io_dtype == DT_FLOAT) ok = by_op<bf16_t, float>(op, ctx);
  if (!ok) ctx->supported = false;
  return ok;
}

}  // namespace ft
