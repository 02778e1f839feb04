// Host-side launch context shared by the runtime and the per-group kernel translation units.
#pragma once
#include <cuda_runtime.h>

#include "../core/comm_dev.h"

#define FT_MAX_THREADS 512
// The NVLS kernel runs 256-thread CTAs: at 512 threads the register cap is 128 and the 8-deep multimem pipeline
// spilled (round 1: 109-179 LDL/STL); 256 threads may use 255 registers, and the switch path does not care whether the
// same number of threads sits in 8 or in 16 CTAs (profiles/tune_n8.md: 594 us either way at 256 MiB on 8 GPUs).
#define FT_NVLS_THREADS 256

namespace ft {

struct LaunchCtx {
  const StagedArgs* args;
  int algo;
  int threads;
  bool cooperative;   // single-GPU emulation: all logical ranks' CTAs must be co-resident
  cudaStream_t stream;
  cudaError_t err;    // out
  bool supported;     // out: false when this (type, op, algo) has no kernel
};

// Returns false when (dtype, io_dtype, op) is not defined at all.
bool cuda_launch(int dtype, int io_dtype, int op, bool vec, LaunchCtx* ctx);
// NVLS in place on registered buffers of the wire type (kernels/nvls_kernel.cu); false = not applicable
bool cuda_launch_nvls_registered(int dtype, int op, LaunchCtx* ctx);
// P2P stages with a thread-block cluster of two CTAs per tile (kernels/cluster_kernel.cu); false = not applicable
bool cuda_launch_cluster(int dtype, int io_dtype, int op, bool vec, LaunchCtx* ctx);
// true when the switch can reduce this (dtype, op) (multimem.ld_reduce): fp32/bf16/fp16 SUM
bool nvls_supported(int dtype, int op);

}  // namespace ft
