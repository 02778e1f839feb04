// Host-side launch context shared by the runtime and the per-group kernel translation units.
#pragma once
#include <cuda_runtime.h>

#include "../core/comm_dev.h"

#define FT_MAX_THREADS 512

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
// true when the switch can reduce this (dtype, op) (multimem.ld_reduce): fp32/bf16/fp16 SUM
bool nvls_supported(int dtype, int op);

}  // namespace ft
