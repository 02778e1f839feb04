#include "launch.h"

#include "../core/dispatch.h"

namespace ft {
bool cuda_launch_g1(int, int, int, bool, LaunchCtx*);
bool cuda_launch_g2(int, int, int, bool, LaunchCtx*);
bool cuda_launch_g4(int, int, int, bool, LaunchCtx*);
bool cuda_launch_g8(int, int, int, bool, LaunchCtx*);
bool cuda_launch_g16(int, int, int, bool, LaunchCtx*);

bool cuda_launch(int dtype, int io_dtype, int op, bool vec, LaunchCtx* ctx) {
  ctx->err = cudaSuccess;
  ctx->supported = false;
  switch (dtype_group(dtype)) {
    case TG_F32: return cuda_launch_g1(dtype, io_dtype, op, vec, ctx);
    case TG_F16: return cuda_launch_g2(dtype, io_dtype, op, vec, ctx);
    case TG_64: return cuda_launch_g4(dtype, io_dtype, op, vec, ctx);
    case TG_I32: return cuda_launch_g8(dtype, io_dtype, op, vec, ctx);
    case TG_SMALL: return cuda_launch_g16(dtype, io_dtype, op, vec, ctx);
  }
  return false;
}

bool nvls_supported(int dtype, int op) {
  return op == OP_SUM && (dtype == DT_FLOAT || dtype == DT_BFLOAT16 || dtype == DT_HALF);
}
}  // namespace ft
