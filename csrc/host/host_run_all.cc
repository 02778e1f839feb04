#include "../core/comm_dev.h"
#include "../core/dispatch.h"
#include "cpu_backend.h"
namespace ft {
bool host_run_g1(int, int, int, bool, const StagedArgs*, int);
#ifdef FT_ONLY_F32  // sanitizer builds compile one slice of the type matrix (+ the 16-bit / fp8 wire slice on request)
#ifdef FT_WITH_F16
bool host_run_g2(int, int, int, bool, const StagedArgs*, int);
#else
static bool host_run_g2(int, int, int, bool, const StagedArgs*, int) { return false; }
#endif
static bool host_run_g4(int, int, int, bool, const StagedArgs*, int) { return false; }
static bool host_run_g8(int, int, int, bool, const StagedArgs*, int) { return false; }
static bool host_run_g16(int, int, int, bool, const StagedArgs*, int) { return false; }
#else
bool host_run_g2(int, int, int, bool, const StagedArgs*, int);
bool host_run_g4(int, int, int, bool, const StagedArgs*, int);
bool host_run_g8(int, int, int, bool, const StagedArgs*, int);
bool host_run_g16(int, int, int, bool, const StagedArgs*, int);
#endif
bool host_run_group(int group, int dtype, int io_dtype, int op, bool vec, const StagedArgs* a, int algo) {
  switch (group) {
    case TG_F32: return host_run_g1(dtype, io_dtype, op, vec, a, algo);
    case TG_F16: return host_run_g2(dtype, io_dtype, op, vec, a, algo);
    case TG_64: return host_run_g4(dtype, io_dtype, op, vec, a, algo);
    case TG_I32: return host_run_g8(dtype, io_dtype, op, vec, a, algo);
    case TG_SMALL: return host_run_g16(dtype, io_dtype, op, vec, a, algo);
  }
  return false;
}
}  // namespace ft

// Tests only: probability (per mille) of a random 0-200 us sleep at every flag publication / acquisition / CTA sync
// of the CPU back end. 0 = off (default).
extern "C" void ft_host_set_jitter(int permille) { ft::host_jitter_permille() = permille < 0 ? 0 : permille; }
