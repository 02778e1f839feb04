// ft_allreduce_emulated_noalloc: the single-GPU (emulated ranks) one-shot allreduce for callers that must
// not allocate. See the declaration in include/flextree.h for the contract.
//
// The regular entry point (ft_allreduce_multi on an ft_comm_create_emulated communicator) keeps its
// scratch - LL receive buffers, per-rank CommDev blocks, the epoch / completion / error words - in
// memory it obtains with cudaMalloc when the communicator is created. Here the same scratch lives in
// one file-scope __device__ array that is part of the loaded module, so nothing is allocated at any
// point: not at set-up, not per call.
//
// Note on the array's name: `g_noalloc_scratch` below is a *device* symbol. In host code that name is
// only a stand-in the runtime uses to look the symbol up; it is not a device address and must never be
// handed to a kernel. The device address is obtained once per device with cudaGetSymbolAddress (with
// that device current) and that pointer is what the kernel receives.
#include <cuda_runtime.h>
#include <string.h>

#include <mutex>
#include <string>

#include "../../include/flextree.h"
#include "../core/comm_dev.h"
#include "../core/types.h"
#include "../kernels/launch.h"
#include "../planner/cost_model.h"

namespace ft {
void set_last_error(const std::string& s);  // planner/c_api.cc
}
using namespace ft;

namespace {

constexpr int kMaxRanks = FT_NOALLOC_MAX_RANKS;
constexpr size_t kMaxBytes = FT_NOALLOC_MAX_BYTES;
// LL receive buffer of one rank, one parity: every source gets room for kMaxBytes of payload, and an LL
// line carries 8 payload bytes in 16.
constexpr size_t kLlBytes = (size_t)kMaxRanks * (kMaxBytes / 8) * 16;
constexpr size_t kCommOff = 0;                                            // CommDev[kMaxRanks]
constexpr size_t kEpochOff = (kCommOff + sizeof(CommDev) * kMaxRanks + 255) & ~(size_t)255;
constexpr size_t kDoneOff = kEpochOff + 256;                              // unsigned[kMaxRanks]
constexpr size_t kErrOff = kDoneOff + 256;                                // unsigned[kMaxRanks]
constexpr size_t kStickyOff = kErrOff + 256;                             // unsigned: first error of any call, kept until read
constexpr size_t kLlOff = kStickyOff + 256;
constexpr size_t kScratchBytes = kLlOff + (size_t)kMaxRanks * 2 * kLlBytes;  // ~4.02 MiB

// Zero-filled when the module is loaded on a device; lives until the process exits.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// Runs in stream order right behind the collective kernel (one warp, one thread works).
//  * Epochs: the collective advances only the epoch words of the logical ranks that took part, and the LL line
//    tag / parity half of a call are derived from a rank's own epoch. A call with fewer ranks would therefore leave
//    the idle ranks one epoch behind and the next larger call would pair lines with different tags. All
//    kMaxRanks words are brought to the same (largest) value here, whatever the call's world size was.
//  * Errors: the per-rank error words are cleared at the start of every call; the first nonzero one is copied
//    to a sticky word that only ft_noalloc_error(clear=1) resets.
__global__ void noalloc_epilogue(unsigned char* base) {
  if (threadIdx.x != 0) return;
  uint64_t* epoch = (uint64_t*)(base + kEpochOff);
  unsigned* err = (unsigned*)(base + kErrOff);
  unsigned* sticky = (unsigned*)(base + kStickyOff);
  uint64_t top = 0;
  for (int l = 0; l < kMaxRanks; ++l) top = epoch[l] > top ? epoch[l] : top;
  for (int l = 0; l < kMaxRanks; ++l) epoch[l] = top;
  for (int l = 0; l < kMaxRanks; ++l)
    if (err[l] && !*sticky) *sticky = err[l];
}

struct DeviceSetup {
  bool ready = false;
  unsigned char* base = nullptr;  // device address of g_noalloc_scratch on this device
};
constexpr int kMaxDevices = 64;
DeviceSetup g_setup[kMaxDevices];
std::mutex g_mu;

bool bad(ft_result_t* rc, ft_result_t code, const char* msg) {
  set_last_error(msg);
  *rc = code;
  return false;
}

// First call on a device: resolve the symbol and write the per-rank CommDev blocks (plain copies into the
// static array - nothing is allocated). The epoch words are left alone on purpose.
bool setup_device(int device, DeviceSetup* st, ft_result_t* rc) {
  void* ptr = nullptr;
  cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
  if (e != cudaSuccess || !ptr) {
    (void)cudaGetLastError();
    return bad(rc, FT_ERR_CUDA, "ft_allreduce_emulated_noalloc: cudaGetSymbolAddress failed for the scratch array");
  }
  unsigned char* base = (unsigned char*)ptr;
  static CommDev host[kMaxRanks];  // static: no heap, and the mutex serialises its use
  for (int l = 0; l < kMaxRanks; ++l) {
    CommDev& d = host[l];
    memset(&d, 0, sizeof d);
    d.rank = l;
    d.nranks = kMaxRanks;
    for (int p = 0; p < kMaxRanks; ++p) d.ll[p] = (char*)(base + kLlOff + (size_t)p * 2 * kLlBytes);
    d.epoch = (uint64_t*)(base + kEpochOff) + l;
    d.done = (unsigned*)(base + kDoneOff) + l;
    d.err = (unsigned*)(base + kErrOff) + l;
    d.ll_bytes = kLlBytes;
  }
  e = cudaMemcpy(base + kCommOff, host, sizeof(CommDev) * kMaxRanks, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return bad(rc, FT_ERR_CUDA, "ft_allreduce_emulated_noalloc: could not initialise the scratch array");
  }
  st->base = base;
  st->ready = true;
  (void)device;
  return true;
}

}  // namespace

extern "C" ft_result_t ft_allreduce_emulated_noalloc(const void* const* sendbufs, void* const* recvbufs, int nranks,
                                                     size_t count, ft_dtype_t dtype, ft_op_t op, void* stream) {
  ft_result_t rc = FT_SUCCESS;
  if (!recvbufs || nranks < 1) {
    bad(&rc, FT_ERR_INVALID, "ft_allreduce_emulated_noalloc: recvbufs is null or nranks < 1");
    return rc;
  }
  if (nranks > kMaxRanks) {
    bad(&rc, FT_ERR_INVALID, "ft_allreduce_emulated_noalloc: at most 8 ranks (FT_NOALLOC_MAX_RANKS); use a communicator for more");
    return rc;
  }
  const int es = dtype_size((int)dtype);
  if (es == 0 || !op_supported((int)op, (int)dtype)) {
    bad(&rc, FT_ERR_UNSUPPORTED, "ft_allreduce_emulated_noalloc: unsupported dtype/op");
    return rc;
  }
  if (count == 0) return FT_SUCCESS;
  if (count > kMaxBytes / (size_t)es) {
    bad(&rc, FT_ERR_INVALID,
        "ft_allreduce_emulated_noalloc: message larger than 16384 bytes per rank (FT_NOALLOC_MAX_BYTES); the static "
        "scratch cannot hold it - use ft_allreduce_multi on a communicator instead");
    return rc;
  }
  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= kMaxDevices) {
    (void)cudaGetLastError();
    bad(&rc, FT_ERR_CUDA, "ft_allreduce_emulated_noalloc: no current CUDA device");
    return rc;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceSetup& st = g_setup[device];
  if (!st.ready && !setup_device(device, &st, &rc)) return rc;

  cudaStream_t s = (cudaStream_t)stream;
  // The array keeps whatever the previous call left behind. Two things must start from zero: the CTA
  // completion counters (a call that was cut short leaves them mid-count) and the sticky error words.
  // The epochs must NOT be cleared: LL lines are validated by an epoch-derived tag, which is what makes
  // stale lines from earlier calls harmless.
  static_assert(kErrOff == kDoneOff + 256 && kStickyOff == kErrOff + 256, "done and err words are cleared together");
  if (cudaMemsetAsync(st.base + kDoneOff, 0, 512, s) != cudaSuccess) {
    (void)cudaGetLastError();
    bad(&rc, FT_ERR_CUDA, "ft_allreduce_emulated_noalloc: cudaMemsetAsync failed");
    return rc;
  }

  StagedArgs a;
  memset(&a, 0, sizeof a);
  int flat = nranks;
  make_plan(&flat, 1, &a.plan);
  bool vec = true;
  for (int l = 0; l < nranks; ++l) {
    if (!recvbufs[l]) {
      bad(&rc, FT_ERR_INVALID, "ft_allreduce_emulated_noalloc: null recvbuf");
      return rc;
    }
    const void* sp = (sendbufs && sendbufs[l] && sendbufs[l] != FT_IN_PLACE) ? sendbufs[l] : recvbufs[l];
    vec = vec && ((uintptr_t)sp % 16 == 0) && ((uintptr_t)recvbufs[l] % 16 == 0);
    a.rc[l].comm = (const CommDev*)(st.base + kCommOff) + l;
    a.rc[l].send = sp;
    a.rc[l].recv = recvbufs[l];
    a.rc[l].win = nullptr;
    a.rc[l].mc = nullptr;
  }
  a.lay = make_layout((long long)count, nranks, es, 4096, vec);
  const int threads = 512;
  const long long lines = ((long long)count * es + 7) / 8;
  int ctas = (int)((lines + threads - 1) / threads);
  if (ctas < 1) ctas = 1;
  a.nlocal = nranks;
  a.ctas = ctas;
  a.mode = MODE_STAGING;
  a.apply_scale = 0;
  a.scale = 1.0;
  a.timeout_ns = 20000000000ull;

  LaunchCtx ctx;
  ctx.args = &a;
  ctx.algo = ALGO_ONESHOT;
  ctx.threads = threads;
  ctx.cooperative = nranks > 1;  // all logical ranks' CTAs wait on one another: they must be co-resident
  ctx.stream = s;
  if (!cuda_launch((int)dtype, (int)dtype, (int)op, vec, &ctx) || !ctx.supported) {
    bad(&rc, FT_ERR_UNSUPPORTED, "ft_allreduce_emulated_noalloc: no kernel for this dtype/op");
    return rc;
  }
  if (ctx.err != cudaSuccess) {
    set_last_error(std::string("ft_allreduce_emulated_noalloc: kernel launch: ") + cudaGetErrorString(ctx.err));
    (void)cudaGetLastError();
    return FT_ERR_CUDA;
  }
  noalloc_epilogue<<<1, 32, 0, s>>>(st.base);
  if (cudaGetLastError() != cudaSuccess) {
    bad(&rc, FT_ERR_CUDA, "ft_allreduce_emulated_noalloc: epilogue launch failed");
    return rc;
  }
  return FT_SUCCESS;
}

extern "C" unsigned ft_noalloc_error(int clear) {
  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess || device < 0 || device >= kMaxDevices) {
    (void)cudaGetLastError();
    return 0xffffffffu;
  }
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceSetup& st = g_setup[device];
  if (!st.ready) return 0;  // never used on this device: nothing can have failed
  unsigned v = 0;
  if (cudaMemcpy(&v, st.base + kStickyOff, sizeof v, cudaMemcpyDeviceToHost) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0xffffffffu;
  }
  if (clear && v) {
    if (cudaMemset(st.base + kStickyOff, 0, sizeof v) != cudaSuccess) (void)cudaGetLastError();
  }
  return v;
}
