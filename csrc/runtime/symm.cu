// Symmetric memory providers (see symm.h).
#include "symm.h"

#include "fd_channel.h"

#include <cuda.h>
#include <errno.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/time.h>
#include <sys/un.h>
#include <unistd.h>

#include <mutex>

namespace ft {

// ---- driver API through the runtime (no link-time libcuda dependency) ---------------------------
namespace drv {
#define FT_DRV_LIST(X)                 \
  X(cuMemCreate)                       \
  X(cuMemRelease)                      \
  X(cuMemExportToShareableHandle)      \
  X(cuMemImportFromShareableHandle)    \
  X(cuMemAddressReserve)               \
  X(cuMemAddressFree)                  \
  X(cuMemMap)                          \
  X(cuMemUnmap)                        \
  X(cuMemSetAccess)                    \
  X(cuMemGetAllocationGranularity)     \
  X(cuMulticastCreate)                 \
  X(cuMulticastAddDevice)              \
  X(cuMulticastBindMem)                \
  X(cuMulticastUnbind)                 \
  X(cuMulticastGetGranularity)         \
  X(cuDeviceGet)                       \
  X(cuDeviceGetAttribute)              \
  X(cuGetErrorString)

#define X(name) static decltype(&::name) p_##name = nullptr;
FT_DRV_LIST(X)
#undef X

static bool g_ok = false;
static std::once_flag g_once;

static bool load() {
  std::call_once(g_once, []() {
    bool ok = true;
#define X(name)                                                                                         \
  {                                                                                                     \
    void* f = nullptr;                                                                                  \
    cudaDriverEntryPointQueryResult q;                                                                  \
    if (cudaGetDriverEntryPoint(#name, &f, cudaEnableDefault, &q) != cudaSuccess || !f) ok = false;     \
    p_##name = (decltype(p_##name))f;                                                                   \
  }
    FT_DRV_LIST(X)
#undef X
    (void)cudaGetLastError();
    g_ok = ok;
  });
  return g_ok;
}

static std::string errstr(CUresult r) {
  const char* s = nullptr;
  if (p_cuGetErrorString) p_cuGetErrorString(r, &s);
  return s ? s : ("CUresult " + std::to_string((int)r));
}
}  // namespace drv

#define FT_CU(call, what)                                                     \
  do {                                                                        \
    CUresult _r = (call);                                                     \
    if (_r != CUDA_SUCCESS) {                                                 \
      if (err) *err = std::string(what) + ": " + drv::errstr(_r);             \
      return false;                                                           \
    }                                                                         \
  } while (0)
#define FT_RT(call, what)                                                     \
  do {                                                                        \
    cudaError_t _r = (call);                                                  \
    if (_r != cudaSuccess) {                                                  \
      if (err) *err = std::string(what) + ": " + cudaGetErrorString(_r);      \
      (void)cudaGetLastError();                                               \
      return false;                                                           \
    }                                                                         \
  } while (0)

bool device_supports_multicast(int device) {
  if (!drv::load()) return false;
  CUdevice dev;
  if (drv::p_cuDeviceGet(&dev, device) != CUDA_SUCCESS) return false;
  int v = 0;
  if (drv::p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev) != CUDA_SUCCESS) return false;
  return v != 0;
}

static bool device_supports_fd_export(int device) {
  if (!drv::load()) return false;
  CUdevice dev;
  if (drv::p_cuDeviceGet(&dev, device) != CUDA_SUCCESS) return false;
  int v = 0;
  if (drv::p_cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED, dev) !=
      CUDA_SUCCESS)
    return false;
  return v != 0;
}

// All ranks must agree on success: allgather a status byte.
static bool all_ok(const Bootstrap& bs, bool mine) {
  char s = mine ? 1 : 0;
  std::vector<char> all((size_t)bs.nranks);
  if (!bs.allgather(&s, all.data(), 1)) return false;
  for (char c : all)
    if (!c) return false;
  return true;
}

static bool upload_table(Segment* seg, std::string* err) {
  FT_RT(cudaMalloc(&seg->d_table, sizeof(char*) * (size_t)seg->nranks), "cudaMalloc(table)");
  FT_RT(cudaMemcpy(seg->d_table, seg->base.data(), sizeof(char*) * (size_t)seg->nranks, cudaMemcpyHostToDevice),
        "cudaMemcpy(table)");
  return true;
}

// ---- VMM provider ------------------------------------------------------------------------------------
// Undo a partially built VMM segment: unmap what was mapped, release every handle (own and imported) and the
// multicast mapping, so that a collective failure leaves nothing behind before the cudaIpc fallback.
static void vmm_release(Segment* seg) {
  if (seg->d_table) { cudaFree(seg->d_table); seg->d_table = nullptr; }
  if (seg->mc_base) {
    drv::p_cuMemUnmap((CUdeviceptr)seg->mc_base, seg->mapped_bytes);
    drv::p_cuMemAddressFree((CUdeviceptr)seg->mc_base, seg->mapped_bytes);
    CUdevice dev;
    if (drv::p_cuDeviceGet(&dev, seg->device) == CUDA_SUCCESS)
      drv::p_cuMulticastUnbind((CUmemGenericAllocationHandle)seg->mc_handle, dev, 0, seg->mapped_bytes);
    drv::p_cuMemRelease((CUmemGenericAllocationHandle)seg->mc_handle);
    seg->mc_base = nullptr;
    seg->mc_handle = 0;
  }
  for (size_t p = 0; p < seg->base.size(); ++p) {
    if (seg->base[p]) {
      drv::p_cuMemUnmap((CUdeviceptr)seg->base[p], seg->mapped_bytes);
      drv::p_cuMemAddressFree((CUdeviceptr)seg->base[p], seg->mapped_bytes);
      seg->base[p] = nullptr;
    }
    if (p < seg->vmm_handles.size() && seg->vmm_handles[p]) {
      drv::p_cuMemRelease((CUmemGenericAllocationHandle)seg->vmm_handles[p]);
      seg->vmm_handles[p] = 0;
    }
  }
  (void)cudaGetLastError();
  seg->kind = SEG_NONE;
}

static bool vmm_create(const Bootstrap& bs, int device, size_t bytes, const SymmOptions& opt, Segment* seg,
                       std::string* err) {
  if (!drv::load()) { if (err) *err = "driver entry points unavailable"; return false; }
  if (!device_supports_fd_export(device)) { if (err) *err = "POSIX fd export unsupported"; return false; }
  CUdevice dev;
  FT_CU(drv::p_cuDeviceGet(&dev, device), "cuDeviceGet");

  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  FT_CU(drv::p_cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "granularity");

  const bool want_mc = opt.try_multicast && device_supports_multicast(device) && bs.nranks > 1;
  size_t mc_gran = 0;
  CUmulticastObjectProp mcp{};
  if (want_mc) {
    mcp.numDevices = (unsigned)bs.nranks;
    mcp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mcp.size = bytes;
    if (drv::p_cuMulticastGetGranularity(&mc_gran, &mcp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS)
      mc_gran = 0;
  }
  size_t align = gran > mc_gran ? gran : mc_gran;
  if (align == 0) align = 2u << 20;
  size_t sz = (bytes + align - 1) / align * align;
  seg->mapped_bytes = sz;
  seg->bytes = bytes;
  seg->nranks = bs.nranks;
  seg->self = bs.rank;
  seg->device = device;
  seg->kind = SEG_VMM;
  seg->base.assign((size_t)bs.nranks, nullptr);
  seg->vmm_handles.assign((size_t)bs.nranks, 0);

  // 1. allocate + export mine
  CUmemGenericAllocationHandle mine = 0;
  bool ok = drv::p_cuMemCreate(&mine, sz, &prop, 0) == CUDA_SUCCESS;
  int my_fd = -1;
  if (ok) ok = drv::p_cuMemExportToShareableHandle(&my_fd, mine, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
  if (!all_ok(bs, ok)) {
    if (mine) drv::p_cuMemRelease(mine);
    if (my_fd >= 0) ::close(my_fd);
    if (err) *err = "cuMemCreate/export failed on some rank";
    return false;
  }
  seg->vmm_handles[(size_t)bs.rank] = mine;

  // 2. open fd channel; token chosen by rank 0
  unsigned long long tok = 0;
  if (bs.rank == 0) {
    timeval tv;
    gettimeofday(&tv, nullptr);
    tok = ((unsigned long long)getpid() << 32) ^ (unsigned long long)tv.tv_usec ^ ((unsigned long long)tv.tv_sec << 20) ^
          (unsigned long long)(uintptr_t)seg;
  }
  std::vector<unsigned long long> toks((size_t)bs.nranks);
  if (!bs.allgather(&tok, toks.data(), sizeof tok)) { if (err) *err = "bootstrap allgather failed"; return false; }
  FdChannel ch;
  std::string cerr_;
  ok = ch.open(toks[0], bs.rank, &cerr_);
  if (!all_ok(bs, ok)) {  // also the "everybody is bound" barrier
    ch.close_();
    drv::p_cuMemRelease(mine);
    ::close(my_fd);
    if (err) *err = "fd channel: " + cerr_;
    return false;
  }

  // 3. exchange allocation fds (all-to-all), import
  ok = true;
  for (int p = 0; p < bs.nranks && ok; ++p)
    if (p != bs.rank) ok = ch.send_fd(p, my_fd, /*tag=*/0, &cerr_);
  for (int i = 0; i < bs.nranks - 1 && ok; ++i) {
    int src, tag, fd;
    ok = ch.recv_fd(&src, &tag, &fd, &cerr_);
    if (!ok) break;
    if (tag != 0 || src < 0 || src >= bs.nranks) { ok = false; cerr_ = "unexpected fd message"; ::close(fd); break; }
    CUmemGenericAllocationHandle h = 0;
    if (drv::p_cuMemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) {
      ok = false;
      cerr_ = "cuMemImportFromShareableHandle failed";
    }
    ::close(fd);
    seg->vmm_handles[(size_t)src] = h;
  }
  ::close(my_fd);
  if (!all_ok(bs, ok)) {
    ch.close_();
    vmm_release(seg);
    if (err) *err = "fd exchange: " + cerr_;
    return false;
  }

  // 4. map everything. A failure on ONE rank must fail the call on EVERY rank (the caller falls back to cudaIpc,
  // whose bootstrap exchanges have a different size and order): accumulate a local flag, then agree.
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  std::string lerr;
  bool ok4 = true;
  for (int p = 0; p < bs.nranks && ok4; ++p) {
    CUdeviceptr va = 0;
    CUresult r = drv::p_cuMemAddressReserve(&va, sz, align, 0, 0);
    if (r != CUDA_SUCCESS) { ok4 = false; lerr = "cuMemAddressReserve: " + drv::errstr(r); break; }
    seg->base[(size_t)p] = (char*)va;
    r = drv::p_cuMemMap(va, sz, 0, seg->vmm_handles[(size_t)p], 0);
    if (r != CUDA_SUCCESS) {
      ok4 = false;
      lerr = "cuMemMap: " + drv::errstr(r);
      drv::p_cuMemAddressFree(va, sz);
      seg->base[(size_t)p] = nullptr;
      break;
    }
    r = drv::p_cuMemSetAccess(va, sz, &acc, 1);
    if (r != CUDA_SUCCESS) { ok4 = false; lerr = "cuMemSetAccess: " + drv::errstr(r); }
  }
  if (ok4 && (cudaMemset(seg->base[(size_t)bs.rank], 0, sz) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess)) {
    ok4 = false;
    lerr = "cudaMemset(segment) failed";
    (void)cudaGetLastError();
  }
  if (!all_ok(bs, ok4)) {
    ch.close_();
    vmm_release(seg);
    if (err) *err = "mapping failed on some rank" + (lerr.empty() ? std::string() : ": " + lerr);
    return false;
  }

  // 5. multicast (optional)
  if (want_mc) {
    bool mok = true;
    CUmemGenericAllocationHandle mch = 0;
    mcp.size = sz;
    if (bs.rank == 0) {
      int mfd = -1;
      mok = drv::p_cuMulticastCreate(&mch, &mcp) == CUDA_SUCCESS &&
            drv::p_cuMemExportToShareableHandle(&mfd, mch, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) == CUDA_SUCCESS;
      // always send something so that receivers do not block: tag 1 = valid, tag 2 = failed
      for (int p = 1; p < bs.nranks; ++p) {
        int fd_to_send = mok ? mfd : ch.sock;
        if (!ch.send_fd(p, fd_to_send, mok ? 1 : 2, &cerr_)) mok = false;
      }
      if (mfd >= 0) ::close(mfd);
    } else {
      int src, tag, fd = -1;
      mok = ch.recv_fd(&src, &tag, &fd, &cerr_);
      if (mok && tag == 1) {
        mok = drv::p_cuMemImportFromShareableHandle(&mch, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) == CUDA_SUCCESS;
      } else {
        mok = false;
      }
      if (fd >= 0) ::close(fd);
    }
    if (all_ok(bs, mok)) mok = drv::p_cuMulticastAddDevice(mch, dev) == CUDA_SUCCESS; else mok = false;
    bool all_added = all_ok(bs, mok);  // every device must be added before any bind
    if (all_added) mok = drv::p_cuMulticastBindMem(mch, 0, mine, 0, sz, 0) == CUDA_SUCCESS;
    bool bound = all_added && all_ok(bs, mok);
    if (bound) {
      CUdeviceptr va = 0;
      mok = drv::p_cuMemAddressReserve(&va, sz, align, 0, 0) == CUDA_SUCCESS &&
            drv::p_cuMemMap(va, sz, 0, mch, 0) == CUDA_SUCCESS && drv::p_cuMemSetAccess(va, sz, &acc, 1) == CUDA_SUCCESS;
      if (all_ok(bs, mok)) {
        seg->mc_base = (char*)va;
        seg->mc_handle = mch;
      }
    }
    if (!seg->mc_base && mch) {  // agreed failure somewhere after the object existed: give it back
      if (bound || all_added) drv::p_cuMulticastUnbind(mch, dev, 0, sz);
      drv::p_cuMemRelease(mch);
    }
    if (!seg->mc_base && opt.require_multicast) {  // the same decision on every rank (mc_base is set after an all_ok)
      ch.close_();
      vmm_release(seg);
      if (err) *err = "multicast requested but could not be set up";
      return false;
    }
  }
  ch.close_();
  const bool up = upload_table(seg, err);
  if (!all_ok(bs, up)) {  // doubles as the exit barrier
    vmm_release(seg);
    if (err && up) *err = "device pointer table upload failed on some rank";
    return false;
  }
  return true;
}

// ---- cudaIpc provider --------------------------------------------------------------------------------
static bool ipc_create(const Bootstrap& bs, int device, size_t bytes, Segment* seg, std::string* err) {
  seg->kind = SEG_IPC;
  seg->bytes = bytes;
  seg->mapped_bytes = bytes;
  seg->nranks = bs.nranks;
  seg->self = bs.rank;
  seg->device = device;
  seg->base.assign((size_t)bs.nranks, nullptr);
  char* mine = nullptr;
  bool ok = cudaMalloc(&mine, bytes) == cudaSuccess && cudaMemset(mine, 0, bytes) == cudaSuccess &&
            cudaDeviceSynchronize() == cudaSuccess;
  cudaIpcMemHandle_t h;
  memset(&h, 0, sizeof h);
  if (ok && bs.nranks > 1) ok = cudaIpcGetMemHandle(&h, mine) == cudaSuccess;
  std::vector<cudaIpcMemHandle_t> all((size_t)bs.nranks);
  if (!bs.allgather(&h, all.data(), sizeof h)) ok = false;
  if (!all_ok(bs, ok)) {
    if (mine) cudaFree(mine);
    (void)cudaGetLastError();
    if (err) *err = "cudaMalloc / cudaIpcGetMemHandle failed on some rank";
    return false;
  }
  seg->base[(size_t)bs.rank] = mine;
  for (int p = 0; p < bs.nranks && ok; ++p) {
    if (p == bs.rank) continue;
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, all[(size_t)p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = false;
    seg->base[(size_t)p] = (char*)ptr;
  }
  if (!all_ok(bs, ok)) {
    (void)cudaGetLastError();
    if (err) *err = "cudaIpcOpenMemHandle failed on some rank";
    return false;
  }
  if (!upload_table(seg, err)) return false;
  return bs.barrier();
}

bool segment_create(const Bootstrap& bs, int device, size_t bytes, const SymmOptions& opt, Segment* seg,
                    std::string* err) {
  FT_RT(cudaSetDevice(device), "cudaSetDevice");
  if (bytes == 0) bytes = 16;
  std::string verr;
  if (opt.try_vmm) {
    Segment s;
    if (vmm_create(bs, device, bytes, opt, &s, &verr)) {
      *seg = s;
      return true;
    }
    if (opt.require_multicast) {
      if (err) *err = verr;
      return false;
    }
  }
  Segment s;
  std::string ierr;
  if (ipc_create(bs, device, bytes, &s, &ierr)) {
    *seg = s;
    return true;
  }
  if (err) *err = "symmetric allocation failed (vmm: " + verr + "; ipc: " + ierr + ")";
  return false;
}

void segment_destroy(Segment* seg) {
  if (!seg || seg->kind == SEG_NONE) return;
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(seg->device);
  if (seg->d_table) cudaFree(seg->d_table);
  if (seg->kind == SEG_VMM && drv::load()) {
    if (seg->mc_base) {
      drv::p_cuMemUnmap((CUdeviceptr)seg->mc_base, seg->mapped_bytes);
      drv::p_cuMemAddressFree((CUdeviceptr)seg->mc_base, seg->mapped_bytes);
      CUdevice dev;
      if (drv::p_cuDeviceGet(&dev, seg->device) == CUDA_SUCCESS)
        drv::p_cuMulticastUnbind((CUmemGenericAllocationHandle)seg->mc_handle, dev, 0, seg->mapped_bytes);
      drv::p_cuMemRelease((CUmemGenericAllocationHandle)seg->mc_handle);
    }
    for (int p = 0; p < seg->nranks; ++p) {
      if (seg->base[(size_t)p]) {
        drv::p_cuMemUnmap((CUdeviceptr)seg->base[(size_t)p], seg->mapped_bytes);
        drv::p_cuMemAddressFree((CUdeviceptr)seg->base[(size_t)p], seg->mapped_bytes);
      }
      if (seg->vmm_handles[(size_t)p]) drv::p_cuMemRelease((CUmemGenericAllocationHandle)seg->vmm_handles[(size_t)p]);
    }
  } else if (seg->kind == SEG_IPC) {
    for (int p = 0; p < seg->nranks; ++p) {
      if (!seg->base[(size_t)p]) continue;
      if (p == seg->self) cudaFree(seg->base[(size_t)p]);
      else cudaIpcCloseMemHandle(seg->base[(size_t)p]);
    }
  }
  // SEG_LOCAL / SEG_LOCAL_VMM: the memory belongs to the group (seg->shared) and goes with its last view
  (void)cudaGetLastError();
  seg->kind = SEG_NONE;
  seg->base.clear();
  seg->d_table = nullptr;
  seg->mc_base = nullptr;
  seg->shared.reset();
  if (prev >= 0) cudaSetDevice(prev);
}

// ---- one process, N devices ---------------------------------------------------------------------------------
namespace {
struct LocalShared {  // what the N views of a one-process segment share; freed with the last view
  SegKind kind = SEG_NONE;
  std::vector<int> devices;
  std::vector<char*> bufs;                        // per rank
  std::vector<unsigned long long> handles;        // SEG_LOCAL_VMM: CUmemGenericAllocationHandle per rank
  unsigned long long mc_handle = 0;
  char* mc_base = nullptr;
  size_t mapped = 0;
  ~LocalShared() {
    int prev = -1;
    cudaGetDevice(&prev);
    // nobody may still be using the memory: wait for every device of the group
    for (int d : devices) {
      cudaSetDevice(d);
      cudaDeviceSynchronize();
    }
    if (kind == SEG_LOCAL) {
      for (size_t i = 0; i < bufs.size(); ++i) {
        if (!bufs[i]) continue;
        cudaSetDevice(devices[i]);
        cudaFree(bufs[i]);
      }
    } else if (kind == SEG_LOCAL_VMM && drv::load()) {
      if (mc_base) {
        drv::p_cuMemUnmap((CUdeviceptr)mc_base, mapped);
        drv::p_cuMemAddressFree((CUdeviceptr)mc_base, mapped);
      }
      if (mc_handle) {
        for (int d : devices) {
          CUdevice dev;
          if (drv::p_cuDeviceGet(&dev, d) == CUDA_SUCCESS) drv::p_cuMulticastUnbind((CUmemGenericAllocationHandle)mc_handle, dev, 0, mapped);
        }
        drv::p_cuMemRelease((CUmemGenericAllocationHandle)mc_handle);
      }
      for (size_t i = 0; i < bufs.size(); ++i) {
        if (bufs[i]) {
          drv::p_cuMemUnmap((CUdeviceptr)bufs[i], mapped);
          drv::p_cuMemAddressFree((CUdeviceptr)bufs[i], mapped);
        }
        if (i < handles.size() && handles[i]) drv::p_cuMemRelease((CUmemGenericAllocationHandle)handles[i]);
      }
    }
    (void)cudaGetLastError();
    if (prev >= 0) cudaSetDevice(prev);
  }
};

bool distinct(const std::vector<int>& v) {
  for (size_t i = 0; i < v.size(); ++i)
    for (size_t j = i + 1; j < v.size(); ++j)
      if (v[i] == v[j]) return false;
  return true;
}

// VMM allocations on every device, each mapped ONCE with access for all devices (one process = one address space),
// optionally bound to one multicast object. Returns false (nothing leaked) if any step is unavailable.
bool local_vmm(const std::vector<int>& devices, size_t bytes, bool want_mc, std::shared_ptr<LocalShared>* out, std::string* err) {
  if (!drv::load()) { if (err) *err = "driver entry points unavailable"; return false; }
  const int n = (int)devices.size();
  auto sh = std::make_shared<LocalShared>();
  sh->kind = SEG_LOCAL_VMM;
  sh->devices = devices;
  sh->bufs.assign((size_t)n, nullptr);
  sh->handles.assign((size_t)n, 0);
  std::vector<CUdevice> cudev((size_t)n);
  for (int i = 0; i < n; ++i) {
    FT_RT(cudaSetDevice(devices[(size_t)i]), "cudaSetDevice");
    FT_RT(cudaFree(0), "context");
    FT_CU(drv::p_cuDeviceGet(&cudev[(size_t)i], devices[(size_t)i]), "cuDeviceGet");
    for (int j = 0; j < n; ++j) {
      if (i == j) continue;
      int can = 0;
      cudaDeviceCanAccessPeer(&can, devices[(size_t)i], devices[(size_t)j]);
      if (!can) { if (err) *err = "no peer access between devices"; return false; }
    }
  }
  want_mc = want_mc && n > 1;
  for (int i = 0; i < n && want_mc; ++i) want_mc = device_supports_multicast(devices[(size_t)i]);
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = devices[0];
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;  // same kind of allocation as the multi-process path
  size_t gran = 0, mc_gran = 0;
  FT_CU(drv::p_cuMemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED), "granularity");
  CUmulticastObjectProp mcp{};
  if (want_mc) {
    mcp.numDevices = (unsigned)n;
    mcp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    mcp.size = bytes;
    if (drv::p_cuMulticastGetGranularity(&mc_gran, &mcp, CU_MULTICAST_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) want_mc = false;
  }
  size_t align = gran > mc_gran ? gran : mc_gran;
  if (align == 0) align = 2u << 20;
  const size_t sz = (bytes + align - 1) / align * align;
  sh->mapped = sz;
  std::vector<CUmemAccessDesc> acc((size_t)n);
  for (int i = 0; i < n; ++i) {
    acc[(size_t)i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc[(size_t)i].location.id = devices[(size_t)i];
    acc[(size_t)i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int i = 0; i < n; ++i) {
    prop.location.id = devices[(size_t)i];
    CUmemGenericAllocationHandle h = 0;
    FT_CU(drv::p_cuMemCreate(&h, sz, &prop, 0), "cuMemCreate");
    sh->handles[(size_t)i] = h;
    CUdeviceptr va = 0;
    FT_CU(drv::p_cuMemAddressReserve(&va, sz, align, 0, 0), "cuMemAddressReserve");
    sh->bufs[(size_t)i] = (char*)va;  // from here on ~LocalShared cleans up
    FT_CU(drv::p_cuMemMap(va, sz, 0, h, 0), "cuMemMap");
    FT_CU(drv::p_cuMemSetAccess(va, sz, acc.data(), (size_t)n), "cuMemSetAccess");
    FT_RT(cudaSetDevice(devices[(size_t)i]), "cudaSetDevice");
    FT_RT(cudaMemset((void*)va, 0, sz), "cudaMemset(segment)");
    FT_RT(cudaDeviceSynchronize(), "sync");
  }
  if (want_mc) {  // failure here only costs the NVLS path
    mcp.size = sz;
    CUmemGenericAllocationHandle mch = 0;
    bool ok = drv::p_cuMulticastCreate(&mch, &mcp) == CUDA_SUCCESS;
    for (int i = 0; i < n && ok; ++i) ok = drv::p_cuMulticastAddDevice(mch, cudev[(size_t)i]) == CUDA_SUCCESS;  // all devices before any bind
    for (int i = 0; i < n && ok; ++i)
      ok = drv::p_cuMulticastBindMem(mch, 0, (CUmemGenericAllocationHandle)sh->handles[(size_t)i], 0, sz, 0) == CUDA_SUCCESS;
    CUdeviceptr va = 0;
    if (ok) ok = drv::p_cuMemAddressReserve(&va, sz, align, 0, 0) == CUDA_SUCCESS;
    if (ok) {
      if (drv::p_cuMemMap(va, sz, 0, mch, 0) == CUDA_SUCCESS && drv::p_cuMemSetAccess(va, sz, acc.data(), (size_t)n) == CUDA_SUCCESS) {
        sh->mc_base = (char*)va;
        sh->mc_handle = mch;
      } else {
        drv::p_cuMemAddressFree(va, sz);
        ok = false;
      }
    }
    if (!ok && mch) {
      if (!sh->mc_handle) {
        for (int i = 0; i < n; ++i) drv::p_cuMulticastUnbind(mch, cudev[(size_t)i], 0, sz);
        drv::p_cuMemRelease(mch);
      }
    }
  }
  *out = sh;
  return true;
}
}  // namespace

bool segment_create_local(const std::vector<int>& devices, size_t bytes, std::vector<Segment>* segs,
                          std::string* err, const SymmOptions* opt) {
  const int n = (int)devices.size();
  if (bytes == 0) bytes = 16;
  segs->assign((size_t)n, Segment());
  int prev = -1;
  cudaGetDevice(&prev);
  std::shared_ptr<LocalShared> sh;
  if (opt && opt->try_vmm && n > 1 && distinct(devices)) {
    std::string verr;
    if (!local_vmm(devices, bytes, opt->try_multicast, &sh, &verr)) {
      sh.reset();
      (void)cudaGetLastError();
      if (opt->require_multicast) { if (err) *err = "multicast requested but VMM set-up failed: " + verr; return false; }
    } else if (opt->require_multicast && !sh->mc_base) {
      if (err) *err = "multicast requested but could not be set up";
      return false;
    }
  }
  if (!sh) {
    sh = std::make_shared<LocalShared>();
    sh->kind = SEG_LOCAL;
    sh->devices = devices;
    sh->bufs.assign((size_t)n, nullptr);
    for (int i = 0; i < n; ++i) {
      FT_RT(cudaSetDevice(devices[(size_t)i]), "cudaSetDevice");
      for (int j = 0; j < n; ++j) {
        if (devices[(size_t)j] == devices[(size_t)i]) continue;
        int can = 0;
        cudaDeviceCanAccessPeer(&can, devices[(size_t)i], devices[(size_t)j]);
        if (!can) { if (err) *err = "no peer access between devices"; return false; }
        cudaError_t e = cudaDeviceEnablePeerAccess(devices[(size_t)j], 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
          if (err) *err = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e);
          return false;
        }
        (void)cudaGetLastError();
      }
      FT_RT(cudaMalloc(&sh->bufs[(size_t)i], bytes), "cudaMalloc(segment)");
      FT_RT(cudaMemset(sh->bufs[(size_t)i], 0, bytes), "cudaMemset(segment)");
      FT_RT(cudaDeviceSynchronize(), "sync");
    }
    sh->mapped = bytes;
  }
  for (int i = 0; i < n; ++i) {
    Segment& s = (*segs)[(size_t)i];
    s.kind = sh->kind;
    s.bytes = bytes;
    s.mapped_bytes = sh->mapped;
    s.nranks = n;
    s.self = i;
    s.device = devices[(size_t)i];
    s.base = sh->bufs;
    s.mc_base = sh->mc_base;
    s.shared = sh;
    FT_RT(cudaSetDevice(s.device), "cudaSetDevice");
    if (!upload_table(&s, err)) return false;
  }
  if (prev >= 0) cudaSetDevice(prev);
  return true;
}

}  // namespace ft
