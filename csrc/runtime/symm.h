// Symmetric memory: one buffer per rank, every rank's buffer mapped into every process, optionally
// bound to an NVLS multicast object. This is the B200-native stand-in for the reference's grow-only
// host staging buffer (flextree_register_the_buffer, allreduce_over_mpi/mpi_mod.hpp:932-950) and for
// the MPI transport underneath handle_send/handle_recv (:663-765): peers are reached with plain
// loads/stores through these mappings.
//
// Three providers:
//   VMM   cuMemCreate + POSIX-fd export, fds passed over abstract unix datagram sockets (SCM_RIGHTS),
//         cuMemMap of every peer, cuMulticastCreate/AddDevice/BindMem when the device supports it
//   IPC   cudaMalloc + cudaIpcGetMemHandle (fallback when VMM export is not permitted)
//   LOCAL one process: cudaMalloc on each device + cudaDeviceEnablePeerAccess (also the single-GPU
//         emulation, where every "rank" is a slice of the same device)
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include <memory>
#include <string>
#include <vector>

#include "../../include/flextree.h"

namespace ft {

// Byte exchange between the ranks of a communicator (collective).
struct Bootstrap {
  int rank = 0, nranks = 1;
  ft_allgather_fn fn = nullptr;
  void* ctx = nullptr;
  bool allgather(const void* send, void* recv, size_t bytes) const {
    if (nranks == 1) {
      memcpy(recv, send, bytes);
      return true;
    }
    return fn && fn(ctx, send, recv, bytes) == 0;
  }
  bool barrier() const {
    char x = 0;
    std::vector<char> all((size_t)nranks);
    return allgather(&x, all.data(), 1);
  }
};

enum SegKind { SEG_NONE = 0, SEG_VMM = 1, SEG_IPC = 2, SEG_LOCAL = 3, SEG_LOCAL_VMM = 4 };

struct Segment {
  SegKind kind = SEG_NONE;
  size_t bytes = 0;              // per-rank size
  int nranks = 0;
  std::vector<char*> base;       // [nranks] address of rank p's buffer as seen from this rank
  char** d_table = nullptr;      // device copy of `base` (what kernels index)
  char* mc_base = nullptr;       // multicast alias (VMM + NVLS only)
  // provider-private
  std::vector<unsigned long long> vmm_handles;  // CUmemGenericAllocationHandle per rank
  unsigned long long mc_handle = 0;
  size_t mapped_bytes = 0;
  int device = 0;
  int self = 0;
  // One-process groups (SEG_LOCAL, SEG_LOCAL_VMM): the N per-rank views alias the same allocations and mappings.
  // They are released when the LAST view is destroyed - a member that goes away early must not pull the memory from
  // under peers whose kernels may still be reading it.
  std::shared_ptr<void> shared;
};

struct SymmOptions {
  bool try_vmm = true;
  bool try_multicast = true;
  bool require_multicast = false;
};

// Multi-process segment (collective over `bs`). zero-fills the local buffer. On failure returns false
// with *err set; all ranks fail or succeed together.
bool segment_create(const Bootstrap& bs, int device, size_t bytes, const SymmOptions& opt, Segment* seg,
                    std::string* err);
void segment_destroy(Segment* seg);

// One-process segments: devices[i] hosts rank i (devices may repeat: single-GPU emulation). Fills
// segs[i] with rank i's view. With distinct devices and opt.try_vmm the buffers come from the VMM API (cuMemCreate,
// one mapping visible to every device) and, if opt.try_multicast and the devices support it, are bound to ONE
// multicast object, so that a single process driving all GPUs gets the NVLS kernels too (no fd passing needed:
// everything happens in one address space). Otherwise cudaMalloc + cudaDeviceEnablePeerAccess.
bool segment_create_local(const std::vector<int>& devices, size_t bytes, std::vector<Segment>* segs,
                          std::string* err, const SymmOptions* opt = nullptr);

bool device_supports_multicast(int device);

}  // namespace ft
