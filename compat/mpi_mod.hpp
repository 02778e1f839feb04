// compat/mpi_mod.hpp - source-level drop-in for the reference's header of the same name.
//
// A program written against Youhe-Jiang/AllReduce-Over-MPI does `#include "mpi_mod.hpp"` and then either calls
// MPI_Allreduce_FT (when built with -DSTANDALONE_TEST, as the reference's benchmark.cpp is) or has its
// MPI_Allreduce calls captured (allreduce_over_mpi/mpi_mod.hpp:1167-1172). Put this
// directory on the include path instead of the reference's and link libflextree_host.so: the same program, not a
// line changed, now runs on this library's CPU data plane (csrc/host/host_comm.cc - shared-memory windows, release/
// acquire flags instead of messages and barriers, the algorithm templates the GPU kernels are built from).
// `make dropin` does exactly that with the reference's unmodified benchmark.cpp; tests/test_dropin.py compares the
// two binaries' CHECK lines; bench/cpu_compare.py times them side by side.
//
// What is kept from the reference's contract (citations: allreduce_over_mpi/mpi_mod.hpp):
//   * signature and "returns 0 always" (:1167-1221); MPI_IN_PLACE (:1194-1215); N <= 1 memcpy fast path (:1181-1188)
//   * FT_TOPO: unset -> {N}; any 1 -> ring; product != N -> "invalid FT_TOPO" + exit(1)  (get_stages, :882-929)
//   * dtypes / ops of handle_reduce (:825-874); anything else prints and exits(1) like the reference does
//   * FlexTree::get_stages, FlexTree::comm_only, FT_enabled() - the other names benchmark.cpp touches
//   * a grow-only staging area (flextree_register_the_buffer, :932-950): the communicator is re-created with a
//     larger window when a larger message arrives, never shrunk, never freed at exit
// What is not: the MPI_Isend/Irecv/Waitall/Barrier data plane. MPI is used for the bootstrap only (a byte
// all-gather built from point-to-point calls, so that it also works on MPI subsets without MPI_Allgather), which
// limits this header to ranks that share one host. Widths are not capped at 20 (MAX_NUM_BLOCKS, :245) but at 32.
#ifndef FLEXTREE_COMPAT_MPI_MOD_HPP_
#define FLEXTREE_COMPAT_MPI_MOD_HPP_

#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <iostream>
#include <map>
#include <vector>

#include "flextree_host.h"
// The GPU half of the library is optional here: a program linked against libflextree.so (the CUDA build) gets its
// DEVICE buffers reduced by the NVLink kernels; linked against the host-only libflextree_host.so these symbols are
// absent (weak, null) and every buffer is treated as host memory.
#pragma weak ft_pointer_device
#pragma weak ft_comm_create
#pragma weak ft_allreduce_blocking
#undef MPI_Allreduce_FT  // flextree.h aliases it to the GPU entry point under STANDALONE_TEST; here it is the function below

static int FT_enabled() { return ft_enabled(); }

#ifdef STANDALONE_TEST
const int INF = 0x3F3F3F3F;  // benchmark.cpp:37 initialises its minimum with it (mpi_mod.hpp:22)
#endif

namespace FlexTree {

static bool comm_only = false;

static std::vector<size_t> get_stages(const size_t& num_nodes) {
  int w[64];
  const int n = ft_parse_topo(nullptr, (int)num_nodes, w, 64);
  if (n < 0) {
    std::cerr << "invalid FT_TOPO (the product of the widths must be " << num_nodes << ")" << std::endl;
    exit(1);
  }
  return std::vector<size_t>(w, w + n);
}

namespace detail {

struct World {
  MPI_Comm comm;
  int rank, size;
};

// Byte all-gather over point-to-point calls (bootstrap only).
inline int allgather_bytes(void* ctx, const void* send, void* recv, size_t bytes) {
  const World* w = (const World*)ctx;
  std::vector<MPI_Request> reqs((size_t)(2 * (w->size - 1)));
  int n = 0;
  memcpy((char*)recv + (size_t)w->rank * bytes, send, bytes);
  for (int p = 0; p < w->size; ++p)
    if (p != w->rank) MPI_Irecv((char*)recv + (size_t)p * bytes, (int)bytes, MPI_UINT8_T, p, 0, w->comm, &reqs[(size_t)n++]);
  for (int p = 0; p < w->size; ++p)
    if (p != w->rank) MPI_Isend(send, (int)bytes, MPI_UINT8_T, p, 0, w->comm, &reqs[(size_t)n++]);
  return MPI_Waitall(n, reqs.data(), MPI_STATUSES_IGNORE) == MPI_SUCCESS ? 0 : 1;
}

struct State {
  World world;
  ft_host_comm_t* hc = nullptr;
  size_t window = 0;
  ft_comm_t* gpu = nullptr;  // created on the first call with device buffers
  int gpu_device = -1;
};

inline long long env_ll(const char* name, long long dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoll(v) : dflt;
}

inline int to_dtype(MPI_Datatype dt) {
  if (dt == MPI_FLOAT) return FT_FLOAT;
  if (dt == MPI_DOUBLE) return FT_DOUBLE;
  if (dt == MPI_INT32_T) return FT_INT32;
  if (dt == MPI_INT64_T) return FT_INT64;
  if (dt == MPI_UINT8_T) return FT_UINT8;
  if (dt == MPI_INT8_T) return FT_INT8;
  if (dt == MPI_UINT16_T) return FT_UINT16;
  if (dt == MPI_INT16_T) return FT_INT16;
  if (dt == MPI_C_BOOL) return FT_BOOL;
  if (dt == MPI_LONG_LONG_INT || dt == MPI_LONG_LONG) return FT_LONG_LONG;
  if (dt == MPI_INT) return FT_INT32;
  if (dt == MPI_UINT32_T) return FT_UINT32;
  if (dt == MPI_UINT64_T) return FT_UINT64;
  return -1;
}

inline int to_op(MPI_Op op) {
  if (op == MPI_SUM) return FT_SUM;
  if (op == MPI_BAND) return FT_BAND;
  if (op == MPI_MAX) return FT_MAX;
  if (op == MPI_MIN) return FT_MIN;
  return -1;
}

// One communicator per MPI communicator, created on first use, re-created (collectively: every rank sees the
// same count) when a message needs a larger window.
inline State& state_for(MPI_Comm comm, size_t msg_bytes) {
  static std::map<MPI_Comm, State> table;
  State& s = table[comm];
  if (!s.hc && !s.window) {
    s.world.comm = comm;
    MPI_Comm_rank(comm, &s.world.rank);
    MPI_Comm_size(comm, &s.world.size);
  }
  // room for the blocks (each padded to 16 bytes) plus the fp8 scale plane the window layout always reserves
  const size_t need = msg_bytes + msg_bytes / 16 + (size_t)s.world.size * 32 + 4096;
  if (!s.hc || need > s.window) {
    size_t win = s.window ? s.window : (size_t)1 << 20;
    while (win < need) win *= 2;
    if (s.hc) ft_host_comm_destroy(s.hc);
    s.hc = nullptr;
    if (ft_host_comm_create(&s.hc, s.world.rank, s.world.size, allgather_bytes, &s.world, win, 0) != FT_SUCCESS) {
      std::cerr << "flextree: cannot create the shared-memory communicator: " << ft_last_error() << std::endl;
      exit(1);
    }
    s.window = ft_host_comm_window_bytes(s.hc);
  }
  return s;
}

// Device buffers: one GPU communicator per MPI communicator (handles exchanged over the same MPI point-to-point
// all-gather), created on first use. The MPI ranks must sit on one NVSwitch box, one GPU each.
inline ft_comm_t* gpu_comm_for(MPI_Comm comm, int device) {
  static std::map<MPI_Comm, State*> table;
  State*& s = table[comm];
  if (!s) {
    s = new State;  // never freed, like the reference's staging buffer (:932-950)
    s->world.comm = comm;
    MPI_Comm_rank(comm, &s->world.rank);
    MPI_Comm_size(comm, &s->world.size);
  }
  if (!s->gpu) {
    if (ft_comm_create(&s->gpu, s->world.rank, s->world.size, device, allgather_bytes, &s->world, nullptr) != FT_SUCCESS) {
      std::cerr << "flextree: cannot create the GPU communicator: " << ft_last_error() << std::endl;
      exit(1);
    }
    s->gpu_device = device;
  } else if (s->gpu_device != device) {
    std::cerr << "flextree: a rank must keep using buffers of one GPU (was " << s->gpu_device << ", now " << device << ")" << std::endl;
    exit(1);
  }
  return s->gpu;
}

}  // namespace detail
}  // namespace FlexTree

inline int MPI_Allreduce_FT(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype, MPI_Op op, MPI_Comm comm) {
  using namespace FlexTree::detail;
  const int dt = to_dtype(datatype), rop = to_op(op);
  if (dt < 0) {
    char name[256] = "?";
    int len = 0;
    MPI_Type_get_name(datatype, name, &len);
    std::cerr << "flextree: unsupported datatype " << name << std::endl;
    exit(1);
  }
  if (rop < 0) {
    std::cerr << "flextree: unsupported reduce op" << std::endl;
    exit(1);
  }
  int tsize = 0, nranks = 1;
  MPI_Type_size(datatype, &tsize);
  MPI_Comm_size(comm, &nranks);
  const bool in_place = sendbuf == MPI_IN_PLACE;
  if (count > 0) {
    // Where do the buffers live? Device memory goes to the NVLink kernels (FT_TOPO / FT_ALGO apply there as well).
    if (ft_pointer_device && ft_comm_create && ft_allreduce_blocking) {
      const int dev = ft_pointer_device(recvbuf);
      const int sdev = in_place ? dev : ft_pointer_device(sendbuf);
      if (dev >= 0 || sdev >= 0) {
        if (dev != sdev) {
          std::cerr << "flextree: sendbuf and recvbuf must both be device memory of the same GPU (or both host memory)" << std::endl;
          exit(1);
        }
        ft_comm_t* g = gpu_comm_for(comm, dev);
        if (ft_allreduce_blocking(in_place ? FT_IN_PLACE : sendbuf, recvbuf, (size_t)count, (ft_dtype_t)dt, (ft_op_t)rop, g) != FT_SUCCESS) {
          std::cerr << "flextree: GPU allreduce failed: " << ft_last_error() << std::endl;
          exit(1);
        }
        return 0;
      }
    }
  }
  if (nranks <= 1 || count <= 0) {
    if (!in_place && count > 0) memcpy(recvbuf, sendbuf, (size_t)count * (size_t)tsize);
    return 0;
  }
  State& s = state_for(comm, (size_t)count * (size_t)tsize);
  const std::vector<size_t> stages = FlexTree::get_stages((size_t)nranks);
  int widths[64];
  for (size_t i = 0; i < stages.size(); ++i) widths[i] = (int)stages[i];
  const bool ring = stages[0] == 1;
  static const int threads = (int)env_ll("FT_HOST_THREADS", 1);
  static const long long tile = env_ll("FT_HOST_TILE_BYTES", 262144);
  static const unsigned long long timeout_ns = (unsigned long long)env_ll("FT_HOST_TIMEOUT_S", 0) * 1000000000ull;
  const long long rc = ft_host_comm_allreduce(s.hc, in_place ? recvbuf : sendbuf, recvbuf, count, dt, dt, rop,
                                              ring ? (int)FT_ALGO_RING : FT_ALGO_STAGES_BULK, widths, (int)stages.size(),
                                              threads, tile, /*mode=*/0, 1.0, 0, timeout_ns, 0, 0);
  if (rc != 0) {
    std::cerr << "flextree: allreduce failed (" << rc << ")" << std::endl;
    exit(1);
  }
  return 0;
}

#ifndef STANDALONE_TEST
// The reference lives inside a modified mpi.h and defines a file-local MPI_Allreduce there (:1170). From a separate
// header the same capture is a macro: every MPI_Allreduce call after this include goes to the function above.
#define MPI_Allreduce MPI_Allreduce_FT
#endif

#endif  // FLEXTREE_COMPAT_MPI_MOD_HPP_
