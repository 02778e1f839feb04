// ft_allreduce_host: the collective for callers whose data lives in HOST memory - which is the reference's whole
// API (MPI_Allreduce on malloc'ed vectors, allreduce_over_mpi/benchmark.cpp:119-159, mpi_mod.hpp:1167-1221).
//
// A host-resident message has to cross PCIe twice, and PCIe (~55 GB/s per direction) is 15-20x slower than the
// NVLink collective. Done naively - copy in, allreduce, copy out - the three steps add up. Here the message is cut
// into chunks that flow through three streams: while chunk i is being reduced over NVLink, chunk i+1 is on its way
// up and chunk i-1 on its way down (PCIe is full duplex), so the call costs about max(H2D, D2H) instead of
// H2D + collective + D2H. The chunks are reduced IN PLACE in symmetric (registered) slots, i.e. on the zero-copy
// NVLS / P2P kernels, never through the staging window.
//
// State per communicator (created lazily by the first call, collectively: ft_symm_alloc is a collective): kSlots
// symmetric chunk buffers, three non-blocking streams, 3 x kSlots events. Released by ft_comm_destroy.
#include <cuda_runtime.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/flextree.h"
#include "../core/plan.h"
#include "../core/types.h"

namespace ft {
void set_last_error(const std::string& s);  // planner/c_api.cc
int comm_nlocal(const ft_comm_t* c);        // runtime/comm.cu
int comm_device(const ft_comm_t* c);
}  // namespace ft
using namespace ft;

namespace {

constexpr int kSlots = 3;

struct HostPipe {
  size_t chunk_bytes = 0;
  int nlocal = 1;
  void* slot[kSlots][FT_MAX_LOCAL];
  cudaStream_t s_in = nullptr, s_ar = nullptr, s_out = nullptr;
  cudaEvent_t e_in[kSlots], e_ar[kSlots], e_out[kSlots];
};

std::mutex g_mu;
std::map<const ft_comm_t*, HostPipe*> g_pipes;

#define HP_RT(call)                                                          \
  do {                                                                       \
    cudaError_t _e = (call);                                                 \
    if (_e != cudaSuccess) {                                                 \
      set_last_error(std::string(#call) + ": " + cudaGetErrorString(_e));    \
      (void)cudaGetLastError();                                              \
      return FT_ERR_CUDA;                                                    \
    }                                                                        \
  } while (0)

ft_result_t get_pipe(ft_comm_t* c, size_t chunk_bytes, HostPipe** out) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_pipes.find(c);
  if (it != g_pipes.end() && it->second->chunk_bytes >= chunk_bytes) {
    *out = it->second;
    return FT_SUCCESS;
  }
  if (it != g_pipes.end()) {
    // a larger chunk was asked for: the old slots stay allocated until the communicator goes (symmetric memory is
    // grow-only here, like the reference's staging buffer, mpi_mod.hpp:932-950)
    delete it->second;
    g_pipes.erase(it);
  }
  auto* p = new HostPipe;
  p->chunk_bytes = chunk_bytes;
  p->nlocal = comm_nlocal(c);
  for (int k = 0; k < kSlots; ++k) {
    void* ptrs[FT_MAX_LOCAL] = {};
    ft_result_t r = ft_symm_alloc(c, chunk_bytes, ptrs);
    if (r != FT_SUCCESS) { delete p; return r; }
    for (int l = 0; l < p->nlocal; ++l) p->slot[k][l] = ptrs[l];
  }
  HP_RT(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
  HP_RT(cudaStreamCreateWithFlags(&p->s_ar, cudaStreamNonBlocking));
  HP_RT(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
  for (int k = 0; k < kSlots; ++k) {
    HP_RT(cudaEventCreateWithFlags(&p->e_in[k], cudaEventDisableTiming));
    HP_RT(cudaEventCreateWithFlags(&p->e_ar[k], cudaEventDisableTiming));
    HP_RT(cudaEventCreateWithFlags(&p->e_out[k], cudaEventDisableTiming));
  }
  g_pipes[c] = p;
  *out = p;
  return FT_SUCCESS;
}

}  // namespace

namespace ft {
// called by ft_comm_destroy (the symmetric slots themselves are freed with the communicator's user segments)
void host_pipe_release(const ft_comm_t* c) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_pipes.find(c);
  if (it == g_pipes.end()) return;
  HostPipe* p = it->second;
  if (p->s_in) {
    cudaStreamDestroy(p->s_in);
    cudaStreamDestroy(p->s_ar);
    cudaStreamDestroy(p->s_out);
    for (int k = 0; k < kSlots; ++k) {
      cudaEventDestroy(p->e_in[k]);
      cudaEventDestroy(p->e_ar[k]);
      cudaEventDestroy(p->e_out[k]);
    }
  }
  (void)cudaGetLastError();
  delete p;
  g_pipes.erase(it);
}
}  // namespace ft

extern "C" ft_result_t ft_allreduce_host_multi(const void* const* sendbufs, void* const* recvbufs, size_t count,
                                               ft_dtype_t dtype, ft_op_t op, double scale, ft_comm_t* comm,
                                               size_t chunk_bytes) {
  if (!comm || !recvbufs) return FT_ERR_INVALID;
  const int es = dtype_size((int)dtype);
  if (es == 0 || (int)dtype == DT_FP8_BLOCK) { set_last_error("ft_allreduce_host: bad dtype"); return FT_ERR_INVALID; }
  if (count == 0) return FT_SUCCESS;
  const int L = comm_nlocal(comm);
  if (chunk_bytes == 0) {
    const char* v = getenv("FT_HOST_CHUNK_BYTES");
    chunk_bytes = (v && *v) ? (size_t)atoll(v) : (size_t)(16u << 20);
  }
  const size_t total = count * (size_t)es;
  if (chunk_bytes > total) chunk_bytes = total;
  chunk_bytes = (chunk_bytes + 4095) & ~(size_t)4095;  // every chunk but the last is a whole number of 16-byte vectors
  HP_RT(cudaSetDevice(comm_device(comm)));
  HostPipe* p = nullptr;
  ft_result_t r = get_pipe(comm, chunk_bytes, &p);
  if (r != FT_SUCCESS) return r;
  const size_t step = p->chunk_bytes > chunk_bytes ? chunk_bytes : p->chunk_bytes;
  const size_t nchunks = (total + step - 1) / step;
  for (size_t i = 0; i < nchunks; ++i) {
    const int k = (int)(i % kSlots);
    const size_t off = i * step, n = total - off < step ? total - off : step;
    if (i >= (size_t)kSlots) HP_RT(cudaStreamWaitEvent(p->s_in, p->e_out[k], 0));  // the slot has been drained
    for (int l = 0; l < L; ++l) {
      const void* src = (sendbufs && sendbufs[l] && sendbufs[l] != FT_IN_PLACE) ? sendbufs[l] : recvbufs[l];
      HP_RT(cudaMemcpyAsync(p->slot[k][l], (const char*)src + off, n, cudaMemcpyHostToDevice, p->s_in));
    }
    HP_RT(cudaEventRecord(p->e_in[k], p->s_in));
    HP_RT(cudaStreamWaitEvent(p->s_ar, p->e_in[k], 0));
    r = ft_allreduce_multi(nullptr, p->slot[k], n / (size_t)es, dtype, dtype, op, scale, comm, p->s_ar);
    if (r != FT_SUCCESS) {
      cudaDeviceSynchronize();
      return r;
    }
    HP_RT(cudaEventRecord(p->e_ar[k], p->s_ar));
    HP_RT(cudaStreamWaitEvent(p->s_out, p->e_ar[k], 0));
    for (int l = 0; l < L; ++l)
      HP_RT(cudaMemcpyAsync((char*)recvbufs[l] + off, p->slot[k][l], n, cudaMemcpyDeviceToHost, p->s_out));
    HP_RT(cudaEventRecord(p->e_out[k], p->s_out));
  }
  HP_RT(cudaStreamSynchronize(p->s_out));  // like MPI_Allreduce: the result is in recvbuf when the call returns
  if (ft_comm_error_peek(comm)) {
    set_last_error("ft_allreduce_host: the device watchdog fired during this call; recvbuf is invalid");
    return FT_ERR_TIMEOUT;
  }
  return FT_SUCCESS;
}

extern "C" ft_result_t ft_allreduce_host(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op,
                                         ft_comm_t* comm) {
  if (!comm) return FT_ERR_INVALID;
  if (comm_nlocal(comm) != 1) {
    set_last_error("emulated communicators take one buffer per logical rank: use ft_allreduce_host_multi");
    return FT_ERR_INVALID;
  }
  const void* s[1] = {sendbuf};
  void* r[1] = {recvbuf};
  return ft_allreduce_host_multi(s, r, count, dtype, op, 1.0, comm, 0);
}
