// Link probe: raw NVLink traffic patterns between the ranks of a communicator, with nothing else in the kernel -
// no flags, no reduction, no dependency between ranks. It answers the questions the link model (planner/cost_model.h)
// and the kernel design depend on, on the box the library runs on:
//   * what does one direction of a GPU's links sustain when only that direction carries data, and what when both do?
//   * is a peer READ (request out, data back) as cheap as a peer WRITE (data out, acknowledgement back)?
//   * register-staged 16-byte accesses vs TMA bulk copies through shared memory
//   * how does the rate depend on the number of peers a GPU talks to at once (fan-in / fan-out)?
// The reference's cost model takes such constants as literals fitted by hand on the authors' cluster
// (cost_model/CostModel.h:3-4, :24, :37); here they are measured (bench/fit_cost_model.py).
//
// Modes: 0 LDG pull (ld.relaxed.sys.v4), 1 STG push (st.global.v4), 2 TMA pull (cp.async.bulk shared <- peer global),
//        3 TMA push (cp.async.bulk peer global <- shared).
// Every CTA works on one peer: peer index (blockIdx.x mod npeers), peers are rank+1, rank+2, ... (mod N). A CTA
// streams its share of `bytes_per_peer` from/to offset region(rank or peer) of the peer's window; contents are
// garbage by design (a pull discards what it read, a push writes whatever is in shared memory / registers).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../core/types.h"
#include "launch.h"
#include "tma_tiles.cuh"

namespace ft {

constexpr int kProbeStages = 4;
constexpr int kProbeChunk = 32 << 10;

__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(512, 1) ft_link_probe_kernel(char* const* win, int rank, int nranks, int npeers,
                                                                long long bytes_per_peer, int mode, unsigned* sink) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int pi = (int)blockIdx.x % npeers;
  const int ctas_per_peer = ((int)gridDim.x - pi + npeers - 1) / npeers;
  const int ci = (int)blockIdx.x / npeers;
  const int p = (rank + 1 + pi) % nranks;
  // my share of the peer's region, in 16-byte vectors
  const long long nvec = bytes_per_peer / 16;
  const long long per = (nvec + ctas_per_peer - 1) / ctas_per_peer;
  const long long v0 = per * ci, v1 = (v0 + per < nvec) ? v0 + per : nvec;
  if (v0 >= v1) return;
  // pull: read the slot the peer keeps for me... any region will do; push: write into the slot the peer keeps for my rank
  char* remote = win[p] + (long long)(mode == 1 || mode == 3 ? rank : p) * bytes_per_peer;
  const int tid = (int)threadIdx.x, nth = (int)blockDim.x;
  if (mode == 0) {
    uint32_t acc = 0;
    const char* src = remote + v0 * 16 + (long long)tid * 16;
    long long left = v1 - v0;
    for (; left >= (long long)nth * 8; left -= (long long)nth * 8, src += (long long)nth * 8 * 16) {
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + (long long)u * nth * 16) : "memory");
#pragma unroll
      for (int u = 0; u < 8; ++u) acc ^= v[u].x ^ v[u].w;
    }
    if (acc == 0x9e3779b9u) sink[0] = acc;  // keeps the loads alive
  } else if (mode == 1) {
    char* dst = remote + v0 * 16 + (long long)tid * 16;
    long long left = v1 - v0;
    const uint4 v = make_uint4((unsigned)tid, (unsigned)rank, 0u, 0u);
    for (; left >= (long long)nth * 8; left -= (long long)nth * 8, dst += (long long)nth * 8 * 16) {
#pragma unroll
      for (int u = 0; u < 8; ++u) *reinterpret_cast<uint4*>(dst + (long long)u * nth * 16) = v;
    }
  } else if (mode == 2) {
    uint64_t* full = (uint64_t*)(smem + kProbeStages * kProbeChunk);
    if (tid == 0) {
      for (int i = 0; i < kProbeStages; ++i) mbar_init(full + i, 1);
      fence_mbarrier_init();
    }
    __syncthreads();
    if (tid == 0) {  // one thread keeps kProbeStages bulk loads in flight; nobody consumes the data
      const long long bytes = (v1 - v0) * 16;
      const char* src = remote + v0 * 16;
      const long long nchunks = (bytes + kProbeChunk - 1) / kProbeChunk;
      for (long long i = 0; i < nchunks + kProbeStages; ++i) {
        const int s = (int)(i % kProbeStages);
        if (i >= kProbeStages) {
          int dummy = 0;
          mbar_wait(full + s, (uint32_t)(((i / kProbeStages) - 1) & 1), &dummy);
        }
        if (i < nchunks) {
          const long long o = i * kProbeChunk;
          const uint32_t len = (uint32_t)(bytes - o < kProbeChunk ? bytes - o : kProbeChunk);
          mbar_expect_tx(full + s, len);
          tma_load_1d(smem + s * kProbeChunk, src + o, len, full + s);
        }
      }
    }
  } else {
    if (tid == 0) {
      const long long bytes = (v1 - v0) * 16;
      char* dst = remote + v0 * 16;
      const long long nchunks = (bytes + kProbeChunk - 1) / kProbeChunk;
      for (long long i = 0; i < nchunks; ++i) {
        const long long o = i * kProbeChunk;
        const uint32_t len = (uint32_t)(bytes - o < kProbeChunk ? bytes - o : kProbeChunk);
        bulk_store(dst + o, smem + (i % kProbeStages) * kProbeChunk, len);
        bulk_commit();
        bulk_wait_read<kProbeStages - 1>();  // at most kProbeStages stores reading shared memory at a time
      }
      bulk_wait<0>();  // all writes performed
    }
  }
}

cudaError_t launch_link_probe(char* const* d_table, int rank, int nranks, int npeers, long long bytes_per_peer, int mode,
                              int ctas, unsigned* sink, cudaStream_t stream) {
  const size_t smem = (mode >= 2) ? (size_t)kProbeStages * kProbeChunk + 128 : 0;
  if (smem > (48u << 10)) {
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
      cudaError_t e = cudaFuncSetAttribute((const void*)ft_link_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      attr_set[dev] = true;
    }
  }
  ft_link_probe_kernel<<<ctas, mode >= 2 ? 32 : 512, smem, stream>>>(d_table, rank, nranks, npeers, bytes_per_peer, mode, sink);
  return cudaGetLastError();
}

}  // namespace ft
