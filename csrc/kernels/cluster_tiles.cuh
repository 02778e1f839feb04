// Thread-block-cluster form of the TMA-staged P2P stages: a CLUSTER of two CTAs (two SMs) owns a tile.
//
// Reduce-scatter: the stage's sources (own slice + w-1 peers) are split between the two CTAs. Each CTA bulk-loads
// ITS half of the sources into its own shared-memory ring (plain cp.async.bulk, no multicast: a remote chunk is
// fetched exactly once, by the CTA that needs it), sums them, and the two partial sums meet through distributed
// shared memory: for every chunk each CTA finalises one half of the vectors and hands the partial sums of the other
// half to its partner with st.shared::cluster + a remote mbarrier arrive; the partner adds them to its own, applies
// the fused scale, rounds and stores the owner's slice. All-gather: the chunk list is dealt out alternately.
// Everything else of the algorithm text (flags, copy loops, epilogues) runs unchanged with the pair acting as one
// 2 x blockDim "CTA": tid()/nthr() span both CTAs and sync() is the hardware cluster barrier.
//
// Numerics: the sum is (sources of CTA 0 in ascending order) + (sources of CTA 1 in ascending order) - deterministic
// and identical on every rank, but NOT the reference's "own first, then peers ascending" order, so fp32 results can
// differ from the non-cluster kernels in the last bit (tests compare with a tolerance).
//
// This is the north-star item "thread-block clusters share reduced tiles across an SM group"; whether it pays on
// B200 is a measurement, see profiles/clusters.md. The reference's counterpart is the k-way reduce_sum over the
// received blocks (allreduce_over_mpi/mpi_mod.hpp:246-452, dispatched from handle_reduce :769-878).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "backend.cuh"
#include "tma_tiles.cuh"

namespace ft {

constexpr int kClStages = 3;                 // ring slots per CTA
constexpr int kClSlotBytes = 48 << 10;
constexpr int kClMaxChunk = 12 << 10;        // per-source chunk: 768 vectors, so half a chunk fits the consumer threads
constexpr int kClXchBytes = (kClMaxChunk / 16 / 2) * 32;  // partial sums of half a chunk: up to 8 fp32 per vector
constexpr int kClRingBytes = kClStages * kClSlotBytes;
constexpr int kClSmemBytes = kClRingBytes + 2 * kClXchBytes + 128;  // + full[3] empty[3] xch[2] mbarriers

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// address of `smem_ptr` (a shared::cta address of THIS CTA) in the shared memory of CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* b, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
      : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
  return ok != 0;
}

struct ClusterRing {
  char* ring;      // kClStages slots
  char* xch;       // 2 exchange buffers (written by the partner CTA through DSMEM)
  uint64_t* full;  // [kClStages]
  uint64_t* empty; // [kClStages]
  uint64_t* xbar;  // [2] "the partner's partial sums for this chunk have landed"
};

// A pair of CTAs presented to the algorithm text as one wide CTA.
struct CudaClusterBE : CudaBE {
  ClusterRing cr_;
  uint32_t crank_;
  uint32_t xit_;  // chunks exchanged so far (same value in every consumer thread of both CTAs)
  __device__ __forceinline__ CudaClusterBE(int* abort_smem, const ClusterRing& cr)
      : CudaBE(abort_smem, TmaRing{cr.ring, cr.full, cr.empty}), cr_(cr), crank_(cluster_ctarank()), xit_(0) {}
  __device__ __forceinline__ int tid() const { return (int)(crank_ * blockDim.x + threadIdx.x); }
  __device__ __forceinline__ int nthr() const { return (int)(2 * blockDim.x); }
  __device__ __forceinline__ void sync() { cluster_sync_all(); }
  __device__ __forceinline__ void finish(const CommDev& cd, int clusters) {
    cluster_sync_all();  // nobody leaves while its partner may still write into its shared memory
    if (crank_ == 0 && threadIdx.x == 0) {
      uint64_t epoch = load_epoch(cd.epoch);
      __threadfence();
      unsigned old = atomicAdd(cd.done, 1u);
      if (old == (unsigned)clusters - 1u) {
        atomicExch(cd.done, 0u);
        __threadfence();
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(cd.epoch), "l"(epoch + 1) : "memory");
      }
    }
  }
  template <typename T, typename IO, class Op, int VE, class Cta>
  __device__ __forceinline__ void rs_tile_tma(Cta& c, int k, int t);
  template <typename T, typename IO, class Op, int VE, class Cta>
  __device__ __forceinline__ void ag_tile_tma(Cta& c, int k, int t);
};

__device__ __forceinline__ void cluster_ring_init(const ClusterRing& R, int consumer_warps, int consumer_threads) {
  if (threadIdx.x == 0) {
    for (int i = 0; i < kClStages; ++i) {
      mbar_init(R.full + i, 1);
      mbar_init(R.empty + i, consumer_warps);
    }
    for (int i = 0; i < 2; ++i) mbar_init(R.xbar + i, consumer_threads);  // every consumer thread of the partner arrives
    fence_mbarrier_init();
  }
  cluster_sync_all();  // both CTAs' barriers exist before anybody arrives remotely
}

// ---- reduce-scatter stage k, tile t, by a CTA pair -------------------------------------------------------------
template <typename T, typename IO, class Op, int VE, class Cta>
__device__ __forceinline__ void CudaClusterBE::rs_tile_tma(Cta& c, int k, int t) {
  using EL = Elem<T>;
  using A = typename EL::acc;
  using PackT = Pack<T, VE>;
  static_assert(sizeof(A) == 4 && VE * 4 <= 32, "exchange slots hold up to 8 32-bit partial sums per vector");
  const ClusterRing& R = cr_;
  const Plan& P = c.a.plan;
  const int w = P.width[k], gn = P.gap[k + 1], nb = P.nranks / gn, rmod = c.r % gn, npeers = w - 1;
  const bool from_user = (k == 0) && (c.a.mode == MODE_STAGING);
  const bool final_stage = (k == P.nstages - 1) && c.a.lonely == 0;
  const bool own_raw = !from_user || Cta::kRawIO;
  // global source list: [own (if it can ride the ring)] + peers ascending; source s belongs to CTA (s & 1)
  const int nsrc = npeers + (own_raw ? 1 : 0);
  const int me = (int)crank_;
  const int mine = (nsrc - me + 1) / 2;          // my sources: me, me+2, ...
  const int half_src = (nsrc + 1) / 2;           // both CTAs size their chunks from the larger half
  int CH = (kClSlotBytes / (half_src > 0 ? half_src : 1)) & ~1023;
  if (CH > kClMaxChunk) CH = kClMaxChunk;
  const char* own_base = from_user ? (const char*)c.rc.send : (const char*)(c.rc.win[c.r] + c.woff);
  char* wr = c.rc.win[c.r] + c.woff;
  const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
  uint32_t it = c.pipe_it;
  auto src_ptr = [&](int s) -> const char* {  // s-th source of the global list
    if (own_raw) return s == 0 ? own_base : c.rs_peer_ptr(k, s - 1);
    return c.rs_peer_ptr(k, s);
  };

  if (warp == 0) {
    if (lane == 0) fence_proxy_async();
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      for (int o = 0; o < bytes; o += CH, ++it) {
        if (lane != 0 || mine == 0) continue;
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kClStages);
        mbar_wait(R.empty + s, ((it / kClStages) & 1) ^ 1, c.abort_flag);
        mbar_expect_tx(R.full + s, (uint32_t)(mine * len));
        char* slot = R.ring + s * kClSlotBytes;
        for (int j = 0; j < mine; ++j) tma_load_1d(slot + j * CH, src_ptr(me + 2 * j) + base_off + o, (uint32_t)len, R.full + s);
      }
    }
  } else {
    const int ctid = (int)threadIdx.x - 32;
    const IO* x = (const IO*)c.rc.send;
    const uint32_t partner = (uint32_t)(me ^ 1);
    uint32_t xit = xit_;
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      for (int o = 0; o < bytes; o += CH, ++it, ++xit) {
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kClStages);
        const int nvec = len >> 4, half_n = (nvec + 1) >> 1;  // CTA 0 finalises [0, half_n), CTA 1 [half_n, nvec)
        const int xb = (int)(xit & 1);
        if (mine > 0) mbar_wait(R.full + s, (it / kClStages) & 1, c.abort_flag);
        const char* slot = R.ring + s * kClSlotBytes;
        // partial sum over MY sources of one vector (ascending source order)
        auto partial = [&](int v, A* acc) {
          bool first = true;
          if (!own_raw && me == 0) {  // the own contribution comes from the user buffer (cast to the wire type first)
            const PackT own = Cta::io_to_wire(Cta::ld_user(x, start + (o >> 4) * VE + (long long)v * VE, VE));
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[i] = EL::up(own.e[i]);
            first = false;
          }
          for (int j = 0; j < mine; ++j) {
            const PackT in = *reinterpret_cast<const PackT*>(slot + j * CH + (v << 4));
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[i] = first ? EL::up(in.e[i]) : Op::apply(acc[i], EL::up(in.e[i]));
            first = false;
          }
          return !first;  // false: this CTA had nothing to contribute
        };
        A keep[VE], give[VE];
        bool have_keep = false, have_give = false;
        const int vk = me == 0 ? ctid : ctid + half_n;  // the vector I finalise, the vector I hand over
        const int vg = me == 0 ? ctid + half_n : ctid;
        const bool k_ok = ctid < half_n && vk < nvec, g_ok = ctid < half_n && vg < nvec;
        if (k_ok) have_keep = partial(vk, keep);
        if (g_ok) have_give = partial(vg, give);
        __syncwarp();
        if (mine > 0 && lane == 0) mbar_arrive(R.empty + s);  // my ring slot may be refilled
        // hand my partial sums of the partner's half to the partner: slot ctid of ITS exchange buffer xb
        if (g_ok) {
          const uint32_t dst = mapa(smem_u32(R.xch + xb * kClXchBytes + ctid * 32), partner);
          float f[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) f[i] = 0.f;
          if (have_give) {
#pragma unroll
            for (int i = 0; i < VE; ++i) f[i] = __uint_as_float(*reinterpret_cast<const uint32_t*>(&give[i]));
          }
          st_cluster_v4(dst, f[0], f[1], f[2], f[3]);
          if (VE > 4) st_cluster_v4(dst + 16, f[4], f[5], f[6], f[7]);
          // whether the partner may use them ("I had sources") is the same decision on both sides: mine_partner > 0
        }
        mbar_arrive_remote(mapa(smem_u32(R.xbar + xb), partner));  // release.cluster: my DSMEM stores are visible first
        // wait for the partner's partial sums of MY half
        while (!mbar_try_wait_cluster(R.xbar + xb, (xit >> 1) & 1)) {
          if (*(volatile const int*)c.abort_flag) break;
        }
        if (k_ok) {
          const int partner_mine = (nsrc - (int)partner + 1) / 2;
          const bool partner_has = partner_mine > 0 || (!own_raw && partner == 0);
          if (partner_has) {
            A other[VE];
            const uint32_t* px = reinterpret_cast<const uint32_t*>(R.xch + xb * kClXchBytes + ctid * 32);
#pragma unroll
            for (int i = 0; i < VE; ++i) *reinterpret_cast<uint32_t*>(&other[i]) = px[i];
            // fixed order on both CTAs: (CTA 0's partial) op (CTA 1's partial)
#pragma unroll
            for (int i = 0; i < VE; ++i) {
              if (!have_keep) keep[i] = other[i];
              else keep[i] = me == 0 ? Op::apply(keep[i], other[i]) : Op::apply(other[i], keep[i]);
            }
          }
          if (final_stage) c.scale_acc(keep);
          PackT out;
#pragma unroll
          for (int i = 0; i < VE; ++i) out.e[i] = EL::down(keep[i]);
          *reinterpret_cast<PackT*>(wr + base_off + o + (vk << 4)) = out;
        }
      }
    }
    xit_ = xit;
  }
  // producer lanes keep xit_ in step (they take no part in the exchange, but the counter is per thread)
  if (warp == 0) {
    uint32_t xit = xit_;
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      const int bytes = n * (int)sizeof(T);
      for (int o = 0; o < bytes; o += CH) ++xit;
    }
    xit_ = xit;
  }
  c.pipe_it = it;
}

// ---- all-gather stage k, tile t: the chunk list is dealt out alternately to the two CTAs -------------------------
template <typename T, typename IO, class Op, int VE, class Cta>
__device__ __forceinline__ void CudaClusterBE::ag_tile_tma(Cta& c, int k, int t) {
  using PackT = Pack<T, VE>;
  const ClusterRing& R = cr_;
  const Plan& P = c.a.plan;
  const int w = P.width[k], g = P.gap[k], gn = P.gap[k + 1], base = stage_base(P, c.r, k), nb = P.nranks / gn;
  const bool last = (k == 0);
  const int my_digit = (c.r - base) / g;
  const int rot = c.a.nlocal > 1 ? 0 : my_digit + 1;
  const bool self_needed = last && !(c.a.mode == MODE_REGISTERED && Cta::kRawIO);
  const int CH = kClSlotBytes & ~1023;
  char* wr = c.rc.win[c.r] + c.woff;
  IO* y = (IO*)c.rc.recv;
  const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
  const int ctid = (int)threadIdx.x - 32, nct = (int)blockDim.x - 32;
  const int me = (int)crank_;
  uint32_t it = c.pipe_it;
  uint32_t deal = 0;  // position in the chunk list (identical in both CTAs)
  if (warp == 0 && lane == 0) fence_proxy_async();
  for (int bi = 0; bi < nb; ++bi) {
    for (int s_ = 0; s_ < w; ++s_) {
      const int p = base + ((rot + s_) % w) * g;
      if (p == c.r && !self_needed) continue;
      long long start;
      const int n = c.tile_valid32((p % gn) + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      const char* src = c.rc.win[p] + c.woff + base_off;
      for (int o = 0; o < bytes; o += CH, ++deal) {
        if ((int)(deal & 1) != me) continue;
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kClStages);
        if (warp == 0) {
          if (lane == 0) {
            mbar_wait(R.empty + s, ((it / kClStages) & 1) ^ 1, c.abort_flag);
            mbar_expect_tx(R.full + s, (uint32_t)len);
            tma_load_1d(R.ring + s * kClSlotBytes, src + o, (uint32_t)len, R.full + s);
          }
        } else {
          mbar_wait(R.full + s, (it / kClStages) & 1, c.abort_flag);
          const char* slot = R.ring + s * kClSlotBytes;
          const int nvec = len >> 4;
          for (int v = ctid; v < nvec; v += nct) {
            const PackT in = *reinterpret_cast<const PackT*>(slot + (v << 4));
            if (last) Cta::st_user(y, start + (long long)((o >> 4) + v) * VE, c.wire_to_io(in), VE);
            else *reinterpret_cast<PackT*>(wr + base_off + o + (v << 4)) = in;
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(R.empty + s);
        }
        ++it;
      }
    }
  }
  c.pipe_it = it;
}

}  // namespace ft
