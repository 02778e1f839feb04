// TMA-staged tiles of the P2P stages (CUDA only; the LDG versions live in core/staged_core.h).
//
// Why: the register-staged pull loops keep at most threads x 8 x 16 B = 64 KiB in flight per SM, so their
// bandwidth grows with the number of SMs spent on communication (measured: 32 CTAs 157 GB/s, 64 CTAs
// 303 GB/s, 148 CTAs 498 GB/s at 2 GPUs). Here ONE elected thread streams the peers' chunks into a
// shared-memory ring with cp.async.bulk (SASS UBLKCP.S.G) tracked by mbarriers, 15 consumer warps
// reduce from shared memory and store the owner's slice: up to (kStages-1) x 48 KiB are in flight per SM
// without holding a single register, and loads, adds and stores overlap.
//
// Ordering: the tile's flag wait (wait_group) has already run for the whole CTA. Data written by peers
// with generic-proxy stores is read here by the async proxy, so the producer issues fence.proxy.async
// after the acquire and before its first bulk copy. Ring slots are recycled through full/empty
// mbarriers (consumers arrive after their last shared-memory read, the producer waits before the next
// cp.async.bulk into the slot). The reference has no counterpart (CPU memcpy inside MPI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../core/staged_core.h"

namespace ft {

constexpr int kTmaStages = 4;
constexpr int kTmaSlotBytes = 48 << 10;
constexpr int kTmaRingBytes = kTmaStages * kTmaSlotBytes;
constexpr int kTmaSmemBytes = kTmaRingBytes + 128;  // + full[4] + empty[4] mbarriers
constexpr int kTmaMaxChunk = 32 << 10;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* b, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
      : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity, const int* abort_flag) {
  while (!mbar_try_wait(b, parity)) {
    if (*(volatile const int*)abort_flag) break;  // a flag watchdog fired somewhere: do not hang on the ring
  }
}
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void fence_mbarrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

struct TmaRing {
  char* ring;
  uint64_t* full;
  uint64_t* empty;
};

// Called by every thread once, before the first tile. Warp 0 is the producer; warps 1.. consume.
__device__ __forceinline__ void tma_ring_init(const TmaRing& R, int consumer_warps) {
  if (threadIdx.x == 0) {
    for (int i = 0; i < kTmaStages; ++i) {
      mbar_init(R.full + i, 1);
      mbar_init(R.empty + i, consumer_warps);
    }
    fence_mbarrier_init();
  }
  __syncthreads();
}

// ---- reduce-scatter stage k, tile t ---------------------------------------------------------------------
template <typename T, typename IO, class Op, int VE, class Cta>
__device__ __forceinline__ void rs_tile_tma_impl(Cta& c, const TmaRing& R, int k, int t) {
  using EL = Elem<T>;
  using A = typename EL::acc;
  using PackT = Pack<T, VE>;
  const Plan& P = c.a.plan;
  const int w = P.width[k], gn = P.gap[k + 1], nb = P.nranks / gn, rmod = c.r % gn, npeers = w - 1;
  const bool from_user = (k == 0) && (c.a.mode == MODE_STAGING);
  const bool final_stage = (k == P.nstages - 1) && c.a.lonely == 0;
  // my own contribution can ride the ring too when it already sits in memory in wire format
  const bool own_raw = !from_user || Cta::kRawIO;
  const int nsrc = npeers + (own_raw ? 1 : 0);
  int CH = (kTmaSlotBytes / nsrc) & ~1023;
  if (CH > kTmaMaxChunk) CH = kTmaMaxChunk;
  const char* own_base = from_user ? (const char*)c.rc.send : (const char*)(c.rc.win[c.r] + c.woff);
  char* wr = c.rc.win[c.r] + c.woff;
  const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
  uint32_t it = c.pipe_it;

  if (warp == 0) {
    if (lane == 0) fence_proxy_async();
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      for (int o = 0; o < bytes; o += CH, ++it) {
        if (lane != 0) continue;
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kTmaStages);
        mbar_wait(R.empty + s, ((it / kTmaStages) & 1) ^ 1, c.abort_flag);
        mbar_expect_tx(R.full + s, (uint32_t)(nsrc * len));
        char* slot = R.ring + s * kTmaSlotBytes;
        if (own_raw) tma_load_1d(slot, own_base + base_off + o, (uint32_t)len, R.full + s);
        for (int j = 0; j < npeers; ++j)
          tma_load_1d(slot + (j + (own_raw ? 1 : 0)) * CH, c.rs_peer_ptr(k, j) + base_off + o, (uint32_t)len, R.full + s);
      }
    }
  } else {
    const int ctid = (int)threadIdx.x - 32, nct = (int)blockDim.x - 32;
    const IO* x = (const IO*)c.rc.send;
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      for (int o = 0; o < bytes; o += CH, ++it) {
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kTmaStages);
        mbar_wait(R.full + s, (it / kTmaStages) & 1, c.abort_flag);
        const char* slot = R.ring + s * kTmaSlotBytes;
        const int nvec = len >> 4;
        for (int v = ctid; v < nvec; v += nct) {
          A acc[VE];
          if (own_raw) {
            const PackT own = *reinterpret_cast<const PackT*>(slot + (v << 4));
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[i] = EL::up(own.e[i]);
          } else {
            const PackT own = Cta::io_to_wire(Cta::ld_user(x, start + (o >> 4) * VE + (long long)v * VE, VE));
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[i] = EL::up(own.e[i]);
          }
          for (int j = 0; j < npeers; ++j) {
            const PackT in = *reinterpret_cast<const PackT*>(slot + (j + (own_raw ? 1 : 0)) * CH + (v << 4));
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[i] = Op::apply(acc[i], EL::up(in.e[i]));
          }
          if (final_stage) c.scale_acc(acc);
          PackT out;
#pragma unroll
          for (int i = 0; i < VE; ++i) out.e[i] = EL::down(acc[i]);
          *reinterpret_cast<PackT*>(wr + base_off + o + (v << 4)) = out;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(R.empty + s);
      }
    }
  }
  c.pipe_it = it;
}

// ---- all-gather stage k, tile t ------------------------------------------------------------------------------
template <typename T, typename IO, class Op, int VE, class Cta>
__device__ __forceinline__ void ag_tile_tma_impl(Cta& c, const TmaRing& R, int k, int t) {
  using EL = Elem<T>;
  using PackT = Pack<T, VE>;
  const Plan& P = c.a.plan;
  const int w = P.width[k], g = P.gap[k], gn = P.gap[k + 1], base = stage_base(P, c.r, k), nb = P.nranks / gn;
  const bool last = (k == 0);
  const int my_digit = (c.r - base) / g;
  // Several logical ranks on ONE GPU (emulation): there are no links to balance, and walking the sources in the
  // same order on every rank lets the readers of a block hit in L2 behind the first one.
  const int rot = c.a.nlocal > 1 ? 0 : my_digit + 1;
  const bool self_needed = last && !(c.a.mode == MODE_REGISTERED && Cta::kRawIO);
  const int CH = kTmaMaxChunk;
  char* wr = c.rc.win[c.r] + c.woff;
  IO* y = (IO*)c.rc.recv;
  const int warp = (int)threadIdx.x >> 5, lane = (int)threadIdx.x & 31;
  const int ctid = (int)threadIdx.x - 32, nct = (int)blockDim.x - 32;
  uint32_t it = c.pipe_it;
  if (warp == 0 && lane == 0) fence_proxy_async();

  for (int bi = 0; bi < nb; ++bi) {
    for (int s_ = 0; s_ < w; ++s_) {
      // Start from my right-hand neighbour in the group and go round: if every rank walked the sources in
      // the same order, all of them would pull from the same GPU at the same time (measured at N=4: flat AG
      // 852 us vs 640 us for the 2x2 plan at 256 MiB before this rotation).
      const int p = base + ((rot + s_) % w) * g;
      if (p == c.r && !self_needed) continue;
      long long start;
      const int n = c.tile_valid32((p % gn) + bi * gn, t, &start);
      const long long base_off = start * (long long)sizeof(T);
      const int bytes = n * (int)sizeof(T);
      const char* src = c.rc.win[p] + c.woff + base_off;
      for (int o = 0; o < bytes; o += CH, ++it) {
        const int len = bytes - o < CH ? bytes - o : CH;
        const int s = (int)(it % kTmaStages);
        if (warp == 0) {
          if (lane == 0) {
            mbar_wait(R.empty + s, ((it / kTmaStages) & 1) ^ 1, c.abort_flag);
            mbar_expect_tx(R.full + s, (uint32_t)len);
            tma_load_1d(R.ring + s * kTmaSlotBytes, src + o, (uint32_t)len, R.full + s);
          }
        } else {
          mbar_wait(R.full + s, (it / kTmaStages) & 1, c.abort_flag);
          const char* slot = R.ring + s * kTmaSlotBytes;
          const int nvec = len >> 4;
          for (int v = ctid; v < nvec; v += nct) {
            const PackT in = *reinterpret_cast<const PackT*>(slot + (v << 4));
            if (last) Cta::st_user(y, start + (long long)((o >> 4) + v) * VE, c.wire_to_io(in), VE);
            else *reinterpret_cast<PackT*>(wr + base_off + o + (v << 4)) = in;
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(R.empty + s);
        }
      }
    }
  }
  c.pipe_it = it;
}

}  // namespace ft
