// CPU model of one CTA: a single thread with the same memory-ordering discipline as the kernel
// (relaxed data accesses, release/acquire flags). Lets the unit tests - and ThreadSanitizer - exercise
// the exact algorithm text of core/staged_core.h on a machine with no GPU.
#pragma once
#include <sched.h>
#include <stdint.h>
#include <string.h>
#include <time.h>

#include "../core/comm_dev.h"
#include "../core/types.h"

namespace ft {

// Stress knob (tests only): with probability permille/1000 a thread sleeps 0-200 us before publishing a flag, after
// acquiring one and at every CTA-wide sync - the schedule perturbation that exposes a missing wait, a buffer reused
// too early or an exit barrier that lets a slow reader be overrun (SURVEY 5.2 "randomized delays").
inline int& host_jitter_permille() {
  static int v = 0;
  return v;
}
inline void host_jitter() {
  const int pm = host_jitter_permille();
  if (pm <= 0) return;
  static thread_local uint64_t x = 0x9E3779B97F4A7C15ull ^ (uint64_t)(uintptr_t)&x;
  x ^= x << 13;
  x ^= x >> 7;
  x ^= x << 17;
  if ((int)(x % 1000) < pm) {
    timespec ts = {0, (long)((x >> 20) % 200000)};
    nanosleep(&ts, nullptr);
  }
}

struct CpuBE {
  static constexpr bool kTma = false;  // no cp.async.bulk here ...
  // ... but the same hook (StagedCta::run<true>) has a CPU meaning: "bulk" tiles. The GPU text gives one
  // 16-byte vector to one thread and pays an index computation per vector, which is free next to an NVLink
  // load and very much not free on a CPU core. The bulk tiles walk each contiguous block-of-a-tile with plain
  // loops over kBulk elements that the compiler vectorises: same flags, same groups, same summation order
  // (own first, peers ascending), same rounding points - results are bit-identical to the per-vector text
  // (tests/test_host_world.py::test_bulk_*). Selected by the host with algo 5, like the TMA kernel on the GPU.
  static constexpr bool kBulkCopyIn = true;
  static constexpr int kBulk = 1024;

  template <typename T, typename IO, class Op, int VE, class Cta> void copy_in_tile_bulk(Cta& c, int t) {
    using EL = Elem<T>;
    using EU = Elem<IO>;
    using A = typename EL::acc;
    const int N = c.a.plan.nranks, g1 = c.a.plan.gap[1];
    char* wr = c.rc.win[c.r] + c.woff;
    for (int b = 0; b < N; ++b) {
      if ((b % g1) == (c.r % g1)) continue;  // blocks I own after stage 0 are read straight from `send`
      long long start;
      const int n = c.tile_valid32(b, t, &start);
      if (n <= 0) continue;
      const IO* __restrict s = (const IO*)c.rc.send + start;
      T* __restrict d = (T*)wr + start;
      if constexpr (Cta::kRawIO) {
        memcpy(d, s, (size_t)n * sizeof(T));
      } else {
        for (int i = 0; i < n; ++i) d[i] = EL::down((A)EU::up(s[i]));
      }
    }
  }

  template <typename T, typename IO, class Op, int VE, class Cta> void rs_tile_tma(Cta& c, int k, int t) {
    using EL = Elem<T>;
    using EU = Elem<IO>;
    using A = typename EL::acc;
    const Plan& P = c.a.plan;
    const int w = P.width[k], gn = P.gap[k + 1], nb = P.nranks / gn, rmod = c.r % gn, npeers = w - 1;
    const bool from_user = (k == 0) && (c.a.mode == MODE_STAGING);
    const bool final_stage = (k == P.nstages - 1) && c.a.lonely == 0;
    char* wr = c.rc.win[c.r] + c.woff;
    const char* peer[FT_MAX_WIDTH];
    for (int j = 0; j < npeers; ++j) peer[j] = c.rs_peer_ptr(k, j);
    A acc[kBulk];
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = c.tile_valid32(rmod + bi * gn, t, &start);
      for (int o = 0; o < n; o += kBulk) {
        const int len = n - o < kBulk ? n - o : kBulk;
        const long long el = start + o;
        if (from_user) {
          const IO* __restrict s = (const IO*)c.rc.send + el;
          if constexpr (Cta::kRawIO) {
            for (int i = 0; i < len; ++i) acc[i] = EL::up(s[i]);
          } else {  // the own contribution is rounded to the wire type first, exactly like io_to_wire
            for (int i = 0; i < len; ++i) acc[i] = EL::up(EL::down((A)EU::up(s[i])));
          }
        } else {
          const T* __restrict s = (const T*)wr + el;
          for (int i = 0; i < len; ++i) acc[i] = EL::up(s[i]);
        }
        for (int j = 0; j < npeers; ++j) {
          const T* __restrict s = (const T*)peer[j] + el;
          for (int i = 0; i < len; ++i) acc[i] = Op::apply(acc[i], EL::up(s[i]));
        }
        if constexpr (EL::is_float) {
          if (final_stage && c.a.apply_scale) {
            const A sc = (A)c.a.scale;
            for (int i = 0; i < len; ++i) acc[i] = (A)(acc[i] * sc);
          }
        }
        T* __restrict d = (T*)wr + el;
        for (int i = 0; i < len; ++i) d[i] = EL::down(acc[i]);
      }
    }
  }

  template <typename T, typename IO, class Op, int VE, class Cta> void ag_tile_tma(Cta& c, int k, int t) {
    using EL = Elem<T>;
    using EU = Elem<IO>;
    const Plan& P = c.a.plan;
    const int w = P.width[k], g = P.gap[k], gn = P.gap[k + 1], base = stage_base(P, c.r, k), nb = P.nranks / gn;
    const bool last = (k == 0);
    const bool self_needed = last && !(c.a.mode == MODE_REGISTERED && Cta::kRawIO);
    const int rot = (c.r - base) / g + 1;  // start at my right-hand neighbour: spreads the readers over the sources
    char* wr = c.rc.win[c.r] + c.woff;
    for (int bi = 0; bi < nb; ++bi) {
      for (int s_ = 0; s_ < w; ++s_) {
        const int p = base + ((rot + s_) % w) * g;
        if (p == c.r && !self_needed) continue;
        long long start;
        const int n = c.tile_valid32((p % gn) + bi * gn, t, &start);
        if (n <= 0) continue;
        const T* __restrict s = (const T*)(c.rc.win[p] + c.woff) + start;
        if (!last) {
          memcpy((T*)wr + start, s, (size_t)n * sizeof(T));
        } else {
          IO* __restrict y = (IO*)c.rc.recv + start;
          if constexpr (Cta::kRawIO && !EL::is_float) {
            memcpy(y, s, (size_t)n * sizeof(T));
          } else {  // same conversions as wire_to_io (the scale was applied when the block became final)
            for (int i = 0; i < n; ++i) y[i] = EU::down((typename EU::acc)EL::up(s[i]));
          }
        }
      }
    }
  }
  int abort_ = 0;
  int tid() const { return 0; }
  int nthr() const { return 1; }
  void sync() { host_jitter(); }
  template <int B> static void ld(void* dst, const void* src, bool /*remote*/) { memcpy(dst, src, B); }
  template <int B> static void st(void* dst, const void* src) { memcpy(dst, src, B); }
  uint64_t flag_ld(const uint64_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
  void flag_st(uint64_t* p, uint64_t v) {
    host_jitter();
    __atomic_store_n(p, v, __ATOMIC_RELEASE);
  }
  void fence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
  void acquire(const uint64_t* p) {
    (void)__atomic_load_n(p, __ATOMIC_ACQUIRE);
    host_jitter();
  }
  uint64_t load_epoch(const uint64_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
  unsigned err_ld(const unsigned* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
  void err_set(const CommDev& cd, unsigned v) {
    unsigned expect = 0;
    if (__atomic_compare_exchange_n(cd.err, &expect, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED) && cd.err_host)
      __atomic_store_n(cd.err_host, v, __ATOMIC_RELEASE);
  }
  uint64_t now_ns() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (uint64_t)ts.tv_sec * 1000000000ull + (uint64_t)ts.tv_nsec;
  }
  void pause() { sched_yield(); }
  int* abort_ptr() { return &abort_; }
  void finish(const CommDev& cd, int ctas) {
    uint64_t epoch = __atomic_load_n(cd.epoch, __ATOMIC_ACQUIRE);
    unsigned old = __atomic_fetch_add(cd.done, 1u, __ATOMIC_ACQ_REL);
    if (old == (unsigned)ctas - 1) {
      __atomic_store_n(cd.done, 0u, __ATOMIC_RELAXED);
      __atomic_store_n(cd.epoch, epoch + 1, __ATOMIC_RELEASE);
    }
  }
  // "switch" emulation: sum the N windows in rank order / write to all N windows
  template <typename T, class Op, int VE, bool EMU = false>
  Pack<T, VE> mc_ld_reduce(const RankCall& rc, long long off, int nranks) {
    using EL = Elem<T>;
    typename EL::acc acc[VE];
    for (int p = 0; p < nranks; ++p) {
      Pack<T, VE> v;
      memcpy(&v, rc.win[p] + off, sizeof(v));
      for (int i = 0; i < VE; ++i) acc[i] = p ? Op::apply(acc[i], EL::up(v.e[i])) : EL::up(v.e[i]);
    }
    Pack<T, VE> o;
    for (int i = 0; i < VE; ++i) o.e[i] = EL::down(acc[i]);
    return o;
  }
  template <typename T, int VE, bool EMU = false>
  void mc_st(const RankCall& rc, long long off, const Pack<T, VE>& v, int nranks) {
    for (int p = 0; p < nranks; ++p) memcpy(rc.win[p] + off, &v, sizeof(v));
  }
  // LL lines: two 8-byte halves {data, flag}, each half single-copy atomic
  void ll_st(char* line, uint32_t d0, uint32_t d1, uint32_t tag) {
    uint64_t* q = (uint64_t*)line;
    __atomic_store_n(q, (uint64_t)d0 | ((uint64_t)tag << 32), __ATOMIC_RELAXED);
    __atomic_store_n(q + 1, (uint64_t)d1 | ((uint64_t)tag << 32), __ATOMIC_RELAXED);
  }
  bool ll_ld(const char* line, uint32_t tag, uint32_t* data) {
    const uint64_t* q = (const uint64_t*)line;
    uint64_t a = __atomic_load_n(q, __ATOMIC_RELAXED), b = __atomic_load_n(q + 1, __ATOMIC_RELAXED);
    if ((uint32_t)(a >> 32) != tag || (uint32_t)(b >> 32) != tag) return false;
    data[0] = (uint32_t)a;
    data[1] = (uint32_t)b;
    return true;
  }
};

}  // namespace ft
