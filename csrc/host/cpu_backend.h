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

struct CpuBE {
  static constexpr bool kTma = false;  // the TMA-staged tiles exist only on the GPU
  template <typename T, typename IO, class Op, int VE, class Cta> void rs_tile_tma(Cta&, int, int) {}
  template <typename T, typename IO, class Op, int VE, class Cta> void ag_tile_tma(Cta&, int, int) {}
  int abort_ = 0;
  int tid() const { return 0; }
  int nthr() const { return 1; }
  void sync() {}
  template <int B> static void ld(void* dst, const void* src, bool /*remote*/) { memcpy(dst, src, B); }
  template <int B> static void st(void* dst, const void* src) { memcpy(dst, src, B); }
  uint64_t flag_ld(const uint64_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
  void flag_st(uint64_t* p, uint64_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
  void fence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
  void acquire(const uint64_t* p) { (void)__atomic_load_n(p, __ATOMIC_ACQUIRE); }
  uint64_t load_epoch(const uint64_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
  unsigned err_ld(const unsigned* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
  void err_set(unsigned* p, unsigned v) {
    unsigned expect = 0;
    __atomic_compare_exchange_n(p, &expect, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED);
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
  template <typename T, class Op, int VE>
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
  template <typename T, int VE>
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
