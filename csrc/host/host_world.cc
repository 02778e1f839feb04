// In-process CPU world: N logical ranks, each "CTA" a std::thread, running the very same algorithm
// templates as the CUDA kernels (core/staged_core.h) through CpuBE. This is the test vehicle for the
// schedule / hazard / flag-protocol logic on machines without a GPU, and a functional CPU data plane
// in its own right (the reference is a CPU library: allreduce_over_mpi/mpi_mod.hpp:953-1221).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <thread>
#include <vector>

#include "../core/dispatch.h"
#include "../core/staged_core.h"
#include "../planner/cost_model.h"
#include "cpu_backend.h"

namespace ft {

struct HostWorld {
  int n = 0;
  size_t win_bytes = 0;   // W: one window; segment = [stage0 | stage1 | registered]
  size_t ll_bytes = 0;
  std::vector<char*> seg;
  std::vector<uint64_t*> pad;
  std::vector<char*> ll;
  std::vector<CommDev> cd;
  std::vector<uint64_t> epoch;
  std::vector<unsigned> done, err;
  std::vector<char*> win_table;  // same table for every rank (one address space)
};

bool host_run_group(int group, int dtype, int io_dtype, int op, bool vec, const StagedArgs* a, int algo);

}  // namespace ft

using namespace ft;

extern "C" {

void* ft_host_world_create(int nranks, size_t window_bytes, size_t ll_bytes) {
  if (nranks < 1 || nranks > FT_MAX_LOCAL) return nullptr;
  auto* w = new HostWorld;
  w->n = nranks;
  w->win_bytes = (window_bytes + 4095) & ~(size_t)4095;
  w->ll_bytes = (ll_bytes + 4095) & ~(size_t)4095;
  w->epoch.assign(nranks, 0);
  w->done.assign(nranks, 0);
  w->err.assign(nranks, 0);
  w->cd.resize(nranks);
  for (int r = 0; r < nranks; ++r) {
    w->seg.push_back((char*)aligned_alloc(4096, 3 * w->win_bytes + 4096));
    memset(w->seg.back(), 0xA5, 3 * w->win_bytes + 4096);  // poison: nothing may rely on zero fill
    w->pad.push_back((uint64_t*)aligned_alloc(4096, FT_PAD_BYTES));
    memset(w->pad.back(), 0, FT_PAD_BYTES);
    w->ll.push_back((char*)aligned_alloc(4096, 2 * w->ll_bytes + 4096));
    memset(w->ll.back(), 0, 2 * w->ll_bytes + 4096);
  }
  w->win_table = w->seg;
  for (int r = 0; r < nranks; ++r) {
    CommDev& c = w->cd[r];
    memset(&c, 0, sizeof c);
    c.rank = r;
    c.nranks = nranks;
    for (int p = 0; p < nranks; ++p) {
      c.pad[p] = w->pad[p];
      c.ll[p] = w->ll[p];
    }
    c.epoch = &w->epoch[r];
    c.done = &w->done[r];
    c.err = &w->err[r];
    c.ll_bytes = w->ll_bytes;
    c.trace = nullptr;
  }
  return w;
}

void ft_host_world_destroy(void* h) {
  auto* w = (HostWorld*)h;
  if (!w) return;
  for (auto p : w->seg) free(p);
  for (auto p : w->pad) free(p);
  for (auto p : w->ll) free(p);
  delete w;
}

// Pointer to the registered (zero-copy) region of rank r: use it as send==recv with mode=1.
void* ft_host_registered(void* h, int rank) {
  auto* w = (HostWorld*)h;
  return w->seg[rank] + 2 * w->win_bytes;
}

unsigned ft_host_error(void* h, int rank) { return ((HostWorld*)h)->err[rank]; }
void ft_host_clear_error(void* h) { for (auto& e : ((HostWorld*)h)->err) e = 0; }
unsigned long long ft_host_epoch(void* h, int rank) { return ((HostWorld*)h)->epoch[rank]; }

// Runs one collective over all logical ranks. Returns 0, or a negative argument error, or the first
// nonzero device error word.
long long ft_host_allreduce(void* h, const void* const* send, void* const* recv, long long count,
                            int dtype, int io_dtype, int op, int algo, const int* widths, int nstages,
                            int ctas, long long tile_bytes, int mode, double scale, int apply_scale,
                            unsigned long long timeout_ns, int skip_rank, int lonely, int phases) {
  auto* w = (HostWorld*)h;
  StagedArgs a;
  memset(&a, 0, sizeof a);
  const bool want_bulk = algo == ALGO_STAGED_TMA;  // staged stages with the CPU's bulk tiles (cpu_backend.h)
  if (want_bulk) algo = ALGO_STAGED;
  if (algo == ALGO_RING || algo == ALGO_NVLS || algo == ALGO_ONESHOT || algo == ALGO_TWOSHOT) {
    int flat = w->n;
    if (!make_plan(&flat, 1, &a.plan)) return -1;
  } else {
    if (!make_plan(widths, nstages, &a.plan)) return -1;
  }
  if (lonely < 0 || a.plan.nranks + lonely != w->n) return -2;
  if (lonely > 0 && (algo != ALGO_STAGED || mode != MODE_STAGING)) return -8;
  a.lonely = lonely;
  a.phases = phases;
  a.nvls_reducers = ctas >= 4 ? ctas / 4 : 0;  // small CTA counts exercise the one-role (skewed) variant
  if (phases != PHASES_ALLREDUCE && (algo != ALGO_STAGED || lonely > 0 || mode != MODE_STAGING || count % w->n)) return -10;
  if (dtype == DT_FP8_BLOCK && (algo != ALGO_STAGED || lonely > 0 || mode != MODE_STAGING)) return -9;
  const int es = dtype_size(dtype), ios = dtype_size(io_dtype);
  if (!es || !ios) return -3;
  bool vec = true;
  for (int r = 0; r < w->n; ++r)
    vec = vec && ((uintptr_t)send[r] % 16 == 0) && ((uintptr_t)recv[r] % 16 == 0);
  if (phases != PHASES_ALLREDUCE && ((count / w->n) * es) % 16) vec = false;  // blocks must not be padded
  a.lay = make_layout(count, a.plan.nranks, es, tile_bytes, vec);
  // same rule as the TMA kernel on the GPU: vector path, whole 16-byte vectors, plain wire, no lonely ranks
  a.use_tma = want_bulk && vec && lonely == 0 && dtype != DT_FP8_BLOCK && count % (16 / es) == 0;
  a.scale_plane_off = ((long long)a.lay.block * a.plan.nranks * es + 255) & ~255LL;
  if ((size_t)(a.scale_plane_off + (a.lay.block * a.plan.nranks) / 16 + 16) > w->win_bytes) return -4;
  if (algo == ALGO_ONESHOT) {
    size_t region = (w->ll_bytes / w->n) & ~(size_t)15;
    size_t lines = (size_t)((count * es + 7) / 8);
    if (lines * 16 > region) return -5;
  }
  if (algo == ALGO_TWOSHOT) {  // [shot 1 | shot 2] x N regions of ceil(lines / N) lines per parity half
    const size_t region2 = (w->ll_bytes / (size_t)(2 * w->n)) & ~(size_t)15;
    const size_t lines2 = (size_t)((count * es + 7) / 8), per_block = (lines2 + (size_t)w->n - 1) / (size_t)w->n;
    if (per_block * 16 > region2) return -5;
  }
  if (algo == ALGO_NVLS && mode == MODE_STAGING && a.lay.ntiles > FT_MAX_CTAS) return -12;  // per-tile flag slots
  if (ctas < 1 || ctas > FT_MAX_CTAS) return -6;
  a.nlocal = w->n;
  a.ctas = ctas;
  a.mode = mode;
  a.apply_scale = apply_scale;
  a.scale = scale;
  a.win_off0 = mode == MODE_REGISTERED ? 2 * (long long)w->win_bytes : 0;
  a.win_stride = mode == MODE_REGISTERED ? 0 : (long long)w->win_bytes;
  a.timeout_ns = timeout_ns;
  for (int r = 0; r < w->n; ++r) {
    a.rc[r].comm = &w->cd[r];
    a.rc[r].send = send[r];
    a.rc[r].recv = recv[r];
    a.rc[r].win = w->win_table.data();
    a.rc[r].mc = nullptr;
  }
  if (skip_rank >= 0) {
    // fault injection: rank `skip_rank` never shows up; the others must time out, not hang
    StagedArgs b = a;
    int j = 0;
    for (int r = 0; r < w->n; ++r)
      if (r != skip_rank) b.rc[j++] = a.rc[r];
    b.nlocal = w->n - 1;
    a = b;
  }
  bool ok = host_run_group(dtype_group(dtype), dtype, io_dtype, op, vec, &a, a.use_tma ? (int)ALGO_STAGED_TMA : algo);
  if (!ok) return -7;
  for (int r = 0; r < w->n; ++r)
    if (w->err[r]) return (long long)w->err[r];
  return 0;
}

// conversions exported for the numerics tests
float ft_bf16_to_float(uint16_t v) { bf16_t b{v}; return to_float(b); }
uint16_t ft_float_to_bf16(float f) { return bf16_from_float(f).v; }
float ft_f16_to_float(uint16_t v) { f16_t b{v}; return to_float(b); }
uint16_t ft_float_to_f16(float f) { return f16_from_float(f).v; }
float ft_e4m3_to_float(unsigned code) { return e4m3_to_float((uint8_t)code); }
unsigned ft_float_to_e4m3(float f) { return float_to_e4m3(f); }
unsigned ft_fp8_block_scale(float amax) { return fp8_block_scale(amax); }

}  // extern "C"
