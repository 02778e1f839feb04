// Multi-process CPU communicator: the same algorithm templates as the GPU kernels, one process per rank,
// windows / flag pads / LL buffers in POSIX shared memory that every rank maps. This is the CPU data
// plane across processes - what the reference is (an MPI_Allreduce over point-to-point messages between
// CPU processes, allreduce_over_mpi/mpi_mod.hpp:1167-1221) without needing an MPI library: the bootstrap is
// the same byte all-gather callback the GPU communicator uses (torch.distributed/gloo in the tests).
#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/time.h>
#include <time.h>
#include <unistd.h>

#include <limits>
#include <memory>
#include <string>
#include <vector>

#include "../../include/flextree.h"
#include "../../include/flextree_host.h"
#include "../core/dispatch.h"
#include "../core/staged_core.h"
#include "../planner/cost_model.h"

namespace ft {
void set_last_error(const std::string& s);
bool host_run_group(int group, int dtype, int io_dtype, int op, bool vec, const StagedArgs* a, int algo);
}  // namespace ft
using namespace ft;

struct ft_host_comm {
  int rank = 0, nranks = 1;
  ft_allgather_fn ag = nullptr;
  void* ctx = nullptr;
  size_t win_bytes = 0, ll_bytes = 0, seg_bytes = 0, total_bytes = 0;
  std::vector<char*> maps;       // whole shared mapping of every rank
  std::vector<char*> win_table;  // window segment base of every rank
  CommDev cd;
  uint64_t epoch = 0, calls = 0;
  size_t board_off = 0;  // agreement board: 2 parities x FT_MAX_RANKS 64-bit slots in the spare page behind the windows
  unsigned done = 0, err = 0;
  // "-k" trees (cost_model/ChooseWidth.h:16-22 lists the trees of N+1 as candidates for N ranks): a tree with k more
  // positions than there are processes. Phantom j (logical rank N+j) is hosted by process j % N: it has its own
  // windows / pad / LL in slot 1 + j / N of that process's segment, its own epoch, and runs as a second logical rank
  // of the host's call with the operation's identity as input.
  int phantoms = 0, slots = 1;
  size_t slot_bytes = 0;
  struct Phantom {
    CommDev cd;
    uint64_t epoch = 0;
    unsigned done = 0, err = 0;
  };
  std::vector<std::unique_ptr<Phantom>> hosted;  // phantoms this process runs, ascending j
  std::vector<int> hosted_id;                    // their j
  std::vector<char> ident, sink;                 // identity input / discarded output of the hosted phantoms
  long long ident_count = -1;
  int ident_dtype = -1, ident_op = -1;
  bool allgather(const void* s, void* r, size_t n) const {
    if (nranks == 1) { memcpy(r, s, n); return true; }
    return ag && ag(ctx, s, r, n) == 0;
  }
};

// Per-call agreement without leaving shared memory: one bit that is AND-ed over the ranks (here: "all my buffers
// are 16-byte aligned") and a 31-bit payload that must be the same on every rank (here: the offset of a registered
// buffer inside the region, in 16-byte units). Every rank owns two 64-bit slots per rank (one per call parity) in
// the spare page behind its windows; it publishes (call number << 32 | payload << 1 | bit) into this call's slot on
// EVERY rank and then waits until its own board shows exactly this call's number in every slot of that parity.
// Why two parities: a call that ends right after the agreement (an argument error every rank sees, e.g. mismatched
// registered offsets) has no data phase to hold a fast rank back, so that rank may already be publishing its NEXT
// call while a slow rank still reads this one - into the other parity. It cannot get two calls ahead: finishing the
// next agreement needs the slow rank's next publication. Replaces a bootstrap all-gather per call (a gloo round trip
// costs more than the whole collective for small messages). Call numbers are compared modulo 2^32.
// Returns 0 = timeout, 1 = agreed, 2 = payloads differ (every rank sees that).
static int agree(ft_host_comm* c, char* bit, uint32_t payload, unsigned long long timeout_ns) {
  const int N = c->nranks;
  const uint64_t seq = ++c->calls;
  const size_t par = (size_t)(seq & 1) * FT_MAX_RANKS;
  const uint64_t mine = (seq << 32) | ((uint64_t)(payload & 0x7fffffffu) << 1) | (uint64_t)(*bit ? 1 : 0);
  for (int p = 0; p < N; ++p)
    __atomic_store_n((uint64_t*)(c->maps[(size_t)p] + c->board_off) + par + c->rank, mine, __ATOMIC_RELEASE);
  const uint64_t* board = (const uint64_t*)(c->maps[(size_t)c->rank] + c->board_off) + par;
  timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  char all = 1;
  bool same = true;
  for (int p = 0; p < N; ++p) {
    uint64_t v;
    unsigned it = 0;
    while ((uint32_t)((v = __atomic_load_n(board + p, __ATOMIC_ACQUIRE)) >> 32) != (uint32_t)seq) {
      if (((++it) & 0x3ff) == 0 && timeout_ns) {
        timespec t1;
        clock_gettime(CLOCK_MONOTONIC, &t1);
        const unsigned long long dt = (unsigned long long)(t1.tv_sec - t0.tv_sec) * 1000000000ull + (unsigned long long)t1.tv_nsec - (unsigned long long)t0.tv_nsec;
        if (dt > timeout_ns) return 0;
      }
      sched_yield();
    }
    all = all && (v & 1);
    same = same && (((v >> 1) & 0x7fffffffu) == (payload & 0x7fffffffu));
  }
  *bit = all;
  return same ? 1 : 2;
}

// The operation's identity in the user's element type: what a phantom rank contributes.
template <typename T> static void fill_n(void* buf, long long n, T v) {
  T* p = (T*)buf;
  for (long long i = 0; i < n; ++i) p[i] = v;
}
static bool fill_identity(void* buf, long long n, int dtype, int op) {
  const size_t bytes = (size_t)n * (size_t)dtype_size(dtype);
  if (op == OP_SUM || op == OP_BOR) { memset(buf, 0, bytes); return true; }
  if (op == OP_BAND) { memset(buf, 0xFF, bytes); return true; }
  if (op != OP_MAX && op != OP_MIN) return false;
  const bool mx = op == OP_MAX;  // identity of MAX is the lowest value, of MIN the highest
  switch (dtype) {
    case DT_UINT8: fill_n<uint8_t>(buf, n, mx ? 0 : 0xFF); return true;
    case DT_BOOL: fill_n<uint8_t>(buf, n, mx ? 0 : 1); return true;
    case DT_INT8: fill_n<int8_t>(buf, n, mx ? INT8_MIN : INT8_MAX); return true;
    case DT_UINT16: fill_n<uint16_t>(buf, n, mx ? 0 : 0xFFFF); return true;
    case DT_INT16: fill_n<int16_t>(buf, n, mx ? INT16_MIN : INT16_MAX); return true;
    case DT_INT32: fill_n<int32_t>(buf, n, mx ? INT32_MIN : INT32_MAX); return true;
    case DT_UINT32: fill_n<uint32_t>(buf, n, mx ? 0u : 0xFFFFFFFFu); return true;
    case DT_INT64: case DT_LONG_LONG: fill_n<int64_t>(buf, n, mx ? INT64_MIN : INT64_MAX); return true;
    case DT_UINT64: fill_n<uint64_t>(buf, n, mx ? 0ull : ~0ull); return true;
    case DT_FLOAT: fill_n<float>(buf, n, mx ? -std::numeric_limits<float>::infinity() : std::numeric_limits<float>::infinity()); return true;
    case DT_DOUBLE: fill_n<double>(buf, n, mx ? -std::numeric_limits<double>::infinity() : std::numeric_limits<double>::infinity()); return true;
    case DT_HALF: fill_n<uint16_t>(buf, n, mx ? 0xFC00 : 0x7C00); return true;      // -inf / +inf
    case DT_BFLOAT16: fill_n<uint16_t>(buf, n, mx ? 0xFF80 : 0x7F80); return true;  // -inf / +inf
  }
  return false;
}

static std::string shm_name(unsigned long long token, int rank) {
  char buf[96];
  snprintf(buf, sizeof buf, "/flextree.%llx.%d", token, rank);
  return buf;
}

extern "C" {

ft_result_t ft_host_comm_create(ft_host_comm** out, int rank, int nranks, ft_allgather_fn ag, void* ctx,
                                size_t window_bytes, size_t ll_bytes) {
  return ft_host_comm_create_ex(out, rank, nranks, ag, ctx, window_bytes, ll_bytes, 0);
}

ft_result_t ft_host_comm_create_ex(ft_host_comm** out, int rank, int nranks, ft_allgather_fn ag, void* ctx,
                                   size_t window_bytes, size_t ll_bytes, int max_phantoms) {
  if (!out || nranks < 1 || nranks > FT_MAX_RANKS || rank < 0 || rank >= nranks || (nranks > 1 && !ag) ||
      max_phantoms < 0 || nranks + max_phantoms > FT_MAX_RANKS) {
    set_last_error("ft_host_comm_create: bad arguments");
    return FT_ERR_INVALID;
  }
  auto* c = new ft_host_comm;
  c->rank = rank;
  c->nranks = nranks;
  c->ag = ag;
  c->ctx = ctx;
  c->win_bytes = ((window_bytes ? window_bytes : (size_t)(4 << 20)) + 4095) & ~(size_t)4095;
  c->ll_bytes = ((ll_bytes ? ll_bytes : (size_t)(1 << 20)) + 4095) & ~(size_t)4095;
  c->seg_bytes = 3 * c->win_bytes + 4096;  // [parity 0 | parity 1 | registered]
  c->board_off = 3 * c->win_bytes;
  const size_t pad_off = c->seg_bytes, ll_off = pad_off + ((FT_PAD_BYTES + 4095) & ~(size_t)4095);
  c->slot_bytes = ll_off + 2 * c->ll_bytes + 4096;
  c->phantoms = max_phantoms;
  c->slots = 1 + (max_phantoms + nranks - 1) / nranks;  // slot 0 = the process itself, then the phantoms it may host
  c->total_bytes = c->slot_bytes * (size_t)c->slots;

  unsigned long long tok = 0;
  if (rank == 0) {
    timeval tv;
    gettimeofday(&tv, nullptr);
    tok = ((unsigned long long)getpid() << 32) ^ ((unsigned long long)tv.tv_sec << 20) ^ (unsigned long long)tv.tv_usec;
  }
  std::vector<unsigned long long> toks((size_t)nranks);
  if (!c->allgather(&tok, toks.data(), sizeof tok)) { set_last_error("bootstrap allgather failed"); delete c; return FT_ERR_INTERNAL; }
  tok = toks[0];

  // create + size + zero my segment
  const std::string mine = shm_name(tok, rank);
  int fd = shm_open(mine.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
  char ok = fd >= 0 && ftruncate(fd, (off_t)c->total_bytes) == 0;
  std::vector<char> oks((size_t)nranks);
  c->allgather(&ok, oks.data(), 1);  // also the "everybody has created its segment" barrier
  for (char o : oks) ok = ok && o;
  if (!ok) {
    if (fd >= 0) { close(fd); shm_unlink(mine.c_str()); }
    set_last_error(std::string("shm_open/ftruncate failed: ") + strerror(errno));
    delete c;
    return FT_ERR_NOMEM;
  }
  c->maps.assign((size_t)nranks, nullptr);
  ok = 1;
  for (int p = 0; p < nranks; ++p) {
    int pfd = p == rank ? fd : shm_open(shm_name(tok, p).c_str(), O_RDWR, 0600);
    void* m = pfd >= 0 ? mmap(nullptr, c->total_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, pfd, 0) : MAP_FAILED;
    if (pfd >= 0 && p != rank) close(pfd);
    if (m == MAP_FAILED) { ok = 0; break; }
    c->maps[(size_t)p] = (char*)m;
  }
  close(fd);
  c->allgather(&ok, oks.data(), 1);  // everybody has mapped everything: names can go
  shm_unlink(mine.c_str());
  for (char o : oks) ok = ok && o;
  if (!ok) { set_last_error("mapping a peer's shared segment failed"); delete c; return FT_ERR_NOMEM; }

  memset(&c->cd, 0, sizeof c->cd);
  c->cd.rank = rank;
  c->cd.nranks = nranks;
  // logical rank L: processes first, then phantom j = L - nranks in slot 1 + j / nranks of process j % nranks
  const int logical = nranks + max_phantoms;
  auto base_of = [&](int L) {
    return L < nranks ? c->maps[(size_t)L] : c->maps[(size_t)((L - nranks) % nranks)] + c->slot_bytes * (size_t)(1 + (L - nranks) / nranks);
  };
  c->win_table.resize((size_t)logical);
  for (int p = 0; p < logical; ++p) {
    c->win_table[(size_t)p] = base_of(p);
    c->cd.pad[p] = (uint64_t*)(base_of(p) + pad_off);
    c->cd.ll[p] = base_of(p) + ll_off;
  }
  for (int j = rank; j < max_phantoms; j += nranks) {
    std::unique_ptr<ft_host_comm::Phantom> ph(new ft_host_comm::Phantom);
    ph->cd = c->cd;  // same tables (rank / counters patched below, after c->cd is complete)
    c->hosted.push_back(std::move(ph));
    c->hosted_id.push_back(j);
  }
  c->cd.epoch = &c->epoch;
  c->cd.done = &c->done;
  c->cd.err = &c->err;
  c->cd.ll_bytes = c->ll_bytes;
  c->cd.trace = nullptr;
  for (size_t i = 0; i < c->hosted.size(); ++i) {
    ft_host_comm::Phantom& ph = *c->hosted[i];
    ph.cd = c->cd;
    ph.cd.rank = nranks + c->hosted_id[i];
    ph.cd.epoch = &ph.epoch;
    ph.cd.done = &ph.done;
    ph.cd.err = &ph.err;
  }
  *out = c;
  return FT_SUCCESS;
}

ft_result_t ft_host_comm_destroy(ft_host_comm* c) {
  if (!c) return FT_SUCCESS;
  char x = 0;
  std::vector<char> all((size_t)c->nranks);
  c->allgather(&x, all.data(), 1);  // nobody unmaps while a peer may still read
  for (char* m : c->maps)
    if (m) munmap(m, c->total_bytes);
  delete c;
  return FT_SUCCESS;
}

// Tests only: jump the per-call agreement counter (same value on every rank), e.g. to just below 2^32.
void ft_host_comm_debug_set_calls(ft_host_comm* c, unsigned long long calls) { c->calls = calls; }
int ft_host_comm_rank(const ft_host_comm* c) { return c->rank; }
int ft_host_comm_size(const ft_host_comm* c) { return c->nranks; }
unsigned ft_host_comm_error(const ft_host_comm* c) { return c->err; }
void* ft_host_comm_registered(ft_host_comm* c) { return c->win_table[(size_t)c->rank] + 2 * c->win_bytes; }
size_t ft_host_comm_window_bytes(const ft_host_comm* c) { return c->win_bytes; }

// One collective for this process's rank. Same argument meaning as ft_host_allreduce (host_world.cc).
long long ft_host_comm_allreduce(ft_host_comm* c, const void* send, void* recv, long long count, int dtype, int io_dtype,
                                 int op, int algo, const int* widths, int nstages, int ctas, long long tile_bytes,
                                 int mode, double scale, int apply_scale, unsigned long long timeout_ns, int lonely,
                                 int phases) {
  StagedArgs a;
  memset(&a, 0, sizeof a);
  const int N = c->nranks;
  const bool want_bulk = algo == ALGO_STAGED_TMA;  // staged stages with the CPU's bulk tiles (cpu_backend.h)
  if (want_bulk) algo = ALGO_STAGED;
  if (send == FT_IN_PLACE || !send) send = recv;
  if (N == 1) {
    if (send != recv) memcpy(recv, send, (size_t)count * (size_t)dtype_size(io_dtype));
    return 0;
  }
  int phant = 0;  // lonely < 0: a "-k" tree, k phantom positions beyond the processes
  if (lonely < 0) {
    phant = -lonely;
    lonely = 0;
  }
  if (algo == ALGO_RING || algo == ALGO_NVLS || algo == ALGO_ONESHOT || algo == ALGO_TWOSHOT) {
    int flat = N;
    if (!make_plan(&flat, 1, &a.plan)) return -1;
    lonely = 0;
    if (phant) return -14;
  } else if (!make_plan(widths, nstages, &a.plan)) {
    return -1;
  }
  if (a.plan.nranks + lonely - phant != N) return -2;
  if (phant && (phant > c->phantoms || mode != MODE_STAGING || phases != PHASES_ALLREDUCE || dtype == DT_FP8_BLOCK)) return -14;
  if (lonely > 0 && (algo != ALGO_STAGED || mode != MODE_STAGING)) return -8;
  if (phases != PHASES_ALLREDUCE && (algo != ALGO_STAGED || lonely > 0 || mode != MODE_STAGING || count % N)) return -10;
  if (dtype == DT_FP8_BLOCK && (algo != ALGO_STAGED || lonely > 0 || mode != MODE_STAGING)) return -9;
  const int es = dtype_size(dtype), ios = dtype_size(io_dtype);
  if (!es || !ios) return -3;
  // every rank must take the same code path: alignment is agreed on collectively
  char vec = ((uintptr_t)send % 16 == 0) && ((uintptr_t)recv % 16 == 0);
  if (phases != PHASES_ALLREDUCE && ((count / N) * es) % 16) vec = 0;
  // registered mode: the buffer is a 16-byte aligned piece of my registered region, at the same offset on every rank
  // (a ragged last vector is written whole: up to 15 bytes behind the message, inside the same 16-byte slot)
  long long reg_off = 0;
  bool reg_bad = false;
  if (mode == MODE_REGISTERED) {
    const char* base = c->win_table[(size_t)c->rank] + 2 * c->win_bytes;
    reg_off = (const char*)recv - base;
    reg_bad = send != recv || reg_off < 0 || reg_off % 16 ||
              (size_t)reg_off + (((size_t)count * (size_t)es + 15) & ~(size_t)15) > c->win_bytes;
  }
  // a rank with a bad buffer still takes part in the agreement (with a payload nobody else can have), so that
  // every rank returns the error instead of some of them waiting for it
  const int agreed = agree(c, &vec, reg_bad ? 0x7fffffffu : (uint32_t)(reg_off >> 4), timeout_ns);
  if (agreed == 0) return -11;
  if (agreed == 2 || reg_bad) return -13;
  a.lonely = lonely;
  a.phases = phases;
  a.nvls_reducers = ctas >= 4 ? ctas / 4 : 0;  // small CTA counts exercise the one-role (skewed) variant
  a.lay = make_layout(count, a.plan.nranks, es, tile_bytes, vec != 0);
  a.use_tma = want_bulk && vec && lonely == 0 && dtype != DT_FP8_BLOCK && count % (16 / es) == 0;
  a.scale_plane_off = ((long long)a.lay.block * a.plan.nranks * es + 255) & ~255LL;
  if ((size_t)(a.scale_plane_off + (a.lay.block * a.plan.nranks) / 16 + 16) > c->win_bytes) return -4;
  if (algo == ALGO_ONESHOT) {
    size_t region = (c->ll_bytes / (size_t)N) & ~(size_t)15;
    if ((size_t)((count * es + 7) / 8) * 16 > region) return -5;
  }
  if (algo == ALGO_TWOSHOT) {  // [shot 1 | shot 2] x N regions of ceil(lines / N) lines per parity half
    const size_t region2 = (c->ll_bytes / (size_t)(2 * N)) & ~(size_t)15;
    const size_t lines2 = (size_t)((count * es + 7) / 8), per_block = (lines2 + (size_t)N - 1) / (size_t)N;
    if (per_block * 16 > region2) return -5;
  }
  if (algo == ALGO_NVLS && mode == MODE_STAGING && a.lay.ntiles > FT_MAX_CTAS) return -12;  // per-tile flag slots
  if (ctas < 1 || ctas > FT_MAX_CTAS) return -6;
  a.nlocal = 1;
  a.ctas = ctas;
  a.mode = mode;
  a.apply_scale = apply_scale;
  a.scale = scale;
  a.win_off0 = mode == MODE_REGISTERED ? 2 * (long long)c->win_bytes + reg_off : 0;
  a.win_stride = mode == MODE_REGISTERED ? 0 : (long long)c->win_bytes;
  a.timeout_ns = timeout_ns;
  a.rc[0].comm = &c->cd;
  a.rc[0].send = send;
  a.rc[0].recv = recv;
  a.rc[0].win = c->win_table.data();
  a.rc[0].mc = nullptr;
  // the phantoms this process hosts ride along as further logical ranks of the same call
  if (phant) {
    const size_t bytes = ((size_t)count * (size_t)ios + 15) & ~(size_t)15;
    size_t live = 0;
    for (int j : c->hosted_id) live += j < phant ? 1 : 0;
    if (live && (c->ident_count != count || c->ident_dtype != io_dtype || c->ident_op != op)) {
      c->ident.assign(bytes + 16, 0);
      if (!fill_identity(c->ident.data() + ((16 - (uintptr_t)c->ident.data() % 16) % 16), count, io_dtype, op)) return -7;
      c->ident_count = count;
      c->ident_dtype = io_dtype;
      c->ident_op = op;
    }
    if (live && c->sink.size() < live * bytes + 16) c->sink.resize(live * bytes + 16);
    size_t k = 0;
    for (size_t i = 0; i < c->hosted.size(); ++i) {
      if (c->hosted_id[i] >= phant) continue;
      RankCall& rc = a.rc[a.nlocal++];
      rc.comm = &c->hosted[i]->cd;
      rc.send = c->ident.data() + ((16 - (uintptr_t)c->ident.data() % 16) % 16);
      rc.recv = c->sink.data() + ((16 - (uintptr_t)c->sink.data() % 16) % 16) + (k++) * bytes;
      rc.win = c->win_table.data();
      rc.mc = nullptr;
    }
  }
  if (!host_run_group(dtype_group(dtype), dtype, io_dtype, op, vec != 0, &a, a.use_tma ? (int)ALGO_STAGED_TMA : algo)) return -7;
  // every logical rank counts every call (flag values and the window parity derive from the epoch): the phantoms
  // that sat this one out keep step with the ranks that ran it
  for (size_t i = 0; i < c->hosted.size(); ++i)
    if (c->hosted_id[i] >= phant) c->hosted[i]->epoch++;
  for (auto& ph : c->hosted)
    if (ph->err) return (long long)ph->err;
  return c->err ? (long long)c->err : 0;
}

}  // extern "C"
