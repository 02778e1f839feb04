// Communicator + collective entry points (the C API of include/flextree.h).
//
// Capability parity: MPI_Allreduce_FT / MPI_Allreduce (allreduce_over_mpi/mpi_mod.hpp:1167-1221):
// context sizing (:216-243), N<=1 copy fast path (:1181-1188), grow-only staging buffer (:932-950,
// :1190), stage vector from FT_TOPO (:882-929, :1191), tree vs ring selection (:1194-1215).
// Additions: algorithm selection by the fitted link model, zero-copy registered buffers, one-shot and
// NVLS paths, fused scale/cast, device watchdog, CUDA-graph-safe epochs.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/flextree.h"
#include "../core/comm_dev.h"
#include "../core/dispatch.h"
#include "../core/types.h"
#include "../kernels/launch.h"
#include "../planner/cost_model.h"
#include "../planner/factorize.h"
#include "symm.h"

namespace ft {
cudaError_t launch_link_probe(char* const* d_table, int rank, int nranks, int npeers, long long bytes_per_peer, int mode,
                              int ctas, unsigned* sink, cudaStream_t stream);  // kernels/link_probe.cu
void set_last_error(const std::string& s);  // planner/c_api.cc
void host_pipe_release(const ft_comm_t* c);  // runtime/host_pipe.cu
}
using namespace ft;

struct ft_link_model { LinkModel m; };

namespace {

enum CommKind { COMM_PROCESS = 0, COMM_LOCAL = 1, COMM_EMULATED = 2 };

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
long long env_ll(const char* name, long long dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoll(v) : dflt;
}
int log_level() {
  static int lvl = [] {
    const char* v = getenv("FT_LOG_LEVEL");
    if (!v) return 0;
    if (!strcmp(v, "debug")) return 2;
    if (!strcmp(v, "info")) return 1;
    return atoi(v);
  }();
  return lvl;
}
#define FT_LOG(lvl, ...)                                   \
  do {                                                     \
    if (log_level() >= (lvl)) {                            \
      fprintf(stderr, "[flextree] " __VA_ARGS__);          \
      fprintf(stderr, "\n");                               \
    }                                                      \
  } while (0)

// State shared by the comms of one ft_comm_create_local() group: segments are (re)allocated for the
// whole group by whichever member asks first.
struct LocalGroup {
  std::vector<int> devices;
  std::vector<struct ft_comm*> members;
  std::mutex mu;
};

}  // namespace

// Resolved launch of one message shape. Everything do_allreduce derives from (count, dtypes, op, scale, buffer
// kind, policy) is kept so that the steady state of a training loop costs one lookup + one kernel launch on
// the host ("communicator: plan cache keyed by (bytes, dtype, op)"; the reference instead rebuilds
// its schedule tables and request arrays on every call, mpi_mod.hpp:964-972).
struct PlanKey {
  size_t count, uoff;
  double scale;
  int io, wire, op, phases, useg;
  bool vec, in_place;
  unsigned long long gen;
  bool operator==(const PlanKey& o) const {
    return count == o.count && uoff == o.uoff && scale == o.scale && io == o.io && wire == o.wire && op == o.op &&
           phases == o.phases && useg == o.useg && vec == o.vec && in_place == o.in_place && gen == o.gen;
  }
};
struct PlanEntry {
  PlanKey key;
  StagedArgs args;  // rc[] is filled per call
  int launch_algo, threads, mode;
  ft_call_info_t info;
  unsigned long long last_use;
};

struct ft_comm {
  std::vector<PlanEntry> plans;          // small LRU
  unsigned long long policy_gen = 1;     // bumped by every setter / (re)allocation: invalidates the cache
  unsigned long long use_clock = 0;
  CommKind kind = COMM_PROCESS;
  int rank = 0, nranks = 1, device = 0;
  int nlocal = 1;  // logical ranks served by this object (N for COMM_EMULATED)
  Bootstrap bs;
  std::shared_ptr<LocalGroup> group;
  SymmOptions sopt;

  // per logical rank (size nlocal): views of the symmetric segments
  std::vector<Segment> pad, ll, win;
  size_t window_bytes = 0;  // per parity
  size_t ll_bytes = 0;      // per parity
  struct UserSeg {
    std::vector<Segment> seg;  // per logical rank
    size_t bytes;
  };
  std::vector<UserSeg> user;
  size_t local_alloc_cursor = 0;  // COMM_LOCAL: which user segment the next ft_symm_alloc returns

  CommDev* d_comm = nullptr;       // [nlocal]
  uint64_t* d_epoch = nullptr;     // [nlocal]
  unsigned* d_done = nullptr;      // [nlocal]
  unsigned* d_err = nullptr;       // [nlocal]
  uint64_t* d_trace = nullptr;     // [nlocal][FT_TRACE_SLOTS], allocated by ft_comm_trace_enable
  unsigned* h_err = nullptr;       // [nlocal] mapped pinned host mirror of d_err (written by the kernel that sets d_err)
  long long nvls_max_tile = 0;  // registered NVLS: tile cap (0: env FT_NVLS_MAX_TILE_BYTES / 4 MiB)
  int nvls_ctas = 0;            // registered NVLS: CTAs for large messages (0: 8)
  // "-k" trees (cost_model/ChooseWidth.h:14-31 lists the trees of N+1 as candidates for N ranks): a tree with k more
  // positions than there are GPUs; phantom j (logical rank N+j) is hosted by process j, see the phantom section below
  struct PhantomWorld;
  PhantomWorld* ph = nullptr;
  int phantom_override = 0;     // widths_override multiplies to nranks + phantom_override
  int cluster = -1;             // P2P stages with a two-CTA cluster per tile: -1 env FT_CLUSTER / off, 0 off, 1 on
  int nvls_variant = -1, nvls_flags = 0;  // ft_comm_set_tuning; -1: env FT_NVLS_VARIANT / default
  int fault_rank = -1;             // ft_comm_inject_fault: this rank plays dead in the next calls (-1 = off)

  // policy
  std::vector<int> widths_override;
  int lonely_override = 0;  // ranks outside the tree when widths_override multiplies to less than the world size
  int algo_override = ALGO_AUTO;
  LinkModel model;
  int ctas_override = 0, threads_override = 0;
  int tma_override = -1;  // -1: env FT_TMA / default
  int max_ctas = 0;       // 0: whole GPU; >0: cap (leave SMs to overlapping compute, e.g. DDP backward)
  long long tile_override = 0;
  unsigned long long timeout_ns = 20000000000ull;
  int sm_count = 148;
  ft_call_info_t last{};
};

namespace {

bool fail(ft_result_t* out, ft_result_t code, const std::string& msg) {
  set_last_error(msg);
  FT_LOG(1, "error: %s", msg.c_str());
  *out = code;
  return false;
}
#define FT_CHECK_RT(call)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (call);                                                                \
    if (_e != cudaSuccess) {                                                                \
      set_last_error(std::string(#call) + ": " + cudaGetErrorString(_e));                   \
      (void)cudaGetLastError();                                                             \
      return FT_ERR_CUDA;                                                                   \
    }                                                                                       \
  } while (0)

void apply_options(ft_comm* c, const ft_comm_options_t* o) {
  ft_comm_options_t d;
  ft_comm_options_default(&d);
  if (o) d = *o;
  c->window_bytes = d.window_bytes ? d.window_bytes : (size_t)env_ll("FT_WINDOW_BYTES", 64ll << 20);
  c->ll_bytes = d.ll_bytes ? d.ll_bytes : (size_t)env_ll("FT_LL_BYTES", 32ll << 20);
  if ((long long)c->window_bytes < (1 << 20)) c->window_bytes = 1 << 20;  // also catches 0 and negative env values
  c->window_bytes = (c->window_bytes + 4095) & ~(size_t)4095;
  c->ll_bytes = (c->ll_bytes + 4095) & ~(size_t)4095;
  int mc = d.use_multicast;
  if (getenv("FT_MULTICAST")) mc = atoi(getenv("FT_MULTICAST"));
  c->sopt.try_multicast = mc != 0;
  c->sopt.require_multicast = mc == 1;
  int vmm = d.use_vmm;
  if (getenv("FT_SYMM")) vmm = strcmp(getenv("FT_SYMM"), "ipc") ? 1 : 0;
  c->sopt.try_vmm = vmm != 0;
  double t = d.timeout_s;
  if (t < 0) t = getenv("FT_TIMEOUT_S") ? atof(getenv("FT_TIMEOUT_S")) : 20.0;
  c->timeout_ns = (unsigned long long)(t * 1e9);
  const char* cm = getenv("FT_COSTMODEL");
  if (cm && *cm) {
    std::string err;
    if (!load_link_model(cm, &c->model, &err)) FT_LOG(1, "FT_COSTMODEL: %s", err.c_str());
  }
  const char* al = getenv("FT_ALGO");
  if (al && *al) {
    int a = algo_from_name(al);
    if (a >= 0) c->algo_override = a;
  }
}

// Device-side CommDev for every logical rank; requires pad/ll segments.
ft_result_t upload_comm_dev(ft_comm* c) {
  const int L = c->nlocal;
  FT_CHECK_RT(cudaSetDevice(c->device));
  FT_CHECK_RT(cudaMalloc(&c->d_comm, sizeof(CommDev) * (size_t)L));
  FT_CHECK_RT(cudaMalloc(&c->d_epoch, sizeof(uint64_t) * (size_t)L));
  FT_CHECK_RT(cudaMalloc(&c->d_done, sizeof(unsigned) * (size_t)L));
  FT_CHECK_RT(cudaMalloc(&c->d_err, sizeof(unsigned) * (size_t)L));
  FT_CHECK_RT(cudaMemset(c->d_epoch, 0, sizeof(uint64_t) * (size_t)L));
  FT_CHECK_RT(cudaMemset(c->d_done, 0, sizeof(unsigned) * (size_t)L));
  FT_CHECK_RT(cudaMemset(c->d_err, 0, sizeof(unsigned) * (size_t)L));
  unsigned* h_err_dev = nullptr;
  if (cudaHostAlloc((void**)&c->h_err, sizeof(unsigned) * (size_t)L, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess &&
      cudaHostGetDevicePointer((void**)&h_err_dev, c->h_err, 0) == cudaSuccess) {
    memset(c->h_err, 0, sizeof(unsigned) * (size_t)L);
  } else {  // no mapped host memory: errors are then only visible through ft_comm_error()
    (void)cudaGetLastError();
    if (c->h_err) cudaFreeHost(c->h_err);
    c->h_err = nullptr;
    h_err_dev = nullptr;
  }
  std::vector<CommDev> h((size_t)L);
  for (int l = 0; l < L; ++l) {
    CommDev& d = h[(size_t)l];
    memset(&d, 0, sizeof d);
    d.rank = c->kind == COMM_EMULATED ? l : c->rank;
    d.nranks = c->nranks;
    for (int p = 0; p < c->nranks; ++p) {
      d.pad[p] = (uint64_t*)c->pad[(size_t)l].base[(size_t)p];
      d.ll[p] = c->ll[(size_t)l].base[(size_t)p];
    }
    d.epoch = c->d_epoch + l;
    d.done = c->d_done + l;
    d.err = c->d_err + l;
    d.err_host = h_err_dev ? h_err_dev + l : nullptr;
    d.ll_bytes = c->ll_bytes;
  }
  FT_CHECK_RT(cudaMemcpy(c->d_comm, h.data(), sizeof(CommDev) * (size_t)L, cudaMemcpyHostToDevice));
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, c->device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
  return FT_SUCCESS;
}

// (Re)allocate one kind of segment for this comm (COMM_PROCESS) or the whole local group / emulation.
bool make_segments(ft_comm* c, size_t bytes, std::vector<Segment>* out, std::string* err) {
  if (c->kind == COMM_PROCESS) {
    out->assign(1, Segment());
    return segment_create(c->bs, c->device, bytes, c->sopt, &(*out)[0], err);
  }
  if (c->kind == COMM_EMULATED) {
    std::vector<int> devs((size_t)c->nranks, c->device);
    return segment_create_local(devs, bytes, out, err);
  }
  return false;  // COMM_LOCAL handled by the group helpers
}

void destroy_segments(std::vector<Segment>* v) {
  for (auto& s : *v) segment_destroy(&s);
  v->clear();
}

ft_result_t grow_window(ft_comm* c, size_t need_per_parity) {
  if (need_per_parity <= c->window_bytes && !c->win.empty()) return FT_SUCCESS;
  c->policy_gen++;
  if (c->group) {
    std::lock_guard<std::mutex> lk(c->group->mu);
    for (auto* m : c->group->members) if (m) m->policy_gen++;
  }
  size_t nb = c->window_bytes < 4096 ? 4096 : c->window_bytes;  // FT_WINDOW_BYTES=0 must not loop forever
  while (nb < need_per_parity) {
    if (nb > ((size_t)1 << 62)) { set_last_error("staging window request too large"); return FT_ERR_NOMEM; }
    nb *= 2;
  }
  nb = (nb + 4095) & ~(size_t)4095;
  FT_LOG(1, "rank %d: staging window -> %zu bytes per parity", c->rank, nb);
  std::string err;
  if (c->kind == COMM_LOCAL) {
    std::lock_guard<std::mutex> lk(c->group->mu);
    bool need = false;
    for (auto* m : c->group->members) need = need || m->win.empty() || m->window_bytes < need_per_parity;
    if (!need) return FT_SUCCESS;
    for (auto* m : c->group->members) {
      cudaSetDevice(m->device);
      cudaDeviceSynchronize();
    }
    std::vector<Segment> segs;
    if (!segment_create_local(c->group->devices, 2 * nb, &segs, &err, &c->sopt)) {
      set_last_error(err);
      return FT_ERR_CUDA;
    }
    for (size_t i = 0; i < c->group->members.size(); ++i) {
      ft_comm* m = c->group->members[i];
      destroy_segments(&m->win);
      m->win.assign(1, segs[i]);
      m->window_bytes = nb;
    }
    cudaSetDevice(c->device);
    return FT_SUCCESS;
  }
  FT_CHECK_RT(cudaSetDevice(c->device));
  FT_CHECK_RT(cudaDeviceSynchronize());
  if (c->kind == COMM_PROCESS && !c->bs.barrier()) {
    set_last_error("bootstrap barrier failed");
    return FT_ERR_INTERNAL;
  }
  std::vector<Segment> segs;
  if (!make_segments(c, 2 * nb, &segs, &err)) {
    set_last_error(err);
    return FT_ERR_CUDA;
  }
  destroy_segments(&c->win);
  c->win = segs;
  c->window_bytes = nb;
  return FT_SUCCESS;
}

// Find the user (registered) segment containing [p, p+bytes) for logical rank l.
int find_user_seg(const ft_comm* c, int l, const void* p, size_t bytes, size_t* off) {
  for (size_t i = 0; i < c->user.size(); ++i) {
    const Segment& s = c->user[i].seg[(size_t)l];
    const char* b = s.base[(size_t)s.self];
    if ((const char*)p >= b && (const char*)p + bytes <= b + s.bytes) {
      *off = (size_t)((const char*)p - b);
      return (int)i;
    }
  }
  return -1;
}

struct Shape {
  int ctas, threads;
  long long tile_bytes;
  int nvls_reducers = 0;  // NVLS with private buffers: CTAs that only reduce (the others copy)
};

// Launch shape heuristics. Measured on 2xB200 (profiles/tune_n2.md): the pull kernels are bound by bytes
// in flight, so bandwidth grows ~linearly with CTAs up to one per SM, and every tile boundary costs a
// system fence + a flag hop, so tiles should be as large as the message allows (512 KiB per block-chunk
// was the sweet spot; 2 MiB starves the grid). Every rank must derive the same shape from the same
// arguments.
Shape pick_shape(const ft_comm* c, int algo, int mode, bool tma, size_t msg_bytes, int nranks, int esize, bool cluster = false) {
  Shape s;
  s.threads = c->threads_override ? c->threads_override : env_int("FT_THREADS", 512);
  if (s.threads > FT_MAX_THREADS) s.threads = FT_MAX_THREADS;
  if (algo == ALGO_NVLS && mode == MODE_REGISTERED && s.threads > FT_NVLS_THREADS) s.threads = FT_NVLS_THREADS;  // launch bound of kernels/nvls_kernel.cu
  if (s.threads < 32) s.threads = 32;
  int max_ctas = c->sm_count;  // one resident CTA per SM at 512 threads x <=128 registers
  const int cap = c->max_ctas > 0 ? c->max_ctas : env_int("FT_MAX_CTAS", 0);
  if (cap > 0 && cap < max_ctas) max_ctas = cap;
  if (max_ctas > FT_MAX_CTAS) max_ctas = FT_MAX_CTAS;
  if (cluster) max_ctas /= 2;  // a "CTA" of the plan is a cluster of two
  if (c->kind == COMM_EMULATED) {
    // cooperative launch: nranks*ctas CTAs must be co-resident
    int per_sm = (s.threads <= 256 && !tma) ? 2 : 1;  // the TMA ring takes the whole shared memory of an SM
    max_ctas = c->sm_count * per_sm / nranks;
    if (cluster) max_ctas /= 2;
    if (max_ctas < 1) max_ctas = 1;
  }
  const long long block_bytes = ((long long)msg_bytes + nranks - 1) / nranks;
  // P2P tiles: a tile's work is ~nranks * tile bytes of loads; 512 KiB was best at N=2 and 64-128 KiB at N=8
  // With TMA-staged tiles the ring drains at every tile boundary, so tiles should be large (measured at
  // N=2: 1-2 MiB best at 256 MiB, 64 KiB x 32 CTAs best at 4 MiB).
  const long long kMinTile = env_ll("FT_MIN_TILE_BYTES", tma ? (64 << 10) : (16 << 10));
  long long kMaxTile = env_ll("FT_MAX_TILE_BYTES", (tma ? (4 << 20) : (1 << 20)) / (nranks < 2 ? 2 : nranks));
  if (kMaxTile < kMinTile) kMaxTile = kMinTile;
  int ctas;
  long long tile;
  if (algo == ALGO_ONESHOT || algo == ALGO_TWOSHOT) {
    long long lines = ((long long)msg_bytes + 7) / 8;  // one LL line (8 payload bytes) per thread
    if (algo == ALGO_TWOSHOT) lines = (lines + nranks - 1) / nranks;  // threads stride over the N blocks
    ctas = (int)((lines + s.threads - 1) / s.threads);
    const int cap_ll = algo == ALGO_TWOSHOT ? 128 : 64;
    if (ctas > cap_ll) ctas = cap_ll;
    tile = kMinTile;
  } else if (algo == ALGO_NVLS) {
    // multimem.ld_reduce/st saturate the switch path from very few SMs (8 CTAs reach 796 GB/s busbw at
    // 256 MiB on 8 GPUs) and every tile boundary costs two system fences with multicast stores in
    // flight: few CTAs, one big tile each (profiles/tune_n8.md).
    if (mode == MODE_REGISTERED) {
      long long want = block_bytes / (32 << 10);
      const int top = c->nvls_ctas > 0 ? c->nvls_ctas : env_int("FT_NVLS_CTAS", 16);  // x 256 threads
      ctas = (int)(want < 2 ? 2 : (want > top ? top : want));
      if (ctas > max_ctas) ctas = max_ctas;
      tile = (block_bytes + ctas - 1) / ctas;
      const long long kNvlsMaxTile = c->nvls_max_tile > 0 ? c->nvls_max_tile : env_ll("FT_NVLS_MAX_TILE_BYTES", 4 << 20);
      if (tile > kNvlsMaxTile) tile = kNvlsMaxTile;
      if (tile < 4096) tile = 4096;
    } else {
      // Private buffers (run_nvls_private): user <-> window copies around the in-switch reduce.
      if ((long long)msg_bytes >= env_ll("FT_NVLS_SPLIT_BYTES", 128ll << 20)) {
        // large: role split - up to 8 CTAs only reduce in the switch, the others copy; tiles small enough that
        // the pipeline fills quickly, large enough that the two fences per tile stay cheap
        tile = block_bytes / env_ll("FT_NVLS_PRIV_TILE_DIV", 64);
        if (tile > env_ll("FT_NVLS_PRIV_TILE_MAX", 256 << 10)) tile = env_ll("FT_NVLS_PRIV_TILE_MAX", 256 << 10);
        if (tile < (16 << 10)) tile = 16 << 10;
        const long long ntiles = (block_bytes + tile - 1) / tile;
        int gr = (int)(ntiles / 4);
        const int gr_top = env_int("FT_NVLS_REDUCERS", 16);
        gr = gr < 1 ? 1 : (gr > gr_top ? gr_top : gr);
        const int cap_env = env_int("FT_NVLS_PRIV_CTAS", 64);
        int cap2 = max_ctas < cap_env ? max_ctas : cap_env;
        int gc = (int)(ntiles < (long long)(cap2 - gr) ? ntiles : (long long)(cap2 - gr));
        if (gc < 1) gc = 1;
        ctas = gr + gc;
        s.nvls_reducers = gr;
      } else {
        // mid-size: every CTA does copy-in / reduce / copy-out on its own tiles, skewed by one tile; several
        // tiles per CTA so that the passes of different CTAs overlap
        long long want = block_bytes / (32 << 10);
        ctas = (int)(want < 2 ? 2 : (want > 64 ? 64 : want));
        if (ctas > max_ctas) ctas = max_ctas;
        tile = block_bytes / ((long long)ctas * 4);
        if (tile > (256 << 10)) tile = 256 << 10;
        if (tile < (16 << 10)) tile = 16 << 10;
        s.nvls_reducers = 0;
      }
    }
  } else {
    // P2P pull kernels: as many CTAs as there are minimum-size tiles, up to the whole GPU; then one tile per
    // CTA until the tile would exceed kMaxTile
    long long want = (block_bytes + kMinTile - 1) / kMinTile;
    ctas = (int)(want < max_ctas ? want : max_ctas);
    if (ctas < 1) ctas = 1;
    tile = (block_bytes + ctas - 1) / ctas;
    if (tile < kMinTile) tile = kMinTile;
    if (tile > kMaxTile) tile = kMaxTile;
  }
  ctas = c->ctas_override ? c->ctas_override : env_int("FT_CTAS", ctas);
  if (ctas < 1) ctas = 1;
  if (s.nvls_reducers >= ctas) s.nvls_reducers = ctas >= 8 ? ctas / 4 : 0;
  if (ctas > FT_MAX_CTAS) ctas = FT_MAX_CTAS;
  if (c->kind == COMM_EMULATED && ctas > max_ctas) ctas = max_ctas;
  s.ctas = ctas;
  long long t2 = c->tile_override ? c->tile_override : env_ll("FT_TILE_BYTES", 0);
  if (t2 > 0) tile = t2;
  tile = (tile + 15) / 16 * 16;
  if (tile < esize) tile = esize;
  s.tile_bytes = tile;
  return s;
}

// NVLS in place on registered buffers of the wire type has a kernel of its own (kernels/nvls_kernel.cu).
bool launch_any(int wire_dtype, int io_dtype, int op, bool vec, int mode, LaunchCtx* ctx) {
  if (ctx->algo == ALGO_STAGED_CLUSTER) {
    if (cuda_launch_cluster(wire_dtype, io_dtype, op, vec, ctx)) return true;
    ctx->algo = ALGO_STAGED_TMA;  // not available for this type / shape: the one-CTA TMA kernel
  }
  if (ctx->algo == ALGO_NVLS && mode == MODE_REGISTERED && wire_dtype == io_dtype && vec) {
    ctx->err = cudaSuccess;
    if (cuda_launch_nvls_registered(wire_dtype, op, ctx)) return true;
  }
  return cuda_launch(wire_dtype, io_dtype, op, vec, ctx);
}

// A device watchdog that fired during an EARLIER call left its error word in the mapped host mirror: refuse to
// queue more work on a communicator whose flags are in an unknown state (the results of that call are garbage).
// Costs a few host loads, synchronises nothing. ft_comm_clear_error() re-arms the communicator.
ft_result_t check_device_error(ft_comm* c) {
  if (!c->h_err) return FT_SUCCESS;
  for (int l = 0; l < c->nlocal; ++l) {
    const unsigned e = *(volatile unsigned*)(c->h_err + l);
    if (e) {
      char buf[320];
      snprintf(buf, sizeof buf,
               "device watchdog fired in an earlier call on this communicator: error word %#x (code %u, phase %u, "
               "waiting for rank %u, on rank %u); its results are invalid - ft_comm_clear_error() to continue",
               e, e & 0xffu, (e >> 8) & 0xffu, (e >> 16) & 0xffu, e >> 24);
      set_last_error(buf);
      return (e & 0xffu) == ERR_TIMEOUT ? FT_ERR_TIMEOUT : FT_ERR_INTERNAL;
    }
  }
  return FT_SUCCESS;
}

ft_result_t phantom_allreduce(ft_comm* c, const void* send, void* recv, size_t count, int io_dtype, int wire_dtype, int op,
                              double scale, cudaStream_t stream, const std::vector<int>& widths, int k, bool vec);
void phantom_release(ft_comm* c);

// The one place a collective is resolved and launched. send/recv: per logical rank.
ft_result_t do_allreduce(ft_comm* c, const void* const* send, void* const* recv, size_t count, int io_dtype,
                         int wire_dtype, int op, double scale, cudaStream_t stream, int phases = PHASES_ALLREDUCE) {
  const int N = c->nranks, L = c->nlocal;
  if (io_dtype < 0 || io_dtype >= DT_COUNT || wire_dtype < 0 || wire_dtype >= DT_COUNT) {
    set_last_error("bad dtype");
    return FT_ERR_INVALID;
  }
  if (!op_supported(op, wire_dtype)) {
    set_last_error(std::string("unsupported op/dtype: ") + op_name(op) + " on " + dtype_name(wire_dtype));
    return FT_ERR_UNSUPPORTED;
  }
  if (wire_dtype == DT_FP8_BLOCK && c->nranks == 1 && io_dtype != DT_FP8_BLOCK) wire_dtype = io_dtype;  // nothing to compress
  const bool fp8_wire = wire_dtype == DT_FP8_BLOCK;
  if (fp8_wire && !((io_dtype == DT_FLOAT || io_dtype == DT_BFLOAT16) && op == OP_SUM)) {
    set_last_error("the block-scaled fp8 wire carries fp32 or bf16 user buffers, SUM only");
    return FT_ERR_UNSUPPORTED;
  }
  if (io_dtype == DT_FP8_BLOCK) {
    set_last_error("fp8-block16 is a wire format, not a buffer dtype");
    return FT_ERR_INVALID;
  }
  if (io_dtype != wire_dtype && !fp8_wire &&
      !(io_dtype == DT_FLOAT && (wire_dtype == DT_BFLOAT16 || wire_dtype == DT_HALF))) {
    set_last_error("mixed formats: fp32 buffers with a bf16/fp16 wire, or fp32/bf16 buffers with the fp8 block-scaled wire");
    return FT_ERR_UNSUPPORTED;
  }
  if (count == 0) return FT_SUCCESS;
  {
    ft_result_t er = check_device_error(c);
    if (er != FT_SUCCESS) return er;
  }
  {
    int cur = -1;
    if (cudaGetDevice(&cur) != cudaSuccess || cur != c->device) FT_CHECK_RT(cudaSetDevice(c->device));
  }
  const int es = dtype_size(wire_dtype), ios = dtype_size(io_dtype);
  const bool is_float = dtype_is_float(wire_dtype);
  const bool apply_scale = is_float && scale != 1.0;
  const size_t msg_bytes = count * (size_t)es;

  const void* s[FT_MAX_LOCAL];
  void* r[FT_MAX_LOCAL];
  bool vec = true, in_place = true;
  for (int l = 0; l < L; ++l) {
    r[(size_t)l] = recv[l];
    s[(size_t)l] = (send[l] == FT_IN_PLACE || send[l] == nullptr) ? recv[l] : send[l];
    if (!r[(size_t)l]) { set_last_error("null recvbuf"); return FT_ERR_INVALID; }
    vec = vec && ((uintptr_t)s[(size_t)l] % 16 == 0) && ((uintptr_t)r[(size_t)l] % 16 == 0);
    in_place = in_place && s[(size_t)l] == r[(size_t)l];
  }

  if (phases != PHASES_ALLREDUCE) {
    // the halves: `count` is the TOTAL element count (nranks * per-rank block); blocks must not be padded
    if (count % (size_t)N) { set_last_error("reduce_scatter/all_gather: total count must be a multiple of the world size"); return FT_ERR_INVALID; }
    if (((count / (size_t)N) * (size_t)es) % 16) vec = false;  // scalar path keeps block == per-rank count exactly
    in_place = false;
    if (N == 1) {
      if (s[0] != r[0]) FT_CHECK_RT(cudaMemcpyAsync(r[0], s[0], count * (size_t)ios, cudaMemcpyDeviceToDevice, stream));
      return FT_SUCCESS;
    }
  }

  // ---- N == 1: copy fast path (mpi_mod.hpp:1181-1188) ---------------------------------------------
  // (a scaled / casting single-rank call still goes through the one-shot kernel below)
  if (N == 1 && !apply_scale && io_dtype == wire_dtype) {
    if (!in_place) FT_CHECK_RT(cudaMemcpyAsync(r[0], s[0], count * (size_t)ios, cudaMemcpyDeviceToDevice, stream));
    c->last = ft_call_info_t{};
    c->last.algo = ALGO_STAGED;
    c->last.nstages = 1;
    c->last.widths[0] = 1;
    return FT_SUCCESS;
  }

  // ---- registered (zero-copy) mode? -------------------------------------------------------------------
  int mode = MODE_STAGING;
  int useg = -1;
  size_t uoff = 0;
  // Zero-copy needs whole 16-byte vectors: the in-place kernels store full vectors into the window, and a
  // ragged tail would spill into whatever follows a sub-view of a symmetric tensor.
  if (in_place && io_dtype == wire_dtype && vec && N > 1 && count % (size_t)(16 / ios) == 0) {
    useg = find_user_seg(c, 0, r[0], count * (size_t)ios, &uoff);
    for (int l = 1; l < L && useg >= 0; ++l) {
      size_t o2 = 0;
      if (find_user_seg(c, l, r[(size_t)l], count * (size_t)ios, &o2) != useg || o2 != uoff) useg = -1;
    }
    if (useg >= 0 && uoff % 16 == 0 && env_int("FT_REGISTERED", 1)) mode = MODE_REGISTERED;
    if (mode != MODE_REGISTERED) { useg = -1; uoff = 0; }
  }

  // ---- plan cache: a message shape seen before under the same policy is one lookup + one launch ----------
  PlanKey key{count, uoff, scale, io_dtype, wire_dtype, op, phases, useg, vec, in_place, c->policy_gen};
  for (PlanEntry& e : c->plans) {
    if (!(e.key == key)) continue;
    StagedArgs a2 = e.args;
    a2.fault_rank_plus1 = c->fault_rank + 1;
    for (int l = 0; l < L; ++l) {
      RankCall& rc = a2.rc[l];
      rc.comm = c->d_comm + l;
      rc.send = s[l];
      rc.recv = r[l];
      const Segment* seg = nullptr;
      if (e.info.algo != ALGO_ONESHOT && e.info.algo != ALGO_TWOSHOT) seg = e.mode == MODE_REGISTERED ? &c->user[(size_t)useg].seg[(size_t)l] : &c->win[(size_t)l];
      rc.win = seg ? seg->d_table : nullptr;
      rc.mc = seg ? seg->mc_base : nullptr;
    }
    LaunchCtx ctx;
    ctx.args = &a2;
    ctx.algo = e.launch_algo;
    ctx.threads = e.threads;
    ctx.cooperative = c->kind == COMM_EMULATED && N > 1;
    ctx.stream = stream;
    if (!launch_any(wire_dtype, io_dtype, op, vec, e.mode, &ctx) || !ctx.supported || ctx.err != cudaSuccess) {
      set_last_error(std::string("kernel launch: ") + cudaGetErrorString(ctx.err));
      (void)cudaGetLastError();
      return FT_ERR_CUDA;
    }
    e.last_use = ++c->use_clock;
    c->last = e.info;
    return FT_SUCCESS;
  }

  // ---- algorithm + widths --------------------------------------------------------------------------------
  int algo = c->algo_override;
  std::vector<int> widths = c->widths_override;
  int lonely = widths.empty() ? 0 : c->lonely_override;
  int phantoms = widths.empty() ? 0 : c->phantom_override;
  if (widths.empty() && getenv("FT_TOPO") && *getenv("FT_TOPO")) {
    // "2,3+1": a tree of 6 plus one lonely rank, "2,4-1": a tree of 8 positions for 7 ranks (extensions; the reference's
    // parser has neither syntax). A product mismatch is an error: the reference exits (mpi_mod.hpp:914-918).
    int buf[FT_MAX_STAGES + 8];
    const int k = ft_parse_topo_ex(nullptr, N, buf, FT_MAX_STAGES + 8, &lonely, &phantoms);
    if (k < 0) return FT_ERR_INVALID;
    widths.assign(buf, buf + k);
  }
  if (phantoms > 0) {
    if (phases != PHASES_ALLREDUCE || c->kind != COMM_PROCESS || phantoms > N) {
      set_last_error("'-k' trees: allreduce on a one-process-per-GPU communicator, k <= world size");
      return FT_ERR_UNSUPPORTED;
    }
    return phantom_allreduce(c, s[0], r[0], count, io_dtype, wire_dtype, op, scale, stream, widths, phantoms, vec);
  }
  if (!widths.empty() && widths[0] == 1 && N > 1) {  // "any 1 means ring" (mpi_mod.hpp:907-910)
    algo = ALGO_RING;
    widths.clear();
  }
  // Is a multicast mapping available for the buffer this call would use? (The staging window is
  // allocated lazily, so before its first use we go by the creation-time probe.)
  bool has_mc;
  if (mode == MODE_REGISTERED) has_mc = c->user[(size_t)useg].seg[0].mc_base != nullptr;
  else if (!c->win.empty()) has_mc = c->win[0].mc_base != nullptr;
  else has_mc = c->model.has_nvls;
  double est = 0;
  if (algo == ALGO_AUTO) {
    if (!widths.empty()) {
      algo = ALGO_STAGED;
    } else {
      LinkModel m = c->model;
      m.has_nvls = m.has_nvls && has_mc && nvls_supported(wire_dtype, op) && vec;
      if (N > 1 && (size_t)((msg_bytes + 7) / 8 * 16) > ((c->ll_bytes / (size_t)N) & ~(size_t)15)) m.max_oneshot_bytes = 0;
      if (N > 1 && (((msg_bytes + 7) / 8 + (size_t)N - 1) / (size_t)N) * 16 > ((c->ll_bytes / (size_t)(2 * N)) & ~(size_t)15)) m.max_twoshot_bytes = 0;
      Choice ch = choose(m, N, (double)msg_bytes, mode == MODE_REGISTERED, FT_MAX_WIDTH);
      algo = ch.algo;
      widths = ch.widths;
      est = ch.est_us;
    }
  }
  if (N == 1) algo = ALGO_ONESHOT;
  if (algo == ALGO_NVLS && !(nvls_supported(wire_dtype, op) && vec)) {
    set_last_error("NVLS path supports fp32/bf16/fp16 SUM on 16-byte aligned buffers only");
    return FT_ERR_UNSUPPORTED;
  }
  if (algo == ALGO_STAGED && widths.empty()) widths = {N};
  if (algo != ALGO_STAGED) { widths = {N}; lonely = 0; }
  if (est == 0) {  // forced algorithm / widths: still report what the model expects (validation tables)
    LinkModel m = c->model;
    m.has_nvls = true;
    est = predict_us(m, algo, widths, N, (double)msg_bytes, mode == MODE_REGISTERED);
  }

  StagedArgs a;
  memset(&a, 0, sizeof a);
  if (lonely > 0) {  // lonely trees run on the staged kernel with private windows
    algo = ALGO_STAGED;
    mode = MODE_STAGING;
  }
  if (phases != PHASES_ALLREDUCE) {  // the halves run on the P2P stage kernels with private windows
    if (lonely > 0 || wire_dtype == DT_FP8_BLOCK) {
      set_last_error("reduce_scatter/all_gather: no lonely trees / fp8 wire");
      return FT_ERR_UNSUPPORTED;
    }
    if (algo != ALGO_STAGED) { algo = ALGO_STAGED; widths = {N}; }
    mode = MODE_STAGING;
  }
  if (wire_dtype == DT_FP8_BLOCK) {  // block-scaled wire: register-staged P2P stages, private windows, vector path
    if (!vec || lonely > 0) {
      set_last_error("the fp8 block-scaled wire needs 16-byte aligned buffers and a full tree");
      return FT_ERR_UNSUPPORTED;
    }
    if (algo != ALGO_STAGED) { algo = ALGO_STAGED; widths = {N}; }
    mode = MODE_STAGING;
  }
  if (!make_plan(widths.data(), (int)widths.size(), &a.plan) || a.plan.nranks + lonely != N) {
    set_last_error("stage widths " + to_string(widths) + " do not multiply to the world size " + std::to_string(N));
    return FT_ERR_INVALID;
  }

  // TMA-staged tiles: vector path and whole 16-byte vectors only (blocks are padded to 16 bytes, so every
  // bulk copy is aligned).
  const int want_threads = c->threads_override ? c->threads_override : env_int("FT_THREADS", 512);
  const bool tma = algo == ALGO_STAGED && lonely == 0 && vec && wire_dtype != DT_FP8_BLOCK && want_threads >= 64 && (count % (size_t)(16 / es) == 0) &&
                   (c->threads_override == 0 || c->threads_override >= 64) &&
                   (c->tma_override >= 0 ? c->tma_override : env_int("FT_TMA", 1)) != 0;
  const int T = N - lonely;  // ranks in the tree = number of blocks
  const bool cluster = tma && want_threads == FT_MAX_THREADS && phases == PHASES_ALLREDUCE &&
                       (c->cluster >= 0 ? c->cluster : env_int("FT_CLUSTER", 0)) != 0 && dtype_is_float(wire_dtype) && es <= 4;
  Shape sh = pick_shape(c, algo, mode, tma, msg_bytes, T, es, cluster);
  a.lay = make_layout((long long)count, T, es, sh.tile_bytes, vec);
  a.lonely = lonely;
  a.phases = phases;
  a.nvls_reducers = sh.nvls_reducers;
  // flag values encode the tile ordinal in FT_EPOCH_SHIFT bits (ring: steps * tiles)
  {
    long long per_cta = (a.lay.ntiles + sh.ctas - 1) / sh.ctas;
    long long steps = algo == ALGO_RING ? 2LL * N : 1;
    while (per_cta * steps >= (1LL << FT_EPOCH_SHIFT) - 2) {
      sh.tile_bytes *= 2;
      a.lay = make_layout((long long)count, T, es, sh.tile_bytes, vec);
      per_cta = (a.lay.ntiles + sh.ctas - 1) / sh.ctas;
    }
  }
  // NVLS with private buffers signals per TILE (slot = tile mod FT_MAX_CTAS): keep every tile in its own slot so
  // that two CTAs can never write one slot out of order.
  while (algo == ALGO_NVLS && mode == MODE_STAGING && a.lay.ntiles > FT_MAX_CTAS) {
    sh.tile_bytes *= 2;
    a.lay = make_layout((long long)count, T, es, sh.tile_bytes, vec);
  }
  if (sh.ctas > a.lay.ntiles && algo != ALGO_ONESHOT && algo != ALGO_TWOSHOT) {
    sh.ctas = a.lay.ntiles;
    if (sh.nvls_reducers >= sh.ctas) sh.nvls_reducers = 0;
  }

  // ---- memory --------------------------------------------------------------------------------------------
  if (algo == ALGO_ONESHOT) {
    size_t region = (c->ll_bytes / (size_t)N) & ~(size_t)15;
    size_t lines = (msg_bytes + 7) / 8;
    if (N > 1 && lines * 16 > region) {
      set_last_error("message too large for the one-shot (LL) buffer; raise ll_bytes / FT_LL_BYTES");
      return FT_ERR_INVALID;
    }
  } else if (algo == ALGO_TWOSHOT) {
    // LL area of a parity half: [shot 1 | shot 2], each N regions by source; a region holds one block of lines
    size_t region = (c->ll_bytes / (size_t)(2 * N)) & ~(size_t)15;
    size_t lines = (msg_bytes + 7) / 8, per_block = (lines + (size_t)N - 1) / (size_t)N;
    if (per_block * 16 > region) {
      set_last_error("message too large for the two-shot (LL) buffer; raise ll_bytes / FT_LL_BYTES");
      return FT_ERR_INVALID;
    }
  } else if (mode == MODE_STAGING) {
    size_t need = (size_t)a.lay.block * (size_t)T * (size_t)es;
    if (wire_dtype == DT_FP8_BLOCK) {
      a.scale_plane_off = ((long long)need + 255) & ~255LL;
      need = (size_t)a.scale_plane_off + need / 16 + 16;
    }
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cap);
    if (need > c->window_bytes || c->win.empty()) {
      if (cap != cudaStreamCaptureStatusNone) {
        set_last_error("staging window too small during stream capture: call ft_comm_reserve() first");
        return FT_ERR_INVALID;
      }
      ft_result_t rr = grow_window(c, need);
      if (rr != FT_SUCCESS) return rr;
    }
  }
  if (algo == ALGO_NVLS) {
    const char* mc = mode == MODE_REGISTERED ? c->user[(size_t)useg].seg[0].mc_base : c->win[0].mc_base;
    if (!mc && c->kind == COMM_EMULATED && c->algo_override == ALGO_NVLS) {
      // one GPU has no switch: the kernel plays it with a loop over the N windows (backend.cuh), which keeps the
      // NVLS control flow - flags, tiles, inner-loop variants, graph replay - testable on a single device
    } else if (!mc) {
      if (c->algo_override == ALGO_NVLS) {
        set_last_error("NVLS requested but this communicator has no multicast mapping");
        return FT_ERR_UNSUPPORTED;
      }
      algo = ALGO_STAGED;  // multicast bind failed when the window was created: use the P2P stages
      c->model.has_nvls = false;
    }
  }

  a.nlocal = L;
  a.ctas = sh.ctas;
  a.mode = mode;
  a.apply_scale = apply_scale ? 1 : 0;
  a.scale = scale;
  a.timeout_ns = c->timeout_ns;
  if (mode == MODE_REGISTERED) {
    a.win_off0 = (long long)uoff;
    a.win_stride = 0;
  } else {
    a.win_off0 = 0;
    a.win_stride = (long long)c->window_bytes;
  }
  for (int l = 0; l < L; ++l) {
    RankCall& rc = a.rc[l];
    rc.comm = c->d_comm + l;
    rc.send = s[l];
    rc.recv = r[l];
    const Segment* seg = nullptr;
    if (algo != ALGO_ONESHOT && algo != ALGO_TWOSHOT) seg = mode == MODE_REGISTERED ? &c->user[(size_t)useg].seg[(size_t)l] : &c->win[(size_t)l];
    rc.win = seg ? seg->d_table : nullptr;
    rc.mc = seg ? seg->mc_base : nullptr;
  }

  a.use_tma = tma ? 1 : 0;
  a.nvls_variant = c->nvls_variant >= 0 ? c->nvls_variant : env_int("FT_NVLS_VARIANT", 0);
  a.nvls_flags = c->nvls_flags | env_int("FT_NVLS_FLAGS", 0);
  a.fault_rank_plus1 = c->fault_rank + 1;

  LaunchCtx ctx;
  ctx.args = &a;
  ctx.algo = tma ? (cluster ? (int)ALGO_STAGED_CLUSTER : (int)ALGO_STAGED_TMA) : algo;
  ctx.threads = sh.threads;
  ctx.cooperative = c->kind == COMM_EMULATED && N > 1;
  ctx.stream = stream;
  if (!launch_any(wire_dtype, io_dtype, op, vec, mode, &ctx) || !ctx.supported) {
    set_last_error(std::string("no kernel for ") + op_name(op) + "/" + dtype_name(wire_dtype) + "/" + algo_name(algo));
    return FT_ERR_UNSUPPORTED;
  }
  if (ctx.err != cudaSuccess) {
    set_last_error(std::string("kernel launch: ") + cudaGetErrorString(ctx.err));
    (void)cudaGetLastError();
    return FT_ERR_CUDA;
  }
  ft_call_info_t& li = c->last;
  li = ft_call_info_t{};
  li.algo = algo;
  li.nstages = a.plan.nstages;
  for (int k = 0; k < a.plan.nstages && k < 8; ++k) li.widths[k] = a.plan.width[k];
  li.ctas = sh.ctas;
  li.threads = sh.threads;
  li.tile_bytes = a.lay.tile * es;
  li.mode = mode;
  li.vectorized = vec;
  li.est_us = est;
  FT_LOG(2, "rank %d allreduce count=%zu %s/%s algo=%s widths=%s ctas=%d tile=%lld mode=%d", c->rank, count,
         dtype_name(wire_dtype), op_name(op), algo_name(algo), to_string(widths).c_str(), sh.ctas, li.tile_bytes, mode);
  if (mode == (useg >= 0 ? MODE_REGISTERED : MODE_STAGING) && key.gen == c->policy_gen) {
    // remember the resolved launch (the generation is unchanged, i.e. nothing was (re)allocated on the way)
    PlanEntry e;
    e.key = key;
    e.args = a;
    e.launch_algo = ctx.algo;
    e.threads = sh.threads;
    e.mode = mode;
    e.info = li;
    e.last_use = ++c->use_clock;
    const size_t kMaxPlans = 32;
    if (c->plans.size() < kMaxPlans) {
      c->plans.push_back(e);
    } else {
      size_t victim = 0;
      for (size_t i = 1; i < c->plans.size(); ++i)
        if (c->plans[i].last_use < c->plans[victim].last_use) victim = i;
      c->plans[victim] = e;
    }
  }
  return FT_SUCCESS;
}

// ---- "-k" trees: phantom ranks -------------------------------------------------------------------------------------
// Capability parity: chooseWidth lists the trees of N+1 nodes, tagged "-1", as candidates for N nodes
// (cost_model/ChooseWidth.h:14-31; worked cases cost_model/README.md:13-27: 2*2*2-1, 2*4-1, 4*2-1 for 7 nodes) and the
// reference's data plane carries the hooks for it (mpi_mod.hpp:172-178, :208-211, :981-987, :1041-1099) without a
// working implementation. Here: a tree of T = N + k positions runs on N GPUs; position N+j is a PHANTOM rank hosted by
// process j. A phantom is a complete logical rank - own flag pad, own window slot, own epoch - whose input is the
// operation's identity and whose output goes to a sink buffer; its host launches TWO logical ranks in one cooperative
// launch (rc[0] = itself, rc[1] = the phantom), exactly like the CPU communicator does (csrc/host/host_comm.cc).
// This is how 7 (or 5, or 3) GPUs get a multi-stage tree with no fold step; the price is that the hosts do double duty.
// P2P stages on private buffers only (the multicast mapping covers slot 0 of every segment, not the phantom slots).
}  // namespace

struct ft_comm::PhantomWorld {
  int k = 0;
  size_t window_bytes = 0;         // per parity, per slot
  Segment pad, win;                // 2 slots per process: [0] the process itself, [1] its phantom
  char** d_win_table = nullptr;    // [T] window base of every logical rank
  CommDev* d_comm = nullptr;       // [2]
  uint64_t* d_epoch = nullptr;     // [2]
  unsigned* d_done = nullptr;      // [2]
  char* ident = nullptr;           // identity input of the hosted phantom
  char* sink = nullptr;            // its discarded output
  size_t io_bytes = 0;
  int ident_dtype = -1, ident_op = -1;
  size_t ident_count = 0;
};

namespace {

__global__ void fill_pattern_kernel(char* p, size_t n, int es, unsigned long long pat) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    for (int b = 0; b < es; ++b) p[i * (size_t)es + b] = (char)((pat >> (8 * b)) & 0xff);
  }
}

// Bit pattern of the identity element of `op` for one element of `dtype`.
unsigned long long identity_pattern(int dtype, int op) {
  const int es = dtype_size(dtype);
  const unsigned long long ones = es >= 8 ? ~0ull : ((1ull << (8 * es)) - 1ull);
  if (op == OP_SUM || op == OP_BOR) return 0;
  if (dtype == DT_BOOL) return (op == OP_BAND || op == OP_MIN) ? 1 : 0;
  if (op == OP_BAND) return ones;
  const bool is_max = op == OP_MAX;
  switch (dtype) {
    case DT_FLOAT: return is_max ? 0xff800000ull : 0x7f800000ull;            // -inf / +inf
    case DT_DOUBLE: return is_max ? 0xfff0000000000000ull : 0x7ff0000000000000ull;
    case DT_HALF: return is_max ? 0xfc00ull : 0x7c00ull;
    case DT_BFLOAT16: return is_max ? 0xff80ull : 0x7f80ull;
    case DT_INT8: case DT_INT16: case DT_INT32: case DT_INT64: case DT_LONG_LONG:
      return is_max ? (1ull << (8 * es - 1)) : (ones >> 1);                  // most negative / most positive
    default: return is_max ? 0 : ones;                                       // unsigned
  }
}

void phantom_free(ft_comm::PhantomWorld* w) {
  if (!w) return;
  segment_destroy(&w->pad);
  segment_destroy(&w->win);
  if (w->d_win_table) cudaFree(w->d_win_table);
  if (w->d_comm) cudaFree(w->d_comm);
  if (w->d_epoch) cudaFree(w->d_epoch);
  if (w->d_done) cudaFree(w->d_done);
  if (w->ident) cudaFree(w->ident);
  if (w->sink) cudaFree(w->sink);
  (void)cudaGetLastError();
  delete w;
}

void phantom_release(ft_comm* c) {
  phantom_free(c->ph);
  c->ph = nullptr;
}

// (Re)build the phantom world for k phantoms and a window of at least `need` bytes per parity. Collective: every
// process reaches it with the same arguments (they derive from the call's arguments).
ft_result_t phantom_build(ft_comm* c, int k, size_t need) {
  const int N = c->nranks, T = N + k;
  size_t wb = c->ph && c->ph->k == k ? c->ph->window_bytes : (size_t)(16u << 20);
  while (wb < need) wb *= 2;
  FT_CHECK_RT(cudaSetDevice(c->device));
  FT_CHECK_RT(cudaDeviceSynchronize());
  if (!c->bs.barrier()) { set_last_error("bootstrap barrier failed"); return FT_ERR_INTERNAL; }
  phantom_release(c);
  auto* w = new ft_comm::PhantomWorld;
  w->k = k;
  w->window_bytes = wb;
  std::string err;
  SymmOptions nomc = c->sopt;
  nomc.try_multicast = false;
  nomc.require_multicast = false;
  if (!segment_create(c->bs, c->device, 2 * FT_PAD_BYTES, nomc, &w->pad, &err) ||
      !segment_create(c->bs, c->device, 2 * 2 * wb, nomc, &w->win, &err)) {
    set_last_error("phantom world: " + err);
    phantom_free(w);
    return FT_ERR_CUDA;
  }
  // logical rank L lives in slot (L < N ? 0 : 1) of process (L < N ? L : L - N)
  auto host_of = [&](int L) { return L < N ? L : L - N; };
  auto slot_of = [&](int L) { return L < N ? 0 : 1; };
  std::vector<char*> wt((size_t)T);
  for (int L = 0; L < T; ++L) wt[(size_t)L] = w->win.base[(size_t)host_of(L)] + (size_t)slot_of(L) * 2 * wb;
  FT_CHECK_RT(cudaMalloc(&w->d_win_table, sizeof(char*) * (size_t)T));
  FT_CHECK_RT(cudaMemcpy(w->d_win_table, wt.data(), sizeof(char*) * (size_t)T, cudaMemcpyHostToDevice));
  FT_CHECK_RT(cudaMalloc(&w->d_comm, sizeof(CommDev) * 2));
  FT_CHECK_RT(cudaMalloc(&w->d_epoch, sizeof(uint64_t) * 2));
  FT_CHECK_RT(cudaMalloc(&w->d_done, sizeof(unsigned) * 2));
  FT_CHECK_RT(cudaMemset(w->d_epoch, 0, sizeof(uint64_t) * 2));
  FT_CHECK_RT(cudaMemset(w->d_done, 0, sizeof(unsigned) * 2));
  CommDev h[2];
  for (int l = 0; l < 2; ++l) {
    CommDev& d = h[l];
    memset(&d, 0, sizeof d);
    d.rank = l == 0 ? c->rank : N + c->rank;
    d.nranks = T;
    for (int L = 0; L < T; ++L) d.pad[L] = (uint64_t*)(w->pad.base[(size_t)host_of(L)] + (size_t)slot_of(L) * FT_PAD_BYTES);
    d.epoch = w->d_epoch + l;
    d.done = w->d_done + l;
    d.err = c->d_err;  // one error word (and its host mirror) per process: a phantom's timeout is its host's
    d.err_host = nullptr;
  }
  {  // reuse the communicator's host mirror for the real rank
    CommDev real;
    if (cudaMemcpy(&real, c->d_comm, sizeof real, cudaMemcpyDeviceToHost) == cudaSuccess) h[0].err_host = h[1].err_host = real.err_host;
    (void)cudaGetLastError();
  }
  FT_CHECK_RT(cudaMemcpy(w->d_comm, h, sizeof h, cudaMemcpyHostToDevice));
  c->ph = w;
  FT_LOG(1, "rank %d: phantom world for %d+%d positions, window %zu bytes per parity", c->rank, N, k, wb);
  return c->bs.barrier() ? FT_SUCCESS : FT_ERR_INTERNAL;
}

ft_result_t phantom_allreduce(ft_comm* c, const void* send, void* recv, size_t count, int io_dtype, int wire_dtype, int op,
                              double scale, cudaStream_t stream, const std::vector<int>& widths, int k, bool vec) {
  const int N = c->nranks, T = N + k;
  if (wire_dtype == DT_FP8_BLOCK) { set_last_error("'-k' trees: no fp8 wire"); return FT_ERR_UNSUPPORTED; }
  StagedArgs a;
  memset(&a, 0, sizeof a);
  if (!make_plan(widths.data(), (int)widths.size(), &a.plan) || a.plan.nranks != T) {
    set_last_error("stage widths " + to_string(widths) + " do not multiply to " + std::to_string(T));
    return FT_ERR_INVALID;
  }
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &cap);
  const int es = dtype_size(wire_dtype), ios = dtype_size(io_dtype);
  const size_t msg_bytes = count * (size_t)es;
  const bool hosting = c->rank < k;
  const int want_threads = c->threads_override ? c->threads_override : env_int("FT_THREADS", 512);
  const bool tma = vec && want_threads >= 64 && (count % (size_t)(16 / es) == 0) &&
                   (c->tma_override >= 0 ? c->tma_override : env_int("FT_TMA", 1)) != 0;
  // hosts run two logical ranks in one cooperative launch: every rank plans with half the CTA budget (the tile -> CTA
  // mapping must be the same everywhere)
  Shape sh = pick_shape(c, ALGO_STAGED, MODE_STAGING, tma, msg_bytes, T, es, /*halve the CTA budget=*/true);
  a.lay = make_layout((long long)count, T, es, sh.tile_bytes, vec);
  {
    long long per_cta = (a.lay.ntiles + sh.ctas - 1) / sh.ctas;
    while (per_cta >= (1LL << FT_EPOCH_SHIFT) - 2) {
      sh.tile_bytes *= 2;
      a.lay = make_layout((long long)count, T, es, sh.tile_bytes, vec);
      per_cta = (a.lay.ntiles + sh.ctas - 1) / sh.ctas;
    }
  }
  if (sh.ctas > a.lay.ntiles) sh.ctas = a.lay.ntiles;
  const size_t need = (size_t)a.lay.block * (size_t)T * (size_t)es;
  if (!c->ph || c->ph->k != k || c->ph->window_bytes < need) {
    if (cap != cudaStreamCaptureStatusNone) { set_last_error("'-k' tree: first use / window growth during stream capture"); return FT_ERR_INVALID; }
    ft_result_t r = phantom_build(c, k, need);
    if (r != FT_SUCCESS) return r;
  }
  ft_comm::PhantomWorld* w = c->ph;
  if (hosting) {
    const size_t iob = count * (size_t)ios + 64;
    if (w->io_bytes < iob) {
      if (cap != cudaStreamCaptureStatusNone) { set_last_error("'-k' tree: buffer growth during stream capture"); return FT_ERR_INVALID; }
      FT_CHECK_RT(cudaDeviceSynchronize());
      if (w->ident) cudaFree(w->ident);
      if (w->sink) cudaFree(w->sink);
      w->ident = w->sink = nullptr;
      FT_CHECK_RT(cudaMalloc(&w->ident, iob));
      FT_CHECK_RT(cudaMalloc(&w->sink, iob));
      w->io_bytes = iob;
      w->ident_dtype = -1;
    }
    if (w->ident_dtype != io_dtype || w->ident_op != op || w->ident_count < count) {
      fill_pattern_kernel<<<256, 256, 0, stream>>>(w->ident, count, ios, identity_pattern(io_dtype, op));
      FT_CHECK_RT(cudaGetLastError());
      w->ident_dtype = io_dtype;
      w->ident_op = op;
      w->ident_count = count;
    }
  }
  const bool apply_scale = dtype_is_float(wire_dtype) && scale != 1.0;
  a.nlocal = hosting ? 2 : 1;
  a.ctas = sh.ctas;
  a.mode = MODE_STAGING;
  a.apply_scale = apply_scale ? 1 : 0;
  a.scale = scale;
  a.timeout_ns = c->timeout_ns;
  a.win_off0 = 0;
  a.win_stride = (long long)w->window_bytes;
  a.use_tma = tma ? 1 : 0;
  a.fault_rank_plus1 = c->fault_rank + 1;
  for (int l = 0; l < a.nlocal; ++l) {
    RankCall& rc = a.rc[l];
    rc.comm = w->d_comm + l;
    rc.send = l == 0 ? send : (const void*)w->ident;
    rc.recv = l == 0 ? recv : (void*)w->sink;
    rc.win = w->d_win_table;
    rc.mc = nullptr;
  }
  LaunchCtx ctx;
  ctx.args = &a;
  ctx.algo = tma ? (int)ALGO_STAGED_TMA : (int)ALGO_STAGED;
  ctx.threads = sh.threads;
  ctx.cooperative = a.nlocal > 1;
  ctx.stream = stream;
  if (!cuda_launch(wire_dtype, io_dtype, op, vec, &ctx) || !ctx.supported) {
    set_last_error("'-k' tree: no kernel for this dtype/op");
    return FT_ERR_UNSUPPORTED;
  }
  if (ctx.err != cudaSuccess) {
    set_last_error(std::string("kernel launch: ") + cudaGetErrorString(ctx.err));
    (void)cudaGetLastError();
    return FT_ERR_CUDA;
  }
  ft_call_info_t& li = c->last;
  li = ft_call_info_t{};
  li.algo = ALGO_STAGED;
  li.nstages = a.plan.nstages;
  for (int i = 0; i < a.plan.nstages && i < 8; ++i) li.widths[i] = a.plan.width[i];
  li.ctas = sh.ctas;
  li.threads = sh.threads;
  li.tile_bytes = a.lay.tile * es;
  li.mode = MODE_STAGING;
  li.vectorized = vec;
  return FT_SUCCESS;
}

ft_result_t finish_create(ft_comm* c) {
  // pads + LL buffers exist now; window is allocated lazily at first use (or ft_comm_reserve)
  return upload_comm_dev(c);
}

}  // namespace

namespace ft {
int comm_nlocal(const ft_comm_t* c) { return c->nlocal; }
int comm_device(const ft_comm_t* c) { return c->device; }
}  // namespace ft

extern "C" {

void ft_comm_options_default(ft_comm_options_t* o) {
  o->window_bytes = 0;
  o->ll_bytes = 0;
  o->use_multicast = -1;
  o->use_vmm = -1;
  o->timeout_s = -1.0;
}

ft_result_t ft_comm_create(ft_comm_t** out, int rank, int nranks, int device, ft_allgather_fn ag, void* ag_ctx,
                           const ft_comm_options_t* opts) {
  ft_result_t rc = FT_SUCCESS;
  if (!out || nranks < 1 || nranks > FT_MAX_RANKS || rank < 0 || rank >= nranks || (nranks > 1 && !ag)) {
    fail(&rc, FT_ERR_INVALID, "ft_comm_create: bad arguments");
    return rc;
  }
  auto* c = new ft_comm;
  c->kind = COMM_PROCESS;
  c->rank = rank;
  c->nranks = nranks;
  c->device = device;
  c->nlocal = 1;
  c->bs.rank = rank;
  c->bs.nranks = nranks;
  c->bs.fn = ag;
  c->bs.ctx = ag_ctx;
  apply_options(c, opts);
  std::string err;
  c->pad.assign(1, Segment());
  c->ll.assign(1, Segment());
  SymmOptions nomc = c->sopt;
  nomc.try_multicast = false;
  nomc.require_multicast = false;
  if (!segment_create(c->bs, device, FT_PAD_BYTES, nomc, &c->pad[0], &err) ||
      !segment_create(c->bs, device, 2 * c->ll_bytes, nomc, &c->ll[0], &err)) {
    fail(&rc, FT_ERR_CUDA, "ft_comm_create: " + err);
    delete c;
    return rc;
  }
  rc = finish_create(c);
  if (rc != FT_SUCCESS) { delete c; return rc; }
  // probe multicast once so that the model knows whether NVLS exists (window itself is lazy)
  c->model.has_nvls = c->sopt.try_multicast && device_supports_multicast(device) && c->pad[0].kind == SEG_VMM && nranks > 1;
  FT_LOG(1, "rank %d/%d on device %d: segments via %s, multicast %s", rank, nranks, device,
         c->pad[0].kind == SEG_VMM ? "VMM" : "cudaIpc", c->model.has_nvls ? "available" : "off");
  *out = c;
  return FT_SUCCESS;
}

ft_result_t ft_comm_create_local(ft_comm_t** comms, int nranks, const int* devices, const ft_comm_options_t* opts) {
  ft_result_t rc = FT_SUCCESS;
  if (!comms || nranks < 1 || nranks > FT_MAX_RANKS || !devices) {
    fail(&rc, FT_ERR_INVALID, "ft_comm_create_local: bad arguments");
    return rc;
  }
  auto group = std::make_shared<LocalGroup>();
  group->devices.assign(devices, devices + nranks);
  std::vector<Segment> pads, lls;
  std::string err;
  std::vector<ft_comm*> made;
  for (int i = 0; i < nranks; ++i) {
    auto* c = new ft_comm;
    c->kind = COMM_LOCAL;
    c->rank = i;
    c->nranks = nranks;
    c->device = devices[i];
    c->nlocal = 1;
    c->group = group;
    apply_options(c, opts);
    made.push_back(c);
  }
  group->members = made;
  // flag pads and LL buffers: VMM when available (same kind of memory as the windows), never multicast-bound
  SymmOptions nomc = made[0]->sopt;
  nomc.try_multicast = false;
  nomc.require_multicast = false;
  if (!segment_create_local(group->devices, FT_PAD_BYTES, &pads, &err, &nomc) ||
      !segment_create_local(group->devices, 2 * made[0]->ll_bytes, &lls, &err, &nomc)) {
    fail(&rc, FT_ERR_CUDA, "ft_comm_create_local: " + err);
    for (auto* c : made) delete c;
    return rc;
  }
  for (int i = 0; i < nranks; ++i) {
    made[(size_t)i]->pad.assign(1, pads[(size_t)i]);
    made[(size_t)i]->ll.assign(1, lls[(size_t)i]);
    rc = finish_create(made[(size_t)i]);
    if (rc != FT_SUCCESS) return rc;
    comms[i] = made[(size_t)i];
  }
  // One process owns all the devices, so the multicast object can be built right here (cuMulticastCreate + AddDevice
  // + BindMem for every device, no fd passing): C programs without a launcher get the NVLS kernels too.
  bool mc = made[0]->sopt.try_multicast && nranks > 1 && pads[0].kind == SEG_LOCAL_VMM;
  for (int i = 0; i < nranks && mc; ++i) mc = device_supports_multicast(devices[i]);
  for (auto* c : made) c->model.has_nvls = mc;
  FT_LOG(1, "local group of %d devices: segments via %s, multicast %s", nranks, pads[0].kind == SEG_LOCAL_VMM ? "VMM" : "cudaMalloc",
         mc ? "available" : "off");
  return FT_SUCCESS;
}

ft_result_t ft_comm_create_emulated(ft_comm_t** out, int nranks, int device, const ft_comm_options_t* opts) {
  ft_result_t rc = FT_SUCCESS;
  if (!out || nranks < 1 || nranks > FT_MAX_LOCAL) {
    fail(&rc, FT_ERR_INVALID, "ft_comm_create_emulated: nranks must be in [1, 32]");
    return rc;
  }
  auto* c = new ft_comm;
  c->kind = COMM_EMULATED;
  c->rank = 0;
  c->nranks = nranks;
  c->device = device;
  c->nlocal = nranks;
  apply_options(c, opts);
  std::string err;
  if (!make_segments(c, FT_PAD_BYTES, &c->pad, &err) || !make_segments(c, 2 * c->ll_bytes, &c->ll, &err)) {
    fail(&rc, FT_ERR_CUDA, "ft_comm_create_emulated: " + err);
    delete c;
    return rc;
  }
  rc = finish_create(c);
  if (rc != FT_SUCCESS) { delete c; return rc; }
  c->model.has_nvls = false;
  // every "link" is local HBM here: no fan-in effect, cheap hops -> the flat plan wins
  c->model.beta_p2p_GBps = c->model.beta_p2p_1peer_GBps = 600.0;
  c->model.alpha_sync_us = 1.5;
  c->model.alpha_peer_us = 0.2;
  *out = c;
  return FT_SUCCESS;
}

ft_result_t ft_comm_destroy(ft_comm_t* c) {
  if (!c) return FT_SUCCESS;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  host_pipe_release(c);
  // every rank has drained its device before anybody unmaps anything: a staging-mode kernel ends without an exit
  // barrier (the next call's window parity protects it), so a peer may still be pulling from my windows right now
  if (c->kind == COMM_PROCESS) c->bs.barrier();
  phantom_release(c);
  destroy_segments(&c->win);
  destroy_segments(&c->pad);
  destroy_segments(&c->ll);
  for (auto& u : c->user) destroy_segments(&u.seg);
  if (c->d_comm) cudaFree(c->d_comm);
  if (c->d_epoch) cudaFree(c->d_epoch);
  if (c->d_done) cudaFree(c->d_done);
  if (c->d_err) cudaFree(c->d_err);
  if (c->d_trace) cudaFree(c->d_trace);
  if (c->h_err) cudaFreeHost(c->h_err);
  (void)cudaGetLastError();
  if (c->group) {
    std::lock_guard<std::mutex> lk(c->group->mu);
    for (auto& m : c->group->members)
      if (m == c) m = nullptr;
    std::vector<ft_comm*> keep;
    for (auto* m : c->group->members)
      if (m) keep.push_back(m);
    c->group->members = keep;
  }
  delete c;
  return FT_SUCCESS;
}

int ft_comm_rank(const ft_comm_t* c) { return c->rank; }
int ft_comm_size(const ft_comm_t* c) { return c->nranks; }
int ft_comm_has_multicast(const ft_comm_t* c) { return c->model.has_nvls ? 1 : 0; }

unsigned ft_comm_error(ft_comm_t* c) {
  std::vector<unsigned> h((size_t)c->nlocal, 0);
  cudaSetDevice(c->device);
  if (cudaMemcpy(h.data(), c->d_err, sizeof(unsigned) * (size_t)c->nlocal, cudaMemcpyDeviceToHost) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0xffffffffu;
  }
  for (unsigned e : h)
    if (e) return e;
  return 0;
}
unsigned ft_comm_error_peek(const ft_comm_t* c) {
  if (!c || !c->h_err) return 0;
  for (int l = 0; l < c->nlocal; ++l) {
    const unsigned e = *(volatile unsigned*)(c->h_err + l);
    if (e) return e;
  }
  return 0;
}
ft_result_t ft_comm_clear_error(ft_comm_t* c) {
  FT_CHECK_RT(cudaSetDevice(c->device));
  FT_CHECK_RT(cudaDeviceSynchronize());  // a kernel that is still spinning would set the words again
  FT_CHECK_RT(cudaMemset(c->d_err, 0, sizeof(unsigned) * (size_t)c->nlocal));
  if (c->h_err) memset(c->h_err, 0, sizeof(unsigned) * (size_t)c->nlocal);
  return FT_SUCCESS;
}
ft_result_t ft_comm_set_tuning(ft_comm_t* c, const char* key, long long value) {
  if (!c || !key) return FT_ERR_INVALID;
  c->policy_gen++;
  const std::string k = key;
  if (k == "nvls_variant") { c->nvls_variant = (int)value; return FT_SUCCESS; }
  if (k == "nvls_flags") { c->nvls_flags = (int)value; return FT_SUCCESS; }
  if (k == "nvls_max_tile_bytes") { c->nvls_max_tile = value; return FT_SUCCESS; }
  if (k == "nvls_ctas") { c->nvls_ctas = (int)value; return FT_SUCCESS; }
  if (k == "cluster") { c->cluster = (int)value; return FT_SUCCESS; }
  set_last_error("ft_comm_set_tuning: unknown key " + k);
  return FT_ERR_INVALID;
}
ft_result_t ft_link_probe(ft_comm_t* c, int mode, int npeers, size_t bytes_per_peer, int ctas, void* stream) {
  if (!c || c->kind == COMM_EMULATED || mode < 0 || mode > 3 || npeers < 1 || npeers >= c->nranks || ctas < 1 ||
      bytes_per_peer < 16 || bytes_per_peer % 16) {
    set_last_error("ft_link_probe: needs a multi-GPU communicator, mode 0-3, 1 <= npeers < nranks, 16-byte multiple");
    return FT_ERR_INVALID;
  }
  const size_t need = bytes_per_peer * (size_t)c->nranks;  // one slot per source rank
  ft_result_t r = grow_window(c, (need + 1) / 2);            // the window has two parity halves: use both
  if (r != FT_SUCCESS) return r;
  FT_CHECK_RT(cudaSetDevice(c->device));
  cudaError_t e = launch_link_probe(c->win[0].d_table, c->rank, c->nranks, npeers, (long long)bytes_per_peer, mode, ctas,
                                    c->d_done /* never written in practice */, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_last_error(std::string("ft_link_probe: ") + cudaGetErrorString(e));
    (void)cudaGetLastError();
    return FT_ERR_CUDA;
  }
  return FT_SUCCESS;
}
ft_result_t ft_comm_inject_fault(ft_comm_t* c, int rank) {
  if (!c || rank >= c->nranks) return FT_ERR_INVALID;
  c->fault_rank = rank < 0 ? -1 : rank;
  return FT_SUCCESS;
}

ft_result_t ft_comm_set_widths(ft_comm_t* c, const int* widths, int nstages) {
  c->policy_gen++;
  if (!widths || nstages <= 0) {
    c->widths_override.clear();
    c->lonely_override = 0;
    c->phantom_override = 0;
    return FT_SUCCESS;
  }
  std::vector<int> w(widths, widths + nstages);
  bool ring = false;
  for (int x : w) ring = ring || x == 1;
  if (ring && c->nranks > 1) {
    c->widths_override = {1};
    c->lonely_override = 0;
    return FT_SUCCESS;
  }
  Plan p;
  if (!make_plan(w.data(), nstages, &p)) {
    set_last_error("invalid widths " + to_string(w) + " for world size " + std::to_string(c->nranks));
    return FT_ERR_INVALID;
  }
  if (p.nranks > c->nranks) {
    // A tree LARGER than the world: the reference's "-k" trees (2*4-1 for 7 nodes, cost_model/README.md:13-27). The k
    // extra positions are phantom ranks hosted by the first k processes (one process per GPU only).
    if (c->kind != COMM_PROCESS || p.nranks > 2 * c->nranks || p.nranks > FT_MAX_RANKS) {
      set_last_error("widths " + to_string(w) + " multiply to more than the world size " + std::to_string(c->nranks) +
                     ": '-k' trees need a one-process-per-GPU communicator and k <= world size");
      return FT_ERR_INVALID;
    }
    c->widths_override = w;
    c->lonely_override = 0;
    c->phantom_override = p.nranks - c->nranks;
    return FT_SUCCESS;
  }
  c->phantom_override = 0;
  // A tree smaller than the world leaves N - T "lonely" ranks (the reference's "+k" trees,
  // cost_model/ChooseWidth.h:16-31): they feed their vector to the tree and gather the result from it.
  c->widths_override = w;
  c->lonely_override = c->nranks - p.nranks;
  return FT_SUCCESS;
}
ft_result_t ft_comm_set_algo(ft_comm_t* c, ft_algo_t algo) {
  c->policy_gen++;
  if ((int)algo < 0 || ((int)algo > ALGO_RING && (int)algo != ALGO_TWOSHOT)) return FT_ERR_INVALID;
  c->algo_override = (int)algo;
  return FT_SUCCESS;
}
ft_result_t ft_comm_set_link_model(ft_comm_t* c, const ft_link_model_t* m) {
  c->policy_gen++;
  const bool probed = c->model.has_nvls;  // what the hardware/bootstrap allows
  c->model = m ? m->m : LinkModel();
  c->model.has_nvls = probed;
  return FT_SUCCESS;
}
ft_result_t ft_comm_set_launch(ft_comm_t* c, int ctas, int threads, long long tile_bytes) {
  c->policy_gen++;
  if (ctas < 0 || ctas > FT_MAX_CTAS || threads < 0 || threads > FT_MAX_THREADS || (threads % 32) || tile_bytes < 0)
    return FT_ERR_INVALID;
  c->ctas_override = ctas;
  c->threads_override = threads;
  c->tile_override = tile_bytes;
  return FT_SUCCESS;
}

// Phase timing (SHOW_TIME parity): device timestamps of CTA 0 at every phase boundary of the staged kernel.
ft_result_t ft_comm_trace_enable(ft_comm_t* c, int on) {
  FT_CHECK_RT(cudaSetDevice(c->device));
  FT_CHECK_RT(cudaDeviceSynchronize());
  const size_t bytes = sizeof(uint64_t) * FT_TRACE_SLOTS * (size_t)c->nlocal;
  if (on && !c->d_trace) {
    FT_CHECK_RT(cudaMalloc(&c->d_trace, bytes));
    FT_CHECK_RT(cudaMemset(c->d_trace, 0, bytes));
  }
  std::vector<CommDev> h((size_t)c->nlocal);
  FT_CHECK_RT(cudaMemcpy(h.data(), c->d_comm, sizeof(CommDev) * (size_t)c->nlocal, cudaMemcpyDeviceToHost));
  for (int l = 0; l < c->nlocal; ++l) h[(size_t)l].trace = on ? c->d_trace + (size_t)l * FT_TRACE_SLOTS : nullptr;
  FT_CHECK_RT(cudaMemcpy(c->d_comm, h.data(), sizeof(CommDev) * (size_t)c->nlocal, cudaMemcpyHostToDevice));
  return FT_SUCCESS;
}
// ns[slot] for local rank `local` of the last traced call; returns the number of valid slots (0 if off).
int ft_comm_trace_read(ft_comm_t* c, int local, unsigned long long* ns, int cap) {
  if (!c->d_trace || local < 0 || local >= c->nlocal) return 0;
  uint64_t h[FT_TRACE_SLOTS];
  cudaSetDevice(c->device);
  if (cudaMemcpy(h, c->d_trace + (size_t)local * FT_TRACE_SLOTS, sizeof h, cudaMemcpyDeviceToHost) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  int n = (int)h[FT_TRACE_SLOTS - 1];
  if (n > FT_TRACE_SLOTS - 1) n = FT_TRACE_SLOTS - 1;
  for (int i = 0; i < n && i < cap; ++i) ns[i] = h[i];
  return n;
}

ft_result_t ft_comm_set_max_ctas(ft_comm_t* c, int max_ctas) {
  c->policy_gen++;
  if (max_ctas < 0) return FT_ERR_INVALID;
  c->max_ctas = max_ctas;
  return FT_SUCCESS;
}

ft_result_t ft_comm_set_tma(ft_comm_t* c, int on) {
  c->policy_gen++;
  c->tma_override = on < 0 ? -1 : (on ? 1 : 0);
  return FT_SUCCESS;
}

ft_result_t ft_comm_reserve(ft_comm_t* c, size_t bytes) {
  // room for the 16-byte block rounding of the widest split
  size_t need = bytes + (size_t)c->nranks * 16 + 64;
  return grow_window(c, need);
}

ft_result_t ft_symm_alloc(ft_comm_t* c, size_t bytes, void** ptr) {
  c->policy_gen++;
  if (!ptr || bytes == 0) return FT_ERR_INVALID;
  size_t padded = ((bytes + 4095) & ~(size_t)4095) + 4096;  // tail pad: full-vector accesses stay inside
  std::string err;
  if (c->kind == COMM_LOCAL) {
    std::lock_guard<std::mutex> lk(c->group->mu);
    if (c->local_alloc_cursor == c->user.size()) {  // first member to ask allocates for everybody
      std::vector<Segment> segs;
      if (!segment_create_local(c->group->devices, padded, &segs, &err, &c->sopt)) {
        set_last_error(err);
        return FT_ERR_CUDA;
      }
      for (size_t i = 0; i < c->group->members.size(); ++i) {
        ft_comm::UserSeg u;
        u.seg.assign(1, segs[i]);
        u.bytes = bytes;
        c->group->members[i]->user.push_back(u);
      }
    }
    ft_comm::UserSeg& u = c->user[c->local_alloc_cursor++];
    if (u.bytes != bytes) {
      set_last_error("ft_symm_alloc: members of a local group must allocate in the same order and size");
      return FT_ERR_INVALID;
    }
    *ptr = u.seg[0].base[(size_t)u.seg[0].self];
    cudaSetDevice(c->device);
    return FT_SUCCESS;
  }
  ft_comm::UserSeg u;
  u.bytes = bytes;
  if (!make_segments(c, padded, &u.seg, &err)) {
    set_last_error(err);
    return FT_ERR_CUDA;
  }
  c->user.push_back(u);
  if (c->kind == COMM_EMULATED) {
    // ptr[0..nranks): one buffer per logical rank
    for (int l = 0; l < c->nlocal; ++l) ptr[l] = u.seg[(size_t)l].base[(size_t)l];
  } else {
    *ptr = u.seg[0].base[(size_t)c->rank];
  }
  return FT_SUCCESS;
}

ft_result_t ft_symm_free(ft_comm_t* c, void* ptr) {
  c->policy_gen++;
  for (size_t i = 0; i < c->user.size(); ++i) {
    Segment& s = c->user[i].seg[0];
    if (s.kind != SEG_NONE && s.base[(size_t)s.self] == (char*)ptr) {
      cudaSetDevice(c->device);
      cudaDeviceSynchronize();
      if (c->kind == COMM_PROCESS) c->bs.barrier();
      destroy_segments(&c->user[i].seg);
      // keep the slot (indices are stable); mark as dead with an empty segment list per rank
      c->user[i].seg.assign((size_t)c->nlocal, Segment());
      for (auto& d : c->user[i].seg) { d.base.assign((size_t)c->nranks, nullptr); d.self = 0; d.bytes = 0; }
      c->user[i].bytes = 0;
      return FT_SUCCESS;
    }
  }
  set_last_error("ft_symm_free: pointer was not returned by ft_symm_alloc");
  return FT_ERR_INVALID;
}

ft_result_t ft_allreduce(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op, ft_comm_t* comm,
                         void* stream) {
  return ft_allreduce_ex(sendbuf, recvbuf, count, dtype, dtype, op, 1.0, comm, stream);
}

ft_result_t ft_allreduce_ex(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_dtype_t wire_dtype,
                            ft_op_t op, double scale, ft_comm_t* comm, void* stream) {
  if (!comm) return FT_ERR_INVALID;
  if (comm->kind == COMM_EMULATED) {
    set_last_error("emulated communicators take one buffer per logical rank: use ft_allreduce_multi");
    return FT_ERR_INVALID;
  }
  const void* s[1] = {sendbuf};
  void* r[1] = {recvbuf};
  return do_allreduce(comm, s, r, count, (int)dtype, (int)wire_dtype, (int)op, scale, (cudaStream_t)stream);
}

// MPI-style blocking form on the default stream: when it returns, recvbuf holds the result (what a captured
// MPI_Allreduce on device memory needs, compat/mpi_mod.hpp).
ft_result_t ft_allreduce_blocking(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op, ft_comm_t* comm) {
  ft_result_t r = ft_allreduce(sendbuf, recvbuf, count, dtype, op, comm, nullptr);
  if (r != FT_SUCCESS) return r;
  FT_CHECK_RT(cudaStreamSynchronize(nullptr));
  const unsigned e = ft_comm_error_peek(comm);
  if (e) {
    set_last_error("device watchdog fired during the collective");
    return FT_ERR_TIMEOUT;
  }
  return FT_SUCCESS;
}

// -1: host (or unknown) memory; otherwise the ordinal of the device the allocation lives on.
int ft_pointer_device(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? a.device : -1;
}

ft_result_t ft_allreduce_multi(const void* const* sendbufs, void* const* recvbufs, size_t count, ft_dtype_t dtype,
                               ft_dtype_t wire_dtype, ft_op_t op, double scale, ft_comm_t* comm, void* stream) {
  if (!comm || !recvbufs) return FT_ERR_INVALID;
  const void* s[FT_MAX_LOCAL];
  for (int l = 0; l < comm->nlocal; ++l) s[l] = sendbufs ? sendbufs[l] : FT_IN_PLACE;
  return do_allreduce(comm, s, recvbufs, count, (int)dtype, (int)wire_dtype, (int)op, scale, (cudaStream_t)stream);
}

// The two halves of the algorithm as collectives (reduce-scatter: recv = my block of the reduction;
// all-gather: send = my block). Emulated communicators pass one pointer per logical rank.
static ft_result_t half_collective(const void* const* sendbufs, void* const* recvbufs, size_t per_rank_count,
                                   ft_dtype_t dtype, ft_op_t op, ft_comm_t* comm, void* stream, int phases) {
  if (!comm || !sendbufs || !recvbufs) return FT_ERR_INVALID;
  return do_allreduce(comm, sendbufs, recvbufs, per_rank_count * (size_t)comm->nranks, (int)dtype, (int)dtype, (int)op, 1.0,
                      (cudaStream_t)stream, phases);
}
ft_result_t ft_reduce_scatter(const void* sendbuf, void* recvbuf, size_t recvcount, ft_dtype_t dtype, ft_op_t op,
                              ft_comm_t* comm, void* stream) {
  if (comm && comm->kind == COMM_EMULATED) { set_last_error("emulated communicator: use ft_reduce_scatter_multi"); return FT_ERR_INVALID; }
  const void* s[1] = {sendbuf};
  void* r[1] = {recvbuf};
  return half_collective(s, r, recvcount, dtype, op, comm, stream, PHASES_REDUCE_SCATTER);
}
ft_result_t ft_all_gather(const void* sendbuf, void* recvbuf, size_t sendcount, ft_dtype_t dtype, ft_comm_t* comm,
                          void* stream) {
  if (comm && comm->kind == COMM_EMULATED) { set_last_error("emulated communicator: use ft_all_gather_multi"); return FT_ERR_INVALID; }
  const void* s[1] = {sendbuf};
  void* r[1] = {recvbuf};
  return half_collective(s, r, sendcount, dtype, FT_SUM, comm, stream, PHASES_ALL_GATHER);
}
ft_result_t ft_reduce_scatter_multi(const void* const* sendbufs, void* const* recvbufs, size_t recvcount, ft_dtype_t dtype,
                                    ft_op_t op, ft_comm_t* comm, void* stream) {
  return half_collective(sendbufs, recvbufs, recvcount, dtype, op, comm, stream, PHASES_REDUCE_SCATTER);
}
ft_result_t ft_all_gather_multi(const void* const* sendbufs, void* const* recvbufs, size_t sendcount, ft_dtype_t dtype,
                                ft_comm_t* comm, void* stream) {
  return half_collective(sendbufs, recvbufs, sendcount, dtype, FT_SUM, comm, stream, PHASES_ALL_GATHER);
}

ft_result_t ft_comm_last_call(const ft_comm_t* c, ft_call_info_t* out) {
  if (!c || !out) return FT_ERR_INVALID;
  *out = c->last;
  return FT_SUCCESS;
}

}  // extern "C"
