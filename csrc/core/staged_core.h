// The generalized multi-stage reduce-scatter + all-gather, written once for two back ends:
//   * CudaBE (kernels/backend.cuh): one CTA of the sm_100a kernel, peers are NVLink-mapped pointers
//   * CpuBE  (host/cpu_backend.h) : one std::thread per "CTA"; used by the CPU tests to check the
//                                   index math, the hazard analysis and the flag protocol without a GPU
//
// Capability parity: FlexTree::tree_allreduce (allreduce_over_mpi/mpi_mod.hpp:953-1111) -
//   reduce-scatter loop :988-1029, all-gather loop :1050-1060, k-way reduce handle_reduce :769-878.
// Design differences (deliberate):
//   * receiver-PULL over peer-mapped memory instead of send/recv messages; the k-way reduction is
//     done in registers in the same loop that loads the peer data (no staging copy, no second pass)
//   * the message is cut into tiles; tile t is a self-contained allreduce over chunk t of every block,
//     always owned by CTA (t mod G) on every rank, so stages pipeline tile by tile and the only
//     cross-rank dependencies are between same-index CTAs (deadlock-free while grids are resident)
//   * stage fences are per-(phase, CTA, source) release/acquire flags instead of world barriers
//   * summation order inside a stage is "own value first, then peers in ascending rank order", the
//     same order handle_reduce uses (:787, :809-817), so fp32 results match the reference bit for bit
//     whenever the block split is the same.
//
// Hazard analysis for the in-place window W (r = this rank, gn = g_{k+1}):
//   RS stage k : r writes W_r[b == r (mod gn)], peers p read W_r[b == p (mod gn)], p != r  -> disjoint
//   AG stage k : r writes W_r[b == p (mod gn)] only after p's flag for a *later* phase of the same
//                tile, which p's CTA releases after its own RS-stage-k reads of W_r         -> ordered
//   next call  : staging mode alternates two windows by epoch parity (any allreduce makes every rank
//                depend on every rank having *started* the call, hence finished the previous one);
//                registered mode ends with an exit barrier instead.
#pragma once
#include "comm_dev.h"
#include "types.h"

#define FT_MAX_THREADS_CORE 512  // == FT_MAX_THREADS (kernels/launch.h): CTA size the streaming NVLS tile of the general kernel is compiled for

namespace ft {

#if defined(__CUDACC__)
#define FT_DEV __device__ __forceinline__
#define FT_NOINLINE __device__ __noinline__
#else
#define FT_DEV inline
#define FT_NOINLINE inline
#endif

template <typename X, typename Y> struct SameType { static constexpr bool value = false; };
template <typename X> struct SameType<X, X> { static constexpr bool value = true; };

template <class BE, typename T, typename IO, class Op, int VE>
struct StagedCta {
  using EL = Elem<T>;
  using EU = Elem<IO>;
  using A = typename EL::acc;
  using PackT = Pack<T, VE>;
  using PackIO = Pack<IO, VE>;
  static constexpr bool kRawIO = SameType<T, IO>::value;

  BE& be;
  const StagedArgs& a;
  const RankCall& rc;
  const CommDev& cd;
  const int r;       // my rank
  const int cta;     // my CTA index within the rank
  uint64_t E;        // flag base of this call
  long long woff;    // byte offset of the window inside the segment
  int* abort_flag;
  bool use_tma_;     // CUDA only: TMA-staged tiles (see kernels/tma_tiles.cuh)
  bool scale_out_;   // wire_to_io applies the scale (one-shot only: the other paths scale when a block becomes final)
  uint32_t pipe_it;  // chunks pushed through the shared-memory ring so far (same value in every thread)

  FT_DEV StagedCta(BE& be_, const StagedArgs& a_, const RankCall& rc_, int cta_)
      : be(be_), a(a_), rc(rc_), cd(*rc_.comm), r(rc_.comm->rank), cta(cta_) {
    uint64_t epoch = be.load_epoch(cd.epoch);
    E = epoch_base(epoch);
    woff = a.win_off0 + (long long)(epoch & 1) * a.win_stride;
    abort_flag = be.abort_ptr();
    // bulk copies need 16-byte multiples: the ragged last vector of an odd-sized message stays on the LDG path
    use_tma_ = BE::kTma && a.use_tma && VE > 1 && (a.lay.count % VE == 0);  // host guarantees this for the TMA kernel
    pipe_it = 0;
    scale_out_ = false;
  }

  // ---- phase timestamps (opt-in) -----------------------------------------------------------------
  FT_DEV void trace(int slot) {
    if (cd.trace && cta == 0 && be.tid() == 0 && slot < FT_TRACE_SLOTS - 1) {
      cd.trace[slot] = be.now_ns();
      cd.trace[FT_TRACE_SLOTS - 1] = (uint64_t)(slot + 1);
    }
  }

  // ---- flags --------------------------------------------------------------------------------------
  FT_DEV void spin(const uint64_t* slot, uint64_t target, int phase, int src) {
    if (*abort_flag) return;
    uint64_t t0 = 0;
    unsigned it = 0;
    while (be.flag_ld(slot) < target) {
      if (((++it) & 0x3ff) == 0) {
        if (be.err_ld(cd.err)) { *abort_flag = 1; return; }
        if (a.timeout_ns) {
          uint64_t now = be.now_ns();
          if (!t0) t0 = now;
          else if (now - t0 > a.timeout_ns) {
            be.err_set(cd, (unsigned)ERR_TIMEOUT | ((unsigned)phase << 8) | ((unsigned)src << 16) |
                                   ((unsigned)r << 24));
            *abort_flag = 1;
            return;
          }
        }
      }
      be.pause();
    }
  }

  // Tell every other member of my stage-k group that `phase` reached `val` on my side.
  // Precondition: be.sync() after the data writes being published.
  FT_DEV void signal_group(int phase, int k, uint64_t val) {
    const int w = a.plan.width[k], g = a.plan.gap[k], base = stage_base(a.plan, r, k);
    for (int j = be.tid(); j < w; j += be.nthr()) {
      int p = base + j * g;
      if (p == r) continue;
      be.fence();
      be.flag_st(cd.pad[p] + pad_slot(phase, cta, r), val);
    }
  }
  // Wait until every other member of my stage-k group has signalled `phase` >= val to me.
  FT_DEV void wait_group(int phase, int k, uint64_t val) {
    const int w = a.plan.width[k], g = a.plan.gap[k], base = stage_base(a.plan, r, k);
    for (int j = be.tid(); j < w; j += be.nthr()) {
      int p = base + j * g;
      if (p == r) continue;
      spin(cd.pad[r] + pad_slot(phase, cta, p), val, phase, p);
      be.acquire(cd.pad[r] + pad_slot(phase, cta, p));
    }
    be.sync();
  }

  // ---- element access -------------------------------------------------------------------------------
  FT_DEV static PackT ld_win(const char* base, long long byte_off, bool remote) {
    PackT v;
    BE::template ld<(int)sizeof(PackT)>(&v, base + byte_off, remote);
    return v;
  }
  FT_DEV static void st_win(char* base, long long byte_off, const PackT& v) {
    BE::template st<(int)sizeof(PackT)>(base + byte_off, &v);
  }
  // nvalid < VE only for the very last vector of the message: user buffers hold exactly `count`
  // elements, so the ragged end is touched element by element.
  FT_DEV static PackIO ld_user(const IO* x, long long elem, int nvalid) {
    PackIO v;
    if (nvalid >= VE) {
      BE::template ld<(int)sizeof(PackIO)>(&v, x + elem, false);
    } else {
      for (int i = 0; i < VE; ++i) {
        if (i < nvalid) BE::template ld<(int)sizeof(IO)>(&v.e[i], x + elem + i, false);
        else v.e[i] = v.e[0];
      }
    }
    return v;
  }
  FT_DEV static void st_user(IO* y, long long elem, const PackIO& v, int nvalid) {
    if (nvalid >= VE) {
      BE::template st<(int)sizeof(PackIO)>(y + elem, &v);
    } else {
      for (int i = 0; i < nvalid; ++i) BE::template st<(int)sizeof(IO)>(y + elem + i, &v.e[i]);
    }
  }
  FT_DEV static PackT io_to_wire(const PackIO& x) {
    PackT o;
#pragma unroll
    for (int i = 0; i < VE; ++i) {
      if constexpr (kRawIO) o.e[i] = x.e[i];
      else o.e[i] = EL::down((A)EU::up(x.e[i]));
    }
    return o;
  }
  FT_DEV PackIO wire_to_io(const PackT& x) const {
    PackIO o;
#pragma unroll
    for (int i = 0; i < VE; ++i) {
      if constexpr (EL::is_float) {
        A v = EL::up(x.e[i]);
        if (scale_out_) v = (A)(v * (A)a.scale);
        o.e[i] = EU::down((typename EU::acc)v);
      } else {
        if constexpr (kRawIO) o.e[i] = x.e[i];
        else o.e[i] = EU::down((typename EU::acc)EL::up(x.e[i]));
      }
    }
    return o;
  }

  // The fused 1/N (or any) scale is applied exactly once per element, at the moment its block becomes final
  // (last reduce-scatter stage / lonely fold / NVLS reduce / last ring reduce step), in the accumulator
  // precision and before the rounding to the wire type. Later phases only copy.
  FT_DEV void scale_acc(A* acc) const {
    if constexpr (EL::is_float) {
      if (a.apply_scale) {
#pragma unroll
        for (int i = 0; i < VE; ++i) acc[i] = (A)(acc[i] * (A)a.scale);
      }
    }
  }

  // Valid element count of block b inside tile t, relative to the tile start (may be <= 0).
  FT_DEV long long tile_valid(long long b, int t, long long* start) const {
    long long lo, hi;
    tile_range(a.lay, b, t, &lo, &hi);
    *start = lo;
    return hi - lo;
  }
  // Same, as a 32-bit count clamped to [0, tile]: jobs are indexed with 32-bit math (64-bit division
  // costs ~40 registers per use and made the reduce loops spill).
  FT_DEV int tile_valid32(int b, int t, long long* start) const {
    long long n = tile_valid(b, t, start);
    return n <= 0 ? 0 : (int)n;
  }

  // Streaming copy of one contiguous run of a tile: n elements starting at element `start` (16-byte aligned in the
  // window; the user side may be ragged only in the very last vector of the message). One base address per batch
  // and no per-vector index arithmetic: the copy passes around the NVLS reduce (private buffers) were ALU-bound on
  // `q / CV` divisions before (819 us per 256 MiB on 8 GPUs with the reduce itself at 575 us).
  //   DIR 0: user send -> my window (io_to_wire)      DIR 1: my window -> user recv (wire_to_io)
  template <int UNR, int DIR> FT_DEV void copy_run(long long start, int n) {
    if (n <= 0) return;
    const IO* x = (const IO*)rc.send;
    IO* y = (IO*)rc.recv;
    char* wr = rc.win[r] + woff;
    const int nth = be.nthr(), tid = be.tid();
    const int nfull = n / VE;  // whole vectors; at most one ragged vector follows
    int v0 = tid;
    for (; v0 + (UNR - 1) * nth < nfull; v0 += UNR * nth) {
      if constexpr (DIR == 0) {
        PackIO v[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) v[u] = ld_user(x, start + (long long)(v0 + u * nth) * VE, VE);
#pragma unroll
        for (int u = 0; u < UNR; ++u) st_win(wr, (start + (long long)(v0 + u * nth) * VE) * (long long)sizeof(T), io_to_wire(v[u]));
      } else {
        PackT v[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) v[u] = ld_win(wr, (start + (long long)(v0 + u * nth) * VE) * (long long)sizeof(T), true);
#pragma unroll
        for (int u = 0; u < UNR; ++u) st_user(y, start + (long long)(v0 + u * nth) * VE, wire_to_io(v[u]), VE);
      }
    }
    for (; v0 * VE < n; v0 += nth) {
      const long long el = start + (long long)v0 * VE;
      const int nv = n - v0 * VE < VE ? n - v0 * VE : VE;
      if constexpr (DIR == 0) st_win(wr, el * (long long)sizeof(T), io_to_wire(ld_user(x, el, nv)));
      else st_user(y, el, wire_to_io(ld_win(wr, el * (long long)sizeof(T), true)), nv);
    }
  }

  // ---- phase IN: private user buffer -> my window (only the blocks some peer will pull) -------------
  template <int UNR> FT_DEV void copy_in_tile(int t) {
    const int N = a.plan.nranks, g1 = a.plan.gap[1];
    for (int b = 0; b < N; ++b) {
      if ((b % g1) == (r % g1)) continue;  // blocks I own after stage 0 are read straight from `send`
      long long start;
      const int n = tile_valid32(b, t, &start);
      copy_run<UNR, 0>(start, n);
    }
  }

  // ---- reduce-scatter stage k on tile t ---------------------------------------------------------------
  // For every block b == r (mod g_{k+1}): W_r[b] = own[b] (+) peers' W_p[b], peers ascending.
  // Each thread works on UNR vectors at a time and pulls PB peers per batch, so UNR*PB 16-byte NVLink
  // loads are in flight per thread before the first add (the pulls are latency-, not issue-bound).
  FT_DEV const char* rs_peer_ptr(int k, int j) const {  // j-th peer in ascending order, self skipped
    const int g = a.plan.gap[k], base = stage_base(a.plan, r, k);
    const int p = base + (j + (base + j * g >= r ? 1 : 0)) * g;
    return rc.win[p] + woff;
  }
  template <int UNR, int PB> FT_DEV void rs_tile(int k, int t) {
    const Plan& P = a.plan;
    const Layout& L = a.lay;
    const int w = P.width[k], gn = P.gap[k + 1];
    const int nb = P.nranks / gn, rmod = r % gn;
    const int CV = (int)((L.tile + VE - 1) / VE);
    const int total = nb * CV;
    const int npeers = w - 1;
    const char* pp[PB];  // first batch of peers kept in registers; later batches are looked up
#pragma unroll
    for (int j = 0; j < PB; ++j) pp[j] = (j < npeers) ? rs_peer_ptr(k, j) : nullptr;
    char* wr = rc.win[r] + woff;
    const IO* x = (const IO*)rc.send;
    const bool own_from_user = (k == 0) && (a.mode == MODE_STAGING);
    const bool final_stage = (k == P.nstages - 1) && a.lonely == 0;  // my blocks become final here
    const int nth = be.nthr(), tid = be.tid();

    for (int q0 = 0; q0 < total; q0 += nth * UNR) {
      PackT in[UNR][PB];
      A acc[UNR][VE];
      long long off[UNR];
      int nv[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int q = q0 + u * nth + tid;
        nv[u] = 0;
        if (q < total) {
          const int bi = q / CV, vi = q - bi * CV;
          long long start;
          const int n = tile_valid32(rmod + bi * gn, t, &start) - vi * VE;
          if (n > 0) {
            nv[u] = n < VE ? n : VE;
            const long long el = start + vi * VE;
            off[u] = el * (long long)sizeof(T);
            PackT own;
            if (own_from_user) own = io_to_wire(ld_user(x, el, nv[u]));
            else own = ld_win(wr, off[u], false);
#pragma unroll
            for (int j = 0; j < PB; ++j)
              if (j < npeers) in[u][j] = ld_win(pp[j], off[u], true);
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[u][i] = EL::up(own.e[i]);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if (!nv[u]) continue;
#pragma unroll
        for (int j = 0; j < PB; ++j) {
          if (j < npeers) {
#pragma unroll
            for (int i = 0; i < VE; ++i) acc[u][i] = Op::apply(acc[u][i], EL::up(in[u][j].e[i]));
          }
        }
      }
      // widths beyond PB+1: further batches of PB peers, again all loads of a batch before its adds
      for (int j0 = PB; j0 < npeers; j0 += PB) {
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          if (!nv[u]) continue;
#pragma unroll
          for (int j = 0; j < PB; ++j)
            if (j0 + j < npeers) in[u][j] = ld_win(rs_peer_ptr(k, j0 + j), off[u], true);
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
          if (!nv[u]) continue;
#pragma unroll
          for (int j = 0; j < PB; ++j) {
            if (j0 + j < npeers) {
#pragma unroll
              for (int i = 0; i < VE; ++i) acc[u][i] = Op::apply(acc[u][i], EL::up(in[u][j].e[i]));
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if (!nv[u]) continue;
        if (final_stage) scale_acc(acc[u]);
        PackT o;
#pragma unroll
        for (int i = 0; i < VE; ++i) o.e[i] = EL::down(acc[u][i]);
        st_win(wr, off[u], o);
      }
    }
  }

  // ---- all-gather stage k on tile t ---------------------------------------------------------------------
  // Pull blocks b == p (mod g_{k+1}) from every group peer p. k > 0: into my window. k == 0: into the
  // user's recv buffer through the scale/cast epilogue, including my own blocks.
  template <int UNR> FT_DEV void ag_tile(int k, int t) {
    const Plan& P = a.plan;
    const Layout& L = a.lay;
    const int w = P.width[k], g = P.gap[k], gn = P.gap[k + 1], base = stage_base(P, r, k);
    const int nb = P.nranks / gn;
    const int CV = (int)((L.tile + VE - 1) / VE);
    const bool last = (k == 0);
    // In registered mode recv aliases my window: my own blocks are already in place unless the
    // epilogue changes them.
    const bool self_needed = last && !(a.mode == MODE_REGISTERED && kRawIO);
    const int total = nb * CV * w;
    char* wr = rc.win[r] + woff;
    IO* y = (IO*)rc.recv;
    const int nth = be.nthr(), tid = be.tid();

    for (int q0 = 0; q0 < total; q0 += nth * UNR) {
      PackT v[UNR];
      long long el[UNR];
      int nv[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int q = q0 + u * nth + tid;
        nv[u] = 0;
        if (q < total) {
          // vector index fastest, then source, then block group: neighbouring CV-runs hit different peers
          const int rest = q / CV, vi = q - rest * CV;
          const int bi = rest / w, s = rest - bi * w;
          const int p = base + s * g;
          if (p != r || self_needed) {
            long long start;
            const int n = tile_valid32((p % gn) + bi * gn, t, &start) - vi * VE;
            if (n > 0) {
              nv[u] = n < VE ? n : VE;
              el[u] = start + vi * VE;
              v[u] = ld_win(rc.win[p] + woff, el[u] * (long long)sizeof(T), p != r);
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if (!nv[u]) continue;
        if (last) st_user(y, el[u], wire_to_io(v[u]), nv[u]);
        else st_win(wr, el[u] * (long long)sizeof(T), v[u]);
      }
    }
  }

  // ---- NVLS (multicast) pieces ---------------------------------------------------------------------------
  // Copy every block of tile t from the user buffer into my window (the switch reads all of them).
  template <int UNR> FT_DEV void copy_in_all_tile(int t) {
    const int N = a.plan.nranks;
    for (int b = 0; b < N; ++b) {
      long long start;
      const int n = tile_valid32(b, t, &start);
      copy_run<UNR, 0>(start, n);
    }
  }

  // In-switch reduction of my block of tile t: one multimem.ld_reduce pulls and sums the N copies,
  // one multimem.st broadcasts the result into every rank's window.
  template <int UNR, bool EMU> FT_DEV void nvls_reduce_tile(int t) {
    long long start;
    const long long n = tile_valid(r, t, &start);
    if (n <= 0) return;
    const long long nvec = (n + VE - 1) / VE;
    const int nth = be.nthr(), tid = be.tid();
    const long long step = (long long)nth * UNR;
    const long long base = woff + start * (long long)sizeof(T);
    // Software pipeline: the next batch of multimem.ld_reduce is issued before the current batch is
    // broadcast, so the inbound (reduced data) and outbound (multimem.st) directions of the links overlap
    // inside one thread instead of alternating.
    PackT cur[UNR], nxt[UNR];
    auto load = [&](PackT* v, long long q0) {
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long q = q0 + (long long)u * nth + tid;
        if (q < nvec) v[u] = be.template mc_ld_reduce<T, Op, VE, EMU>(rc, base + q * (long long)sizeof(PackT), a.plan.nranks);
      }
    };
    load(cur, 0);
    for (long long q0 = 0; q0 < nvec; q0 += step) {
      const bool more = q0 + step < nvec;
      if (more) load(nxt, q0 + step);
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long q = q0 + (long long)u * nth + tid;
        if (q >= nvec) continue;
        scale_pack(cur[u]);
        be.template mc_st<T, VE, EMU>(rc, base + q * (long long)sizeof(PackT), cur[u], a.plan.nranks);
      }
      if (more) {
#pragma unroll
        for (int u = 0; u < UNR; ++u) cur[u] = nxt[u];
      }
    }
  }

  // Same work without the software pipeline: UNR loads, then UNR stores (general path: private buffers, casts,
  // single-GPU emulation).
  enum { NVLS_KEEP_END = 4, NVLS_DEEP_COPY = 8 };
  template <int UNR, bool EMU> FT_DEV void nvls_reduce_tile_simple(int t) {
    long long start;
    const long long n = tile_valid(r, t, &start);
    if (n <= 0) return;
    const long long nvec = (n + VE - 1) / VE;
    const int nth = be.nthr(), tid = be.tid();
    const long long step = (long long)nth * UNR;
    const long long base = woff + start * (long long)sizeof(T);
    for (long long q0 = 0; q0 < nvec; q0 += step) {
      PackT v[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long q = q0 + (long long)u * nth + tid;
        if (q < nvec) v[u] = be.template mc_ld_reduce<T, Op, VE, EMU>(rc, base + q * (long long)sizeof(PackT), a.plan.nranks);
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const long long q = q0 + (long long)u * nth + tid;
        if (q >= nvec) continue;
        scale_pack(v[u]);
        be.template mc_st<T, VE, EMU>(rc, base + q * (long long)sizeof(PackT), v[u], a.plan.nranks);
      }
    }
  }
  FT_DEV void scale_pack(PackT& v) const {
    if constexpr (EL::is_float) {
      if (a.apply_scale) {  // the switch returned the final sum: scale it once, before the broadcast
        A acc[VE];
#pragma unroll
        for (int i = 0; i < VE; ++i) acc[i] = EL::up(v.e[i]);
        scale_acc(acc);
#pragma unroll
        for (int i = 0; i < VE; ++i) v.e[i] = EL::down(acc[i]);
      }
    }
  }

  // Streaming form for the dedicated kernel (fixed NTH threads per CTA). What it fixes, measured on 2 GPUs with the
  // loops above at 256 MiB: 4.8 ms for "UNR loads, then UNR stores" whatever UNR was (4, 8, 16), i.e. ~2.3 us per
  // multimem.ld_reduce per thread - one full switch round trip each, the loads of a thread did not overlap at all.
  // The SASS showed why: every LDGMC of a thread took its address from the SAME register pair, recomputed between
  // the loads, and a multimem load holds its address registers until it has been sent off, so the next address
  // computation (a write-after-read on that pair) waited for the previous load. Here a batch is addressed as ONE base
  // pointer per thread plus compile-time offsets (u * NTH * 16 bytes): the base is not written while the batch is
  // in flight, the full batches carry no bounds checks, and only the ragged end of a tile goes through the checked
  // loop. MODE: 0 = reduce + broadcast; 1 = measurement, reduce only (result stays local); 2 = measurement,
  // broadcast only (of the unreduced local block). DEPTH 2 = software pipeline (next batch's loads before this
  // batch's stores).
  template <int UNR, int DEPTH, int MODE, int NTH, bool EMU> FT_DEV void nvls_tile_stream(int t) {
    long long start;
    const long long n = tile_valid(r, t, &start);
    if (n <= 0) return;
    const long long nvec = (n + VE - 1) / VE;
    const int tid = be.tid();
    constexpr long long kBatch = (long long)NTH * UNR;               // vectors per batch per CTA
    constexpr long long kStride = (long long)NTH * (long long)sizeof(PackT);  // bytes between a thread's vectors
    const long long full = nvec / kBatch;
    const int nr = a.plan.nranks;
    char* wr = rc.win[r];
    long long off = woff + start * (long long)sizeof(T) + (long long)tid * (long long)sizeof(PackT);
    auto ld = [&](PackT* v, long long o) {
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if constexpr (MODE == 2) v[u] = ld_win(wr, o + u * kStride, false);
        else v[u] = be.template mc_ld_reduce<T, Op, VE, EMU>(rc, o + u * kStride, nr);
      }
    };
    auto st = [&](PackT* v, long long o) {
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        scale_pack(v[u]);
        if constexpr (MODE == 1) st_win(wr, o + u * kStride, v[u]);
        else be.template mc_st<T, VE, EMU>(rc, o + u * kStride, v[u], nr);
      }
    };
    if constexpr (DEPTH == 2) {
      PackT cur[UNR], nxt[UNR];
      if (full > 0) ld(cur, off);
      for (long long b = 0; b < full; ++b) {
        const long long o = off + b * kBatch * (long long)sizeof(PackT);
        if (b + 1 < full) ld(nxt, o + kBatch * (long long)sizeof(PackT));
        st(cur, o);
        if (b + 1 < full) {
#pragma unroll
          for (int u = 0; u < UNR; ++u) cur[u] = nxt[u];
        }
      }
    } else {
      for (long long b = 0; b < full; ++b) {
        const long long o = off + b * kBatch * (long long)sizeof(PackT);
        PackT v[UNR];
        ld(v, o);
        st(v, o);
      }
    }
    // ragged end of the tile: fewer than kBatch vectors
    for (long long q = full * kBatch + tid; q < nvec; q += NTH) {
      const long long o = woff + start * (long long)sizeof(T) + q * (long long)sizeof(PackT);
      PackT v;
      if constexpr (MODE == 2) v = ld_win(wr, o, false);
      else v = be.template mc_ld_reduce<T, Op, VE, EMU>(rc, o, nr);
      scale_pack(v);
      if constexpr (MODE == 1) st_win(wr, o, v);
      else be.template mc_st<T, VE, EMU>(rc, o, v, nr);
    }
  }

  // The in-switch reduce of one tile. Inner-loop variants (ft_comm_set_tuning "nvls_variant"; each one is its own
  // kernel in kernels/nvls_kernel.cu, with its own register allocation):
  //   0  streaming, 8 vectors per batch            1  streaming, 16 vectors per batch
  //   2  streaming, 8 vectors, software-pipelined  3  streaming, 4 vectors, software-pipelined
  //   4  round-1 loop (pipelined, 8 vectors, per-vector addresses and bounds checks) - kept for A/B
  //   5  warp-specialised: half the warps run multimem.ld_reduce into a shared-memory ring, the other half drain it
  //      with multimem.st (kernels/nvls_ws.cuh)
  //   6  plain loop (8 loads then 8 stores, per-vector addresses) - the form that serialises, kept as evidence
  //   7  MEASUREMENT: in-switch reduce only    8  MEASUREMENT: multicast store only   (wrong sums by design)
  FT_DEV void nvls_tile(int t) {
    // EMU: no multicast mapping (single-GPU emulation) - the back end plays the switch with a loop over the N
    // windows. A template flag, not a branch inside the loops: the multi-GPU code carries no trace of it.
    constexpr int U = (VE == 1) ? 1 : 4;
    if (!rc.mc) {
      nvls_reduce_tile_simple<U, true>(t);
    } else {
      // few reducer CTAs carry the whole switch traffic of a rank: they need deep batches (8 CTAs x 256 threads x 4
      // vectors in flight ran at 727 us per 256 MiB on 8 GPUs against 575 us with twice the bytes in flight)
      if constexpr (VE > 1) {
        if (be.nthr() == FT_MAX_THREADS_CORE) { nvls_tile_stream<8, 1, 0, FT_MAX_THREADS_CORE, false>(t); return; }
      }
      nvls_reduce_tile_simple<U, false>(t);
    }
  }
  // One variant, chosen at compile time (the dedicated registered-buffer kernel, kernels/nvls_kernel.cu).
  template <int VARIANT, int NTH, bool EMU> FT_DEV void nvls_tile_variant(int t) {
    if constexpr (VE == 1) nvls_reduce_tile_simple<1, EMU>(t);
    else if constexpr (VARIANT == 1) nvls_tile_stream<16, 1, 0, NTH, EMU>(t);
    else if constexpr (VARIANT == 2) nvls_tile_stream<8, 2, 0, NTH, EMU>(t);
    else if constexpr (VARIANT == 3) nvls_tile_stream<4, 2, 0, NTH, EMU>(t);
    else if constexpr (VARIANT == 4) nvls_reduce_tile<8, EMU>(t);
    else if constexpr (VARIANT == 5 && BE::kTma) be.template nvls_tile_ws<T, IO, Op, VE, EMU>(*this, t);
    else if constexpr (VARIANT == 6) nvls_reduce_tile_simple<8, EMU>(t);
    else if constexpr (VARIANT == 7) nvls_tile_stream<8, 1, 1, NTH, EMU>(t);
    else if constexpr (VARIANT == 8) nvls_tile_stream<8, 1, 2, NTH, EMU>(t);
    else nvls_tile_stream<8, 1, 0, NTH, EMU>(t);
  }
  // NVLS on registered (zero-copy) buffers of the wire type: the whole call for this CTA. Same protocol as the
  // registered branch of run_nvls, without the code of the private-buffer roles around it.
  template <int VARIANT, int NTH, bool EMU> FT_DEV void run_nvls_registered() {
    const int G = a.ctas, nt = a.lay.ntiles;
    const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;
    signal_group(PH_IN, 0, E + mine);
    for (int j = 0; j < mine; ++j) {
      wait_group(PH_IN, 0, E + j + 1);
      nvls_tile_variant<VARIANT, NTH, EMU>(cta + j * G);
      be.sync();
      signal_group(PH_RS0, 0, E + j + 1);
    }
    for (int j = 0; j < mine; ++j) wait_group(PH_RS0, 0, E + j + 1);
    if (a.nvls_flags & NVLS_KEEP_END) {  // see run_nvls: the exit barrier is redundant
      be.sync();
      signal_group(PH_END, 0, E + 1);
      wait_group(PH_END, 0, E + 1);
    }
    be.finish(cd, a.ctas);
  }

  // Local copy of every block of tile t from my window to the user buffer through the epilogue.
  template <int UNR> FT_DEV void out_tile(int t) {
    const int N = a.plan.nranks;
    for (int b = 0; b < N; ++b) {
      long long start;
      const int n = tile_valid32(b, t, &start);
      copy_run<UNR, 1>(start, n);  // written by the peers' multimem.st: read around L1
    }
  }

  // NVLS with private (non-symmetric) user buffers. Three passes per tile - user -> window (HBM), in-switch
  // reduce (NVLink), window -> user (HBM) - that must overlap to reach the link rate. CTAs are therefore split
  // by ROLE: the last `nvls_reducers` CTAs of every rank only run multimem.ld_reduce/st, all the others only
  // copy (a few tiles ahead in, then out as results arrive). Roles meet through per-TILE flags (slot =
  // tile mod FT_MAX_CTAS, value grows with tile / FT_MAX_CTAS) that every rank also sends to itself.
  // Measured on 8 GPUs before the split: 1104 us (phase-major) / 922 us (tile-major, one role) per 256 MiB,
  // against 579 us for zero-copy buffers and 720 us for NCCL.
  FT_DEV uint64_t tile_flag(int t) const { return E + (uint64_t)(t / FT_MAX_CTAS) + 1; }
  FT_DEV void signal_all_tile(int phase, int t) {  // to every rank, myself included
    const int N = a.plan.nranks, slot = t % FT_MAX_CTAS;
    for (int p = be.tid(); p < N; p += be.nthr()) {
      be.fence();
      be.flag_st(cd.pad[p] + pad_slot(phase, slot, r), tile_flag(t));
    }
  }
  FT_DEV void wait_all_tile(int phase, int t) {
    const int N = a.plan.nranks, slot = t % FT_MAX_CTAS;
    for (int p = be.tid(); p < N; p += be.nthr()) {
      spin(cd.pad[r] + pad_slot(phase, slot, p), tile_flag(t), phase, p);
      be.acquire(cd.pad[r] + pad_slot(phase, slot, p));
    }
    be.sync();
  }
  FT_DEV void run_nvls_private() {
    const int G = a.ctas, nt = a.lay.ntiles;
    constexpr int CU = (VE == 1) ? 1 : 8;
    constexpr int CD = (VE == 1) ? 1 : 16;           // deep copies (a.nvls_flags & NVLS_DEEP_COPY): twice the bytes in flight per CTA
    const bool deep = (a.nvls_flags & NVLS_DEEP_COPY) != 0;
    const int Gr = a.nvls_reducers > 0 && a.nvls_reducers < G ? a.nvls_reducers : 0;
    const int Gc = G - Gr;
    if (Gr == 0) {
      // Every CTA plays both roles on its own tiles (t = cta + j*G), skewed by one tile so that each flag wait
      // sits behind a tile's worth of work. Best for mid-size messages (measured on 8 GPUs: 80 us at 16 MiB,
      // 232 us at 64 MiB against 115 / 279 us for the role split, which wins from ~128 MiB up).
      const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;
      if (mine > 0) {
        copy_in_all_tile<CU>(cta);
        be.sync();
        signal_all_tile(PH_IN, cta);
      }
      for (int j = 0; j < mine; ++j) {
        const int t = cta + j * G;
        if (j + 1 < mine) {
          copy_in_all_tile<CU>(t + G);
          be.sync();
          signal_all_tile(PH_IN, t + G);
        }
        wait_all_tile(PH_IN, t);
        nvls_tile(t);
        be.sync();
        signal_all_tile(PH_RS0, t);
        if (j > 0) {
          wait_all_tile(PH_RS0, t - G);
          out_tile<CU>(t - G);
        }
      }
      if (mine > 0) {
        wait_all_tile(PH_RS0, cta + (mine - 1) * G);
        out_tile<CU>(cta + (mine - 1) * G);
      }
    } else if (cta < Gc) {
      // copy role: tiles cta, cta + Gc, ...; stay up to kAhead tiles ahead with the copy-in
      const int mine = cta < nt ? (nt - cta + Gc - 1) / Gc : 0;
      constexpr int kAhead = 2;
      int jin = 0, jout = 0;
      while (jout < mine) {
        if (jin < mine && jin - jout < kAhead) {
          const int t = cta + jin * Gc;
          if (deep) copy_in_all_tile<CD>(t);
          else copy_in_all_tile<CU>(t);
          be.sync();
          signal_all_tile(PH_IN, t);
          ++jin;
        } else {
          const int t = cta + jout * Gc;
          wait_all_tile(PH_RS0, t);
          if (deep) out_tile<CD>(t);
          else out_tile<CU>(t);
          ++jout;
        }
      }
    } else {
      // reduce role: tiles q, q + Gr, ...
      const int q = cta - Gc;
      for (int t = q; t < nt; t += Gr) {
        wait_all_tile(PH_IN, t);
        nvls_tile(t);
        be.sync();
        signal_all_tile(PH_RS0, t);
      }
    }
    be.finish(cd, a.ctas);
  }

  // NVLS allreduce. Plan must be the flat {N} plan (stage-0 group == everybody).
  FT_DEV void run_nvls() {
    const int G = a.ctas, nt = a.lay.ntiles;
    const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;
    constexpr int CU = (VE == 1) ? 1 : 8;
    if (a.mode == MODE_STAGING) {
      run_nvls_private();
      return;
    }
    // zero-copy (registered) buffers
    signal_group(PH_IN, 0, E + mine);
    for (int j = 0; j < mine; ++j) {
      wait_group(PH_IN, 0, E + j + 1);
      nvls_tile(cta + j * G);
      be.sync();
      signal_group(PH_RS0, 0, E + j + 1);
    }
    const bool need_out = !kRawIO;
    for (int j = 0; j < mine; ++j) {
      wait_group(PH_RS0, 0, E + j + 1);
      if (need_out) out_tile<CU>(cta + j * G);
    }
    // No exit barrier: a peer reads my buffer only while it reduces its own block, and it signals PH_RS0 for a tile
    // after that tile's reads (and broadcast stores) are done. Once every CTA of mine has passed its PH_RS0 waits -
    // which is when this kernel can end - every peer has finished with my buffer. (The pairwise redundancy of the
    // registered-NVLS waits was found by the wait-removal campaign on the CPU back end; NVLS_KEEP_END restores
    // the barrier for A/B timing.)
    if (a.nvls_flags & NVLS_KEEP_END) {
      be.sync();
      signal_group(PH_END, 0, E + 1);
      wait_group(PH_END, 0, E + 1);
    }
    be.finish(cd, a.ctas);
  }

  // ---- one-shot LL ----------------------------------------------------------------------------------------
  // Every rank pushes its whole (small) message to every peer as 16-byte lines {data, flag, data, flag}
  // and then reduces the N-1 received copies with its own; a line is valid when both flags carry this
  // call's tag, so there is no separate barrier: latency is one NVLink store hop.
  // (No counterpart in the reference, which has no latency-optimised path.)
  static constexpr int LE = 8 / (int)sizeof(T) > 0 ? 8 / (int)sizeof(T) : 1;  // elements per line
  // my payload of LL line i in wire format
  FT_DEV Pack<T, LE> ll_own(long long e0, int nv) const {
    const IO* x = (const IO*)rc.send;
    Pack<T, LE> own;
    if (VE > 1 && nv == LE) {
      Pack<IO, LE> raw;
      BE::template ld<(int)sizeof(Pack<IO, LE>)>(&raw, x + e0, false);
#pragma unroll
      for (int q = 0; q < LE; ++q) {
        if constexpr (kRawIO) own.e[q] = raw.e[q];
        else own.e[q] = EL::down((A)EU::up(raw.e[q]));
      }
    } else {
      for (int q = 0; q < LE; ++q) {
        IO t0;
        if (q < nv) BE::template ld<(int)sizeof(IO)>(&t0, x + e0 + q, false);
        else BE::template ld<(int)sizeof(IO)>(&t0, x + e0, false);
        if constexpr (kRawIO) own.e[q] = t0;
        else own.e[q] = EL::down((A)EU::up(t0));
      }
    }
    return own;
  }
  FT_DEV void run_oneshot() {
    scale_out_ = true;
    const int N = a.plan.nranks, G = a.ctas;
    const long long count = a.lay.count;
    const long long nlines = (count + LE - 1) / LE;
    const uint32_t tag = (uint32_t)(E >> FT_EPOCH_SHIFT);  // epoch + 1, never 0 for 2^32 calls
    const size_t region = (cd.ll_bytes / (size_t)N) & ~(size_t)15;
    const size_t par_off = ((E >> FT_EPOCH_SHIFT) & 1) ? cd.ll_bytes : 0;
    IO* y = (IO*)rc.recv;
    const long long stride = (long long)G * be.nthr();
    const long long first_line = (long long)cta * be.nthr() + be.tid();
    // pass 1: push ALL of my lines to every peer before polling anything, so a thread that owns several lines
    // pays one NVLink round trip in total instead of one per line
    for (long long i = first_line; i < nlines; i += stride) {
      const long long e0 = i * LE;
      const int nv = (count - e0) < LE ? (int)(count - e0) : LE;
      const Pack<T, LE> own = ll_own(e0, nv);
      uint32_t d[2] = {0, 0};
      memcpy(d, &own, sizeof(own) < 8 ? sizeof(own) : 8);
      for (int p = 0; p < N; ++p) {
        if (p == r) continue;
        be.ll_st(cd.ll[p] + par_off + (size_t)r * region + (size_t)i * 16, d[0], d[1], tag);
      }
    }
    // pass 2: poll the N-1 copies of each line, reduce in ascending rank order (identical on every rank)
    for (long long i = first_line; i < nlines; i += stride) {
      const long long e0 = i * LE;
      const int nv = (count - e0) < LE ? (int)(count - e0) : LE;
      const Pack<T, LE> own = ll_own(e0, nv);
      A acc[LE] = {};
      bool first = true;
      for (int q = 0; q < N; ++q) {
        Pack<T, LE> in;
        if (q == r) {
          in = own;
        } else {
          uint32_t g[2];
          if (!ll_wait(cd.ll[r] + par_off + (size_t)q * region + (size_t)i * 16, tag, g, q)) { g[0] = g[1] = 0; }
          memcpy(&in, g, sizeof(in) < 8 ? sizeof(in) : 8);
        }
#pragma unroll
        for (int e = 0; e < LE; ++e) acc[e] = first ? EL::up(in.e[e]) : Op::apply(acc[e], EL::up(in.e[e]));
        first = false;
      }
      Pack<IO, LE> out;
#pragma unroll
      for (int e = 0; e < LE; ++e) {
        if constexpr (EL::is_float) {
          A v = acc[e];
          if (a.apply_scale) v = (A)(v * (A)a.scale);
          out.e[e] = EU::down((typename EU::acc)v);
        } else {
          out.e[e] = EU::down((typename EU::acc)acc[e]);
        }
      }
      if (VE > 1 && nv == LE) BE::template st<(int)sizeof(Pack<IO, LE>)>(y + e0, &out);
      else for (int q = 0; q < nv; ++q) BE::template st<(int)sizeof(IO)>(y + e0 + q, &out.e[q]);
    }
    be.finish(cd, a.ctas);
  }
  // ---- two-shot LL -----------------------------------------------------------------------------------------
  // The one-shot pushes the whole message to every peer: (N-1) * S * 2 bytes per rank on the wire (a 16-byte line
  // carries 8 bytes of data), which is past its sweet spot around 1 MiB. Two shots over the same LL lines move
  // 2 * (N-1)/N * S * 2 instead: shot 1 sends block p of my message to rank p only, every rank reduces its own block
  // (ascending rank order, like the one-shot), shot 2 broadcasts the reduced block. Still no barrier and no flag
  // words: a line is its own flag. Costs one more store hop of latency than the one-shot.
  // LL area of a parity half: [shot 1: N regions by source | shot 2: N regions by source], region = ll_bytes/(2N).
  // (No counterpart in the reference.)
  FT_DEV void run_twoshot() {
    const int N = a.plan.nranks, G = a.ctas;
    const long long count = a.lay.count;
    const long long nlines = (count + LE - 1) / LE;
    const long long Bl = (nlines + N - 1) / N;  // lines per block; block b = lines [b*Bl, min((b+1)*Bl, nlines))
    const uint32_t tag = (uint32_t)(E >> FT_EPOCH_SHIFT);
    const size_t region = (cd.ll_bytes / (size_t)(2 * N)) & ~(size_t)15;
    const size_t par_off = ((E >> FT_EPOCH_SHIFT) & 1) ? cd.ll_bytes : 0;
    const size_t shot2 = par_off + (size_t)N * region;
    IO* y = (IO*)rc.recv;
    const long long stride = (long long)G * be.nthr();
    const long long first_line = (long long)cta * be.nthr() + be.tid();
    // shot 1: every line that is not in my block goes to the rank that owns its block
    for (long long i = first_line; i < nlines; i += stride) {
      const int p = (int)(i / Bl);
      if (p == r) continue;
      const long long e0 = i * LE;
      const int nv = (count - e0) < LE ? (int)(count - e0) : LE;
      const Pack<T, LE> own = ll_own(e0, nv);
      uint32_t d[2] = {0, 0};
      memcpy(d, &own, sizeof(own) < 8 ? sizeof(own) : 8);
      be.ll_st(cd.ll[p] + par_off + (size_t)r * region + (size_t)(i - (long long)p * Bl) * 16, d[0], d[1], tag);
    }
    // my block: reduce the N copies in ascending rank order, round to the wire type (the block is final: the fused
    // scale is applied here, once), keep it and push it to everybody
    const long long b0 = (long long)r * Bl, b1 = (b0 + Bl < nlines) ? b0 + Bl : nlines;
    for (long long i = b0 + first_line; i < b1; i += stride) {
      const long long e0 = i * LE, li = i - b0;
      const int nv = (count - e0) < LE ? (int)(count - e0) : LE;
      const Pack<T, LE> own = ll_own(e0, nv);
      A acc[LE] = {};
      bool first = true;
      for (int q = 0; q < N; ++q) {
        Pack<T, LE> in;
        if (q == r) {
          in = own;
        } else {
          uint32_t g[2];
          if (!ll_wait(cd.ll[r] + par_off + (size_t)q * region + (size_t)li * 16, tag, g, q)) { g[0] = g[1] = 0; }
          memcpy(&in, g, sizeof(in) < 8 ? sizeof(in) : 8);
        }
#pragma unroll
        for (int e = 0; e < LE; ++e) acc[e] = first ? EL::up(in.e[e]) : Op::apply(acc[e], EL::up(in.e[e]));
        first = false;
      }
      Pack<T, LE> res;
#pragma unroll
      for (int e = 0; e < LE; ++e) {
        if constexpr (EL::is_float) {
          A v = acc[e];
          if (a.apply_scale) v = (A)(v * (A)a.scale);
          res.e[e] = EL::down(v);
        } else {
          res.e[e] = EL::down(acc[e]);
        }
      }
      uint32_t d[2] = {0, 0};
      memcpy(d, &res, sizeof(res) < 8 ? sizeof(res) : 8);
      for (int p = 0; p < N; ++p) {
        if (p == r) continue;
        be.ll_st(cd.ll[p] + shot2 + (size_t)r * region + (size_t)li * 16, d[0], d[1], tag);
      }
      twoshot_out(y, e0, nv, res);
    }
    // shot 2: the other blocks arrive reduced
    for (long long i = first_line; i < nlines; i += stride) {
      const int q = (int)(i / Bl);
      if (q == r) continue;
      const long long e0 = i * LE;
      const int nv = (count - e0) < LE ? (int)(count - e0) : LE;
      uint32_t g[2];
      if (!ll_wait(cd.ll[r] + shot2 + (size_t)q * region + (size_t)(i - (long long)q * Bl) * 16, tag, g, q)) { g[0] = g[1] = 0; }
      Pack<T, LE> res;
      memcpy(&res, g, sizeof(res) < 8 ? sizeof(res) : 8);
      twoshot_out(y, e0, nv, res);
    }
    be.finish(cd, a.ctas);
  }
  FT_DEV static void twoshot_out(IO* y, long long e0, int nv, const Pack<T, LE>& res) {
    Pack<IO, LE> out;
#pragma unroll
    for (int e = 0; e < LE; ++e) {
      if constexpr (kRawIO) out.e[e] = res.e[e];
      else out.e[e] = EU::down((typename EU::acc)EL::up(res.e[e]));
    }
    if (VE > 1 && nv == LE) BE::template st<(int)sizeof(Pack<IO, LE>)>(y + e0, &out);
    else for (int q = 0; q < nv && q < LE; ++q) BE::template st<(int)sizeof(IO)>(y + e0 + q, &out.e[q]);
  }
  FT_DEV bool ll_wait(const char* line, uint32_t tag, uint32_t* data, int src) {
    if (*abort_flag) return false;
    uint64_t t0 = 0;
    unsigned it = 0;
    while (!be.ll_ld(line, tag, data)) {
      if (((++it) & 0x3ff) == 0) {
        if (be.err_ld(cd.err)) { *abort_flag = 1; return false; }
        if (a.timeout_ns) {
          uint64_t now = be.now_ns();
          if (!t0) t0 = now;
          else if (now - t0 > a.timeout_ns) {
            be.err_set(cd, (unsigned)ERR_TIMEOUT | (0xfeu << 8) | ((unsigned)src << 16) | ((unsigned)r << 24));
            *abort_flag = 1;
            return false;
          }
        }
      }
      be.pause();
    }
    return true;
  }

  // ---- ring ----------------------------------------------------------------------------------------------
  // Capability parity: FlexTree::ring_allreduce (mpi_mod.hpp:1113-1163): N-1 reduce-scatter steps then N-1
  // all-gather steps around the ring, here as pulls from the left neighbour gated by one progress
  // counter per (CTA, neighbour). Progress value for "step s, tile j done" is E + s*mine + j + 1 with
  // s = 0 for the input, 1..N-1 for RS, N..2N-2 for AG (monotone because loops are step-major).
  template <int UNR> FT_DEV void ring_step_tile(int t, int blk, bool reduce, bool to_user, bool own_from_user) {
    const int N = a.plan.nranks, left = (r + N - 1) % N;
    long long start, n = tile_valid(blk, t, &start);
    if (n <= 0) return;
    const long long nvec = (n + VE - 1) / VE;
    const char* wl = rc.win[left] + woff;
    char* wr = rc.win[r] + woff;
    const IO* x = (const IO*)rc.send;
    IO* y = (IO*)rc.recv;
    const int nth = be.nthr(), tid = be.tid();
    for (long long q0 = 0; q0 < nvec; q0 += (long long)nth * UNR) {
      PackT in[UNR], own[UNR];
      int nv[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        long long q = q0 + (long long)u * nth + tid;
        nv[u] = 0;
        if (q < nvec) {
          long long rem = n - q * VE;
          nv[u] = rem < VE ? (int)rem : VE;
          long long el = start + q * VE;
          in[u] = ld_win(wl, el * (long long)sizeof(T), true);
          if (reduce) {
            if (own_from_user) own[u] = io_to_wire(ld_user(x, el, nv[u]));
            else own[u] = ld_win(wr, el * (long long)sizeof(T), false);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        if (!nv[u]) continue;
        long long el = start + (q0 + (long long)u * nth + tid) * VE;
        PackT o = in[u];
        if (reduce) {
          // the partial sum arriving from the left is accumulated first, then my contribution
          A acc[VE];
#pragma unroll
          for (int i = 0; i < VE; ++i) acc[i] = Op::apply(EL::up(in[u].e[i]), EL::up(own[u].e[i]));
          if (to_user) scale_acc(acc);  // last reduce step: the block is final
#pragma unroll
          for (int i = 0; i < VE; ++i) o.e[i] = EL::down(acc[i]);
        }
        st_win(wr, el * (long long)sizeof(T), o);
        if (to_user) st_user(y, el, wire_to_io(o), nv[u]);
      }
    }
  }
  template <int UNR> FT_DEV void ring_copy_tile(int t, int blk, bool in_dir) {
    // in_dir: user -> window (input block); else window -> user (my fully reduced block)
    long long start, n = tile_valid(blk, t, &start);
    if (n <= 0) return;
    const long long nvec = (n + VE - 1) / VE;
    char* wr = rc.win[r] + woff;
    const int nth = be.nthr(), tid = be.tid();
    for (long long q = tid; q < nvec; q += nth) {
      long long rem = n - q * VE, el = start + q * VE;
      int nv = rem < VE ? (int)rem : VE;
      if (in_dir) st_win(wr, el * (long long)sizeof(T), io_to_wire(ld_user((const IO*)rc.send, el, nv)));
      else st_user((IO*)rc.recv, el, wire_to_io(ld_win(wr, el * (long long)sizeof(T), false)), nv);
    }
  }
  FT_DEV void ring_signal(uint64_t val) {  // to my right neighbour
    const int N = a.plan.nranks, right = (r + 1) % N;
    if (be.tid() == 0) {
      be.fence();
      be.flag_st(cd.pad[right] + pad_slot(PH_RS0, cta, r), val);
    }
  }
  FT_DEV void ring_wait(uint64_t val) {  // on my left neighbour
    const int N = a.plan.nranks, left = (r + N - 1) % N;
    if (be.tid() == 0) {
      spin(cd.pad[r] + pad_slot(PH_RS0, cta, left), val, PH_RS0, left);
      be.acquire(cd.pad[r] + pad_slot(PH_RS0, cta, left));
    }
    be.sync();
  }
  FT_DEV void run_ring() {
    const int N = a.plan.nranks, G = a.ctas, nt = a.lay.ntiles;
    const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;
    constexpr int CU = (VE == 1) ? 1 : 4;
    const bool staging = a.mode == MODE_STAGING;
    // step 0: my raw block r must be in my window for the right neighbour
    for (int j = 0; j < mine; ++j) {
      if (staging) ring_copy_tile<CU>(cta + j * G, r, true);
      be.sync();
      ring_signal(E + (uint64_t)j + 1);
    }
    // RS steps 1..N-1: pull block (r - s) from the left (it finished step s-1 on it), add mine
    for (int s = 1; s <= N - 1; ++s) {
      const int blk = ((r - s) % N + N) % N;
      for (int j = 0; j < mine; ++j) {
        ring_wait(E + (uint64_t)(s - 1) * mine + j + 1);
        // after the last RS step block (r+1) is complete on this rank
        ring_step_tile<CU>(cta + j * G, blk, true, s == N - 1, staging);
        be.sync();
        ring_signal(E + (uint64_t)s * mine + j + 1);
      }
    }
    // AG steps N..2N-2: pull the completed block (r + 1 - (s - N + 1)) from the left
    for (int s = N; s <= 2 * N - 2; ++s) {
      const int blk = ((r - (s - N)) % N + N) % N;
      for (int j = 0; j < mine; ++j) {
        ring_wait(E + (uint64_t)(s - 1) * mine + j + 1);
        ring_step_tile<CU>(cta + j * G, blk, false, true, false);
        be.sync();
        ring_signal(E + (uint64_t)s * mine + j + 1);
      }
    }
    if (a.mode == MODE_REGISTERED) {  // exit barrier with both neighbours' readers: the right one reads me
      const int right = (r + 1) % N, left = (r + N - 1) % N;
      be.sync();
      if (be.tid() == 0) {
        be.fence();
        be.flag_st(cd.pad[left] + pad_slot(PH_END, cta, r), E + 1);
        spin(cd.pad[r] + pad_slot(PH_END, cta, right), E + 1, PH_END, right);
        be.acquire(cd.pad[r] + pad_slot(PH_END, cta, right));
      }
      be.sync();
    }
    be.finish(cd, a.ctas);
  }

  // Register-staged reduce-scatter tile: pick the (vectors per thread, peers per batch) shape so that ~8 remote
  // 16-byte loads are in flight per thread whatever the stage width (more than that spills at 512 threads).
  FT_DEV void rs_dispatch(int k, int t) {
    if constexpr (VE == 1) {
      rs_tile<1, 4>(k, t);  // scalar path (unaligned user buffers): correctness first
    } else {
      if (a.plan.width[k] == 2) rs_tile<4, 1>(k, t);  // one peer: more vectors per thread instead
      else rs_tile<2, 4>(k, t);
    }
  }

  // ---- block-scaled fp8 wire (T == fp8_t, opt-in) -----------------------------------------------------------
  // Same stages as above, but a pack is 16 e4m3 values that share one power-of-two scale stored in the
  // window's scale plane (a.scale_plane_off + element/16). Every stage dequantises its inputs, reduces in
  // fp32 and re-quantises its output block with a fresh scale; the own contribution of stage 0 is taken
  // from the user buffer unquantised. Register-staged loads only (no TMA / NVLS / LL variants).
  static constexpr bool kBlockScaled = SameType<T, fp8_t>::value;
  FT_DEV long long scale_off(long long el) const { return a.scale_plane_off + (el >> 4); }
  FT_DEV static void q_dequant(const PackT& p, uint8_t s, float* acc, bool first) {
    const float sc = fp8_scale_value(s);
#if defined(__CUDA_ARCH__)
    static_assert(VE != 16 || sizeof(PackT) == 16, "a block-scaled pack is 16 e4m3 bytes");
    if constexpr (VE == 16) {  // four words of four e4m3 each, converted in hardware (no byte indexing, no local memory)
      const uint4 w = *reinterpret_cast<const uint4*>(&p);
      const uint32_t ws[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float f[4];
        e4m3x4_to_float4(ws[q], f);
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[4 * q + i] = first ? f[i] * sc : Op::apply(acc[4 * q + i], f[i] * sc);
      }
      return;
    }
#endif
#pragma unroll
    for (int i = 0; i < VE; ++i) {
      const float v = e4m3_to_float(reinterpret_cast<const uint8_t*>(&p)[i]) * sc;
      acc[i] = first ? v : Op::apply(acc[i], v);
    }
  }
  FT_DEV static PackT q_quant(const float* acc, uint8_t* s_out) {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < VE; ++i) {
      const float m = acc[i] < 0 ? -acc[i] : acc[i];
      amax = m > amax ? m : amax;
    }
    const uint8_t s = fp8_block_scale(amax);
    const float inv = fp8_scale_inverse(s);
    PackT o;
    *s_out = s;
#if defined(__CUDA_ARCH__)
    if constexpr (VE == 16) {
      uint4 w;
      w.x = float4_to_e4m3x4(acc[0] * inv, acc[1] * inv, acc[2] * inv, acc[3] * inv);
      w.y = float4_to_e4m3x4(acc[4] * inv, acc[5] * inv, acc[6] * inv, acc[7] * inv);
      w.z = float4_to_e4m3x4(acc[8] * inv, acc[9] * inv, acc[10] * inv, acc[11] * inv);
      w.w = float4_to_e4m3x4(acc[12] * inv, acc[13] * inv, acc[14] * inv, acc[15] * inv);
      *reinterpret_cast<uint4*>(&o) = w;
      return o;
    }
#endif
#pragma unroll
    for (int i = 0; i < VE; ++i) reinterpret_cast<uint8_t*>(&o)[i] = float_to_e4m3(acc[i] * inv);
    return o;
  }
  FT_DEV static void q_load_user(const IO* x, long long el, int nvalid, float* acc) {
    PackIO u = ld_user(x, el, nvalid);
#pragma unroll
    for (int i = 0; i < VE; ++i) acc[i] = i < nvalid ? (float)EU::up(u.e[i]) : 0.f;
  }
  FT_DEV void q_store_user(IO* y, long long el, int nvalid, const float* acc) const {
    PackIO o;
#pragma unroll
    for (int i = 0; i < VE; ++i) o.e[i] = EU::down((typename EU::acc)acc[i]);
    st_user(y, el, o, nvalid);
  }
  // The three tile functions walk one block at a time with a plain vector index (no per-vector divisions) and issue
  // every load of a vector - the own contribution, the packs of ALL peers and their scale bytes - before the first
  // conversion, like rs_tile<UNR, PB> does for the unquantised wire.
  FT_DEV void q_copy_in_tile(int t) {
    const int N = a.plan.nranks, g1 = a.plan.gap[1];
    char* wr = rc.win[r] + woff;
    for (int b = 0; b < N; ++b) {
      if ((b % g1) == (r % g1)) continue;
      long long start;
      const int n = tile_valid32(b, t, &start);
      const int nvec = (n + VE - 1) / VE;
      for (int v = be.tid(); v < nvec; v += be.nthr()) {
        const long long el = start + (long long)v * VE;
        const int nv = n - v * VE < VE ? n - v * VE : VE;
        float acc[VE];
        q_load_user((const IO*)rc.send, el, nv, acc);
        uint8_t sc;
        PackT p = q_quant(acc, &sc);
        st_win(wr, el, p);
        BE::template st<1>(wr + scale_off(el), &sc);
      }
    }
  }
  static constexpr int kQPeers = 8;  // peers whose packs are in flight together (7 cover a stage of width 8)
  FT_DEV void q_rs_tile(int k, int t) {
    const Plan& P = a.plan;
    const int w = P.width[k], gn = P.gap[k + 1], nb = P.nranks / gn, rmod = r % gn, npeers = w - 1;
    const bool own_from_user = (k == 0) && (a.mode == MODE_STAGING);
    const bool final_stage = (k == P.nstages - 1) && a.lonely == 0;
    char* wr = rc.win[r] + woff;
    const char* pw[kQPeers];
#pragma unroll
    for (int j = 0; j < kQPeers; ++j) pw[j] = j < npeers ? rs_peer_ptr(k, j) : nullptr;
    for (int bi = 0; bi < nb; ++bi) {
      long long start;
      const int n = tile_valid32(rmod + bi * gn, t, &start);
      const int nvec = (n + VE - 1) / VE;
      for (int v = be.tid(); v < nvec; v += be.nthr()) {
        const long long el = start + (long long)v * VE;
        const int nv = n - v * VE < VE ? n - v * VE : VE;
        // all loads first
        PackT in[kQPeers];
        uint8_t sj[kQPeers];
#pragma unroll
        for (int j = 0; j < kQPeers; ++j) {
          if (j < npeers) {
            in[j] = ld_win(pw[j], el, true);
            BE::template ld<1>(&sj[j], pw[j] + scale_off(el), true);
          }
        }
        float acc[VE];
        if (own_from_user) {
          q_load_user((const IO*)rc.send, el, nv, acc);
        } else {
          uint8_t s0;
          PackT p0 = ld_win(wr, el, false);
          BE::template ld<1>(&s0, wr + scale_off(el), false);
          q_dequant(p0, s0, acc, true);
        }
#pragma unroll
        for (int j = 0; j < kQPeers; ++j)
          if (j < npeers) q_dequant(in[j], sj[j], acc, false);
        for (int j = kQPeers; j < npeers; ++j) {  // widths beyond 9: one at a time
          const char* pj = rs_peer_ptr(k, j);
          uint8_t s1;
          PackT p1 = ld_win(pj, el, true);
          BE::template ld<1>(&s1, pj + scale_off(el), true);
          q_dequant(p1, s1, acc, false);
        }
        if (final_stage && a.apply_scale) {
#pragma unroll
          for (int i = 0; i < VE; ++i) acc[i] *= (float)a.scale;
        }
        uint8_t sc;
        PackT o = q_quant(acc, &sc);
        st_win(wr, el, o);
        BE::template st<1>(wr + scale_off(el), &sc);
      }
    }
  }
  FT_DEV void q_ag_tile(int k, int t) {
    const Plan& P = a.plan;
    const int w = P.width[k], g = P.gap[k], gn = P.gap[k + 1], base = stage_base(P, r, k), nb = P.nranks / gn;
    const bool last = (k == 0);
    char* wr = rc.win[r] + woff;
    for (int bi = 0; bi < nb; ++bi) {
      for (int s_ = 0; s_ < w; ++s_) {
        const int p = base + s_ * g;
        if (p == r && !last) continue;
        long long start;
        const int n = tile_valid32((p % gn) + bi * gn, t, &start);
        const int nvec = (n + VE - 1) / VE;
        const char* pw = rc.win[p] + woff;
        for (int v0 = be.tid(); v0 < nvec; v0 += 2 * be.nthr()) {  // two vectors per thread in flight
          const int v1 = v0 + be.nthr();
          const long long e0 = start + (long long)v0 * VE, e1 = start + (long long)v1 * VE;
          uint8_t s0 = 0, s1 = 0;
          PackT p0 = ld_win(pw, e0, p != r), p1;
          BE::template ld<1>(&s0, pw + scale_off(e0), p != r);
          if (v1 < nvec) {
            p1 = ld_win(pw, e1, p != r);
            BE::template ld<1>(&s1, pw + scale_off(e1), p != r);
          }
          if (last) {
            float acc[VE];
            q_dequant(p0, s0, acc, true);
            q_store_user((IO*)rc.recv, e0, n - v0 * VE < VE ? n - v0 * VE : VE, acc);
            if (v1 < nvec) {
              q_dequant(p1, s1, acc, true);
              q_store_user((IO*)rc.recv, e1, n - v1 * VE < VE ? n - v1 * VE : VE, acc);
            }
          } else {
            st_win(wr, e0, p0);
            BE::template st<1>(wr + scale_off(e0), &s0);
            if (v1 < nvec) {
              st_win(wr, e1, p1);
              BE::template st<1>(wr + scale_off(e1), &s1);
            }
          }
        }
      }
    }
  }

  // ---- the two halves as collectives of their own (ZeRO / FSDP style users) ------------------------------
  // The reference's algorithm IS a reduce-scatter followed by an all-gather (mpi_mod.hpp:988-1029 /
  // :1050-1060); exposing the halves costs two small copy loops.
  // reduce-scatter epilogue: my fully reduced block r, tile t -> recv[0 .. B)
  FT_DEV void half_rs_out_tile(int t) {
    long long start;
    const int n = tile_valid32(r, t, &start);
    if (n <= 0) return;
    const long long blk0 = (long long)r * a.lay.block;
    const char* wr = rc.win[r] + woff;
    IO* y = (IO*)rc.recv;
    const int nvec = (n + VE - 1) / VE;
    for (int q = be.tid(); q < nvec; q += be.nthr()) {
      const long long el = start + (long long)q * VE;
      const int nv = n - q * VE < VE ? n - q * VE : VE;
      st_user(y, el - blk0, wire_to_io(ld_win(wr, el * (long long)sizeof(T), false)), nv);
    }
  }
  // all-gather prologue: send[0 .. B) -> my block r of my window, tile t
  FT_DEV void half_ag_in_tile(int t) {
    long long start;
    const int n = tile_valid32(r, t, &start);
    if (n <= 0) return;
    const long long blk0 = (long long)r * a.lay.block;
    char* wr = rc.win[r] + woff;
    const IO* x = (const IO*)rc.send;
    const int nvec = (n + VE - 1) / VE;
    for (int q = be.tid(); q < nvec; q += be.nthr()) {
      const long long el = start + (long long)q * VE;
      const int nv = n - q * VE < VE ? n - q * VE : VE;
      st_win(wr, el * (long long)sizeof(T), io_to_wire(ld_user(x, el - blk0, nv)));
    }
  }

  // ---- lonely ranks (world size = tree ranks + L, the reference's "+1" trees) ------------------------------
  // Capability parity: the lonely ops of Send_Ops/Recv_Ops (mpi_mod.hpp:172-178, :208-211) and the lonely
  // branches of tree_allreduce (:981-987, :1019-1026, :1041-1099), which the reference ships as stubs
  // ("must be modified before use"). Here: a lonely rank publishes its whole vector in its window; tree
  // rank r folds block r of every lonely rank into its fully reduced block right after the last
  // reduce-scatter stage (lonely ranks in ascending order, after all tree contributions); the lonely
  // ranks then gather block b straight from tree rank b.
  template <int UNR> FT_DEV void lonely_fold_tile(int t) {
    const int Tn = a.plan.nranks, L = a.lonely;
    long long start;
    const int n = tile_valid32(r, t, &start);
    if (n <= 0) return;
    const int nvec = (n + VE - 1) / VE;
    char* wr = rc.win[r] + woff;
    const int nth = be.nthr(), tid = be.tid();
    for (int q = tid; q < nvec; q += nth) {
      const long long off = (start + (long long)q * VE) * (long long)sizeof(T);
      PackT own = ld_win(wr, off, false);
      A acc[VE];
#pragma unroll
      for (int i = 0; i < VE; ++i) acc[i] = EL::up(own.e[i]);
      for (int l = 0; l < L; ++l) {
        PackT in = ld_win(rc.win[Tn + l] + woff, off, true);
#pragma unroll
        for (int i = 0; i < VE; ++i) acc[i] = Op::apply(acc[i], EL::up(in.e[i]));
      }
      scale_acc(acc);
      PackT o;
#pragma unroll
      for (int i = 0; i < VE; ++i) o.e[i] = EL::down(acc[i]);
      st_win(wr, off, o);
    }
  }
  // a lonely rank's gather: block b of tile t comes from tree rank b
  template <int UNR> FT_DEV void lonely_out_tile(int t) {
    const Layout& L_ = a.lay;
    const int Tn = a.plan.nranks;
    const int CV = (int)((L_.tile + VE - 1) / VE);
    const int total = Tn * CV;
    IO* y = (IO*)rc.recv;
    const int nth = be.nthr(), tid = be.tid();
    for (int q0 = 0; q0 < total; q0 += nth * UNR) {
      PackT v[UNR];
      long long el[UNR];
      int nv[UNR];
#pragma unroll
      for (int u = 0; u < UNR; ++u) {
        const int q = q0 + u * nth + tid;
        nv[u] = 0;
        if (q < total) {
          const int b = q / CV, vi = q - b * CV;
          long long start;
          const int n = tile_valid32(b, t, &start) - vi * VE;
          if (n > 0) {
            nv[u] = n < VE ? n : VE;
            el[u] = start + vi * VE;
            v[u] = ld_win(rc.win[b] + woff, el[u] * (long long)sizeof(T), true);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNR; ++u)
        if (nv[u]) st_user(y, el[u], wire_to_io(v[u]), nv[u]);
    }
  }
  // flags between the tree and the lonely ranks (any-to-any inside the two sets)
  FT_DEV void signal_ranks(int phase, int first, int count, uint64_t val) {
    for (int j = be.tid(); j < count; j += be.nthr()) {
      be.fence();
      be.flag_st(cd.pad[first + j] + pad_slot(phase, cta, r), val);
    }
  }
  FT_DEV void wait_ranks(int phase, int first, int count, uint64_t val) {
    for (int j = be.tid(); j < count; j += be.nthr()) {
      spin(cd.pad[r] + pad_slot(phase, cta, first + j), val, phase, first + j);
      be.acquire(cd.pad[r] + pad_slot(phase, cta, first + j));
    }
    be.sync();
  }
  FT_DEV void run_lonely_rank() {
    const int G = a.ctas, nt = a.lay.ntiles, Tn = a.plan.nranks;
    const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;
    constexpr int CU = (VE == 1) ? 1 : 8;
    for (int j = 0; j < mine; ++j) {
      copy_in_all_tile<CU>(cta + j * G);
      be.sync();
      signal_ranks(PH_IN, 0, Tn, E + j + 1);
    }
    for (int j = 0; j < mine; ++j) {
      wait_ranks(PH_LONELY, 0, Tn, E + j + 1);
      lonely_out_tile<CU>(cta + j * G);
    }
    be.finish(cd, a.ctas);
  }

  // ---- the whole call, for this CTA ---------------------------------------------------------------------
  // TMA = true is the kernel whose tiles are staged through shared memory by cp.async.bulk (CUDA only; the
  // host selects it only for 16-byte-multiple messages on the vector path).
  template <bool TMA = false> FT_DEV void run() {
    const int G = a.ctas, m = a.plan.nstages, nt = a.lay.ntiles;
    const int mine = cta < nt ? (nt - cta + G - 1) / G : 0;  // my tiles: t = cta + j*G
    constexpr int CU = (VE == 1) ? 1 : 8;
    if (a.lonely > 0 && r >= a.plan.nranks) {  // I am outside the tree
      run_lonely_rank();
      return;
    }

    trace(0);
    if (a.phases == PHASES_ALL_GATHER) {
      // my block is already "reduced": place it in my window and announce it like a finished last RS stage
      for (int j = 0; j < mine; ++j) {
        half_ag_in_tile(cta + j * G);
        be.sync();
        signal_group(PH_RS0 + m - 1, m - 1, E + j + 1);
      }
    }
    // IN: publish my window to my stage-0 group
    if (a.phases == PHASES_ALL_GATHER) {
    } else if (a.mode == MODE_STAGING) {
      for (int j = 0; j < mine; ++j) {
        if constexpr (kBlockScaled) q_copy_in_tile(cta + j * G);
        else if constexpr (TMA && BE::kBulkCopyIn) be.template copy_in_tile_bulk<T, IO, Op, VE>(*this, cta + j * G);
        else copy_in_tile<CU>(cta + j * G);
        be.sync();
        signal_group(PH_IN, 0, E + j + 1);
      }
    } else {
      signal_group(PH_IN, 0, E + mine);  // data was complete before the kernel started
    }
    trace(1);

    // reduce-scatter, stage-major so that a wait is almost always for a long-finished tile
    for (int k = 0; k < m && a.phases != PHASES_ALL_GATHER; ++k) {
      for (int j = 0; j < mine; ++j) {
        wait_group(k == 0 ? (int)PH_IN : PH_RS0 + k - 1, k, E + j + 1);
        if constexpr (kBlockScaled) q_rs_tile(k, cta + j * G);
        else if constexpr (TMA) be.template rs_tile_tma<T, IO, Op, VE>(*this, k, cta + j * G);
        else rs_dispatch(k, cta + j * G);
        be.sync();
        if (a.lonely > 0 && k == m - 1) {  // my block is reduced over the tree: add the lonely ranks' copies
          wait_ranks(PH_IN, a.plan.nranks, a.lonely, E + j + 1);
          lonely_fold_tile<1>(cta + j * G);
          be.sync();
          signal_ranks(PH_LONELY, a.plan.nranks, a.lonely, E + j + 1);
        }
        // consumers: RS stage k+1 pulls from its stage-(k+1) group; after the last RS stage the AG of
        // the same stage pulls
        signal_group(PH_RS0 + k, k + 1 < m ? k + 1 : k, E + j + 1);
      }
      trace(2 + k);
    }

    if (a.phases == PHASES_REDUCE_SCATTER) {
      for (int j = 0; j < mine; ++j) half_rs_out_tile(cta + j * G);  // my own writes, same CTA: no flag needed
    }
    // all-gather, stages unwound in reverse
    for (int k = m - 1; k >= 0 && a.phases != PHASES_REDUCE_SCATTER; --k) {
      for (int j = 0; j < mine; ++j) {
        wait_group(k == m - 1 ? PH_RS0 + k : PH_AG0 + k + 1, k, E + j + 1);
        if constexpr (kBlockScaled) q_ag_tile(k, cta + j * G);
        else if constexpr (TMA) be.template ag_tile_tma<T, IO, Op, VE>(*this, k, cta + j * G);
        else ag_tile<CU>(k, cta + j * G);
        if (k > 0) {
          be.sync();
          signal_group(PH_AG0 + k, k - 1, E + j + 1);
        }
      }
      trace(2 + m + (m - 1 - k));
    }

    // registered mode: nobody may leave (and let the user overwrite the buffer) while a peer still reads it
    if (a.mode == MODE_REGISTERED) {
      be.sync();
      for (int k = 0; k < m; ++k) signal_group(PH_END, k, E + 1);
      for (int k = 0; k < m; ++k) wait_group(PH_END, k, E + 1);
    }
    trace(2 + 2 * m);
    be.finish(cd, a.ctas);
  }
};

}  // namespace ft
