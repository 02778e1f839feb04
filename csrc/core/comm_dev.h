// Device-visible communicator state, flag-pad layout and kernel argument blocks.
//
// What this replaces in the reference: MPI_Isend/Irecv request arrays + MPI_Waitall + one
// MPI_Barrier per stage (allreduce_over_mpi/mpi_mod.hpp:969-972, :1003, :1027-1028, :1058-1059) and
// the grow-only host staging buffer (:932-950). Here a stage boundary is a 64-bit flag per
// (phase, CTA, source rank) that the producer pushes into the consumer's pad with a system-scope
// release store; flags carry a device-resident epoch so that nothing is ever reset from the host
// (CUDA-graph replays just keep counting).
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "plan.h"

namespace ft {

// Flag phases. RS stage k uses PH_RS0 + k, AG stage k uses PH_AG0 + k.
enum Phase {
  PH_IN = 0,                          // "my window holds this tile's input"
  PH_RS0 = 1,
  PH_AG0 = 1 + FT_MAX_STAGES,
  PH_END = 1 + 2 * FT_MAX_STAGES,     // "I have finished reading everybody's window" (registered mode)
  PH_LONELY = 2 + 2 * FT_MAX_STAGES,  // tree rank -> lonely ranks: "my block of this tile is final"
  PH_COUNT = 3 + 2 * FT_MAX_STAGES
};

FT_HD size_t pad_slot(int phase, int cta, int src) {
  return ((size_t)phase * FT_MAX_CTAS + (size_t)cta) * FT_MAX_RANKS + (size_t)src;
}
static const size_t FT_PAD_WORDS = (size_t)PH_COUNT * FT_MAX_CTAS * FT_MAX_RANKS;
static const size_t FT_PAD_BYTES = FT_PAD_WORDS * sizeof(uint64_t);

// Flag value for "tile j of the call with this epoch is done": strictly increasing across calls.
static const int FT_EPOCH_SHIFT = 20;  // up to 2^20 - 1 tiles per CTA per call
FT_HD uint64_t epoch_base(uint64_t epoch) { return (epoch + 1) << FT_EPOCH_SHIFT; }

// Error word layout (nonzero == failed): code | phase << 8 | src << 16 | rank << 24
enum ErrCode { ERR_NONE = 0, ERR_TIMEOUT = 1, ERR_PEER = 2 };

// One per (logical) rank, resident in that rank's device memory.
struct CommDev {
  int rank;
  int nranks;
  uint64_t* pad[FT_MAX_RANKS];  // flag pad of every rank as mapped into this rank's address space
  char* ll[FT_MAX_RANKS];       // LL (one-shot) receive buffers of every rank
  uint64_t* epoch;              // local: number of collectives completed on this communicator
  unsigned* done;               // local: CTAs of the current kernel that have finished
  unsigned* err;                // local: sticky error word
  unsigned* err_host;           // mirror of `err` in mapped host memory (the host peeks it without a sync), or null
  size_t ll_bytes;              // capacity of one parity half of an LL buffer
  uint64_t* trace;              // local, optional: FT_TRACE_SLOTS phase timestamps of CTA 0 (ns), null = off
};

// Phase-boundary timestamps (parity with the reference's SHOW_TIME logging, mpi_mod.hpp:33-38 and the
// TIME_LOG_IF sites in tree_allreduce): slot 0 = start, 1 = window published, 2+k = RS stage k done,
// 2+m+i = i-th AG stage done (stage m-1-i), last used slot = exit. Slot FT_TRACE_SLOTS-1 holds the count.
static const int FT_TRACE_SLOTS = 24;

enum Phases { PHASES_ALLREDUCE = 0, PHASES_REDUCE_SCATTER = 1, PHASES_ALL_GATHER = 2 };

enum Mode {
  MODE_STAGING = 0,     // user buffers are private: copy-in to the symmetric window, parity double-buffer
  MODE_REGISTERED = 1   // the user buffer IS the symmetric window (in place): zero copy + exit barrier
};

// Per logical rank, per call.
struct RankCall {
  const CommDev* comm;
  const void* send;   // user input (local memory of this rank)
  void* recv;         // user output
  char* const* win;   // device table [nranks]: base of the window segment of every rank
  char* mc;           // multicast alias of the window segment base (NVLS), or null
};

struct StagedArgs {
  Plan plan;
  Layout lay;
  int nlocal;  // logical ranks served by this launch (1 in multi-process / multi-device mode)
  int ctas;    // CTAs per logical rank
  int mode;    // Mode
  int apply_scale;
  double scale;
  long long win_off0;    // window byte offset inside the segment ...
  long long win_stride;  // ... plus (epoch & 1) * win_stride (0 in registered mode)
  unsigned long long timeout_ns;  // 0 = spin forever
  int use_tma;  // CUDA back end: stage peer chunks through shared memory with cp.async.bulk (TMA)
  int lonely;   // ranks plan.nranks .. plan.nranks+lonely-1 are outside the tree (N = tree + lonely)
  long long scale_plane_off;  // fp8 block-scaled wire: byte offset of the scale plane inside the window
  int phases;   // 0 = allreduce, 1 = reduce-scatter only (recv gets my block), 2 = all-gather only (send is my block)
  int nvls_reducers;  // NVLS with private buffers: CTAs per rank that only run the in-switch reduce (rest copy)
  int nvls_variant;   // inner loop of the in-switch reduce, see StagedCta::nvls_tile (tuning; 0 = default)
  int nvls_flags;     // 4: keep the exit barrier of the registered path (A/B timing)
  int fault_rank_plus1;  // fault injection (tests): rank + 1 of a rank that plays dead in this call, 0 = none
  RankCall rc[FT_MAX_LOCAL];
};

}  // namespace ft
