/* flextree.h - C API of the B200-native generalized allreduce library.
 *
 * Capability map to the reference (Youhe-Jiang/AllReduce-Over-MPI; citations are reference paths):
 *   ft_allreduce            <-> MPI_Allreduce_FT / MPI_Allreduce override  allreduce_over_mpi/mpi_mod.hpp:1167-1221
 *   FT_IN_PLACE             <-> MPI_IN_PLACE handling                     mpi_mod.hpp:959-962, 1194-1215
 *   env FT_TOPO, ft_comm_set_widths <-> get_stages                        mpi_mod.hpp:882-929
 *   ft_dtype_t / ft_op_t    <-> MPI_Datatype / MPI_Op dispatch            mpi_mod.hpp:825-874
 *   ft_enumerate_widths ... <-> cost_model/GetWidth.h, ChooseWidth.h, CostModel.h, topo_count/factor_count.py
 * Everything that takes a `stream` is asynchronous and CUDA-graph capturable once the communicator
 * has been warmed up for that message size (see ft_comm_reserve).
 */
#ifndef FLEXTREE_H_
#define FLEXTREE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FT_VERSION_MAJOR 0
#define FT_VERSION_MINOR 1

typedef enum {
  FT_SUCCESS = 0,
  FT_ERR_INVALID = 1,     /* bad argument (also: widths whose product is not the world size) */
  FT_ERR_CUDA = 2,        /* a CUDA runtime/driver call failed; see ft_last_error() */
  FT_ERR_UNSUPPORTED = 3, /* (op, dtype) or algorithm not available */
  FT_ERR_TIMEOUT = 4,     /* device watchdog fired: a peer never signalled */
  FT_ERR_NOMEM = 5,
  FT_ERR_INTERNAL = 6
} ft_result_t;

/* Numeric values match ft::DType / ft::RedOp (csrc/core/types.h). The first ten are the reference's
 * list (mpi_mod.hpp:827-837); the rest are additions. */
typedef enum {
  FT_UINT8 = 0, FT_INT8 = 1, FT_UINT16 = 2, FT_INT16 = 3, FT_INT32 = 4, FT_INT64 = 5,
  FT_FLOAT = 6, FT_DOUBLE = 7, FT_BOOL = 8, FT_LONG_LONG = 9,
  FT_UINT32 = 10, FT_UINT64 = 11, FT_HALF = 12, FT_BFLOAT16 = 13,
  /* wire_dtype of ft_allreduce_ex only (opt-in): e4m3 values + one power-of-two scale byte per 16 elements,
   * re-quantised at every reduce-scatter stage; carries fp32 or bf16 buffers, SUM, P2P stages. */
  FT_FP8_BLOCK16 = 14
} ft_dtype_t;

typedef enum { FT_SUM = 0, FT_BAND = 1, FT_MAX = 2, FT_MIN = 3, FT_BOR = 4 } ft_op_t;

/* TWOSHOT: reduce-scatter + all-gather over LL lines (no flags, no barrier) for the sizes between the one-shot's
 * sweet spot and the bandwidth kernels; value 5 is internal (the TMA flavour of STAGES). */
typedef enum { FT_ALGO_AUTO = 0, FT_ALGO_ONESHOT = 1, FT_ALGO_STAGES = 2, FT_ALGO_NVLS = 3, FT_ALGO_RING = 4,
               FT_ALGO_TWOSHOT = 6 } ft_algo_t;

#define FT_IN_PLACE ((const void*)(intptr_t)-1)

const char* ft_last_error(void);      /* thread-local description of the last failure */
const char* ft_version_string(void);  /* "flextree-b200 x.y (git describe, date)" */
int ft_enabled(void);                 /* prints "FlexTree enabled", returns 0 (FT_enabled(), mpi_mod.hpp:8-12) */

/* ------------------------------------------------------------------------------------------------
 * Planner (pure CPU, no CUDA needed)
 * ---------------------------------------------------------------------------------------------- */

/* H(n): number of ordered factorizations of n (topo_count/factor_count.py:1-11). */
uint64_t ft_count_factorizations(uint64_t n);

/* Ordered factorizations of n, flattened as [len, w0..w(len-1), len, ...] into out[cap]. legacy_flat=1
 * lists the single-stage tree the way cost_model/GetWidth.h:15-23 does ({1,n} and {n,1}).
 * Returns the number of ints written, or -(needed) if cap is too small; *n_lists gets the list count. */
long long ft_enumerate_widths(int n, int legacy_flat, int* out, long long cap, long long* n_lists);

/* Reference chooseWidth(n) listing (cost_model/ChooseWidth.h:8-38) as text lines "2*4", "3*3-1", "7*1+1". */
long long ft_candidates_text(int n, char* out, long long cap);

int ft_prime_factors(int n, int* out, int cap); /* returns count */
int ft_is_prime(int n);
int ft_is_prime_legacy(int n);

/* Reference cost formula (cost_model/CostModel.h). terms[3] = latency, memory, bandwidth (may be NULL). */
double ft_legacy_cost(const int* widths, int nstages, int total_nodes, double chunk, double* terms);
/* arg-min over ft_enumerate_widths(n, legacy) - what cost_model/main.cpp prints as the optimum.
 * Returns the number of stages written to widths_out (<= cap), 0 if there is no candidate. */
int ft_legacy_choose(int n, double chunk, int* widths_out, int cap, double* cost_out);

/* Block-granular traffic of one rank for a plan, as the reference would send it (one message per
 * block, mpi_mod.hpp:673-703). */
int ft_plan_traffic(const int* widths, int nstages, int rank, long long count, long long* messages,
                    long long* elements);
/* ASCII dump of the send/recv tables of one rank (print_ops, mpi_mod.hpp:107-144). */
long long ft_print_ops(const int* widths, int nstages, int rank, int lonely, char* out, long long cap);

/* Link model (alpha-beta-gamma, fitted on the box). json may be NULL -> built-in defaults /
 * env FT_COSTMODEL. */
typedef struct ft_link_model ft_link_model_t;
ft_link_model_t* ft_link_model_create(const char* json_path);
void ft_link_model_destroy(ft_link_model_t*);
long long ft_link_model_dump(const ft_link_model_t*, char* out, long long cap);
int ft_link_model_set(ft_link_model_t*, const char* key, double value);
double ft_predict_us(const ft_link_model_t*, int algo, const int* widths, int nstages, int nranks,
                     double bytes, int registered);
/* Picks algorithm + widths for a message; returns nstages written. */
int ft_choose(const ft_link_model_t*, int nranks, double bytes, int registered, int* algo_out,
              int* widths_out, int cap, double* est_us);

/* FT_TOPO semantics of the reference (mpi_mod.hpp:882-929): unset/empty -> {n}; any 1 -> ring ({1});
 * product != n -> error. text may be NULL to read the environment. Returns nstages or -1. */
int ft_parse_topo(const char* text, int nranks, int* widths_out, int cap);
/* The same with the tree-size suffixes of the reference's listings (PrintTreeStructure.h: "a*b+1", "a*b-1"):
 * "2,3+1" = tree of nranks-1 plus one lonely rank, "2,4-1" = tree of nranks+1 with one phantom position. */
int ft_parse_topo_ex(const char* text, int nranks, int* widths_out, int cap, int* lonely_out, int* phantoms_out);

/* ------------------------------------------------------------------------------------------------
 * Communicators
 * ---------------------------------------------------------------------------------------------- */
typedef struct ft_comm ft_comm_t;

/* Byte all-gather used once at creation to exchange memory handles: every rank contributes `bytes`
 * and receives nranks*bytes in rank order. Must be a collective over the same ranks. */
typedef int (*ft_allgather_fn)(void* ctx, const void* send, void* recv, size_t bytes);

typedef struct {
  size_t window_bytes;   /* initial staging window per parity (grow-only, like mpi_mod.hpp:932-950); 0 = default */
  size_t ll_bytes;       /* LL (one-shot) buffer per parity; 0 = default */
  int use_multicast;     /* -1 auto, 0 never, 1 require */
  int use_vmm;           /* -1 auto (VMM + POSIX fd, fall back to cudaIpc), 0 cudaIpc only */
  double timeout_s;      /* device watchdog; <0 = default (env FT_TIMEOUT_S or 20 s); 0 = off */
} ft_comm_options_t;
void ft_comm_options_default(ft_comm_options_t*);

/* One process per GPU (what torch.distributed / torchrun gives you). */
ft_result_t ft_comm_create(ft_comm_t** out, int rank, int nranks, int device, ft_allgather_fn ag,
                           void* ag_ctx, const ft_comm_options_t* opts);
/* One process per GPU started by the launcher `ftrun -np N prog args...` (build/ftrun; the stand-in for the reference's
 * mpirun, benchmark.cpp:48-52): rank, size and a bootstrap token come from the environment (FT_RANK, FT_SIZE,
 * FT_LOCAL_RANK, FT_BOOT_TOKEN) and the handle exchange runs over unix sockets - no Python, no MPI, and the same
 * VMM + multicast segments as the torch path. device < 0: use FT_LOCAL_RANK. The ft_env_* helpers give a C program
 * the few collectives it needs around the library (rank-ordered printing, max over ranks of a timing). */
ft_result_t ft_comm_create_from_env(ft_comm_t** out, int device, const ft_comm_options_t* opts);
int ft_env_rank(void);        /* -1 when not started by ftrun */
int ft_env_size(void);
int ft_env_local_rank(void);
int ft_env_allgather(const void* send, void* recv, size_t bytes); /* 0 = ok */
int ft_env_barrier(void);
/* One process, one rank per listed device (cudaDeviceEnablePeerAccess); comms[i] is rank i. */
ft_result_t ft_comm_create_local(ft_comm_t** comms, int nranks, const int* devices,
                                 const ft_comm_options_t* opts);
/* One process, ONE device, nranks logical ranks served by a single cooperative launch: the kernels'
 * protocol and arithmetic with every "peer" in local HBM. For tests and the 1-GPU benchmark. */
ft_result_t ft_comm_create_emulated(ft_comm_t** out, int nranks, int device, const ft_comm_options_t* opts);
ft_result_t ft_comm_destroy(ft_comm_t*);

int ft_comm_rank(const ft_comm_t*);
int ft_comm_size(const ft_comm_t*);
int ft_comm_has_multicast(const ft_comm_t*);
/* Sticky device error word (0 = healthy): code | phase<<8 | src<<16 | rank<<24.
 * ft_comm_error copies it from the device: it BLOCKS until the work queued on the device has drained, so after it
 * returns 0 every collective issued so far has completed cleanly. ft_comm_error_peek reads the kernel-written
 * mirror in mapped host memory instead: no synchronisation, may lag a running kernel. Every collective call
 * peeks first and returns FT_ERR_TIMEOUT (with the decoded word in ft_last_error()) once an EARLIER call on the
 * communicator has tripped the watchdog - the output of that earlier call is garbage - until
 * ft_comm_clear_error() (which waits for the device) re-arms it.
 * (The reference has no failure path at all: an absent peer hangs MPI_Waitall, mpi_mod.hpp:1003.) */
unsigned ft_comm_error(ft_comm_t*);
unsigned ft_comm_error_peek(const ft_comm_t*);
ft_result_t ft_comm_clear_error(ft_comm_t*);
/* Tuning knobs that have no place in the stable API (launch-shape scans, kernel-variant A/B runs). Keys:
 * "nvls_variant" (inner loop of the in-switch reduce on registered buffers, 0-6, see StagedCta::nvls_tile_variant;
 * 7 / 8 are measurement kernels that leave out the multicast store / the in-switch load and give wrong sums),
 * "nvls_flags" (4: keep the exit barrier),
 * "nvls_ctas", "nvls_max_tile_bytes" (shape of the registered NVLS path). */
ft_result_t ft_comm_set_tuning(ft_comm_t*, const char* key, long long value);
/* Raw link traffic for the cost-model fit (kernels/link_probe.cu): this rank streams `bytes_per_peer` bytes from
 * (mode 0: 16-byte loads, 2: TMA bulk loads) or to (1: 16-byte stores, 3: TMA bulk stores) each of its next `npeers`
 * ranks' staging windows with `ctas` CTAs. No synchronisation between ranks and no meaningful data: time it with
 * events while the other ranks run the pattern you want to measure (collective only in that the first call may
 * grow the window). */
ft_result_t ft_link_probe(ft_comm_t*, int mode, int npeers, size_t bytes_per_peer, int ctas, void* stream);
/* Fault injection for tests and drills: logical rank `rank` plays dead in every following collective (its kernel
 * exits without touching a buffer or signalling anybody), so its peers run into the watchdog. rank < 0 = off. */
ft_result_t ft_comm_inject_fault(ft_comm_t*, int rank);

/* Overrides (NULL/0 restores automatic choice). Same meaning as env FT_TOPO / FT_ALGO. */
ft_result_t ft_comm_set_widths(ft_comm_t*, const int* widths, int nstages);
ft_result_t ft_comm_set_algo(ft_comm_t*, ft_algo_t algo);
ft_result_t ft_comm_set_link_model(ft_comm_t*, const ft_link_model_t*);
/* Launch shape overrides: CTAs per rank, threads per CTA, tile bytes (0 = automatic). */
ft_result_t ft_comm_set_launch(ft_comm_t*, int ctas, int threads, long long tile_bytes);

/* Phase timing (the reference's SHOW_TIME switch, mpi_mod.hpp:33-38): when enabled, CTA 0 of the staged
 * kernel records %globaltimer at every phase boundary: start, window published, each reduce-scatter
 * stage, each all-gather stage, exit. ft_comm_trace_read returns the slots of the last call (ns). */
ft_result_t ft_comm_trace_enable(ft_comm_t*, int on);
int ft_comm_trace_read(ft_comm_t*, int local_rank, unsigned long long* ns, int cap);

/* Upper bound on the CTAs (= SMs) a collective may occupy; 0 = no bound. Use it when the collective
 * overlaps compute (DDP backward): NVLS needs 8 CTAs anyway, the TMA P2P stages reach most of their
 * bandwidth from ~32. Same as env FT_MAX_CTAS. */
ft_result_t ft_comm_set_max_ctas(ft_comm_t*, int max_ctas);

/* P2P stage kernels: 1 = stage peer chunks through shared memory with TMA bulk copies (cp.async.bulk +
 * mbarrier ring), 0 = register-staged 16-byte loads, -1 = default (env FT_TMA). */
ft_result_t ft_comm_set_tma(ft_comm_t*, int on);

/* Collective: make sure a message of `bytes` fits the staging window (allocates + exchanges handles
 * if it has to grow). Call it before stream capture. */
ft_result_t ft_comm_reserve(ft_comm_t*, size_t bytes);

/* Symmetric (registered) memory: collective allocation of `bytes` on every rank. Buffers from here
 * can be all-reduced in place with zero copies. */
ft_result_t ft_symm_alloc(ft_comm_t*, size_t bytes, void** ptr);
ft_result_t ft_symm_free(ft_comm_t*, void* ptr);

/* ------------------------------------------------------------------------------------------------
 * Collectives. `stream` is a cudaStream_t.
 * ---------------------------------------------------------------------------------------------- */
ft_result_t ft_allreduce(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op,
                         ft_comm_t* comm, void* stream);

/* Blocking form on the default stream (MPI semantics: recvbuf is ready when the call returns) and a pointer
 * classifier (-1 = host memory, else the device ordinal): together they let compat/mpi_mod.hpp route a captured
 * MPI_Allreduce on cudaMalloc'ed buffers to the GPU kernels (mpi_mod.hpp:1167-1171 captures every MPI_Allreduce). */
ft_result_t ft_allreduce_blocking(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op,
                                  ft_comm_t* comm);
int ft_pointer_device(const void* ptr);

/* Extended form used by the torch path: fp32 user buffers with a 16-bit wire (wire_dtype), and a
 * fused `out = cast(sum * scale)` epilogue (the DDP 1/N). wire_dtype == dtype and scale == 1 give
 * ft_allreduce. */
ft_result_t ft_allreduce_ex(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype,
                            ft_dtype_t wire_dtype, ft_op_t op, double scale, ft_comm_t* comm, void* stream);

/* Emulated communicators: one send/recv pointer per logical rank. */
ft_result_t ft_allreduce_multi(const void* const* sendbufs, void* const* recvbufs, size_t count,
                               ft_dtype_t dtype, ft_dtype_t wire_dtype, ft_op_t op, double scale,
                               ft_comm_t* comm, void* stream);

/* Host-resident data (what the reference's MPI_Allreduce takes: benchmark.cpp:119-159 allreduces a malloc'ed
 * vector). Synchronous like MPI_Allreduce: when the call returns, recvbuf (host memory) holds the result.
 * The message is cut into chunks (env FT_HOST_CHUNK_BYTES, default 16 MiB) that are pipelined over three streams -
 * H2D of chunk i+1, in-place zero-copy allreduce of chunk i, D2H of chunk i-1 - so the call costs about
 * max(H2D, D2H), not their sum plus the collective. Pass page-locked memory (cudaHostAlloc / cudaHostRegister /
 * torch pin_memory) for full PCIe speed; pageable memory works but copies synchronously.
 * Collective: the first call (and any call with a larger chunk size) allocates symmetric chunk slots.
 * sendbuf may be FT_IN_PLACE. `_multi`: one pointer per logical rank (emulated communicators), fused scale,
 * explicit chunk size (0 = default). */
ft_result_t ft_allreduce_host(const void* sendbuf, void* recvbuf, size_t count, ft_dtype_t dtype, ft_op_t op,
                              ft_comm_t* comm);
ft_result_t ft_allreduce_host_multi(const void* const* sendbufs, void* const* recvbufs, size_t count,
                                    ft_dtype_t dtype, ft_op_t op, double scale, ft_comm_t* comm, size_t chunk_bytes);

/* The two halves of the algorithm as collectives of their own (the reference's tree_allreduce is a
 * reduce-scatter loop followed by an all-gather loop, mpi_mod.hpp:988-1029 / :1050-1060).
 * ft_reduce_scatter: sendbuf holds nranks*recvcount elements, recvbuf receives this rank's reduced block.
 * ft_all_gather: sendbuf holds sendcount elements (this rank's block), recvbuf nranks*sendcount. */
ft_result_t ft_reduce_scatter(const void* sendbuf, void* recvbuf, size_t recvcount, ft_dtype_t dtype, ft_op_t op,
                              ft_comm_t* comm, void* stream);
ft_result_t ft_all_gather(const void* sendbuf, void* recvbuf, size_t sendcount, ft_dtype_t dtype, ft_comm_t* comm,
                          void* stream);
ft_result_t ft_reduce_scatter_multi(const void* const* sendbufs, void* const* recvbufs, size_t recvcount,
                                    ft_dtype_t dtype, ft_op_t op, ft_comm_t* comm, void* stream);
ft_result_t ft_all_gather_multi(const void* const* sendbufs, void* const* recvbufs, size_t sendcount,
                                ft_dtype_t dtype, ft_comm_t* comm, void* stream);

/* What the last call on this communicator resolved to (for logs, tests and the benchmark). */
typedef struct {
  int algo;
  int nstages;
  int widths[8];
  int ctas, threads;
  long long tile_bytes;
  int mode;        /* 0 staging, 1 registered */
  int vectorized;
  double est_us;
} ft_call_info_t;
ft_result_t ft_comm_last_call(const ft_comm_t*, ft_call_info_t*);

/* Allreduce across `nranks` buffers that all live on the CURRENT device (the single-GPU "emulated ranks"
 * form of ft_allreduce_multi, one-shot algorithm), for callers that are not allowed to allocate.
 * No counterpart in the reference (which has neither a GPU path nor an allocation-free mode).
 *
 * Who needs this: code that runs where a device allocation is off limits - a plugin or callback invoked
 * from inside another framework's execution loop (the host framework owns all device memory and forbids
 * cudaMalloc/cudaFree there, or treats their implicit device-wide synchronisation as a bug), a service
 * that has handed the whole GPU to a pool allocator at start-up, or a latency-critical loop that must not
 * risk an out-of-memory failure mid-run. Such callers cannot create an ft_comm (its creation allocates
 * the scratch buffers), so this entry point keeps ALL of its scratch - receive buffers, per-rank control
 * blocks, epoch/completion/error words - in one static __device__ array inside the library
 * (about 4 MiB, resident on every GPU that loads the library, until the process exits). The call itself
 * performs no cudaMalloc, cudaFree, host heap allocation or any other allocating routine.
 *
 * Limits (rejected with FT_ERR_INVALID and a message in ft_last_error()):
 *   nranks <= FT_NOALLOC_MAX_RANKS (8), count * sizeof(dtype) <= FT_NOALLOC_MAX_BYTES (16384) per rank.
 * sendbufs[i] / recvbufs[i]: device pointers of logical rank i (sendbufs == NULL or an entry equal to
 * FT_IN_PLACE means in place). Every recvbuf receives the same reduced vector, bit-identical to what
 * ft_allreduce_multi produces with the one-shot algorithm.
 *
 * Set-up: the FIRST call on each device (with that device current) looks up the device address of the
 * scratch array and writes the control blocks into it with a synchronous copy. Make that first call
 * before any stream capture is started on the device.
 * Sharing: every call on a device uses the same array, and the array keeps the previous call's contents.
 * The call clears what has to start from zero (completion counters, error words) in stream order before
 * the kernel; everything else is protected by per-call tags. Two calls whose kernels overlap on the GPU
 * (different streams, or different host threads without ordering) race on the array: serialise them.
 * World sizes may change freely from call to call (4 ranks, then 8, then 2 ...): a one-thread kernel queued
 * behind every collective keeps the epoch words of all FT_NOALLOC_MAX_RANKS logical ranks in step.
 * Failure: the return code only covers the launch. If the device watchdog fires (20 s without a peer's
 * line, e.g. because two calls overlapped) the recvbufs of that call hold garbage; the error word of the
 * first such call on the device is kept until ft_noalloc_error(1) reads and clears it. */
#define FT_NOALLOC_MAX_RANKS 8
#define FT_NOALLOC_MAX_BYTES 16384
ft_result_t ft_allreduce_emulated_noalloc(const void* const* sendbufs, void* const* recvbufs, int nranks,
                                          size_t count, ft_dtype_t dtype, ft_op_t op, void* stream);
/* Sticky error word of the allocation-free entry point on the CURRENT device (0 = every call so far completed;
 * layout as ft_comm_error). Waits for the device (one 4-byte copy), allocates nothing; clear != 0 resets it.
 * No counterpart in the reference. */
unsigned ft_noalloc_error(int clear);

#ifdef STANDALONE_TEST
/* Source-level alias for code written against the reference benchmark (benchmark.cpp:153). */
#define MPI_Allreduce_FT(send, recv, count, dtype, op, comm) \
  ((int)ft_allreduce((send), (recv), (count), (dtype), (op), (comm), 0))
#endif

#ifdef __cplusplus
}
#endif
#endif /* FLEXTREE_H_ */
