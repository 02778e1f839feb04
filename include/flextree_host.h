/* flextree_host.h - C API of the CPU communicator (csrc/host/host_comm.cc): one process per rank on one host,
 * windows / flag pads in POSIX shared memory, the same algorithm templates as the GPU kernels.
 *
 * This is the data plane a program written against the reference moves to when its buffers live in host
 * memory: the reference IS an allreduce between CPU processes (allreduce_over_mpi/mpi_mod.hpp:1167-1221).
 * compat/mpi_mod.hpp wraps it in the reference's own signature. Needs only libflextree_host.so (g++, no CUDA).
 */
#ifndef FLEXTREE_HOST_H_
#define FLEXTREE_HOST_H_

#include "flextree.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ft_host_comm ft_host_comm_t;

/* Collective over all ranks. `ag` is a byte all-gather over the same ranks (used at creation and destruction
 * only; calls agree through shared memory). window_bytes: capacity for one message in the wire type, per
 * parity (0 = 4 MiB; fixed for the life of the communicator - create a larger one to grow). ll_bytes: one-shot
 * (LL) buffer, 0 = 1 MiB. */
ft_result_t ft_host_comm_create(ft_host_comm_t** out, int rank, int nranks, ft_allgather_fn ag, void* ag_ctx,
                                size_t window_bytes, size_t ll_bytes);
/* Same, with room for up to max_phantoms phantom ranks: "-k" trees (cost_model/ChooseWidth.h:16-22 proposes the trees
 * of N+1 for N ranks), run with lonely = -k in ft_host_comm_allreduce. Phantom j is hosted by process j % nranks as a
 * second logical rank that contributes the operation's identity. Costs one more window set per hosting process. */
ft_result_t ft_host_comm_create_ex(ft_host_comm_t** out, int rank, int nranks, ft_allgather_fn ag, void* ag_ctx,
                                   size_t window_bytes, size_t ll_bytes, int max_phantoms);
ft_result_t ft_host_comm_destroy(ft_host_comm_t* comm); /* collective */
int ft_host_comm_rank(const ft_host_comm_t*);
int ft_host_comm_size(const ft_host_comm_t*);
size_t ft_host_comm_window_bytes(const ft_host_comm_t*);
unsigned ft_host_comm_error(const ft_host_comm_t*); /* sticky watchdog word, 0 = healthy */
/* This rank's registered region (window_bytes long): data placed here is reduced in place with mode 1, zero copy. */
void* ft_host_comm_registered(ft_host_comm_t*);

/* One collective. dtype = wire type, io_dtype = type of the user buffers (equal unless a 16-bit / fp8 wire carries
 * fp32 data). algo: ft_algo_t, or 5 = stages with the CPU's bulk tiles (falls back to 2 where they do not apply).
 * widths[nstages]: the tree (product + lonely = world size); ignored by ring / one-shot / nvls. ctas = threads used
 * by this rank; tile_bytes = pipelining grain. mode: 0 = private buffers, 1 = `send == recv ==` registered region.
 * apply_scale: multiply the result by `scale` when a block becomes final (fused average). timeout_ns: watchdog,
 * 0 = wait forever. lonely: k > 0 = the last k ranks are outside the tree ("+k"), k < 0 = the tree has -k phantom
 * positions beyond the ranks ("-k", needs create_ex). phases: 0 allreduce, 1 reduce-scatter (recv = my count/N block),
 * 2 all-gather.
 * Returns 0, a negative argument error, or the watchdog word. */
long long ft_host_comm_allreduce(ft_host_comm_t* comm, const void* send, void* recv, long long count, int dtype,
                                 int io_dtype, int op, int algo, const int* widths, int nstages, int ctas,
                                 long long tile_bytes, int mode, double scale, int apply_scale,
                                 unsigned long long timeout_ns, int lonely, int phases);

#define FT_ALGO_STAGES_BULK 5

#ifdef __cplusplus
}
#endif
#endif /* FLEXTREE_HOST_H_ */
