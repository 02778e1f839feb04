/* One process per GPU from plain C, no Python and no MPI in the process:
 *
 *   make cuda tools
 *   gcc -O2 -Iinclude -I/usr/local/cuda/include examples/c_ftrun_allreduce.c -o build/c_ftrun_allreduce \
 *       -Lallreduce-over-mpi_b200/lib -lflextree -Wl,-rpath,$PWD/allreduce-over-mpi_b200/lib -L/usr/local/cuda/lib64 -lcudart
 *   build/ftrun -np 8 build/c_ftrun_allreduce
 *
 * What `mpirun -np 8 ./allreduce_over_mpi` was to the reference (allreduce_over_mpi/benchmark.cpp:48-52, :149-159):
 * the launcher hands every process its rank, ft_comm_create_from_env() finds the peers over unix sockets and maps
 * their memory (VMM + NVLS multicast where the box has it), ft_allreduce() is the MPI_Allreduce-shaped call. */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "flextree.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA: %s\n", cudaGetErrorString(e_)); exit(1); } } while (0)
#define FT(x) do { ft_result_t r_ = (x); if (r_ != FT_SUCCESS) { fprintf(stderr, "flextree: %s\n", ft_last_error()); exit(1); } } while (0)

int main(void) {
  const int rank = ft_env_rank(), size = ft_env_size();
  if (rank < 0) { fprintf(stderr, "start me with: ftrun -np N %s\n", "c_ftrun_allreduce"); return 2; }
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const int dev = ft_env_local_rank() % ndev;
  CK(cudaSetDevice(dev));
  ft_comm_t* comm = NULL;
  FT(ft_comm_create_from_env(&comm, dev, NULL));

  const size_t count = 1 << 24;                      /* 64 MiB of fp32 */
  float* buf = NULL;                                 /* symmetric memory: the zero-copy (NVLS / in-place P2P) path */
  FT(ft_symm_alloc(comm, count * sizeof(float), (void**)&buf));
  float* host = (float*)malloc(count * sizeof(float));
  for (size_t i = 0; i < count; ++i) host[i] = (float)(i % 1000) + (float)rank;
  CK(cudaMemcpy(buf, host, count * sizeof(float), cudaMemcpyHostToDevice));

  FT(ft_allreduce(FT_IN_PLACE, buf, count, FT_FLOAT, FT_SUM, comm, NULL));   /* default stream; asynchronous */
  CK(cudaDeviceSynchronize());
  if (ft_comm_error(comm)) { fprintf(stderr, "rank %d: device watchdog fired\n", rank); return 1; }

  CK(cudaMemcpy(host, buf, 16 * sizeof(float), cudaMemcpyDeviceToHost));
  const float want9 = (float)size * 9.0f + (float)(size * (size - 1) / 2);
  ft_call_info_t info;
  ft_comm_last_call(comm, &info);
  for (int r = 0; r < size; ++r) {                   /* rank-ordered output, like the reference's CHECK loop */
    ft_env_barrier();
    if (r == rank) printf("rank %d/%d on GPU %d: buf[9] = %g (want %g), algorithm id %d, multicast %d\n", rank, size, dev, host[9], want9,
                          info.algo, ft_comm_has_multicast(comm));
    fflush(stdout);
  }
  /* host-resident data, as the reference's benchmark has it: chunk-pipelined H2D / allreduce / D2H */
  for (size_t i = 0; i < count; ++i) host[i] = 1.0f;
  FT(ft_allreduce_host(FT_IN_PLACE, host, count, FT_FLOAT, FT_SUM, comm));
  if (host[12345] != (float)size) { fprintf(stderr, "rank %d: host path wrong: %g\n", rank, host[12345]); return 1; }
  ft_symm_free(comm, buf);
  ft_comm_destroy(comm);
  free(host);
  return 0;
}
