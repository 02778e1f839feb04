/* One process, every visible GPU is a rank (ft_comm_create_local), in-place fp32 sum with the reference's
 * call shape. Build:  nvcc -Iinclude examples/c_local_allreduce.c -Lallreduce-over-mpi_b200/lib -lflextree -o build/c_local
 * (plain C: the API is MPI_Allreduce(sendbuf, recvbuf, count, datatype, op, comm) plus a stream). */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include "flextree.h"

int main(void) {
  int n = 0, devs[64];
  ft_comm_t* comms[64];
  float* buf[64];
  const size_t count = 1 << 20;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n < 1) { fprintf(stderr, "no GPU\n"); return 1; }
  for (int i = 0; i < n; ++i) devs[i] = i;
  if (ft_comm_create_local(comms, n, devs, NULL) != FT_SUCCESS) { fprintf(stderr, "%s\n", ft_last_error()); return 1; }
  float* host = (float*)malloc(count * sizeof(float));
  for (size_t i = 0; i < count; ++i) host[i] = (float)i;
  for (int r = 0; r < n; ++r) {
    cudaSetDevice(r);
    cudaMalloc((void**)&buf[r], count * sizeof(float));
    cudaMemcpy(buf[r], host, count * sizeof(float), cudaMemcpyHostToDevice);
  }
  for (int r = 0; r < n; ++r)  /* launches are asynchronous: issue all ranks, then synchronise */
    if (ft_allreduce(FT_IN_PLACE, buf[r], count, FT_FLOAT, FT_SUM, comms[r], 0) != FT_SUCCESS) {
      fprintf(stderr, "%s\n", ft_last_error());
      return 1;
    }
  for (int r = 0; r < n; ++r) {
    cudaSetDevice(r);
    cudaDeviceSynchronize();
    if (ft_comm_error(comms[r])) { fprintf(stderr, "device error %#x\n", ft_comm_error(comms[r])); return 2; }
  }
  cudaSetDevice(0);
  cudaMemcpy(host, buf[0], 20 * sizeof(float), cudaMemcpyDeviceToHost);
  printf("CHECK 0:");
  for (int i = 9; i < 20; ++i) printf(" %g", host[i]);  /* = n * i, like the reference's CHECK line */
  printf("\n");
  for (int r = 0; r < n; ++r) ft_comm_destroy(comms[r]);
  free(host);
  return 0;
}
