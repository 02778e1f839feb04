// A plain MPI program with DEVICE buffers: it includes the drop-in header and calls MPI_Allreduce on cudaMalloc'ed
// memory - the header classifies the pointers (ft_pointer_device) and sends them to the NVLink kernels, host buffers
// in the same program still take the CPU data plane. Contract being exercised: "include the header, every
// MPI_Allreduce is captured" (allreduce_over_mpi/mpi_mod.hpp:1167-1171), here for GPU memory.
// Built by tests/test_multigpu.py with nvcc against the in-repo MPI subset; run as `ftmpirun -np 2 dropin_gpu_user`.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "mpi_mod.hpp"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA: %s (%s:%d)\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char** argv) {
  int provided = 0, rank = 0, size = 1, ndev = 0;
  MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  CK(cudaGetDeviceCount(&ndev));
  CK(cudaSetDevice(rank % ndev));
  int bad = 0;
  for (int count : {35, 1000003, 1 << 22}) {
    // out of place, float SUM
    std::vector<float> h((size_t)count);
    for (int i = 0; i < count; ++i) h[(size_t)i] = (float)((i % 1000) + rank);
    float *d_in = nullptr, *d_out = nullptr;
    CK(cudaMalloc(&d_in, sizeof(float) * (size_t)count));
    CK(cudaMalloc(&d_out, sizeof(float) * (size_t)count));
    CK(cudaMemcpy(d_in, h.data(), sizeof(float) * (size_t)count, cudaMemcpyHostToDevice));
    MPI_Allreduce(d_in, d_out, count, MPI_FLOAT, MPI_SUM, MPI_COMM_WORLD);
    std::vector<float> got((size_t)count);
    CK(cudaMemcpy(got.data(), d_out, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost));
    for (int i = 0; i < count; ++i) {
      const float want = (float)(size * (i % 1000)) + (float)(size * (size - 1) / 2);
      if (got[(size_t)i] != want) { if (!bad) fprintf(stderr, "rank %d: float[%d] = %g, want %g\n", rank, i, got[(size_t)i], want); ++bad; }
    }
    // in place, int MAX
    std::vector<int> hi((size_t)count);
    for (int i = 0; i < count; ++i) hi[(size_t)i] = (i * 7 + rank * 13) % 1001;
    int* d_i = nullptr;
    CK(cudaMalloc(&d_i, sizeof(int) * (size_t)count));
    CK(cudaMemcpy(d_i, hi.data(), sizeof(int) * (size_t)count, cudaMemcpyHostToDevice));
    MPI_Allreduce(MPI_IN_PLACE, d_i, count, MPI_INT32_T, MPI_MAX, MPI_COMM_WORLD);
    CK(cudaMemcpy(hi.data(), d_i, sizeof(int) * (size_t)count, cudaMemcpyDeviceToHost));
    for (int i = 0; i < count; ++i) {
      int want = 0;
      for (int r = 0; r < size; ++r) { int v = (i * 7 + r * 13) % 1001; want = v > want ? v : want; }
      if (hi[(size_t)i] != want) { if (!bad) fprintf(stderr, "rank %d: int[%d] = %d, want %d\n", rank, i, hi[(size_t)i], want); ++bad; }
    }
    CK(cudaFree(d_in));
    CK(cudaFree(d_out));
    CK(cudaFree(d_i));
  }
  // host memory in the same program: CPU data plane
  double x[4] = {1.0 * rank, 2.0, 3.0, -1.0 * rank};
  MPI_Allreduce(MPI_IN_PLACE, x, 4, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
  if (x[0] != size * (size - 1) / 2.0 || x[1] != 2.0 * size || x[3] != -x[0]) ++bad;
  for (int r = 0; r < size; ++r) {
    MPI_Barrier(MPI_COMM_WORLD);
    if (r == rank) { printf("DROPIN_GPU %d/%d %s\n", rank, size, bad ? "FAILED" : "OK"); fflush(stdout); }
  }
  MPI_Finalize();
  return bad ? 1 : 0;
}
