// A "user program of the reference" for tests/test_dropin.py: plain MPI code that includes mpi_mod.hpp WITHOUT
// STANDALONE_TEST, so that its MPI_Allreduce calls are captured (allreduce_over_mpi/mpi_mod.hpp:1170). Built against
// compat/mpi_mod.hpp + the in-repo MPI subset; run under ftmpirun. Prints DROPIN_USER_OK on rank 0.
#include <mpi.h>
#include <stdint.h>
#include <stdio.h>

#include <vector>

#include "mpi_mod.hpp"

template <typename T> static int check(const char* what, const std::vector<T>& got, const std::vector<T>& want) {
  for (size_t i = 0; i < want.size(); ++i)
    if (got[i] != want[i]) {
      printf("MISMATCH %s at %zu\n", what, i);
      return 1;
    }
  return 0;
}

int main(int argc, char** argv) {
  int provided = 0, rank = 0, n = 1, bad = 0;
  MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &n);

  // small, then large (the window has to grow), then small again; in place and out of place
  for (int count : {35, 7, 3000000, 1001}) {
    std::vector<float> x((size_t)count), out((size_t)count, -1.f), want((size_t)count);
    for (int i = 0; i < count; ++i) {
      x[(size_t)i] = (float)((i % 13) * (rank + 1));
      want[(size_t)i] = (float)((i % 13) * (n * (n + 1) / 2));
    }
    std::vector<float> keep = x;
    MPI_Allreduce(x.data(), out.data(), count, MPI_FLOAT, MPI_SUM, MPI_COMM_WORLD);
    bad |= check("float out-of-place", out, want);
    bad |= check("sendbuf preserved", x, keep);
    MPI_Allreduce(MPI_IN_PLACE, x.data(), count, MPI_FLOAT, MPI_SUM, MPI_COMM_WORLD);
    bad |= check("float in-place", x, want);
  }
  {  // the other dtypes / ops of the reference's dispatch table
    const int count = 4099;
    std::vector<int32_t> a((size_t)count), wa((size_t)count);
    std::vector<uint8_t> b((size_t)count), wb((size_t)count);
    std::vector<double> d((size_t)count), wd((size_t)count);
    std::vector<long long> l((size_t)count), wl((size_t)count);
    for (int i = 0; i < count; ++i) {
      a[(size_t)i] = i * (rank + 1);
      wa[(size_t)i] = i * (n * (n + 1) / 2);
      b[(size_t)i] = (uint8_t)(0xF0 | (1u << (rank % 4)) | (i & 1));
      uint8_t w = 0xFF;
      for (int r = 0; r < n; ++r) w &= (uint8_t)(0xF0 | (1u << (r % 4)) | (i & 1));
      wb[(size_t)i] = w;
      d[(size_t)i] = 0.5 * i + rank;
      wd[(size_t)i] = 0;
      for (int r = 0; r < n; ++r) wd[(size_t)i] += 0.5 * i + r;
      l[(size_t)i] = (long long)i * 1000000007LL + rank;
      wl[(size_t)i] = (long long)i * 1000000007LL * n + (long long)n * (n - 1) / 2;
    }
    MPI_Allreduce(MPI_IN_PLACE, a.data(), count, MPI_INT32_T, MPI_SUM, MPI_COMM_WORLD);
    bad |= check("int32 sum", a, wa);
    MPI_Allreduce(MPI_IN_PLACE, b.data(), count, MPI_UINT8_T, MPI_BAND, MPI_COMM_WORLD);
    bad |= check("uint8 band", b, wb);
    MPI_Allreduce(MPI_IN_PLACE, d.data(), count, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD);
    bad |= check("double sum", d, wd);
    MPI_Allreduce(MPI_IN_PLACE, l.data(), count, MPI_LONG_LONG, MPI_SUM, MPI_COMM_WORLD);
    bad |= check("long long sum", l, wl);
  }
  int all_bad = bad;
  MPI_Allreduce(MPI_IN_PLACE, &all_bad, 1, MPI_INT32_T, MPI_SUM, MPI_COMM_WORLD);  // count < N
  MPI_Barrier(MPI_COMM_WORLD);
  if (rank == 0) printf(all_bad ? "DROPIN_USER_FAILED\n" : "DROPIN_USER_OK\n");
  MPI_Finalize();
  return all_bad ? 1 : 0;
}
