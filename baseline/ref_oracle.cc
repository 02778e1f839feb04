// ref_oracle: runs the UNMODIFIED reference allreduce (mpi_mod.hpp, included as-is) on caller-supplied
// data so that tests can compare this library against the reference bit for bit.
//   ftmpirun -np N ref_oracle <dir> <count> <dtype: f32|i32|f64> <op: sum|band> [inplace]
// rank r reads <dir>/in.<r>.bin and writes <dir>/out.<r>.bin. FT_TOPO selects the widths (reference semantics).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include <mpi.h>
#include <glog/logging.h>
#define STANDALONE_TEST
#include "mpi_mod.hpp"

int main(int argc, char** argv) {
  int provided, rank, size;
  MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  if (argc < 5) return 2;
  const std::string dir = argv[1];
  const int count = atoi(argv[2]);
  const std::string dt = argv[3], op = argv[4];
  const bool inplace = argc > 5;
  const size_t es = dt == "f64" ? 8 : 4;
  MPI_Datatype mdt = dt == "f32" ? MPI_FLOAT : dt == "f64" ? MPI_DOUBLE : MPI_INT32_T;
  MPI_Op mop = op == "band" ? MPI_BAND : MPI_SUM;
  char* in = new char[(size_t)count * es + 4096];   // new[]: lands in the shim's shared arena
  char* out = new char[(size_t)count * es + 4096];
  FILE* f = fopen((dir + "/in." + std::to_string(rank) + ".bin").c_str(), "rb");
  if (!f || fread(in, es, (size_t)count, f) != (size_t)count) return 3;
  fclose(f);
  if (inplace) {
    MPI_Allreduce_FT(MPI_IN_PLACE, in, count, mdt, mop, MPI_COMM_WORLD);
    memcpy(out, in, (size_t)count * es);
  } else {
    MPI_Allreduce_FT(in, out, count, mdt, mop, MPI_COMM_WORLD);
  }
  f = fopen((dir + "/out." + std::to_string(rank) + ".bin").c_str(), "wb");
  fwrite(out, es, (size_t)count, f);
  fclose(f);
  MPI_Finalize();
  return 0;
}
